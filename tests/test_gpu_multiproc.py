"""The REAL multi-GPU path -- one process per GPU under torchrun, CUDA IPC peer mappings of the accumulation buffers,
the NCCL rendezvous around `reduce_resolve_peers_kernel` writing into rank 0's image -- with the assembled image
CHECKED (bench.py's `parity` block): against a rank-ordered host merge of the ranks' own buffers and against one
context rendering the same global frames.  What is sharded is the accumulation of CPUPathTracer.cpp:77-80; SURVEY 8e
promises bit-identity under the tile partition.  Skipped on a box with fewer than two GPUs (tests/test_gpu_fused_tail.py
covers the kernel with three contexts on one device, tests/test_multi_rank_cpu.py the host logic under gloo)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def run_ranks(n, *args, port=29533):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", str(n), "--steps", "1", "--warmup", "3",
           "--no-cpu-baseline", *args]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert out.returncode == 0 and len(lines) == 1, (out.stdout[-1500:], out.stderr[-3000:])
    return json.loads(lines[0])


@pytest.mark.parametrize("partition,tail", [("tiles", "fused"), ("samples", "fused"), ("samples", "nccl")])
def test_two_ranks_assemble_a_checked_image(built, partition, tail):
    if n_gpus() < 2:
        pytest.skip("needs two GPUs")
    j = run_ranks(2, "--workload", "c3", "--spp", "32", "--partition", partition, "--tail", tail)
    p = j["parity"]
    print(json.dumps(p))
    assert j["n_gpus"] == 2 and p["ok"] and p["frames_per_pixel"] == 64
    if tail == "fused":
        assert p["float_sums_bit_identical_to_rank_ordered_host_merge"] and p["rgba8_equal_to_resolve_of_host_merge"] == 1.0
    if partition == "tiles":
        assert p["pixels_bit_identical_to_single_context"] == 1.0 and p["rgba8_equal_to_single_context"] == 1.0
    else:
        assert p["max_rel_diff_vs_single_context"] <= 2e-5 and p["rgba8_equal_to_single_context"] >= 0.999 and p["rgba8_max_channel_difference_lsb"] <= 1


def test_two_ranks_cornell_with_nee_tiles(built):
    """The closed box: every pixel hits, shadow rays, 4 rays per lane -- the tile partition on the other integrator path."""
    if n_gpus() < 2:
        pytest.skip("needs two GPUs")
    j = run_ranks(2, "--workload", "c2", "--spp", "32", "--spp-pass", "32", "--partition", "tiles", port=29541)
    assert j["parity"]["ok"] and j["parity"]["pixels_bit_identical_to_single_context"] == 1.0

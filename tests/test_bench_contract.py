"""bench.py's output contract, as far as it can be checked without a GPU: the reference arm (`--impl reference`, the
reference's own CPU code of the path timed on the host cores) prints ONE JSON line with the keys the driver reads, and
the GPU arm refuses to run without a device instead of falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600, env=e)


def test_reference_arm_json_line(built):
    out = run_bench("--impl", "reference", "--workload", "c1", "--steps", "1", "--warmup", "0")
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in j, k
    assert j["impl"] == "reference" and j["unit"] == "Msamples/s" and j["higher_is_better"] is True and j["vs_baseline"] is None
    assert j["value"] > 0 and j["steps"] == 1 and j["data"] == "synthetic"
    assert j["config"]["name"] == "c1" and (j["config"]["width"], j["config"]["height"], j["config"]["spp_per_gpu"]) == (640, 360, 16)
    cb = j["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_render.so")):
        assert cb["kind"] == "reference" and cb["cores"] == 1     # the reference's render loop is single-threaded


def test_reference_arm_on_other_ranks_prints_nothing(built):
    """Under torchrun only rank 0 runs the CPU arm; the other ranks exit 0 without work."""
    out = run_bench("--impl", "reference", "--workload", "c1", "--gpus", "2", "--steps", "1", "--warmup", "0",
                    env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29517"})
    assert out.returncode == 0, out.stderr[-2000:]
    assert not [l for l in out.stdout.splitlines() if l.startswith("{")]


def test_gpu_arm_has_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = run_bench("--workload", "c1", "--steps", "1", "--warmup", "0")
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)
    assert not [l for l in out.stdout.splitlines() if l.startswith("{")]

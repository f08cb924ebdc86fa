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
    # the CPU arm runs the GPU arm's configuration, the whole image every step, and says how many steps it timed
    assert j["config"]["same_config_as_gpu_arm"] is True and j["config"]["share_of_image_per_step"] == 1.0 and j["steps_timed"] == 1
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


def test_mesh_cpu_sample_runs_the_stated_configuration(built):
    """cpu_sample() for a mesh workload: the twin at the workload's own resolution and sample count on an interleaved
    tile share; `full=True` (the reference arm) renders the whole image.  Checked on a shrunken copy of c2."""
    import importlib
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    scenes = importlib.import_module("software-path-tracer_b200.scenes")
    wl = dict(bench.WORKLOADS["c2"], name="c2", w=128, h=72, spp=8)
    v, t, m, mats = scenes.cornell_box(sphere_res=8)
    scene = dict(v=v, t=t, m=m, mats=mats)
    part = bench.cpu_sample(wl, scene, budget_s=1e-4)       # a budget nothing fits: the smallest share, 1/256
    full = bench.cpu_sample(wl, scene, full=True)
    assert full["share"] == 1.0 and full["n"] == 128 * 72 * 8 and "the whole image" in full["sample"]
    assert part["share"] == 1.0 / 256 and 0 < part["n"] <= full["n"] // 64 and "128x72, 8 spp, depth 8" in part["sample"]
    assert part["kind"] == full["kind"] == "port" and full["rays"] > full["n"]


def test_roofline_capture_stamp(built, tmp_path, monkeypatch):
    """bench.trace_build_hash() stamps profiles/trace_kernel_ncu.json; it must change when the traversal kernel's sources
    change (so that a stale ncu capture is withheld) and must not depend on anything else."""
    import importlib
    import shutil
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    h0 = bench.trace_build_hash()
    assert len(h0) == 16 and h0 == bench.trace_build_hash()
    fake = tmp_path / "repo"
    (fake / "software-path-tracer_b200").mkdir(parents=True)
    shutil.copytree(os.path.join(ROOT, "software-path-tracer_b200", "csrc"), fake / "software-path-tracer_b200" / "csrc",
                    ignore=shutil.ignore_patterns("*.o", "*.log"))
    shutil.copy(os.path.join(ROOT, "Makefile"), fake / "Makefile")
    monkeypatch.setattr(bench, "ROOT", str(fake))
    assert bench.trace_build_hash() == h0
    with open(fake / "software-path-tracer_b200" / "csrc" / "post.cu", "a") as f:
        f.write("// unrelated kernel\n")
    assert bench.trace_build_hash() == h0
    with open(fake / "software-path-tracer_b200" / "csrc" / "bvh8_traverse.cuh", "a") as f:
        f.write("// the traversal changed\n")
    assert bench.trace_build_hash() != h0
    cap = os.path.join(ROOT, "profiles", "trace_kernel_ncu.json")
    if os.path.exists(cap):
        j = json.load(open(cap))
        assert "build_hash" in j and "commit" in j and "c3" in j and j["c3"]["dram_bytes_per_ray"] > 0

"""The C++ host side: CUDAPathTracer behind the reference's unchanged render::PathTracer interface, driven by a
headless stand-in for src/App.cpp next to the reference's own CPUPathTracer (tests/cpp/headless_compare.cpp).
The binary is built where /root/reference is mounted (`make host`) and travels to the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "headless_compare")


@pytest.mark.gpu
def test_cuda_backend_is_a_drop_in_for_the_reference_backend():
    if not os.path.exists(EXE):
        pytest.skip("build/headless_compare not built (needs /root/reference at build time)")
    out = subprocess.run([EXE, "640", "360", "16"], capture_output=True, text=True, timeout=600)
    print(out.stdout, out.stderr[-2000:])
    assert out.returncode == 0 and "OK" in out.stdout
    assert "CPU Path Tracer (Embree)" in out.stdout and "CUDA Path Tracer (sm_100a)" in out.stdout
    # both backends came out of the REGISTERED factory (PathTracer.cpp:11-21 + the GPU_CUDA case of
    # host/reference_glue/register_gpu_cuda.patch): enumerators 0 and 3, the default branch still throws
    assert "factory: create_path_tracer(CPU_EMBREE) -> type 0, create_path_tracer(GPU_CUDA) -> type 3, GPU_OPTIX throws: yes" in out.stdout


@pytest.mark.gpu
def test_adapter_options_equal_the_same_render_through_the_c_abi(pt, scenes):
    """CUDAPathTracer::Options (honour RenderSettings: 8 bounces, roulette depth 3, exposure 2; 4 samples per
    render() call): the accumulation and the image the adapter returns are those of the same settings applied
    through the C ABI directly."""
    import re
    import numpy as np
    if not os.path.exists(EXE):
        pytest.skip("build/headless_compare not built (needs /root/reference at build time)")
    out = subprocess.run([EXE, "256", "144", "2"], capture_output=True, text=True, timeout=600)
    m = re.search(r"options: (\d+)x(\d+), (\d+) render\(\) calls -> (\d+) frames \(device (\d+)\), accum fnv ([0-9a-f]{16}), image fnv ([0-9a-f]{16})", out.stdout)
    assert m, out.stdout + out.stderr[-2000:]
    W, H, calls, frames, dev_frames = (int(m.group(k)) for k in range(1, 6))
    assert (W, H, calls, frames, dev_frames) == (128, 72, 3, 12, 12)

    def fnv(a):
        h = 1469598103934665603
        for b in np.ascontiguousarray(a).view(np.uint8).ravel().tolist():
            h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return f"{h:016x}"

    sph = scenes.app_spheres()
    ids = np.arange(2, 2 + len(sph), dtype=np.uint32)
    with pt.PathTracer(pt.default_config(max_bounces=8, rr_start=3)) as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(sph, ids)
        tr.set_postprocess(pt.default_postprocess(exposure=2.0))
        for _ in range(3):
            tr.render(4)
        acc, fc = tr.read_accum()
        img = tr.resolve_rgba8()
    assert fc == 12 and fnv(acc) == m.group(6) and fnv(img) == m.group(7)
    # and the options do change the picture: the reference's literals (4 bounces, exposure 1) give another image
    with pt.PathTracer() as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(sph, ids)
        tr.render(12)
        assert fnv(tr.resolve_rgba8()) != m.group(7)


def test_factory_registration_patch_is_the_two_line_delta():
    """CPU check (needs the reference tree, i.e. runs in the build container): the patch applies cleanly to the
    reference's PathTracer.h / PathTracer.cpp and adds exactly the enumerator, the include and the case."""
    ref = "/root/reference/libs/render"
    patch = os.path.join(ROOT, "software-path-tracer_b200", "host", "reference_glue", "register_gpu_cuda.patch")
    text = open(patch).read()
    added = [l[1:].strip() for l in text.splitlines() if l.startswith("+") and not l.startswith("+++")]
    removed = [l[1:].strip() for l in text.splitlines() if l.startswith("-") and not l.startswith("---")]
    assert added == ["GPU_METAL,", "GPU_CUDA", '#include "engines/pathtracer/backends/cuda/CUDAPathTracer.h"',
                     "case BackendType::GPU_CUDA:", "return std::make_unique<render::CUDAPathTracer>();"]
    assert removed == ["GPU_METAL"]
    if not os.path.isdir(ref):
        pytest.skip("reference tree not mounted")
    r = subprocess.run(["patch", "-p1", "--dry-run", "-d", "/root/reference", "-i", patch], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_adapter_sources_mirror_the_reference_interface():
    """CPU check: the adapter overrides exactly the eight pure virtuals of render::PathTracer."""
    hdr = open(os.path.join(ROOT, "software-path-tracer_b200", "host", "CUDAPathTracer.h")).read()
    for name in ["void render() override", "set_scene(", "set_settings(", "get_scene() const override",
                 "get_settings() const override", "get_backend_name() const override",
                 "get_backend_type() const override", "get_render_result() override"]:
        assert name in hdr, name


@pytest.mark.gpu
def test_cli_writes_the_reference_image(built, pt, oracle, scenes, tmp_path):
    """tools/cudapt_cli.c (plain C over the C ABI): the App scene at 96x54x3 spp must give the reference's image."""
    import subprocess
    exe = os.path.join(ROOT, "build", "cudapt_cli")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", ROOT, "build/cudapt_cli"], check=True, capture_output=True)
    out = tmp_path / "app.ppm"
    r = subprocess.run([exe, "--size", "96x54", "--spp", "3", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = out.read_bytes()
    head = b"P6\n96 54\n255\n"
    assert raw.startswith(head)
    import numpy as np
    rgb = np.frombuffer(raw[len(head):], np.uint8).reshape(54, 96, 3)
    want, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(96, 54), scenes.app_spheres(), 3)
    w = oracle.twin_resolve_rgba8(want, 3)
    want_rgb = np.stack([(w >> 24) & 0xFF, (w >> 16) & 0xFF, (w >> 8) & 0xFF], -1).astype(np.uint8)
    assert (rgb == want_rgb).all(axis=-1).mean() >= 0.9999
    # an OBJ through the mesh integrator, tone-mapped
    obj = tmp_path / "quad.obj"
    obj.write_text("v -2 -2 5\nv 2 -2 5\nv 2 2 5\nv -2 2 5\nf 1 2 3 4\n")
    out2 = tmp_path / "quad.pfm"
    r = subprocess.run([exe, "--obj", str(obj), "--size", "32x18", "--spp", "4", "--out", str(out2)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert out2.read_bytes().startswith(b"PF\n32 18\n-1.0\n")

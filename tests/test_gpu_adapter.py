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


def test_adapter_sources_mirror_the_reference_interface():
    """CPU check: the adapter overrides exactly the eight pure virtuals of render::PathTracer."""
    hdr = open(os.path.join(ROOT, "software-path-tracer_b200", "host", "CUDAPathTracer.h")).read()
    for name in ["void render() override", "set_scene(", "set_settings(", "get_scene() const override",
                 "get_settings() const override", "get_backend_name() const override",
                 "get_backend_type() const override", "get_render_result() override"]:
        assert name in hdr, name

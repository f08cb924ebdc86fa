"""The C-ABI library loads and exports every symbol include/cudapt.h declares (no compute calls)."""
import ctypes
import os
import re
import subprocess

import pytest


def header_symbols(path):
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(cudapt_[a-z0-9_]+)\s*\(", txt)))


def test_header_matches_export_list(pt):
    assert header_symbols(pt.HEADER_PATH) == sorted(pt.EXPORTS)


def test_library_exports_every_symbol(pt):
    lib = pt.load_library()
    for name in pt.EXPORTS:
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", pt.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (cudapt_[a-z0-9_]+)", out))
    assert exported == set(pt.EXPORTS)
    assert lib.cudapt_abi_version() == 1


def test_struct_layouts_match_header(pt):
    """ctypes mirrors vs the C structs: compile a tiny C program that prints sizeof/offsetof."""
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "cudapt.h"
int main(void) {
  printf("%zu %zu %zu %zu\n", sizeof(cudapt_config), sizeof(cudapt_camera), sizeof(cudapt_stats), sizeof(cudapt_material));
  printf("%zu %zu %zu\n", offsetof(cudapt_config, samples_per_pass), offsetof(cudapt_stats, frame_count), offsetof(cudapt_material, param));
  return 0;
}'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.run(["gcc", "-std=c99", "-I", os.path.dirname(pt.HEADER_PATH), c, "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()
    sizes = [int(x) for x in out]
    assert sizes[0] == ctypes.sizeof(pt.Config)
    assert sizes[1] == ctypes.sizeof(pt.Camera)
    assert sizes[2] == ctypes.sizeof(pt.Stats)
    assert sizes[3] == pt.MATERIAL_DTYPE.itemsize == 32
    assert sizes[4] == pt.Config.samples_per_pass.offset
    assert sizes[5] == pt.Stats.frame_count.offset
    assert sizes[6] == 28


def test_defaults_are_the_reference_literals(pt):
    c = pt.default_config()
    assert (c.rng_mode, c.max_bounces, c.rr_start) == (0, 4, 2)       # CPT.cpp:199,264
    import numpy as np
    assert np.float32(c.tnear) == np.float32(0.001) and np.float32(c.eps) == np.float32(1e-4)
    assert np.float32(c.albedo) == np.float32(0.7) and c.samples_per_pass == 1
    cam = pt.default_camera()
    assert list(cam.origin) == [0, 0, 0] and list(cam.forward) == [0, 0, 1] and cam.tan_half_fov == 1.0


def test_no_cpu_fallback(pt):
    """Without a GPU, creating a context must fail loudly (status CUDA), never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pt.CudaptError) as e:
        pt.PathTracer()
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)


def test_product_never_touches_the_oracle():
    """No source under the product package includes, imports, links or dlopens anything under oracle/."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "software-path-tracer_b200")
    bad = re.compile(r'#include\s*[<"][^>"]*oracle|import\s+oracle|from\s+oracle|libtwin|libref_render|oracle/_ref|twin\.h')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not bad.search(txt), f
    mk = open(os.path.join(root, "Makefile")).read()
    link = [l for l in mk.splitlines() if "-shared" in l]
    assert link and all("oracle" not in l for l in link)


def test_look_at_reproduces_the_default_camera_and_rejects_degenerate_input(pt):
    """cudapt_camera_look_at mirrors the reference's placeholder Camera (Camera.h:14-20,33-36); its defaults are the
    hard-coded camera of CPT.cpp:62-73.  Pure host arithmetic: no GPU needed."""
    import numpy as np
    d = pt.default_camera()
    c = pt.look_at((0, 0, 0), (0, 0, 1), (0, 1, 0), 90.0)
    for f in ("origin", "right", "up", "forward"):
        assert list(getattr(c, f)) == list(getattr(d, f))
    assert c.tan_half_fov == d.tan_half_fov == 1.0
    c = pt.look_at((1, 2, 3), (4, -2, 9), (0, 1, 0), 60.0)
    r, u, f = (np.array(list(getattr(c, k)), np.float64) for k in ("right", "up", "forward"))
    assert abs(np.linalg.norm(r) - 1) < 1e-6 and abs(np.linalg.norm(u) - 1) < 1e-6 and abs(np.linalg.norm(f) - 1) < 1e-6
    assert abs(r @ u) < 1e-6 and abs(r @ f) < 1e-6 and abs(u @ f) < 1e-6
    assert np.allclose(np.cross(u, f), r, atol=1e-6)  # right-handed like the default (right = up x forward)
    assert abs(c.tan_half_fov - np.tan(np.radians(30.0))) < 1e-6
    import pytest
    for bad in (((0, 0, 0), (0, 0, 0), (0, 1, 0), 90.0), ((0, 0, 0), (0, 2, 0), (0, 1, 0), 90.0), ((0, 0, 0), (0, 0, 1), (0, 1, 0), 180.0)):
        with pytest.raises(ValueError):
            pt.look_at(*bad)

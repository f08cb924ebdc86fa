"""Generates tests/golden/c2_cornell_64x36.npz from the Tier-2 oracle (the CPU twin, oracle/twin/twin.cpp).

The reference cannot render triangles (CPUPathTracer.cpp:367-398), so these vectors are NOT pinned by the reference:
they freeze the twin's mesh specification so that an accidental change of the twin or of the CUDA
path shows up against committed numbers.  Run:  python tests/golden/make_golden_mesh.py
Scene = scenes.cornell_box(sphere_res=6) (BASELINE.json configs[1] at reduced tessellation), 64x36, Philox.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

scenes = importlib.import_module("software-path-tracer_b200.scenes")

W, H = 64, 36
KW = dict(rng_mode=1, max_bounces=6, jitter=1, nee=1, sky_scale=0.0, seed=17, cam_tan_half_fov=0.36)
oracle.build()
v, t, m, mats = scenes.cornell_box(sphere_res=6)
mesh = oracle.TwinMesh(v, t, m, mats)
acc4, idx, pt, c4 = mesh.render(oracle.twin_params(W, H, **KW), 4, want_primary=True)
acc32, _, _, c32 = mesh.render(oracle.twin_params(W, H, **KW), 32)
lens32, _, _, _ = mesh.render(oracle.twin_params(W, H, cam_lens_radius=0.05, cam_focus_distance=4.0, **KW), 32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c2_cornell_64x36.npz"), vertices=v, triangles=t, material_id=m,
                    materials=mats, width=W, height=H, accum4=acc4, accum32=acc32, accum32_lens=lens32, primary_idx=idx,
                    primary_t=pt, rays4=np.uint64(c4.rays), rays32=np.uint64(c32.rays))
print("wrote mesh golden: rays", c4.rays, c32.rays)

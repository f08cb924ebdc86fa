"""Generates tests/golden/c1_spheres_96x54.npz from the Tier-1 oracle, i.e. from the reference's own
CPUPathTracer.cpp compiled unmodified against the Embree stand-in (oracle/_ref/libref_render.so).

Needs /root/reference (only present in the build container).  Run:  python tests/golden/make_golden.py
Scene = src/App.cpp:101-122 (38 spheres); 96x54 pixels (the 16:9 aspect of config 1); frames 1..4 and 1..16.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import importlib

import oracle  # noqa: E402

scenes = importlib.import_module("software-path-tracer_b200.scenes")

W, H = 96, 54
oracle.build()
sph = scenes.app_spheres()
out = {"spheres": sph, "width": W, "height": H}
s = oracle.RefSession(sph, W, H)
s.render(4)
acc4, rgba4, fc = s.read()
assert fc == 4
idx, t = s.primary()
s.render(12)
acc16, rgba16, fc = s.read()
assert fc == 16
out.update(accum4=acc4, rgba4=rgba4, accum16=acc16, rgba16=rgba16, primary_idx=idx, primary_t=t,
           rays16=np.uint64(s.rays))
# RNG known answers: raw PCG words of CPT.cpp:294-301 for a few (x, y, W, frame)
keys = [(0, 0, 640, 1), (17, 5, 640, 1), (639, 359, 640, 16), (3, 2, 96, 7)]
words = np.stack([oracle.pcg_stream(*k, 10)[0] for k in keys])
floats = np.stack([oracle.pcg_stream(*k, 10)[1] for k in keys])
out.update(pcg_keys=np.array(keys, np.uint32), pcg_words=words, pcg_floats=floats)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c1_spheres_96x54.npz"), **out)
print("wrote golden: rays16 =", s.rays)

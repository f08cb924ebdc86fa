#!/usr/bin/env python
"""Phases of the device BVH build on a bench scene (CUDAPT_BUILD_DEBUG=1 prints them); run on the GPU box."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CUDAPT_BUILD_DEBUG"] = "1"
pt = importlib.import_module("software-path-tracer_b200")
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "c3"
wl = dict(bench.WORKLOADS[name], name=name)
scenes = importlib.import_module("software-path-tracer_b200.scenes")
scene = bench.build_scene(scenes, wl)
tr = pt.PathTracer(pt.default_config(**wl["cfg"]))
tr.set_resolution(320, 180)
tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"], build=False)
for k in range(3):
    t0 = time.perf_counter()
    tr.build_accel_device()
    print(f"build {k}: {(time.perf_counter() - t0) * 1e3:.1f} ms", flush=True)

#!/usr/bin/env python
"""Does cudapt_stage_scene overlap the upload with rendering?  (run on the GPU box)"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pt = importlib.import_module("software-path-tracer_b200")
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
wl = dict(bench.WORKLOADS[name], name=name)
scenes = importlib.import_module("software-path-tracer_b200.scenes")
scene = bench.build_scene(scenes, wl)
tr = pt.PathTracer(pt.default_config(**wl["cfg"]))
tr.set_resolution(wl["w"], wl["h"])
tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"])
spp = wl["spp"]
def t(f, n=3):
    f(); tr.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    tr.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
def render_only():
    tr.reset(); tr.render(spp); tr.synchronize()
def reup_render():
    tr.reupload_scene(); tr.render(spp); tr.synchronize()
def render_stage():
    tr.render(spp); tr.stage_scene(); tr.synchronize()
def stage_only():
    tr.stage_scene(); tr.synchronize()
    tr.render(1); tr.synchronize()
print(name, "render only %.2f ms" % t(render_only))
print(name, "reupload + render %.2f ms" % t(reup_render))
print(name, "render + staged upload %.2f ms" % t(render_stage))
t0 = time.perf_counter(); tr.stage_scene(); t1 = time.perf_counter(); tr.synchronize(); 
print("stage call returns after %.2f ms" % ((t1 - t0) * 1e3))

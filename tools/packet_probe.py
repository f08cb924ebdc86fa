#!/usr/bin/env python
"""Camera-ray packet statistics on a bench workload: node / triangle steps per packet, share of packets handed to the
per-ray kernel, and the render time for a list of budgets (run on the GPU box)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pt = importlib.import_module("software-path-tracer_b200")
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 32
budgets = sys.argv[3].split(",") if len(sys.argv) > 3 else ["off", "160"]
wl = dict(bench.WORKLOADS[name], name=name)
scenes = importlib.import_module("software-path-tracer_b200.scenes")
scene = bench.build_scene(scenes, wl)
for b in budgets:
    if b == "off":
        os.environ["CUDAPT_PRIMARY_PACKETS"] = "0"
    else:
        os.environ["CUDAPT_PRIMARY_PACKETS"] = "1"
        os.environ["CUDAPT_PACKET_BUDGET"] = b
    tr = pt.PathTracer(pt.default_config(**dict(wl["cfg"], samples_per_pass=spp)))
    tr.set_camera(pt.default_camera(tan_half_fov=wl["fov"]))
    tr.set_resolution(wl["w"], wl["h"])
    tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"], device_build=os.environ.get("PROBE_DEVICE_BUILD") == "1")
    print(f"   accel build {tr.stats().accel_build_ms:.1f} ms, {tr.stats().n_nodes} nodes", flush=True)
    for counting in (False, True):
        tr.set_counting(counting)
        for _ in range(2):
            tr.reset()
            tr.render(spp)
        st = tr.stats()
        if not counting:
            print(f"{name} spp {spp} budget {b}: render {st.last_render_ms:.3f} ms, trace launches {st.last_trace_ms:.3f} ms, packets {st.last_packet_ms:.3f} ms", flush=True)
            print(f"closest-hit rays traced by the per-ray kernel in one pass: {st.rays - st.shadow_rays - st.packet_rays}", flush=True)
        else:
            npk = max(st.packets + st.packet_bails, 1)
            print(f"   per packet: nodes {st.packet_nodes / npk:.2f} tris {st.packet_tris / npk:.2f} bails {st.packet_bails / npk:.4f}"
                  f" | per-ray kernel: nodes/ray {st.nodes_visited / max(st.rays, 1):.2f} tris/ray {st.tris_tested / max(st.rays, 1):.2f}", flush=True)
    tr.close()

#!/usr/bin/env python
"""Small passes (1..16 samples per render() call -- the viewer's use, and a strong-scaled rank's): render time of a bench
workload with the multi-pixel camera-ray packets on and off (CUDAPT_PACKET_MINI), plus the packet counters."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pt = importlib.import_module("software-path-tracer_b200")
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
wl = dict(bench.WORKLOADS[name], name=name)
scenes = importlib.import_module("software-path-tracer_b200.scenes")
scene = bench.build_scene(scenes, wl)
for S in (1, 4, 8, 16):
    for mini in ("0", "1"):
        os.environ["CUDAPT_PACKET_MINI"] = mini
        tr = pt.PathTracer(pt.default_config(**dict(wl["cfg"], samples_per_pass=S)))
        tr.set_camera(pt.default_camera(tan_half_fov=wl["fov"]))
        tr.set_resolution(wl["w"], wl["h"])
        tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"])
        ms = []
        for k in range(6):
            tr.render(S)            # progressive: no reset, as a viewer would
            tr.synchronize()
            ms.append(tr.stats().last_render_ms)
        st = tr.stats()
        tr.set_counting(True)
        tr.reset(); tr.render(S); tr.synchronize()
        c = tr.stats()
        npk = max(c.packets + c.packet_bails, 1)
        print(f"{name} {S:2d} spp/call mini={mini}: render {min(ms[2:]):7.3f} ms ({wl['w'] * wl['h'] * S / min(ms[2:]) / 1e3:7.1f} Msamples/s), "
              f"packets {st.last_packet_ms:6.3f} ms, trace {st.last_trace_ms:6.3f} ms; per packet: nodes {c.packet_nodes / npk:5.2f} "
              f"tris {c.packet_tris / npk:5.2f} handed over {c.packet_bails / npk:.4f}", flush=True)
        tr.close()

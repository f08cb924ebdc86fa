#!/usr/bin/env python
""""size a pass so that its live path state fits the 126 MB L2".  The compacted tile partition
(csrc/kernels.h PassParams::owned_tiles) gives exactly such passes on ONE GPU: share r of `world` interleaved 32x32-tile
shares of the C3 image holds 1/world of the pixels x 64 samples of state.  This probe renders a few shares of every
`world` and extrapolates the step (all shares back to back), next to the state size of one share."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pt = importlib.import_module("software-path-tracer_b200")
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
wl = dict(bench.WORKLOADS[name], name=name)
scenes = importlib.import_module("software-path-tracer_b200.scenes")
scene = bench.build_scene(scenes, wl)
W, H, spp = wl["w"], wl["h"], wl["spp"]
tr = pt.PathTracer(pt.default_config(**wl["cfg"]))
tr.set_camera(pt.default_camera(tan_half_fov=wl["fov"]))
tr.set_resolution(W, H)
tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"])
print(f"{name}: {W}x{H}x{spp} spp; path state 148 B per path", flush=True)
for world in (1, 2, 4, 16, 64, 256):
    shares = list(range(min(world, 6)))
    ms, launches = [], []
    for r in shares:
        tr.set_tile_partition(r, world, 32, 32)
        for k in range(3):          # the third is timed (state allocation, L2 warm as in a steady band loop)
            tr.reset()
            l0 = tr.stats().kernel_launches
            tr.render(spp)
            tr.synchronize()
        st = tr.stats()
        ms.append(st.last_render_ms)
        launches.append(st.kernel_launches - l0)
    state_mb = (W * H / world) * spp * 148 / 1e6
    print(f"world {world:4d}: state of one band {state_mb:9.1f} MB, one band {np.mean(ms):7.3f} ms ({int(np.mean(launches))} launches), "
          f"whole image = {world} bands: {np.mean(ms) * world:8.2f} ms", flush=True)
tr.close()

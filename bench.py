#!/usr/bin/env python
"""bench.py -- Msamples/s (paths/s) of the path-tracing hot path on N B200s, next to the CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4] [--impl b200|reference]

A "step" = one full progressive render of the workload (reset, spp frames of 1 sample/pixel each) on every
rank; at N > 1 each rank renders its own sample range (rank g: global frames g+1, g+1+N, ...; weak scaling:
per-GPU work fixed) and the float accumulation buffers are summed to rank 0 with one NCCL reduce at the
read-back boundary, inside the step.  One JSON line is printed by rank 0 (contract in the task statement).

Timing: per-step CUDA-event pairs recorded on the stream libcudapt launches on; L2 is flushed (256 MiB memset)
between steps, outside the event pairs; the reported time is the max over ranks.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (description, width, height, spp, config overrides, camera tan_half_fov)
    "c1": dict(desc="configs[0]: App.cpp 38-sphere scene, 640x360, 16 spp, reference constants (PCG, 4 bounces)",
               w=640, h=360, spp=16, cfg=dict(), fov=1.0, kind="spheres"),
    "c2": dict(desc="configs[1]: Cornell box mesh (diffuse+mirror+dielectric), 1280x720, 256 spp, area-light NEE, depth 8",
               w=1280, h=720, spp=256, cfg=dict(rng_mode=1, max_bounces=8, sky_scale=0.0, jitter=1, nee=1, seed=2,
                                                samples_per_pass=256), fov=0.36, kind="mesh"),
    "c3": dict(desc="configs[2]: 1,002,528-triangle fBm height field, GGX-coated, 1920x1080, 64 spp, depth 8",
               w=1920, h=1080, spp=64, cfg=dict(rng_mode=1, max_bounces=8, jitter=1, seed=3, samples_per_pass=64),
               fov=1.0, kind="mesh"),
    "c4": dict(desc="configs[3]: 20,000,002-triangle flattened rock field (BVH > L2), 1920x1080, 128 spp, depth 16",
               w=1920, h=1080, spp=128, cfg=dict(rng_mode=1, max_bounces=16, jitter=1, seed=4, samples_per_pass=128),
               fov=1.0, kind="mesh"),
    "c5": dict(desc="configs[4]: the 1,002,528-triangle scene at 3840x2160, 1024 spp IN TOTAL sharded over the ranks "
                    "(sample ranges), accumulation combined at the read-back boundary",
               w=3840, h=2160, spp=1024, cfg=dict(rng_mode=1, max_bounces=8, jitter=1, seed=3, samples_per_pass=64),
               fov=1.0, kind="mesh", strong=True),
}
DEFAULT_WORKLOAD = "c3"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def build_scene(scenes, wl):
    if wl["kind"] == "spheres":
        return dict(spheres=scenes.app_spheres())
    name = wl["name"]
    if name == "c2":
        v, t, m, mats = scenes.cornell_box()
    elif name in ("c3", "c5"):
        v, t, m, mats = scenes.heightfield()
    else:
        v, t, m, mats = scenes.rock_field()
    return dict(v=v, t=t, m=m, mats=mats)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                mx.append(float(f[1]))
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[0]))
                    for n, val in zip(names, f[3:7]):
                        if val.lower().startswith("active"):
                            reasons.add(n)
            except ValueError:
                continue
        if not sm and self.rows:  # region shorter than the sampling period: take the nearest sample
            try:
                sm = [float(self.rows[-1][1].split(",")[0])]
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU implementation of the path (oracle/_ref) or, for the mesh workloads the
# reference cannot render (CPUPathTracer.cpp:367-398: spheres only), the CPU twin on all host threads.
# ------------------------------------------------------------------------------------------------------
def cpu_sample(wl, scene, budget_s=12.0, full=False):
    """One CPU sample of the workload AT THE WORKLOAD'S OWN CONFIGURATION (scene, resolution, samples per pixel, depth,
    materials): the whole image when `full` (the reference arm), else the largest interleaved-tile share of it that
    the probe predicts to fit `budget_s`.  Returns a dict: samples, seconds, rays, kind, cores, sample (description),
    share (fraction of the image's pixels that was rendered)."""
    oracle = importlib.import_module("oracle")
    W, H = wl["w"], wl["h"]
    if wl["kind"] == "spheres":
        if oracle.ref_available():
            s = oracle.RefSession(scene["spheres"], W, H)
            s.render(1)  # first call builds the scene (rebuild_scene); not timed, as on the GPU side
            r0 = s.rays
            frames = wl["spp"]
            sec = s.render(frames, True)  # render() + get_render_result() per frame, as App::run does
            n, rays = W * H * frames, s.rays - r0
            s.close()
            return dict(n=n, sec=sec, rays=rays, kind="reference", cores=1, share=1.0,
                        sample=(f"unmodified reference CPUPathTracer + Embree stand-in, {W}x{H}, {frames} frames incl. "
                                "get_render_result() each, 1 thread (the reference render loop is single-threaded)"))
        p = oracle.twin_params(W, H, threads=1)
        t0 = time.time()
        _, _, _, c = oracle.twin_render_spheres(p, scene["spheres"], wl["spp"])
        return dict(n=W * H * wl["spp"], sec=time.time() - t0, rays=c.rays, kind="port", cores=1, share=1.0,
                    sample=f"CPU twin, {W}x{H}, {wl['spp']} frames, 1 thread")
    # Mesh workloads: the reference cannot render triangles (CPUPathTracer.cpp:367-398), so the CPU arm is the twin --
    # the reference's loop around the mesh spec -- on all host threads, at the workload's resolution and sample count.
    cores = os.cpu_count() or 1
    mesh = scene.get("_twin_mesh")
    if mesh is None:
        mesh = oracle.TwinMesh(scene["v"], scene["t"], scene["m"], scene["mats"])
        scene["_twin_mesh"] = mesh
    spp = wl["spp"]
    cfg = dict(wl["cfg"])
    cfg.pop("samples_per_pass", None)

    def run(world):
        p = oracle.twin_params(W, H, cam_tan_half_fov=wl["fov"], part_rank=0, part_world=world, tile_w=8, tile_h=8, **cfg)
        t0 = time.time()
        _, _, _, c = mesh.render(p, spp)
        return c, time.time() - t0

    world = 1
    if not full:
        c, probe = run(256)                              # 1/256 of the 8x8 tiles, spread over the whole image
        full_s = probe * (W * H * spp) / max(c.paths, 1)
        while world < 256 and full_s / world > budget_s:
            world *= 2
    c, sec = run(world)
    return dict(n=int(c.paths), sec=sec, rays=int(c.rays), kind="port", cores=cores, share=1.0 / world,
                sample=(f"CPU twin (the reference's loop, CPUPathTracer.cpp:43-85,197-284, around the mesh spec; own binary BVH), same "
                        f"scene, {W}x{H}, {spp} spp, depth {cfg.get('max_bounces', 4)}: " +
                        ("the whole image" if world == 1 else f"1/{world} of the image's 8x8 tiles (interleaved over the whole image)") +
                        f", {cores} threads"))


def run_reference_arm(args, wl, scene, rank, world):
    """The reference's CPU path on this box's host cores, on the GPU arm's configuration: every step renders the WHOLE
    image at the workload's sample count.  Such a step takes seconds (c3) to a minute (c4), so the number of steps
    actually timed is bounded by a wall-clock budget and reported as `steps_timed`."""
    if rank != 0:
        return
    budget_s = float(os.environ.get("BENCH_REFERENCE_BUDGET_S", "150"))
    t_begin = time.time()
    last = None
    if args.warmup > 0:
        last = cpu_sample(wl, scene, full=True)       # one untimed warm-up step (threads, page cache, BVH build)
    vals, secs = [], []
    for i in range(max(1, args.steps)):
        last = cpu_sample(wl, scene, full=True)
        vals.append(last["n"])
        secs.append(last["sec"])
        mean = sum(secs) / len(secs)
        if time.time() - t_begin + mean > budget_s:
            break
    value = sum(vals) / sum(secs) / 1e6
    line = {
        "impl": "reference", "metric": "Msamples/s (paths/s)", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "steps_timed": len(secs), "ms_per_step": 1e3 * sum(secs) / len(secs),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "name": wl["name"], "width": wl["w"], "height": wl["h"], "spp_per_gpu": wl["spp"],
                   "same_config_as_gpu_arm": True, "share_of_image_per_step": last["share"],
                   "partition": "host CPU only (rank 0)",
                   "note": f"whole steps of the stated configuration; {len(secs)} of the {args.steps} requested steps timed "
                           f"(wall-clock budget {budget_s:.0f} s, BENCH_REFERENCE_BUDGET_S)"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": last["cores"], "kind": last["kind"], "sample": last["sample"],
                         "mrays_per_s": last["rays"] / last["sec"] / 1e6},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true", help="N > 1: skip the check of the assembled image after the timed regions")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (debug)")
    ap.add_argument("--spp-pass", type=int, default=0, help="override samples per wavefront pass (debug)")
    ap.add_argument("--partition", default="tiles", choices=["samples", "tiles", "combined"],
                    help="N > 1: how a step's work is split over the ranks (SURVEY 8e): interleaved 32x32 tiles (A, default: "
                         "per-pixel sums in single-GPU order, so the assembled image is BIT-IDENTICAL to one GPU's; the slots of "
                         "a pass are compacted to the rank's own tiles, csrc/kernels.h), sample ranges (B: every rank renders "
                         "the whole image for its own frames), or tile groups x sample ranges")
    ap.add_argument("--tile", type=int, default=32, help="N > 1, tile partitions: edge of the square interleaved tiles in pixels (multiple of 8)")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling for any workload: the workload's samples per pixel are the TOTAL over the ranks (c5 always is)")
    ap.add_argument("--tail", default="fused", choices=["fused", "nccl"],
                    help="N > 1 read-back tail: 'fused' = one reduce-scatter+resolve+gather kernel per rank over NVLink "
                         "peer memory (cudapt_reduce_resolve_peers); 'nccl' = NCCL reduce of the float buffer to rank 0")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = dict(WORKLOADS[args.workload], name=args.workload)
    if args.spp:
        wl["spp"] = args.spp
    if args.spp_pass:
        wl["cfg"] = dict(wl["cfg"], samples_per_pass=args.spp_pass)
    strong = bool(wl.get("strong")) or args.strong
    if strong:  # total work fixed: every rank renders its share of the samples
        wl["spp"] = max(1, wl["spp"] // world)
    scenes = importlib.import_module("software-path-tracer_b200.scenes")

    if args.impl == "reference":
        if rank == 0:
            run_reference_arm(args, wl, build_scene(scenes, wl), rank, world)
        return

    import torch
    import torch.distributed as dist
    pt = importlib.import_module("software-path-tracer_b200")
    part = importlib.import_module("software-path-tracer_b200.partition")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene = build_scene(scenes, wl)
    W, H, spp = wl["w"], wl["h"], wl["spp"]
    tile_rank, tile_world, group, n_groups = part.layout(rank, world, args.partition)
    if tile_world > 1 and wl["kind"] == "mesh":
        # a tile rank renders tile_world x more frames of 1/tile_world of the pixels: with the slots compacted to its own
        # tiles (csrc/kernels.h) the same memory holds tile_world x more samples per pass
        per_pass = int(wl["cfg"].get("samples_per_pass", 1))
        wl["cfg"] = dict(wl["cfg"], samples_per_pass=min(1024, per_pass * tile_world))
    cfg = pt.default_config(device=local_rank, **wl["cfg"])
    tr = pt.PathTracer(cfg)
    tr.set_camera(pt.default_camera(tan_half_fov=wl["fov"]))
    tr.set_resolution(W, H)
    if wl["kind"] == "spheres":
        tr.upload_spheres(scene["spheres"])
        scene_bytes = scene["spheres"].nbytes + 4 * len(scene["spheres"])
    else:
        tr.upload_mesh(scene["v"], scene["t"], scene["m"], scene["mats"])
        scene_bytes = int(tr.stats().accel_bytes)
    # Work split.  Every mode gives a rank W*H*spp paths per step; they differ in WHICH (pixel, frame) pairs.
    #   samples : rank g renders frames g+1, g+1+G, ... of every pixel                      -> render(spp)
    #   tiles   : rank g renders all G*spp frames of the 32x32 tiles it owns               -> render(spp * G)
    #   combined: T tile ranks x G/T sample groups                                          -> render(spp * T)
    frames_per_render = spp * tile_world
    if world > 1:
        if n_groups > 1:
            tr.set_frame_range(*part.frame_range(group, n_groups))
        if tile_world > 1:
            tr.set_tile_partition(tile_rank, tile_world, args.tile, args.tile)

    stream = torch.cuda.ExternalStream(tr.stream_handle, device=local_rank)
    n_px = W * H
    accum_t = None
    if world > 1:
        # zero-copy view of libcudapt's accumulation buffer for the NCCL reduce
        class _Cai:
            __cuda_array_interface__ = {"shape": (n_px * 4,), "typestr": "<f4", "data": (tr.accum_device_ptr, False),
                                        "version": 2}
        accum_t = torch.as_tensor(_Cai(), device=f"cuda:{local_rank}")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local_rank}")

    # fused tail: exchange CUDA IPC handles once, map every peer's accumulation buffer and the root's outputs
    fused = world > 1 and args.tail == "fused"
    peer_accum, root_rgba, root_accum, tick = None, None, None, None
    if fused:
        handles = [None] * world
        dist.all_gather_object(handles, tr.ipc_export())
        peer_accum = [tr.accum_device_ptr if g == rank else tr.ipc_open(handles[g][0]) for g in range(world)]
        root_rgba = tr.rgba_device_ptr if rank == 0 else tr.ipc_open(handles[0][1])
        root_accum = peer_accum[0]
        tick = torch.zeros(1, dtype=torch.int32, device=f"cuda:{local_rank}")

    def tail():
        """Read-back boundary of a step, enqueued on libcudapt's stream (no host synchronisation)."""
        if world == 1:
            return
        with torch.cuda.stream(stream):
            if fused:
                dist.all_reduce(tick)   # stream-ordered rendezvous: every rank's render is complete
                tr.reduce_resolve_peers(peer_accum, rank, spp * world, root_rgba, None)  # spp * world frames per pixel in every mode; image only: the merged floats stay distributed
                dist.all_reduce(tick)   # every rank's slice has landed in the root's buffers
            else:
                part.reduce_accum(accum_t, dst=0)

    def step():
        tr.reset()
        tr.render(frames_per_render)
        tail()

    def sync_all():
        tr.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def check_multi_gpu_image():
        """Renders one more step and compares what the N ranks assembled on rank 0 with (a) a rank-ordered HOST merge of
        the ranks' own float buffers (the fused tail adds in rank order: bit-identical; NCCL's order is its own:
        allclose) and (b) ONE context on rank 0 rendering the same global frames of the whole image: bit-identical for
        the tile partition (every pixel is summed by one rank in single-GPU order), equal up to float re-association
        for sample ranges.  Returns the `parity` block of the JSON line."""
        total = spp * world
        bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
        tr.reset()
        tr.render(frames_per_render)
        sync_all()
        mine, _ = tr.read_accum()                                   # this rank's own buffer, before any merge
        tail()
        sync_all()
        rgba = merged = None
        if rank == 0:
            if fused:
                rgba = tr.download_rgba8()
            else:
                tr.set_frame_count(total)
                rgba = tr.resolve_rgba8()
                merged, _ = tr.read_accum()
        if fused:  # the float sums too: the same kernel with root_accum = rank 0's buffer (slices are disjoint)
            with torch.cuda.stream(stream):
                dist.all_reduce(tick)
                tr.reduce_resolve_peers(peer_accum, rank, total, root_rgba, root_accum)
                dist.all_reduce(tick)
            sync_all()
            if rank == 0:
                merged, _ = tr.read_accum()
                rgba2 = tr.download_rgba8()
        mine_t = torch.from_numpy(mine).to(f"cuda:{local_rank}")
        parts = [torch.empty_like(mine_t) for _ in range(world)] if rank == 0 else None
        dist.gather(mine_t, parts, dst=0)
        torch.cuda.synchronize()
        out = None
        if rank == 0:
            host = parts[0].cpu().numpy().copy()
            for g in range(1, world):
                host += parts[g].cpu().numpy()                       # float32, rank order, as the fused kernel adds
            del parts
            same_merge = bool(np.array_equal(bits(merged), bits(host)))
            close_merge = bool(np.allclose(merged, host, rtol=2e-5, atol=1e-6))
            # the image of the host merge through the library's own resolve (tests/ pin it to the reference's arithmetic)
            tr.set_tile_partition(0, 1, 32, 32)
            tr.set_frame_range(1, 1)
            tr.write_accum(host, total)
            rgba_host = tr.resolve_rgba8()
            img_vs_merge = float((rgba == rgba_host).mean())
            if fused:
                img_vs_merge = min(img_vs_merge, float((rgba2 == rgba_host).mean()))
            # one context, the whole image, global frames 1..total
            tr.reset()
            tr.render(total)
            single, fc = tr.read_accum()
            rgba_single = tr.resolve_rgba8()
            same_single = float(np.all(bits(merged) == bits(single), axis=-1).mean())
            denom = np.maximum(np.abs(single), 1e-3)
            max_rel = float((np.abs(merged - single) / denom).max())
            img_vs_single = float((rgba == rgba_single).mean())
            chan = lambda w: np.stack([(w >> sh) & 0xFF for sh in (24, 16, 8, 0)], -1).astype(np.int32)
            max_lsb = int(np.abs(chan(rgba) - chan(rgba_single)).max())
            tiles_only = n_groups == 1
            ok = (same_merge and img_vs_merge == 1.0) if fused else (close_merge and img_vs_merge >= 0.9999)
            if tiles_only:
                ok = ok and same_single == 1.0 and img_vs_single == 1.0
            else:
                ok = ok and max_rel <= 2e-5 and img_vs_single >= 0.999 and max_lsb <= 1
            out = {"mode": args.partition, "tail": args.tail, "frames_per_pixel": int(total), "ok": bool(ok),
                   "float_sums_bit_identical_to_rank_ordered_host_merge": same_merge,
                   "rgba8_equal_to_resolve_of_host_merge": img_vs_merge,
                   "pixels_bit_identical_to_single_context": same_single, "max_rel_diff_vs_single_context": max_rel,
                   "rgba8_equal_to_single_context": img_vs_single, "rgba8_max_channel_difference_lsb": max_lsb,
                   "bar": ("tiles: float sums and RGBA8 bit-identical to ONE context rendering the same frames" if tiles_only else
                           "sample ranges: float sums within 2e-5 relative of ONE context rendering the same frames (re-association of "
                           f"{total} adds), RGBA8 within 1 LSB per channel everywhere and equal on >= 99.9 % of the words (a sum that lands "
                           "within 4e-6 relative of an 8-bit truncation boundary may flip it)") +
                          ("; fused tail bit-identical to the rank-ordered host merge" if fused else "; NCCL reduce allclose to the host merge")}
        sync_all()
        return out

    if wl["kind"] == "mesh":
        # node/triangle visit counts for the roofline, from the instrumented traversal, outside the timed region
        tr.reset()
        tr.set_counting(True)
        tr.render(min(spp, int(wl["cfg"].get("samples_per_pass", 1))))
        cs = tr.stats()
        # nodes_visited / tris_tested are counted by the per-ray kernel only: closest-hit rays it traced + shadow rays
        per_ray = max(cs.rays - cs.packet_rays, 1)
        n_pk = max(cs.packets + cs.packet_bails, 1)
        tr._counted = {"nodes_per_ray": cs.nodes_visited / per_ray, "tris_per_ray": cs.tris_tested / per_ray,
                       "packet_nodes": cs.packet_nodes / n_pk, "packet_tris": cs.packet_tris / n_pk,
                       "packet_bail_frac": cs.packet_bails / n_pk if cs.packet_rays or cs.packet_bails else None,
                       "rays_per_packet": cs.packet_rays / max(cs.packets, 1)}
        tr.set_counting(False)
    for _ in range(args.warmup):
        step()
    sync_all()
    launches0 = tr.stats().kernel_launches
    sampler = ClockSampler(local_rank) if rank == 0 else None
    time.sleep(0.25) if sampler else None
    sync_all()  # ranks enter the timed region together (rank 0 has just spent 0.25 s starting the clock sampler)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_wall0 = time.time()
    for k in range(args.steps):
        flush.zero_()                      # L2 flush, on torch's stream, outside the event pair
        torch.cuda.synchronize()
        ev[k][0].record(stream)
        step()
        ev[k][1].record(stream)
    sync_all()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    if os.environ.get("BENCH_DEBUG"):
        print(f"[rank {rank}] step_ms {step_ms}", file=sys.stderr, flush=True)
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    st = tr.stats()
    launches = st.kernel_launches - launches0
    # dominant-kernel time from the library's own event pair around cudapt_render (same stream)
    render_ms = st.last_render_ms
    paths_per_step = n_px * spp
    value = world * paths_per_step * args.steps / (total_ms * 1e-3) / 1e6
    rays_per_path = st.rays / max(st.paths, 1)

    # ---- e2e through the C ABI with host buffers: scene H2D + render + resolve + RGBA8 D2H every step ----
    # The caller's result buffers: page-locked (as the C++ adapter pins RenderResult::image_buffer) and TWO of them,
    # because the read-back is double-buffered -- cudapt_resolve_rgba8_async / cudapt_download_rgba8_async copy image
    # k to the host on the library's download stream while step k+1 renders, as a viewer that shows frame k while
    # frame k+1 renders would.  Every step's image still crosses to the host inside the timed region; the host waits
    # for copy k-1 before it starts copy k, so it never runs more than one step ahead of the GPU.
    host_rgba = [np.empty((H, W), np.uint32) for _ in range(2)]
    for b in host_rgba:
        tr.host_register(b)

    def e2e_step(k=0, first=True, last=True):
        # The step's scene goes host -> device inside the timed region.  Mesh scenes are double-buffered: the copy for
        # step k+1 (cudapt_stage_scene, the library's copy stream) is started as soon as step k's kernels are queued and
        # overlaps them; step k+1's render waits for it and swaps the buffers.  The first step uploads in line.
        if wl["kind"] == "spheres":
            tr.upload_spheres(scene["spheres"])     # H2D of the step's scene; also restarts accumulation
        elif first:
            tr.reupload_scene()
        tr.render(frames_per_render)
        if wl["kind"] == "mesh" and not last:
            tr.stage_scene()                         # next step's scene: H2D concurrent with this step's rendering
        if fused and rank == 0:
            tr.stream_wait_download()                # the peers rewrite the root's image after the rendezvous in tail()
        tail()
        if world > 1:
            tr.set_frame_count(spp * world if rank == 0 else spp)
        if fused:
            if rank == 0:
                tr.wait_download()                             # image k-1 is on the host
                tr.download_rgba8_async(host_rgba[k & 1])      # the image the ranks assembled in the root's buffer
        else:
            tr.wait_download()
            tr.resolve_rgba8_async(host_rgba[k & 1])          # resolve kernel + D2H, not waited for here

    # The e2e loop is timed on the HOST clock, so one scheduling hiccup of the host (a short step makes K steps a few
    # tens of ms) can swamp it.  The K-step loop is therefore run three times when it is short (< 2 s) and the MEDIAN
    # run is reported, with all runs listed in `e2e.runs_ms`; a long loop (c4, c5) runs once.
    e2e_step()
    tr.wait_download()
    e2e_runs = []
    while True:
        sync_all()
        t0 = time.perf_counter()
        e2e_marks = []
        for k in range(args.steps):
            e2e_step(k, first=(k == 0), last=(k == args.steps - 1))   # exactly one scene upload per step
            e2e_marks.append(time.perf_counter())
        tr.wait_download()                                            # the last image is on the host
        sync_all()
        if os.environ.get("BENCH_DEBUG"):
            print(f"[rank {rank}] e2e step s {[round(b - a, 4) for a, b in zip([t0] + e2e_marks[:-1], e2e_marks)]}", file=sys.stderr, flush=True)
        e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=f"cuda:{local_rank}")
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)   # every rank sees the same time, so all leave the loop together
        e2e_runs.append(float(e2e_s.item()))
        if len(e2e_runs) >= 3 or e2e_runs[0] >= 2.0:
            break
    e2e_value = world * paths_per_step * args.steps / sorted(e2e_runs)[len(e2e_runs) // 2] / 1e6

    # the read-back tail alone (N > 1): CUDA events on the library's stream, max over ranks
    tail_info = None
    if world > 1:
        sync_all()
        te = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for a, b in te:
            a.record(stream)
            tail()
            b.record(stream)
        sync_all()
        tms = torch.tensor([sorted(a.elapsed_time(b) for a, b in te)[len(te) // 2]], dtype=torch.float64,
                           device=f"cuda:{local_rank}")
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        tail_info = {"kind": args.tail, "median_ms": float(tms.item()), "accum_bytes_per_rank": int(n_px * 16),
                     "note": "fused: 2 stream-ordered 4-byte all-reduces (rendezvous) + one reduce-scatter/resolve/gather "
                             "kernel per rank over NVLink peer memory" if fused else
                             "ncclReduce(sum) of the float accumulation buffer to rank 0 (resolve not included)"}

    # ---- N > 1: the image the ranks assembled is a CHECKED result (CPUPathTracer.cpp:77-80 is what the sharding replaces) ----
    parity = None
    if world > 1 and not args.no_parity_check:
        parity = check_multi_gpu_image()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, peak_src = load_peaks()
    roof = roofline(tr, pt, wl, scene, st, render_ms, hbm_peak, peak_src)
    line = {
        "metric": "Msamples/s (paths/s)", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "name": wl["name"], "width": W, "height": H, "spp_per_gpu": spp,
                   "l2": "flushed between steps (256 MiB memset outside the per-step CUDA-event pairs)",
                   "partition": ("one process per GPU; " + {"samples": "sample ranges per rank", "tiles": f"interleaved {args.tile}x{args.tile} tiles per rank",
                                                            "combined": f"{tile_world} tile ranks x {n_groups} sample groups"}[args.partition] + "; " +
                                 ("fused tail every step: one reduce-scatter + resolve + gather kernel per rank over "
                                  "NVLink peer memory (CUDA IPC), rank-ordered sums" if fused else
                                  "NCCL reduce(sum) of the float accumulation buffer to rank 0 at the end of every step"))
                   if world > 1 else "single GPU"},
        "mrays_per_s": value * rays_per_path, "rays_per_path": rays_per_path,
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(scene_bytes),
                "d2h_bytes_per_step": int(n_px * 4), "runs_ms": [round(1e3 * t, 3) for t in e2e_runs],
                "note": f"median of {len(e2e_runs)} host-clock timed runs of {args.steps} steps each"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
    }
    if tail_info:
        line["tail"] = tail_info
    if parity is not None:
        line["parity"] = parity
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_sample(wl, scene)
        line["cpu_baseline"] = {"value": cb["n"] / cb["sec"] / 1e6, "unit": "Msamples/s", "cores": cb["cores"], "kind": cb["kind"],
                                "sample": cb["sample"], "share_of_image": cb["share"], "mrays_per_s": cb["rays"] / cb["sec"] / 1e6}
    print(json.dumps(line), flush=True)
    for b in host_rgba:
        tr.host_unregister(b)
    tr.close()
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit("bench.py: the image assembled by the ranks FAILED its parity check (see the 'parity' block)")


def roofline(tr, pt, wl, scene, st, render_ms, hbm_peak, peak_src):
    """Algorithmic bytes per launch / CUDA-event duration of the render launch (DESIGN.md 'Roofline')."""
    W, H, spp = wl["w"], wl["h"], wl["spp"]
    n_px = W * H
    if wl["kind"] == "spheres":
        # one launch renders all spp frames: accumulation read + write once per pixel (32 B) + the scene once per CTA
        alg_bytes = 32.0 * n_px
        note = ("sphere megakernel: scene is in shared memory, the only compulsory HBM traffic is the 32 B/pixel "
                "accumulation RMW per launch; the kernel is FP32/FP64-issue bound, not HBM bound -- see 'compute'")
        rays = st.rays / max(st.frame_count, 1) * spp if st.frame_count else 0
        n_sph = len(scene["spheres"])
        flops = (st.rays / max(st.paths, 1)) * (25.0 * n_sph + 12.0) + (st.shade_invocations / max(st.paths, 1)) * 90.0
        achieved = alg_bytes / (render_ms * 1e-3) / 1e9
        # counters of render_spheres_kernel from an `ncu --set full` capture of this build (tools/run_c1.sh ->
        # tools/make_trace_ncu_json.py), stamped with a hash of the kernel's sources like the traversal kernel's
        issue, traffic = None, None
        spath = os.path.join(ROOT, "profiles", "spheres_kernel_ncu.json")
        if os.path.exists(spath):
            cap = json.load(open(spath))
            fresh = cap.get("build_hash") == kernel_build_hash(SPHERES_SOURCES)
            if fresh and cap.get("c1"):
                issue = dict(cap["c1"], capture="profiles/spheres_kernel_ncu.json", commit=cap.get("commit"), build_hash_matches=True)
                traffic = cap["c1"].get("dram_bytes_per_launch")
            else:
                issue = {"capture": "profiles/spheres_kernel_ncu.json", "commit": cap.get("commit"), "build_hash_matches": bool(fresh),
                         "note": "STALE: the kernel's sources changed since the capture; figures withheld"}
        return {"bound": "issue", "kernel": "render_spheres_kernel", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "issue": issue, "peak_source": peak_src,
                "note": note + ". 'achieved' / 'peak' / 'frac' are the contract's algorithmic bytes over the HBM copy peak (0.3 %: "
                               "nothing here is memory bound); 'issue' holds the counters that bind it",
                "compute": {"flops_per_path": flops, "achieved_tflops": flops * n_px * spp / (render_ms * 1e-3) / 1e12,
                            "fp32_peak_tflops": 148 * 128 * 2 * 1.965e9 / 1e12}}
    paths = max(st.paths, 1)
    R = st.rays / paths
    B = st.shade_invocations / paths
    counted = getattr(tr, "_counted", None)
    Nn = counted["nodes_per_ray"] if counted else None
    Nt = counted["tris_per_ray"] if counted else None
    if Nn is None:
        return {"bound": "issue", "achieved": None, "peak": hbm_peak, "unit": "GB/s", "frac": None, "traffic": None,
                "peak_source": peak_src, "note": "node/triangle counters unavailable"}
    # Dominant kernel = trace_kernel<closest> (per-ray BVH traversal).  Its launches of the last render are timed
    # inside libcudapt with CUDA-event pairs on the launch stream (stats.last_trace_ms = their sum).
    # Algorithmic bytes of those launches: per ray 36 B in (ray + queue id) + 8 B out (hit) + Nn nodes x 80 B
    # + Nt triangles x 48 B, over the closest-hit rays the kernel traced = all rays - shadow rays - camera rays
    # finished by packets (primary_packet.cu).  Reported per launch: both divided by the number of trace launches.
    closest_rays = float(st.rays - st.shadow_rays - st.packet_rays)
    per_ray_rays = float(st.rays - st.packet_rays)
    n_packets = float(st.packets)
    pk_bytes = n_packets * (counted["packet_nodes"] * 80.0 + counted["packet_tris"] * 48.0) + st.packet_rays * 24.0
    bytes_per_path = (per_ray_rays * (Nn * 80.0 + Nt * 48.0) + pk_bytes) / paths + B * 160.0 + 32.0
    trace_ms = float(st.last_trace_ms)
    # one closest-hit launch per bounce and pass; a render() call runs ceil(frames / samples_per_pass) passes
    frames = int(st.frame_count) if st.frame_count else spp
    n_launches = max(1, -(-frames // int(wl["cfg"].get("samples_per_pass", 1))) * int(wl["cfg"].get("max_bounces", 4)))
    trace_bytes = closest_rays * (44.0 + Nn * 80.0 + Nt * 48.0)
    achieved = trace_bytes / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else None
    # What ncu measured on THIS build's traversal kernel (profiles/trace_kernel_ncu.json, written by
    # tools/make_trace_ncu_json.py from `ncu --set full` captures): DRAM bytes per ray, issue-slot utilisation, lanes per
    # instruction.  The file is stamped with a hash of the kernel's sources; a stamp that does not match the tree the
    # bench runs from is reported as stale and its figures are withheld.
    traffic, issue = None, None
    tpath = os.path.join(ROOT, "profiles", "trace_kernel_ncu.json")
    if os.path.exists(tpath):
        cap = json.load(open(tpath))
        fresh = cap.get("build_hash") == trace_build_hash()
        t = cap.get(wl["name"] if wl["name"] != "c5" else "c3")
        if t and fresh:
            traffic = t["dram_bytes_per_ray"] * closest_rays / n_launches
            issue = dict(t, capture="profiles/trace_kernel_ncu.json", commit=cap.get("commit"), build_hash_matches=True)
        else:
            issue = {"capture": "profiles/trace_kernel_ncu.json", "commit": cap.get("commit"), "build_hash_matches": bool(fresh),
                     "note": "no capture of this workload" if fresh else "STALE: the traversal kernel's sources changed since the capture; figures withheld"}
    # what the memory system gives a per-thread 16-B gather (tools/microbench_l2.cu run on this pool's B200): the
    # ceiling the algorithmic bytes of an L2-resident BVH can be fetched at, beside the HBM copy peak above
    alg = {"value": achieved, "hbm_copy_peak_GBps": hbm_peak, "over_hbm_copy_peak": achieved / hbm_peak if achieved else None}
    lpath = os.path.join(ROOT, "profiles", "l2_bandwidth.json")
    if os.path.exists(lpath):
        rows = json.load(open(lpath))["rows"]
        def row(mib, pat):
            return next((r["GBps"] for r in rows if r["set_MiB"] == mib and r["pattern"] == pat), None)
        g = row(64, "node80_per_thread")
        alg.update({"l2_record_gather_GBps": g, "over_l2_record_gather": (achieved / g) if achieved and g else None,
                    "l2_stream_GBps": row(64, "stream"),
                    "note": "two denominators for the SAME algorithmic bytes: the HBM copy peak (MEASURED_PEAKS.json) and what an "
                            "L2-resident 80-B record gather with 16-B loads reaches (profiles/l2_bandwidth.json, tools/microbench_l2.cu). "
                            "'over_l2_record_gather' is a RATIO, not a roofline fraction: rays of one warp share nodes through L1, so it "
                            "can exceed 1 -- i.e. no memory roofline binds this kernel"})
    packets = None
    if st.packet_rays:
        packets = {"kernel": "primary_packet_kernel", "share_of_step": st.last_packet_ms / render_ms if render_ms else None,
                   "ms_per_step": st.last_packet_ms, "camera_rays_finished": int(st.packet_rays),
                   "rays_per_packet": counted["rays_per_packet"], "node_steps_per_packet": counted["packet_nodes"], "triangle_steps_per_packet": counted["packet_tris"],
                   "handed_to_per_ray_kernel": counted["packet_bail_frac"],
                   "algorithmic_GBps": pk_bytes / (st.last_packet_ms * 1e-3) / 1e9 if st.last_packet_ms else None,
                   "note": "one warp = 32 samples of a pixel: fused ray generation + shared BVH walk; issue bound "
                           "(profiles/), bytes given for completeness"}
    return {"bound": "issue", "kernel": "trace_kernel<closest>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
            "frac": achieved / hbm_peak if achieved else None, "traffic": traffic,
            "issue": issue, "algorithmic_GBps": alg,
            "algorithmic_bytes_per_launch": trace_bytes / n_launches, "launch_ms": trace_ms / n_launches,
            "launches_per_step": n_launches, "kernel_share_of_step": trace_ms / render_ms if render_ms else None,
            "peak_source": peak_src, "bytes_per_ray": 44.0 + Nn * 80.0 + Nt * 48.0, "bytes_per_path_whole_step": bytes_per_path,
            "whole_step_GBps": bytes_per_path * paths / (render_ms * 1e-3) / 1e9,
            "R": R, "Nn": Nn, "Nt": Nt, "B": B, "rays_traced_by_this_kernel_per_step": closest_rays, "packets": packets,
            "note": "bound = instruction issue (alu pipe) on every mesh workload, the 1.3 GB BVH of c4 included: DRAM carries a few "
                    "per cent of the algorithmic bytes ('traffic' vs 'algorithmic_bytes_per_launch'), the BVH is served by L1/L2. "
                    "'achieved' / 'peak' / 'frac' are the contract's algorithmic-bytes figure over the HBM copy peak and do NOT say "
                    "the kernel is HBM bound; 'issue' holds the counters that do bind it (ncu capture of this build). Nn/Nt: nodes / "
                    "triangles per ray of the per-ray kernel (camera rays finished by packets excluded)"}


TRACE_SOURCES = ("trace_bvh.cu", "bvh8_traverse.cuh", "bvh8.h", "common.cuh", "kernels.h")
SPHERES_SOURCES = ("spheres.cu", "common.cuh", "kernels.h")


def kernel_build_hash(files):
    """sha256 over the sources that make a kernel (and the flags it is compiled with): the stamp of an ncu capture under
    profiles/."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "software-path-tracer_b200", "csrc")
    for f in files:
        h.update(open(os.path.join(csrc, f), "rb").read())
    for line in open(os.path.join(ROOT, "Makefile")):
        if line.startswith("NVFLAGS") or line.startswith("            -"):
            h.update(line.encode())
    return h.hexdigest()[:16]


def trace_build_hash():
    """Stamp of profiles/trace_kernel_ncu.json: the per-ray traversal kernel's sources and flags."""
    return kernel_build_hash(TRACE_SOURCES)


if __name__ == "__main__":
    main()

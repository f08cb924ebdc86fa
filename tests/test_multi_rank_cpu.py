"""world_size-2 (and 3) runs of the N>1 host logic on CPU with the gloo backend: each rank renders its share with
the CPU twin (standing in for its GPU), the accumulation buffers are reduced as bench.py does, and rank 0
compares with the single-rank render."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, mode, out_path):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle = importlib.import_module("oracle")
    part = importlib.import_module("software-path-tracer_b200.partition")
    scenes = importlib.import_module("software-path-tracer_b200.scenes")
    W, H, per = 64, 36, 3
    sph = scenes.app_spheres()
    if mode == "combined":  # bench.py --partition combined: tile ranks inside a group x sample ranges across groups
        tile_rank, tile_world, group, n_groups = part.layout(rank, world, "combined")
        first, stride = part.frame_range(group, n_groups)
        p = oracle.twin_params(W, H, threads=1, part_rank=tile_rank, part_world=tile_world, tile_w=16, tile_h=8)
        acc, _, _, _ = oracle.twin_render_spheres(p, sph, per, first_frame=first, frame_stride=stride)
    elif mode == "frames":
        first, stride = part.frame_range(rank, world)
        p = oracle.twin_params(W, H, threads=1)
        acc, _, _, _ = oracle.twin_render_spheres(p, sph, per, first_frame=first, frame_stride=stride)
    else:
        p = oracle.twin_params(W, H, threads=1, part_rank=rank, part_world=world, tile_w=16, tile_h=8)
        acc, _, _, _ = oracle.twin_render_spheres(p, sph, per)
        own = np.array([[part.owns_tile(rank, world, x, y, W, 16, 8) for x in range(W)] for y in range(H)])
        assert np.array_equal(own, acc[..., 3] > 0)  # python rule == twin rule (== kernel rule, tested on GPU)
    t = torch.from_numpy(acc)
    part.reduce_accum(t, dst=0)
    if rank == 0:
        np.save(out_path, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,mode", [(2, "frames"), (2, "tiles"), (3, "tiles"), (4, "combined")])
def test_reduce_of_rank_shares_matches_single_rank(tmp_path, oracle, scenes, world, mode):
    out = str(tmp_path / "acc.npy")
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, mode, out), nprocs=world, join=True)
    got = np.load(out)
    part = importlib.import_module("software-path-tracer_b200.partition")
    W, H, per = 64, 36, 3
    n = per * 2 if mode == "combined" else part.total_frame_count(per, world, mode)  # 4 ranks = 2 tile ranks x 2 groups
    want, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H, threads=1), scenes.app_spheres(), n)
    assert np.all(got[..., 3] == n)
    if mode == "tiles":
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))  # bit-identical: x + 0 = x
    else:
        assert np.allclose(got, want, rtol=1e-6, atol=1e-5)  # float re-association only


def test_layouts_cover_every_pixel_frame_pair_once():
    part = importlib.import_module("software-path-tracer_b200.partition")
    for world in (1, 2, 4, 8):
        for mode in ("samples", "tiles", "combined"):
            seen = {}
            for r in range(world):
                tr_, tw, g, ng = part.layout(r, world, mode)
                assert 0 <= tr_ < tw and 0 <= g < ng and tw * ng == world
                seen[(tr_, g)] = seen.get((tr_, g), 0) + 1
            assert len(seen) == world and set(seen.values()) == {1}
    assert part.layout(5, 8, "combined") == (1, 4, 1, 2) and part.layout(3, 4, "combined") == (1, 2, 1, 2)


def test_frame_ranges_partition_the_frames():
    part = importlib.import_module("software-path-tracer_b200.partition")
    for world in (1, 2, 4, 8):
        frames = sorted(f for r in range(world) for f in part.frames_of_rank(r, world, 5))
        assert frames == list(range(1, 5 * world + 1))
    with pytest.raises(ValueError):
        part.frame_range(3, 2)

"""GPU vs the CPU twin AT THE SIZES OF BASELINE.json's configs[1..3] -- the configurations every published number of
the mesh path is quoted on -- plus configs[3]'s scene at a reduced size with every camera-ray code path forced.

The reference renders spheres only (CPUPathTracer.cpp:367-398), so the oracle is the twin (oracle/twin/twin.cpp: the
loop of CPUPathTracer.cpp:43-85,197-284 around the mesh spec of DESIGN.md 5): "parity unpinned by the reference", pinned
by closed forms in tests/test_gpu_analytic.py.  A whole image at these sizes costs the scalar twin minutes, so the twin
renders ONE interleaved-tile share of the image (the `part_rank / part_world / tile_w / tile_h` partition it shares
with cudapt_set_tile_partition: 8x8 tiles spread over the whole image, every depth and material of the scene in the
sample) at the config's full resolution and full sample count; the GPU renders the whole image through the code path
the bench runs (one pass per step: camera-ray packets + bounce-0 hit list + per-ray kernel) and every owned pixel is
compared.  Bar: EVERY compared pixel bit-identical -- primary triangle id, primary t bits, float accumulation.  (The
mesh spec uses only correctly rounded + - * / sqrt and explicit fma in a fixed order, and the closest hit does not
depend on the acceleration structure, so there is no allowance to state; the sphere path's double-precision sincos
allowance does not apply here.)
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def tile_mask(W, H, rank, world, tw, th):
    """Pixels owned under cudapt_set_tile_partition's rule: (tile + tile_row) % world == rank."""
    tiles_x = (W + tw - 1) // tw
    ty, tx = np.meshgrid(np.arange(H) // th, np.arange(W) // tw, indexing="ij")
    return (ty * tiles_x + tx + ty) % world == rank


def gpu_render(pt, scene, W, H, N, S, fov, cfg, counting=False):
    with pt.PathTracer(pt.default_config(samples_per_pass=S, **cfg)) as tr:
        tr.set_camera(pt.default_camera(tan_half_fov=fov))
        tr.set_resolution(W, H)
        tr.upload_mesh(*scene)
        if counting:
            tr.set_counting(True)
        tr.render(N)
        acc, fc = tr.read_accum()
        prim, t = tr.read_primary_hits()
        st = tr.stats()
    assert fc == N
    return acc, prim, t, st


def compare_share(oracle, twin_mesh, acc, prim, t, W, H, N, fov, cfg, share, label):
    """Twin renders tile share `share` = (rank, world, tile_w, tile_h) of the W x H image at N spp; every owned pixel
    of the GPU image must equal it bit for bit.  Returns the number of pixels compared."""
    rank, world, tw, th = share
    p = oracle.twin_params(W, H, cam_tan_half_fov=fov, part_rank=rank, part_world=world, tile_w=tw, tile_h=th, **cfg)
    want, widx, wt, c = twin_mesh.render(p, N, want_primary=True)
    own = tile_mask(W, H, rank, world, tw, th)
    assert c.paths == int(own.sum()) * N                     # the twin rendered exactly the share
    assert np.all(want[~own] == 0) and np.all(want[own][:, 3] == N)
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    ok_prim = (got == widx)[own] & (bits(t) == bits(wt))[own]
    same = np.all(bits(acc) == bits(want), axis=-1)[own]
    hit_frac = float((widx[own] >= 0).mean())
    print(f"{label}: {own.sum()} pixels x {N} spp against the twin: primary identical {ok_prim.sum()}/{ok_prim.size}, "
          f"accumulation bit-identical {same.sum()}/{same.size}, camera rays hitting {hit_frac:.3f}, twin rays/path {c.rays / c.paths:.3f}")
    assert ok_prim.all(), f"{label}: {(~ok_prim).sum()} primary hits differ"
    assert same.all(), f"{label}: {(~same).sum()} of {same.size} pixels differ in the float accumulation"
    return int(own.sum())


# ---- configs[3] scene generator at a reduced size, depth 16: every camera-ray code path ---------------------------
C4_CFG = dict(rng_mode=1, max_bounces=16, jitter=1, seed=4)


@pytest.fixture(scope="module")
def small_rocks(scenes, oracle):
    scene = scenes.rock_field(n_instances=2_000)
    assert len(scene[1]) == 200_002
    return scene, oracle.TwinMesh(*scene)


def test_c4_reduced_packets_fallback_hitlist(pt, oracle, small_rocks, monkeypatch):
    """scenes.rock_field (the generator of configs[3]) with 2 000 rocks = 200 002 triangles, 320x180, depth 16, one
    32-sample pass: packets + hit list, with the default budget and with a budget small enough that a sizeable part of
    the packets is handed to the per-ray kernel (the fall-back queue matters most on C4)."""
    scene, twin = small_rocks
    W, H, N = 320, 180, 32
    want, widx, wt, c = twin.render(oracle.twin_params(W, H, **C4_CFG), N, want_primary=True)
    assert c.paths == W * H * N
    bails = {}
    for budget in ("400", "60"):
        monkeypatch.setenv("CUDAPT_PACKET_BUDGET", budget)
        acc, prim, t, st = gpu_render(pt, scene, W, H, N, 32, 1.0, C4_CFG, counting=True)
        assert st.packet_rays + 32 * st.packet_bails == W * H * N and st.paths == W * H * N
        bails[budget] = st.packet_bails / (W * H)
        got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
        assert np.array_equal(got, widx) and np.array_equal(bits(t), bits(wt))
        same = np.all(bits(acc) == bits(want), axis=-1)
        print(f"budget {budget}: packets handed to the per-ray kernel {bails[budget]:.4f}, accumulation bit-identical {same.sum()}/{same.size}, "
              f"rays GPU {st.rays} twin {c.rays}")
        assert same.all()
        assert st.rays == c.rays and st.shade_invocations == c.shade_invocations
    assert bails["60"] > 0.02 and bails["60"] > bails["400"]  # the small budget really exercised the fall-back queue


def test_c4_reduced_per_ray_pass(pt, oracle, small_rocks):
    """The same scene through raygen + the per-ray kernel only (a 3-sample pass cannot form packets), two passes."""
    scene, twin = small_rocks
    W, H, N = 320, 180, 6
    want, widx, wt, c = twin.render(oracle.twin_params(W, H, **C4_CFG), N, want_primary=True)
    acc, prim, t, st = gpu_render(pt, scene, W, H, N, 3, 1.0, C4_CFG)
    assert st.packet_rays == 0
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    assert np.array_equal(got, widx) and np.array_equal(bits(t), bits(wt))
    assert np.all(bits(acc) == bits(want)) and st.rays == c.rays


# ---- configs[3] at its full size ---------------------------------------------------------------------------------
def test_c4_full_size_tile_share(pt, oracle, scenes):
    """20 000 002 triangles (1.3 GB BVH, > L2), 1920x1080, depth 16: ONE 128-sample pass exactly as `bench.py --workload
    c4` runs it (4 rays per lane, hit list, fall-back queue), and a 3-sample per-ray pass, each against a 1/256
    interleaved-tile share rendered by the twin at the same sample counts."""
    scene = scenes.rock_field()
    assert len(scene[1]) == 20_000_002
    twin = oracle.TwinMesh(*scene)
    W, H = 1920, 1080
    acc, prim, t, st = gpu_render(pt, scene, W, H, 128, 128, 1.0, C4_CFG)
    assert st.packet_rays > 0.9 * W * H * 128 and st.paths == W * H * 128
    n = compare_share(oracle, twin, acc, prim, t, W, H, 128, 1.0, C4_CFG, (0, 256, 8, 8), "c4 1920x1080 128 spp (packets)")
    assert n >= 8000
    acc, prim, t, st = gpu_render(pt, scene, W, H, 3, 3, 1.0, C4_CFG)   # 3 samples: no packet shape -> raygen + per-ray kernel
    assert st.packet_rays == 0
    compare_share(oracle, twin, acc, prim, t, W, H, 3, 1.0, C4_CFG, (101, 256, 8, 8), "c4 1920x1080 3 spp (per-ray)")
    twin.close()


# ---- configs[2] at its full size ---------------------------------------------------------------------------------
C3_CFG = dict(rng_mode=1, max_bounces=8, jitter=1, seed=3)


def test_c3_full_size_tile_share(pt, oracle, scenes):
    """1 002 528 triangles, 1920x1080: the 64-sample step of the default bench (2 rays per lane, hit list) and a
    3-sample per-ray pass and a 16-sample pass of multi-pixel packets, against 1/64 and 1/256 tile shares of the twin."""
    scene = scenes.heightfield()
    assert len(scene[1]) == 1_002_528
    twin = oracle.TwinMesh(*scene)
    W, H = 1920, 1080
    acc, prim, t, st = gpu_render(pt, scene, W, H, 64, 64, 1.0, C3_CFG)
    assert st.packet_rays > 0.99 * W * H * 64
    n = compare_share(oracle, twin, acc, prim, t, W, H, 64, 1.0, C3_CFG, (5, 64, 8, 8), "c3 1920x1080 64 spp (packets)")
    assert n >= 32000
    acc, prim, t, st = gpu_render(pt, scene, W, H, 3, 3, 1.0, C3_CFG)   # 3 samples: no packet shape -> raygen + per-ray kernel
    assert st.packet_rays == 0
    compare_share(oracle, twin, acc, prim, t, W, H, 3, 1.0, C3_CFG, (17, 256, 8, 8), "c3 1920x1080 3 spp (per-ray)")
    acc, prim, t, st = gpu_render(pt, scene, W, H, 16, 16, 1.0, C3_CFG)  # 16 samples: packets of 2 pixels x 16 samples
    assert st.packet_rays > 0.9 * W * H * 16
    compare_share(oracle, twin, acc, prim, t, W, H, 16, 1.0, C3_CFG, (33, 256, 8, 8), "c3 1920x1080 16 spp (multi-pixel packets)")
    twin.close()


# ---- configs[1] at its full size ---------------------------------------------------------------------------------
C2_CFG = dict(rng_mode=1, max_bounces=8, sky_scale=0.0, jitter=1, nee=1, seed=2)


def test_c2_full_size_tile_share(pt, oracle, scenes):
    """Cornell box (diffuse + mirror + dielectric, area-light NEE), 1280x720, 256 spp in ONE pass as `bench.py
    --workload c2` runs it (4 rays per lane; shadow rays through the any-hit kernel), against a 1/64 tile share."""
    scene = scenes.cornell_box()
    twin = oracle.TwinMesh(*scene)
    W, H = 1280, 720
    acc, prim, t, st = gpu_render(pt, scene, W, H, 256, 256, 0.36, C2_CFG)
    assert st.packet_rays > 0.99 * W * H * 256 and st.shadow_rays > 0
    n = compare_share(oracle, twin, acc, prim, t, W, H, 256, 0.36, C2_CFG, (9, 64, 8, 8), "c2 1280x720 256 spp (packets, NEE)")
    assert n >= 14000
    twin.close()

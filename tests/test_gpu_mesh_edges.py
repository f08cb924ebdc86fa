"""Edge cases of the mesh path (GPU vs the CPU twin, through the C ABI), and size-independent properties at the full
size of BASELINE.json's configs[2].  The reference has no triangle path (CPUPathTracer.cpp:367-398), so the oracle is
the twin ("parity unpinned by the reference"); bar as in tests/test_gpu_mesh.py: every pixel bit-identical.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_frac(a, b):
    return float(np.all(bits(a) == bits(b), axis=-1).mean())


def quad(z, x0=-2.0, x1=2.0, y0=-2.0, y1=2.0):
    v = np.array([[x0, y0, z], [x1, y0, z], [x1, y1, z], [x0, y1, z]], np.float32)
    t = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)
    return v, t


def pair(pt, oracle, v, t, W, H, N=2, m=None, mats=None, **kw):
    cfg = pt.default_config(rng_mode=1, **kw)
    with pt.PathTracer(cfg) as tr:
        tr.set_resolution(W, H)
        tr.upload_mesh(v, t, m, mats)
        tr.render(N)
        acc, fc = tr.read_accum()
        prim, tt = tr.read_primary_hits()
    kw.pop("samples_per_pass", None)
    want, widx, wt, _ = oracle.TwinMesh(v, t, m, mats).render(oracle.twin_params(W, H, rng_mode=1, **kw), N, want_primary=True)
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    return acc, want, got, widx, tt, wt


def test_coincident_triangles_tie_goes_to_smallest_index(pt, oracle):
    """Three copies of the same quad at the same depth: every hit is an exact tie in t; the closest-hit rule picks the
    smallest caller index, whatever the BVH looks like."""
    v1, t1 = quad(4.0)
    v = np.concatenate([v1, v1, v1])
    t = np.concatenate([t1 + 8, t1 + 4, t1])  # caller order deliberately not the geometric duplicate order
    acc, want, got, widx, tt, wt = pair(pt, oracle, v, t, 96, 54, max_bounces=3, seed=1)
    assert np.array_equal(got, widx) and np.array_equal(bits(tt), bits(wt))
    hit = got >= 0
    assert hit.any() and set(np.unique(got[hit])) <= {0, 1}  # only the first copy's two triangles
    assert same_frac(acc, want) == 1.0


def test_degenerate_triangles_are_inert(pt, oracle):
    v1, t1 = quad(5.0)
    extra_v = np.array([[0, 0, 3], [0, 0, 3], [0, 0, 3],        # a point
                        [-1, 0, 3], [0, 0, 3], [1, 0, 3]], np.float32)  # a segment
    v = np.concatenate([v1, extra_v])
    t = np.concatenate([np.array([[4, 5, 6], [7, 8, 9]], np.uint32), t1])
    acc, want, got, widx, tt, wt = pair(pt, oracle, v, t, 80, 45, max_bounces=4, seed=2)
    assert np.isfinite(acc).all()
    assert np.array_equal(got, widx) and not np.isin(got, [0, 1]).any()
    assert same_frac(acc, want) == 1.0


@pytest.mark.parametrize("size", [(1, 1), (3, 2), (2, 7), (33, 17)])
def test_single_triangle_tiny_images(pt, oracle, size):
    W, H = size
    v = np.array([[-3, -3, 2], [3, -3, 2], [0, 3, 2]], np.float32)
    t = np.array([[0, 1, 2]], np.uint32)
    acc, want, got, widx, tt, wt = pair(pt, oracle, v, t, W, H, N=3, max_bounces=2, jitter=1, seed=3)
    assert np.array_equal(got, widx) and same_frac(acc, want) == 1.0


def test_axis_parallel_rays(pt, oracle):
    """Odd width/height with corner sampling: the centre column/row have d.x == 0 / d.y == 0 exactly, and an
    axis-aligned wall has zero-thickness boxes -- the slab test's 1/0 handling."""
    v, t = quad(6.0, -8, 8, -8, 8)
    v2, t2 = quad(9.0, -20, 20, -20, 20)
    vv = np.concatenate([v, v2])
    tt_ = np.concatenate([t, t2 + 4])
    acc, want, got, widx, tt, wt = pair(pt, oracle, vv, tt_, 64, 36, max_bounces=3, seed=4)
    assert np.array_equal(got, widx) and np.array_equal(bits(tt), bits(wt))
    assert (got[18, 32] >= 0)  # the centre pixel (u = 0.5, v = 0.5 -> d = (0, 0, 1)) hits the wall
    assert same_frac(acc, want) == 1.0


def test_large_coordinates_and_thin_slivers(pt, oracle, scenes):
    """The same height field pushed 4096 units away from the origin (camera moved along): coordinate magnitudes where
    one quantisation step of the root node is ~1e-2 and float spacing ~5e-4."""
    v, t, m, mats = scenes.heightfield(n_quads=32)
    off = np.array([4096.0, -2048.0, 8192.0], np.float32)
    v2 = (v + off).astype(np.float32)
    kw = dict(max_bounces=4, jitter=1, seed=6)
    cfg = pt.default_config(rng_mode=1, **kw)
    cam = dict(origin=tuple(float(x) for x in off))
    with pt.PathTracer(cfg) as tr:
        tr.set_camera(pt.default_camera(**cam))
        tr.set_resolution(96, 54)
        tr.upload_mesh(v2, t, m, mats)
        tr.render(2)
        acc, _ = tr.read_accum()
        prim, tt = tr.read_primary_hits()
    want, widx, wt, _ = oracle.TwinMesh(v2, t, m, mats).render(
        oracle.twin_params(96, 54, rng_mode=1, cam_origin=cam["origin"], **kw), 2, want_primary=True)
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    ok = (got == widx) & (bits(tt) == bits(wt))
    print(f"far-away scene: primary identical {ok.mean():.5f}, accumulation {same_frac(acc, want):.5f}")
    assert ok.all() and same_frac(acc, want) == 1.0


def test_deep_paths_max_bounces_64(pt, oracle, scenes):
    """A closed mirror-ish box keeps paths alive: 64 bounces, roulette off until late -> long queues at every bounce."""
    v, t, m, mats = scenes.cornell_box(sphere_res=6)
    mats = mats.copy()
    mats["albedo"][:] = np.minimum(mats["albedo"] * 1.3, 0.98)
    kw = dict(max_bounces=64, rr_start=40, sky_scale=0.0, jitter=1, nee=1, seed=8)
    cfg = pt.default_config(rng_mode=1, samples_per_pass=2, **kw)
    with pt.PathTracer(cfg) as tr:
        tr.set_camera(pt.default_camera(tan_half_fov=0.36))
        tr.set_resolution(64, 36)
        tr.upload_mesh(v, t, m, mats)
        tr.render(2)
        acc, _ = tr.read_accum()
        st = tr.stats()
    want, _, _, c = oracle.TwinMesh(v, t, m, mats).render(oracle.twin_params(64, 36, rng_mode=1, cam_tan_half_fov=0.36, **kw), 2)
    assert st.paths == 64 * 36 * 2 and st.rays == c.rays
    assert same_frac(acc, want) == 1.0


def test_full_size_properties_of_config3(pt, scenes):
    """BASELINE.json configs[2] at its full size (1,002,528 triangles, 1920x1080), 2 spp: properties that do not need
    the CPU twin -- alpha counts samples, values finite and non-negative, interleaved-tile partition over 3 ranks sums
    to the single-context image bit for bit, pass size is invisible, a refit with unchanged vertices is the identity."""
    scene = scenes.heightfield()
    v, t, m, mats = scene
    assert len(t) == 1_002_528
    W, H, N = 1920, 1080, 2
    cfg = dict(rng_mode=1, max_bounces=8, jitter=1, seed=3)

    def mk(spp_pass):
        tr = pt.PathTracer(pt.default_config(samples_per_pass=spp_pass, **cfg))
        tr.set_resolution(W, H)
        tr.upload_mesh(*scene)
        return tr

    full = mk(2)
    full.render(N)
    acc, fc = full.read_accum()
    assert fc == N and np.all(acc[..., 3] == N) and np.isfinite(acc).all() and (acc[..., :3] >= 0).all()
    st = full.stats()
    assert st.paths == W * H * N and st.rays >= st.paths and st.n_triangles == 1_002_528
    total = np.zeros_like(acc)
    for r in range(3):
        full.set_tile_partition(r, 3, 32, 32)
        full.reset()
        full.render(N)
        total += full.read_accum()[0]
    assert np.array_equal(bits(total), bits(acc))
    full.set_tile_partition(0, 1, 32, 32)
    full.update_vertices(v)  # identity refit
    full.render(N)
    assert np.array_equal(bits(full.read_accum()[0]), bits(acc))
    full.close()
    one = mk(1)
    one.render(N)
    assert np.array_equal(bits(one.read_accum()[0]), bits(acc))
    one.close()

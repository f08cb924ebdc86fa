"""GPU parity of the mesh wavefront integrator against the CPU twin (oracle/twin), through the C ABI.

The reference renders spheres only (CPUPathTracer.cpp:367-398), so this path is PARITY UNPINNED by the
reference: the oracle is the twin, which implements DESIGN.md "Mesh integrator numerics" independently.
Bar: EVERY pixel -- primary-hit triangle id, primary t bits, float accumulation bits.  The spec uses only
correctly rounded + - * / sqrt and explicitly written fma in a fixed order on both sides, and the closest hit
(smallest t, ties to the smallest caller index) does not depend on the acceleration structure: the twin's
binary BVH and the GPU's quantised BVH8 are both conservative, so there is no allowance to state.  (The
99.99 % allowance of the sphere path exists for its double-precision libm sincos and does not apply here.)
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_frac(a, b):
    return float(np.all(bits(a) == bits(b), axis=-1).mean())


def run_pair(pt, oracle, scene, W, H, N, fov=1.0, spp_pass=1, **kw):
    v, t, m, mats = scene
    cfg = pt.default_config(rng_mode=1, samples_per_pass=spp_pass, **kw)
    tr = pt.PathTracer(cfg)
    tr.set_camera(pt.default_camera(tan_half_fov=fov))
    tr.set_resolution(W, H)
    tr.upload_mesh(v, t, m, mats)
    tr.render(N)
    mesh = oracle.TwinMesh(v, t, m, mats)
    want, widx, wt, c = mesh.render(oracle.twin_params(W, H, rng_mode=1, cam_tan_half_fov=fov, **kw), N, want_primary=True)
    return tr, want, widx, wt, c


def check_primary(tr, widx, wt, min_frac=1.0):
    prim, t = tr.read_primary_hits()
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    ok = (got == widx) & (bits(t) == bits(wt))
    print(f"primary hits identical (id and t bits): {ok.sum()}/{ok.size}")
    assert ok.mean() >= min_frac


def test_cornell_nee_matches_twin(pt, oracle, scenes):
    scene = scenes.cornell_box(sphere_res=16)
    W, H, N = 160, 90, 6
    kw = dict(max_bounces=8, sky_scale=0.0, jitter=1, nee=1, seed=5)
    tr, want, widx, wt, c = run_pair(pt, oracle, scene, W, H, N, fov=0.36, spp_pass=3, **kw)
    acc, fc = tr.read_accum()
    assert fc == N
    check_primary(tr, widx, wt)
    f = same_frac(acc, want)
    print(f"cornell accumulation bit-identical pixels: {f:.5f}")
    assert f == 1.0
    assert abs(acc[..., :3].mean() - want[..., :3].mean()) <= 0.02 * want[..., :3].mean() + 1e-6
    st = tr.stats()
    assert st.paths == W * H * N
    assert st.rays == c.rays and st.shade_invocations == c.shade_invocations
    assert np.array_equal(tr.resolve_rgba8(), oracle.twin_resolve_rgba8(acc, N))
    tr.close()


def test_heightfield_ggx_matches_twin(pt, oracle, scenes):
    scene = scenes.heightfield(n_quads=96)
    W, H, N = 192, 108, 4
    kw = dict(max_bounces=8, jitter=1, seed=9)
    tr, want, widx, wt, c = run_pair(pt, oracle, scene, W, H, N, spp_pass=4, **kw)
    acc, _ = tr.read_accum()
    check_primary(tr, widx, wt)
    f = same_frac(acc, want)
    print(f"heightfield accumulation bit-identical pixels: {f:.5f}")
    assert f == 1.0
    tr.close()


def test_all_material_types_small(pt, oracle, scenes):
    """Mirror / dielectric / Lambert / GGX on the same geometry, no NEE, with sky."""
    v, t, m, mats = scenes.cornell_box(sphere_res=8)
    mats = mats.copy()
    mats[0]["type"] = scenes.GGX_COATED
    mats[0]["param"] = 0.25
    tr, want, widx, wt, c = run_pair(pt, oracle, (v, t, m, mats), 128, 72, 5, fov=0.36, spp_pass=2, max_bounces=10, jitter=0,
                                     seed=1, nee=0, sky_scale=0.5)
    acc, _ = tr.read_accum()
    assert same_frac(acc, want) == 1.0
    tr.close()


def test_samples_per_pass_is_invisible(pt, scenes):
    """Tracing 1, 3 or 5 samples per pass changes batching only: accumulation is added in sample order."""
    scene = scenes.cornell_box(sphere_res=8)
    out = []
    for spp_pass in (1, 3, 5):
        tr = pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=spp_pass, max_bounces=6, nee=1, jitter=1, sky_scale=0.0))
        tr.set_camera(pt.default_camera(tan_half_fov=0.36))
        tr.set_resolution(96, 54)
        tr.upload_mesh(*scene)
        tr.render(7)
        out.append(tr.read_accum()[0])
        tr.close()
    assert np.array_equal(bits(out[0]), bits(out[1])) and np.array_equal(bits(out[0]), bits(out[2]))


def test_mesh_tile_partition_and_resume(pt, scenes):
    scene = scenes.heightfield(n_quads=48)
    W, H, N, G = 128, 72, 4, 3
    cfg = dict(rng_mode=1, max_bounces=5, jitter=1, samples_per_pass=2)

    def mk():
        tr = pt.PathTracer(pt.default_config(**cfg))
        tr.set_resolution(W, H)
        tr.upload_mesh(*scene)
        return tr

    full = mk()
    full.render(N)
    facc, _ = full.read_accum()
    total = np.zeros_like(facc)
    for r in range(G):
        tr = mk()
        tr.set_tile_partition(r, G, 16, 8)
        tr.render(N)
        total += tr.read_accum()[0]
        tr.close()
    assert np.array_equal(bits(total), bits(facc))
    # resume: 2 frames, read, new context continues with frames 3..4
    a = mk()
    a.render(2)
    part, fc = a.read_accum()
    b = mk()
    b.write_accum(part, fc)
    b.render(2)
    assert np.array_equal(bits(b.read_accum()[0]), bits(facc))
    # re-upload of the packed scene from pinned host memory restarts accumulation and reproduces the image
    full.reupload_scene()
    full.render(N)
    assert np.array_equal(bits(full.read_accum()[0]), bits(facc))
    for x in (full, a, b):
        x.close()


def test_counting_mode_and_stats(pt, scenes):
    scene = scenes.heightfield(n_quads=64)
    tr = pt.PathTracer(pt.default_config(rng_mode=1, max_bounces=4))
    tr.set_resolution(160, 90)
    tr.upload_mesh(*scene)
    tr.render(2)
    plain, _ = tr.read_accum()
    tr.reset()
    tr.set_counting(True)
    tr.render(2)
    counted, _ = tr.read_accum()
    st = tr.stats()
    assert np.array_equal(bits(plain), bits(counted))
    assert st.nodes_visited > st.rays > 0 and st.tris_tested > 0
    assert st.n_triangles == 2 * 64 * 64 and st.n_nodes > 0 and st.accel_bytes == st.n_nodes * 80 + st.n_triangles * 48
    tr.close()


def test_mesh_requires_philox_and_scene(pt, scenes):
    tr = pt.PathTracer()
    tr.set_resolution(32, 32)
    with pytest.raises(pt.CudaptError):
        tr.build_accel()
    tr.upload_mesh(*scenes.heightfield(n_quads=8))
    with pytest.raises(pt.CudaptError) as e:
        tr.render(1)  # default config is REFERENCE_PCG; the mesh integrator is Philox-only
    assert "PHILOX" in str(e.value)
    tr.close()


def test_medium_heightfield_primary_hits_at_config_resolution(pt, oracle, scenes):
    """131 k triangles at 640x360: primary-hit parity (north_star: ids bit-exact, t within 1e-4 relative --
    here required bit-exact) plus size-independent properties of the full accumulation."""
    scene = scenes.heightfield(n_quads=256)
    W, H = 640, 360
    tr, want, widx, wt, c = run_pair(pt, oracle, scene, W, H, 1, spp_pass=1, max_bounces=8, jitter=1, seed=3)
    check_primary(tr, widx, wt)
    acc, _ = tr.read_accum()
    assert np.all(acc[..., 3] == 1) and np.isfinite(acc).all() and (acc[..., :3] >= 0).all()
    assert same_frac(acc, want) == 1.0
    tr.close()


def test_update_vertices_refit_equals_fresh_build(pt, oracle, scenes):
    """moving the vertices and refitting (no rebuild) gives the image of a fresh build of the moved
    mesh, bit for bit (the closest-hit rule does not depend on the acceleration structure), and equals the twin."""
    W, H, N = 128, 72, 3
    v, t, m, mats = scenes.heightfield(n_quads=48)
    rng = np.random.default_rng(3)
    v2 = (v + rng.normal(0.0, 0.05, v.shape)).astype(np.float32)
    kw = dict(max_bounces=5, jitter=1, seed=21)
    cfg = pt.default_config(rng_mode=1, samples_per_pass=3, **kw)
    with pt.PathTracer(cfg) as tr:
        tr.set_resolution(W, H)
        tr.upload_mesh(v, t, m, mats)
        tr.render(N)
        before, _ = tr.read_accum()
        tr.update_vertices(v2)
        assert tr.stats().frame_count == 0  # accumulation restarted, like a scene change (CPT.cpp:122-131)
        tr.render(N)
        refit, fc = tr.read_accum()
        prim_refit, t_refit = tr.read_primary_hits()
        with pytest.raises(pt.CudaptError):
            tr.update_vertices(v2[:-1])  # topology change is not an update
    with pt.PathTracer(cfg) as tr:
        tr.set_resolution(W, H)
        tr.upload_mesh(v2, t, m, mats)
        tr.render(N)
        fresh, _ = tr.read_accum()
        prim_fresh, t_fresh = tr.read_primary_hits()
    assert fc == N
    assert np.array_equal(bits(refit), bits(fresh))
    assert np.array_equal(prim_refit, prim_fresh) and np.array_equal(bits(t_refit), bits(t_fresh))
    assert same_frac(before, refit) < 0.9
    want, _, _, _ = oracle.TwinMesh(v2, t, m, mats).render(oracle.twin_params(W, H, rng_mode=1, **kw), N)
    assert same_frac(refit, want) == 1.0


def test_mesh_golden_fixture(pt, golden_mesh):
    """GPU vs the COMMITTED mesh vectors (tests/golden/c2_cornell_64x36.npz, generated by the twin): per-ray kernel
    (4 spp, passes of 2), packet path (one 32-sample pass) and the thin lens."""
    g = golden_mesh
    W, H = int(g["width"]), int(g["height"])
    kw = dict(rng_mode=1, max_bounces=6, jitter=1, nee=1, sky_scale=0.0, seed=17)

    def run(n, spp_pass, **cam):
        with pt.PathTracer(pt.default_config(samples_per_pass=spp_pass, **kw)) as tr:
            tr.set_camera(pt.default_camera(tan_half_fov=0.36, **cam))
            tr.set_resolution(W, H)
            tr.upload_mesh(g["vertices"], g["triangles"], g["material_id"], g["materials"])
            tr.render(n)
            acc, _ = tr.read_accum()
            prim, t = tr.read_primary_hits()
            st = tr.stats()
        return acc, prim, t, st

    acc, prim, t, st = run(4, 2)
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    assert np.array_equal(got, g["primary_idx"]) and np.array_equal(bits(t), bits(g["primary_t"]))
    assert same_frac(acc, g["accum4"]) == 1.0 and int(st.rays) == int(g["rays4"])
    acc, _, _, st = run(32, 32)
    assert st.packet_rays > 0 and same_frac(acc, g["accum32"]) == 1.0
    acc, _, _, st = run(32, 32, lens_radius=0.05, focus_distance=4.0)
    assert st.packet_rays == 0 and same_frac(acc, g["accum32_lens"]) == 1.0

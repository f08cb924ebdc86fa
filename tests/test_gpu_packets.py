"""Camera-ray packets (csrc/primary_packet.cu): a pass of 32k samples per pixel traces its camera rays as one packet
per (pixel, 32 samples).  The per-ray closest hit is a minimum over triangles and does not depend on which superset
of candidates was tested, so the results must be BIT-IDENTICAL to the per-ray kernel's (CUDAPT_PRIMARY_PACKETS=0) and
equal the CPU twin's within the mesh parity bar; a tiny budget forces every packet through the fall-back queue."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def render(pt, scene, W, H, N, S, cam=None, part=None, **kw):
    with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=S, **kw)) as tr:
        if cam:
            tr.set_camera(cam if isinstance(cam, pt.Camera) else pt.default_camera(**cam))
        tr.set_resolution(W, H)
        tr.upload_mesh(*scene)
        if part:
            tr.set_tile_partition(*part)
        tr.render(N)
        acc, _ = tr.read_accum()
        prim, t = tr.read_primary_hits()
        st = tr.stats()
    return acc, prim, t, st


CASES = {
    "cornell": dict(W=96, H=54, kw=dict(max_bounces=5, jitter=1, nee=1, sky_scale=0.0, seed=5), cam=dict(tan_half_fov=0.36)),
    "heightfield": dict(W=128, H=72, kw=dict(max_bounces=4, jitter=1, seed=7), cam=None),
    "heightfield_nojitter": dict(W=64, H=36, kw=dict(max_bounces=3, jitter=0, seed=7), cam=None),
}


def look(origin, forward, up_hint=(0.0, 1.0, 0.0)):
    f = np.asarray(forward, np.float64)
    f /= np.linalg.norm(f)
    r = np.cross(np.asarray(up_hint, np.float64), f)
    r /= np.linalg.norm(r)
    u = np.cross(f, r)
    return dict(origin=origin, right=tuple(r), up=tuple(u), forward=tuple(f))


def scene_of(scenes, name):
    return scenes.cornell_box(sphere_res=8) if name == "cornell" else scenes.heightfield(n_quads=96)


@pytest.mark.parametrize("name", list(CASES))
def test_packets_equal_per_ray_kernel_and_twin(pt, oracle, scenes, name, monkeypatch):
    c = CASES[name]
    scene = scene_of(scenes, name)
    N = 32
    monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", "1")
    monkeypatch.delenv("CUDAPT_PACKET_BUDGET", raising=False)
    a1, p1, t1, s1 = render(pt, scene, c["W"], c["H"], N, 32, cam=c["cam"], **c["kw"])
    monkeypatch.setenv("CUDAPT_PACKET_BUDGET", "0")  # every packet gives up after its first node
    a2, p2, t2, s2 = render(pt, scene, c["W"], c["H"], N, 32, cam=c["cam"], **c["kw"])
    monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", "0")
    a0, p0, t0, s0 = render(pt, scene, c["W"], c["H"], N, 32, cam=c["cam"], **c["kw"])
    # (the bounce-0 hit list of the packets adds its scan and expand launches)
    assert s1.kernel_launches in (s0.kernel_launches, s0.kernel_launches + 2) and s1.rays == s0.rays == s2.rays and s1.paths == s0.paths
    for a, p, t in ((a1, p1, t1), (a2, p2, t2)):
        assert np.array_equal(p, p0) and np.array_equal(bits(t), bits(t0))
        assert np.array_equal(bits(a), bits(a0))
    tw = {("cam_" + k): v for k, v in (c["cam"] or {}).items()}
    want, widx, wt, _ = oracle.TwinMesh(*scene).render(oracle.twin_params(c["W"], c["H"], rng_mode=1, **c["kw"], **tw), N, want_primary=True)
    got = np.where(p1 == 0xFFFFFFFF, -1, p1.astype(np.int64))
    ok = (got == widx) & (bits(t1) == bits(wt))
    same = float(np.all(bits(a1) == bits(want), axis=-1).mean())
    print(f"{name}: packets vs twin: primary {ok.mean():.5f}, accumulation {same:.5f}")
    assert ok.all() and same == 1.0


@pytest.mark.parametrize("rays", [1, 2, 4])
def test_packets_of_32_64_128_samples(pt, scenes, rays, monkeypatch):
    """128-sample passes: 1, 2 or 4 rays per lane (CUDAPT_PACKET_RAYS) -- the same image as the per-ray kernel."""
    scene = scenes.heightfield(n_quads=48)
    kw = dict(max_bounces=3, jitter=1, seed=13)
    monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", "0")
    a0, p0, t0, s0 = render(pt, scene, 64, 36, 128, 128, **kw)
    monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", "1")
    monkeypatch.setenv("CUDAPT_PACKET_RAYS", str(rays))
    a1, p1, t1, s1 = render(pt, scene, 64, 36, 128, 128, **kw)
    assert s1.packet_rays > 0.9 * 64 * 36 * 128 and s1.rays == s0.rays
    assert np.array_equal(p1, p0) and np.array_equal(bits(t1), bits(t0)) and np.array_equal(bits(a1), bits(a0))


def test_packets_rotated_camera_and_partition(pt, scenes, monkeypatch):
    """Camera looking down -Y/-X-ish (dominant axis and sign differ from the default +Z), 3-rank tile partition."""
    scene = scenes.heightfield(n_quads=64)
    kw = dict(max_bounces=3, jitter=1, seed=11)
    cams = (look((0.0, 6.0, 10.0), (0.0, -1.0, 0.05), up_hint=(0.0, 0.0, 1.0)), look((12.0, 1.0, 10.0), (-1.0, -0.3, 0.1)),
            pt.look_at((-10.0, 3.0, 10.0), (0.0, -2.0, 10.0), (0.0, 1.0, 0.0), 50.0))  # +X dominant, narrow lens
    for cam in cams:
        out = {}
        for on in ("1", "0"):
            monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", on)
            out[on] = render(pt, scene, 80, 48, 64, 64, cam=cam, **kw)
        assert np.array_equal(bits(out["1"][0]), bits(out["0"][0])) and np.array_equal(out["1"][1], out["0"][1])
        assert (out["1"][1] != 0xFFFFFFFF).mean() > 0.05  # the camera does see the mesh
    total = None
    monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", "1")
    full = render(pt, scene, 80, 48, 32, 32, **kw)[0]
    for r in range(3):
        a = render(pt, scene, 80, 48, 32, 32, part=(r, 3, 16, 8), **kw)[0]
        total = a if total is None else total + a
    assert np.array_equal(bits(total), bits(full))


def test_packets_full_size_config3_bit_identical(pt, scenes, monkeypatch):
    """BASELINE.json configs[2] at its full size (1,002,528 triangles, 1920x1080), one 64-sample pass: the packet path
    (2 rays per lane + fall-back queue) and the per-ray kernel produce the same accumulation buffer bit for bit, the
    same primary-hit AOV and the same ray count; an environment map and a refit do not disturb that."""
    scene = scenes.heightfield()
    v = scene[0]
    assert len(scene[1]) == 1_002_528
    W, H, N = 1920, 1080, 64
    env = scenes.octahedral_environment(64, sun_dir=(0.3, 0.8, 0.5), sun_power=20.0, seed=4)
    out = {}
    for on in ("1", "0"):
        monkeypatch.setenv("CUDAPT_PRIMARY_PACKETS", on)
        with pt.PathTracer(pt.default_config(rng_mode=1, max_bounces=8, jitter=1, seed=3, samples_per_pass=64)) as tr:
            tr.set_resolution(W, H)
            tr.upload_mesh(*scene)
            tr.render(N)
            a = tr.read_accum()[0]
            prim, t = tr.read_primary_hits()
            st = tr.stats()
            tr.upload_environment(env)
            w2 = v.copy()
            w2[:, 1] += (0.05 * np.sin(2.0 * v[:, 0])).astype(np.float32)
            tr.update_vertices(w2)
            tr.render(N)
            b = tr.read_accum()[0]
        out[on] = (a, prim, t, st, b)
    a1, p1, t1, s1, b1 = out["1"]
    a0, p0, t0, s0, b0 = out["0"]
    assert s1.packet_rays > 0.99 * W * H * N and s0.packet_rays == 0
    assert s1.rays == s0.rays and s1.paths == s0.paths == W * H * N
    assert np.array_equal(p1, p0) and np.array_equal(bits(t1), bits(t0))
    assert np.array_equal(bits(a1), bits(a0))
    assert np.array_equal(bits(b1), bits(b0)) and not np.array_equal(bits(b1), bits(a1))


@pytest.mark.parametrize("name", ["cornell", "heightfield", "heightfield_two_passes", "heightfield_env_aov"])
def test_hit_list_bit_identical(pt, scenes, name, monkeypatch):
    """Bounce-0 hit list (primary_packet.cu, kList): all-sky packets are finished by the packet kernel and never reach
    the shade kernel; tiles of 64 packets are claimed dynamically.  Accumulation, primary-hit AOVs and counters must
    equal those of the identity queue (CUDAPT_PACKET_HITLIST=0), with a budget that sends packets to the per-ray kernel
    as well."""
    if name == "cornell":
        scene, W, H, N, S, cam, kw, env = scenes.cornell_box(sphere_res=8), 96, 54, 32, 32, dict(tan_half_fov=0.36), dict(max_bounces=5, jitter=1, nee=1, sky_scale=0.0, seed=5), None
    elif name == "heightfield":
        scene, W, H, N, S, cam, kw, env = scenes.heightfield(n_quads=96), 200, 113, 64, 64, None, dict(max_bounces=4, jitter=1, seed=7), None
    elif name == "heightfield_two_passes":
        scene, W, H, N, S, cam, kw, env = scenes.heightfield(n_quads=64), 97, 67, 256, 128, None, dict(max_bounces=3, jitter=1, seed=3), None
    else:
        scene, W, H, N, S, cam, kw, env = scenes.heightfield(n_quads=64), 128, 72, 32, 32, None, dict(max_bounces=4, jitter=0, seed=9, sky_scale=0.6), scenes.octahedral_environment(16, sun_dir=(-0.3, 0.5, 0.4))
    out = {}
    for budget in ("400", "12"):
        monkeypatch.setenv("CUDAPT_PACKET_BUDGET", budget)
        for mode in ("0", "1"):
            monkeypatch.setenv("CUDAPT_PACKET_HITLIST", mode)
            with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=S, **kw)) as tr:
                if cam:
                    tr.set_camera(pt.default_camera(**cam))
                tr.set_resolution(W, H)
                tr.upload_mesh(*scene)
                if env is not None:
                    tr.upload_environment(env)
                tr.render(N)
                acc, _ = tr.read_accum()
                prim, t = tr.read_primary_hits()
                st = tr.stats()
            out[(budget, mode)] = (acc, prim, t, st)
        (a0, p0, t0, s0), (a1, p1, t1, s1) = out[(budget, "0")], out[(budget, "1")]
        assert np.array_equal(bits(a0), bits(a1)) and np.array_equal(p0, p1) and np.array_equal(bits(t0), bits(t1))
        assert (s0.paths, s0.rays, s0.shade_invocations, s0.packet_rays, s0.packets) == (s1.paths, s1.rays, s1.shade_invocations, s1.packet_rays, s1.packets)
        assert s0.paths == W * H * N and s1.kernel_launches > s0.kernel_launches
    assert np.array_equal(bits(out[("400", "1")][0]), bits(out[("12", "1")][0]))


@pytest.mark.parametrize("case", ["cornell", "heightfield_ragged", "many_ranks"])
def test_compacted_tile_partition_keeps_packets_and_hit_list(pt, scenes, case, monkeypatch):
    """Interleaved-tile partition with tiles that are multiples of 8x8 pixels: a pass whose camera rays go as packets
    compacts its slots to the rank's own slot tiles (csrc/kernels.h PassParams::owned_tiles), so it keeps the bounce-0
    hit list and holds world x more samples in the same memory.  The ranks' buffers must still sum to the one-context
    image BIT FOR BIT (zeros elsewhere), AOVs and counters must add up, and tiles that are NOT multiples of 8 (the
    sparse form: explicit queue, one atomic per packet) must give the same image."""
    monkeypatch.delenv("CUDAPT_PACKET_HITLIST", raising=False)
    monkeypatch.delenv("CUDAPT_PRIMARY_PACKETS", raising=False)
    if case == "cornell":
        scene, W, H, N, S, cam, kw, G, tile = scenes.cornell_box(sphere_res=8), 96, 54, 64, 64, dict(tan_half_fov=0.36), dict(max_bounces=5, jitter=1, nee=1, sky_scale=0.0, seed=5), 3, (16, 8)
    elif case == "heightfield_ragged":  # width and height not multiples of 8 or of the tile: padded slots at the borders
        scene, W, H, N, S, cam, kw, G, tile = scenes.heightfield(n_quads=64), 203, 117, 96, 32, None, dict(max_bounces=4, jitter=1, seed=7), 2, (32, 32)
    else:  # more ranks than tile columns: some ranks own very little
        scene, W, H, N, S, cam, kw, G, tile = scenes.heightfield(n_quads=48), 72, 40, 32, 32, None, dict(max_bounces=3, jitter=1, seed=3), 8, (24, 16)
    full, fprim, ft, fst = render(pt, scene, W, H, N, S, cam=cam, **kw)
    for tw, th in (tile, (tile[0] + 4, tile[1] + 2)):  # compacted form, then a tile size that forces the sparse form
        total = np.zeros_like(full)
        prim = np.full((H, W), 0xFFFFFFFF, np.uint32)
        tt = np.full((H, W), np.inf, np.float32)
        paths = rays = packet_rays = 0
        for r in range(G):
            a, p_, t_, st = render(pt, scene, W, H, N, S, cam=cam, part=(r, G, tw, th), **kw)
            ty, tx = np.meshgrid(np.arange(H) // th, np.arange(W) // tw, indexing="ij")
            own = (ty * ((W + tw - 1) // tw) + tx + ty) % G == r
            assert np.all(a[~own] == 0) and np.all(a[own][:, 3] == N)
            total += a
            prim[own], tt[own] = p_[own], t_[own]
            paths, rays, packet_rays = paths + st.paths, rays + st.rays, packet_rays + st.packet_rays
        assert np.array_equal(bits(total), bits(full))
        assert np.array_equal(prim, fprim) and np.array_equal(bits(tt), bits(ft))
        assert (paths, rays) == (fst.paths, fst.rays) and packet_rays > 0.9 * W * H * N


def test_pass_of_512_samples_under_a_tile_partition(pt, scenes):
    """samples_per_pass up to 1024: what a tile rank of an 8-GPU run uses (512 frames of its own pixels in ONE pass, 4 rays
    per lane).  Same image, bit for bit, as 32-sample passes of one context over the whole image."""
    scene = scenes.heightfield(n_quads=64)
    W, H, N, G = 128, 72, 512, 8
    kw = dict(max_bounces=4, jitter=1, seed=19)
    full, fprim, ft, fst = render(pt, scene, W, H, N, 32, **kw)
    total = np.zeros_like(full)
    paths = 0
    for r in range(G):
        a, _, _, st = render(pt, scene, W, H, N, 512, part=(r, G, 32, 32), **kw)
        total += a
        paths += st.paths
        assert st.packet_rays > 0 or st.paths == 0
    assert np.array_equal(bits(total), bits(full)) and paths == fst.paths == W * H * N
    with pytest.raises(pt.CudaptError):
        pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=1025))


@pytest.mark.parametrize("S", [1, 2, 4, 8, 16])
def test_multi_pixel_packets_equal_per_ray_path(pt, oracle, scenes, S, monkeypatch):
    """Passes of 1, 2, 4, 8 or 16 samples trace their camera rays as packets of 32 / S neighbouring pixels x S samples
    (primary_packet.cu kMini, slot-major path ids) -- what a viewer that renders one sample per call gets.  Accumulation,
    primary-hit AOVs and counters must equal the raygen + per-ray path (CUDAPT_PACKET_MINI=0) bit for bit, with the hit
    list on and off, with a budget that sends every packet to the per-ray kernel, on an image whose size is not a
    multiple of the 8x8 slot tiles, and the twin's image."""
    scene = scenes.heightfield(n_quads=64)
    W, H, N = 203, 117, 3 * S          # three passes
    kw = dict(max_bounces=4, jitter=1, seed=23)
    monkeypatch.setenv("CUDAPT_PACKET_MINI", "0")
    a0, p0, t0, s0 = render(pt, scene, W, H, N, S, **kw)
    assert s0.packet_rays == 0
    monkeypatch.setenv("CUDAPT_PACKET_MINI", "1")
    for hitlist, budget in (("1", "400"), ("0", "400"), ("1", "0")):
        monkeypatch.setenv("CUDAPT_PACKET_HITLIST", hitlist)
        monkeypatch.setenv("CUDAPT_PACKET_BUDGET", budget)
        a1, p1, t1, s1 = render(pt, scene, W, H, N, S, **kw)
        assert np.array_equal(bits(a1), bits(a0)) and np.array_equal(p1, p0) and np.array_equal(bits(t1), bits(t0)), (hitlist, budget)
        assert (s1.paths, s1.rays, s1.shade_invocations) == (s0.paths, s0.rays, s0.shade_invocations)
        if budget != "0":
            assert s1.packet_rays > 0.9 * W * H * N
    monkeypatch.delenv("CUDAPT_PACKET_HITLIST")
    monkeypatch.delenv("CUDAPT_PACKET_BUDGET")
    want, widx, wt, _ = oracle.TwinMesh(*scene).render(oracle.twin_params(W, H, rng_mode=1, **kw), N, want_primary=True)
    got = np.where(p0 == 0xFFFFFFFF, -1, p0.astype(np.int64))
    assert np.array_equal(got, widx) and np.array_equal(bits(t0), bits(wt)) and np.array_equal(bits(a0), bits(want))


def test_multi_pixel_packets_tiny_images_partition_and_cornell(pt, scenes, monkeypatch):
    """Images of one tile or less (the hit list's scratch does not fit: identity queue), a compacted 3-rank partition with
    4-sample passes, and the closed Cornell box with NEE at 1 sample per pass."""
    hf = scenes.heightfield(n_quads=32)
    for W, H in ((1, 1), (7, 3), (8, 8), (33, 17)):
        out = {}
        for on in ("0", "1"):
            monkeypatch.setenv("CUDAPT_PACKET_MINI", on)
            out[on] = render(pt, hf, W, H, 4, 2, max_bounces=3, jitter=1, seed=5)
        assert np.array_equal(bits(out["0"][0]), bits(out["1"][0])) and np.array_equal(out["0"][1], out["1"][1]), (W, H)
    monkeypatch.setenv("CUDAPT_PACKET_MINI", "1")
    kw = dict(max_bounces=4, jitter=1, seed=7)
    full, fprim, ft, fst = render(pt, hf, 203, 117, 8, 4, **kw)
    total = np.zeros_like(full)
    packet_rays = 0
    for r in range(3):
        a, _, _, st = render(pt, hf, 203, 117, 8, 4, part=(r, 3, 16, 8), **kw)
        total += a
        packet_rays += st.packet_rays
    assert np.array_equal(bits(total), bits(full)) and packet_rays > 0.9 * 203 * 117 * 8
    box = scenes.cornell_box(sphere_res=8)
    ckw = dict(max_bounces=5, jitter=1, nee=1, sky_scale=0.0, seed=5)
    out = {}
    for on in ("0", "1"):
        monkeypatch.setenv("CUDAPT_PACKET_MINI", on)
        out[on] = render(pt, box, 96, 54, 6, 1, cam=dict(tan_half_fov=0.36), **ckw)
    assert np.array_equal(bits(out["0"][0]), bits(out["1"][0])) and out["1"][3].packet_rays > 0 and out["0"][3].packet_rays == 0

"""Closed-form pins of the mesh integrator (materials, NEE, Russian roulette, emission bookkeeping).

The reference has no materials, no lights and no triangles (CPUPathTracer.cpp:260 is a constant 0.7, :367-398 spheres
only), so the mesh path's oracle -- the twin -- and the CUDA kernels implement the same builder-written spec (DESIGN.md
5): "bit-identical to the twin" proves consistency, not correctness.  These tests pin BOTH against facts that do not
come from that spec: radiometry that any unbiased path tracer must reproduce.

  * white furnace: a closed Lambert box of albedo a that emits Le everywhere shows Le * sum_k a^k (geometric series cut
    at the depth limit) in every direction -- with cosine sampling every path carries exactly that value, with Russian
    roulette or next-event estimation only the mean does;
  * direct irradiance under a rectangular Lambertian emitter: E = pi * Le * F with the closed-form form factor of a
    differential element to a parallel rectangle (e.g. Howell, "A Catalog of Radiation Heat Transfer Configuration
    Factors", B-4), radiance a/pi * E; complete occlusion gives exactly 0;
  * a dielectric slab at normal incidence: Schlick's Fresnel term is r0 = ((1-n)/(1+n))^2 there, and the series over the
    internal reflections gives T = (1-r0)/(1+r0), R = 2 r0/(1+r0);
  * the GGX-coated material under a uniform environment: the pixel mean is the directional albedo of
    f = F D G / (4 cos_o cos_i) + albedo (1 - F(cos_o)) / pi, computed here by plain quadrature over the hemisphere
    (no importance sampling, no code shared with the integrator), and never exceeds 1 for a white base.

Each test runs on the twin (CPU, `-m "not gpu"`) and on the CUDA path through the C ABI (`-m gpu`).
"""
import math

import numpy as np
import pytest

IMPLS = ["twin", pytest.param("gpu", marks=pytest.mark.gpu)]


def quad(p0, p1, p2, p3):
    return np.array([p0, p1, p2, p3], np.float64), np.array([[0, 2, 1], [0, 3, 2]], np.int64)


class Builder:
    def __init__(self):
        self.v, self.t, self.m = [], [], []

    def add(self, vt, mat):
        v, t = vt
        base = sum(len(x) for x in self.v)
        self.v.append(v)
        self.t.append(t + base)
        self.m.append(np.full(len(t), mat, np.uint32))

    def mesh(self):
        return (np.concatenate(self.v).astype(np.float32), np.concatenate(self.t).astype(np.uint32), np.concatenate(self.m))


def render(impl, request, scene, mats, W, H, N, cam=None, env=None, **cfg):
    """Mean radiance per pixel (H, W, 3) and the primary-hit triangle ids, from the twin or from the CUDA path."""
    v, t, m = scene
    cam = dict(cam or {})
    if impl == "twin":
        oracle = request.getfixturevalue("oracle")
        kw = {("cam_" + k): val for k, val in cam.items()}
        if env is not None:
            kw["env"] = env
        p = oracle.twin_params(W, H, rng_mode=1, **cfg, **kw)
        acc, idx, _, _ = oracle.TwinMesh(v, t, m, mats).render(p, N, want_primary=True)
    else:
        pt = request.getfixturevalue("pt")
        with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=min(N, 256), **cfg)) as tr:
            tr.set_camera(pt.default_camera(**cam))
            tr.set_resolution(W, H)
            tr.upload_mesh(v, t, m, mats)
            if env is not None:
                tr.upload_environment(env)
            tr.render(N)
            acc, fc = tr.read_accum()
            prim, _ = tr.read_primary_hits()
        assert fc == N
        idx = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    assert np.all(acc[..., 3] == N)
    return acc[..., :3].astype(np.float64) / N, idx


def closed_box(mat, lo=(-2.0, -2.0, -2.0), hi=(2.0, 2.0, 6.0)):
    """Six quads with the geometric normal (cross(e1, e2)) pointing INTO the box: the camera at the origin is inside."""
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    b = Builder()
    b.add(quad([x0, y0, z0], [x1, y0, z0], [x1, y0, z1], [x0, y0, z1]), mat)  # floor, normal +y
    b.add(quad([x0, y1, z0], [x0, y1, z1], [x1, y1, z1], [x1, y1, z0]), mat)  # ceiling, -y
    b.add(quad([x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1]), mat)  # back, -z
    b.add(quad([x0, y0, z0], [x0, y1, z0], [x1, y1, z0], [x1, y0, z0]), mat)  # behind the camera, +z
    b.add(quad([x0, y0, z0], [x0, y0, z1], [x0, y1, z1], [x0, y1, z0]), mat)  # left, +x
    b.add(quad([x1, y0, z0], [x1, y1, z0], [x1, y1, z1], [x1, y0, z1]), mat)  # right, -x
    v, t, m = b.mesh()
    n = np.cross(v[t[:, 1]] - v[t[:, 0]], v[t[:, 2]] - v[t[:, 0]])
    centre = (np.array(lo) + np.array(hi)) / 2
    assert np.all(((centre - v[t[:, 0]]) * n).sum(1) > 0)  # every face looks at the inside
    return v, t, m


# ---- white furnace ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_white_furnace_lambert_every_path_carries_the_series(impl, request, scenes):
    """No roulette, no NEE: cosine sampling has weight = albedo exactly, so EVERY sample equals Le * sum_{k<B} a^k."""
    a, Le, B = 0.6, np.array([0.5, 0.25, 1.0]), 6
    mats = np.array([scenes.material((a, a, a), emission=tuple(Le))], scenes.MATERIAL_DTYPE)
    img, idx = render(impl, request, closed_box(0), mats, 16, 10, 8, max_bounces=B, rr_start=1000, nee=0, jitter=1, seed=1, sky_scale=1.0)
    want = Le * (1 - a ** B) / (1 - a)
    assert (idx >= 0).all()  # closed: no camera ray escapes (the sky, scale 1, would show)
    assert np.allclose(img, want, rtol=2e-6, atol=0), (img.min(axis=(0, 1)), img.max(axis=(0, 1)), want)


@pytest.mark.parametrize("impl", IMPLS)
def test_white_furnace_with_russian_roulette_is_unbiased(impl, request, scenes):
    """Roulette from the second bounce on (survivors divided by p): the MEAN is the series, now almost complete."""
    a, Le, B = 0.7, np.array([1.0, 0.5, 0.25]), 48
    mats = np.array([scenes.material((a, a, a), emission=tuple(Le))], scenes.MATERIAL_DTYPE)
    img, _ = render(impl, request, closed_box(0), mats, 16, 9, 1024, max_bounces=B, rr_start=1, nee=0, jitter=1, seed=2, sky_scale=0.0)
    want = Le * (1 - a ** B) / (1 - a)
    got = img.mean(axis=(0, 1))
    # 147 456 paths; one path's value has a relative standard deviation below 1, so the mean's is below 0.3 %
    assert np.allclose(got, want, rtol=0.012), (got, want)
    assert img.max() > 1.05 * want.max() and img.min() < 0.95 * want.min()  # and roulette did introduce variance


@pytest.mark.parametrize("impl", IMPLS)
def test_furnace_next_event_estimation_adds_one_term(impl, request, scenes):
    """NEE on, depth 1: the camera hit shows its own emission (Le) plus the light sample, whose mean is the direct
    light a/pi * E with E = pi * Le inside a uniform furnace -> Le * (1 + a).  Pins the light pick (1/n), the area
    pdf, cos cos / d^2 and the 1/pi of the Lambert BRDF.  Pixels in the middle of the back wall only (near the edges
    the 1/d^2 estimator has a heavy tail)."""
    a, Le = 0.5, np.array([0.8, 0.4, 0.2])
    mats = np.array([scenes.material((a, a, a), emission=tuple(Le))], scenes.MATERIAL_DTYPE)
    img, idx = render(impl, request, closed_box(0), mats, 8, 8, 4096, cam=dict(tan_half_fov=0.08), max_bounces=1, rr_start=1000, nee=1, jitter=1,
                      seed=3, sky_scale=0.0)
    assert np.isin(idx, [4, 5]).all()  # the back wall
    got = img.mean(axis=(0, 1))
    assert np.allclose(got, Le * (1 + a), rtol=0.01), (got, Le * (1 + a))


# ---- direct irradiance under a rectangular emitter -------------------------------------------------------------------
def corner_form_factor(a, b, h):
    """Differential element -> rectangle a x b in a parallel plane at distance h, the element below one corner."""
    if a <= 0 or b <= 0:
        return 0.0
    A, Bq = math.hypot(a, h), math.hypot(b, h)
    return (a / A * math.atan(b / A) + b / Bq * math.atan(a / Bq)) / (2 * math.pi)


def rect_form_factor(px, pz, x0, x1, z0, z1, h):
    """Superposition of four corner rectangles (signed) for a point at (px, pz) below the rectangle [x0,x1] x [z0,z1]."""
    def F(xe, ze):
        sx, sz = math.copysign(1.0, xe - px), math.copysign(1.0, ze - pz)
        return sx * sz * corner_form_factor(abs(xe - px), abs(ze - pz), h)
    return F(x1, z1) - F(x0, z1) - F(x1, z0) + F(x0, z0)


def light_scene(blocker=False):
    b = Builder()
    b.add(quad([-6, -1, 1], [6, -1, 1], [6, -1, 13], [-6, -1, 13]), 0)                 # floor y = -1, normal +y
    b.add(quad([-0.5, 1, 3.5], [-0.5, 1, 5.0], [1.0, 1, 5.0], [1.0, 1, 3.5]), 1)      # emitter y = +1, normal -y
    if blocker:
        b.add(quad([-3, 0.9, 1], [-3, 0.9, 8], [4, 0.9, 8], [4, 0.9, 1]), 2)          # black sheet just below it
    return b.mesh()


@pytest.mark.parametrize("impl", IMPLS)
def test_direct_light_of_a_rectangular_emitter_matches_the_form_factor(impl, request, scenes):
    a, Le = 0.8, np.array([10.0, 6.0, 2.0])
    mats = np.array([scenes.material((a, a, a)), scenes.material((0, 0, 0), emission=tuple(Le)), scenes.material((0, 0, 0))],
                    scenes.MATERIAL_DTYPE)
    W, H, N, tan = 12, 8, 4096, 0.5
    img, idx = render(impl, request, light_scene(), mats, W, H, N, cam=dict(tan_half_fov=tan), max_bounces=1, rr_start=1000, nee=1, jitter=0, seed=4,
                      sky_scale=0.0)
    checked = 0
    for y in range(H):
        for x in range(W):
            if idx[y, x] not in (0, 1):
                continue
            # the reference camera, CPUPathTracer.cpp:62-73: corner sampling, v = 1 - y/H
            ux = ((x / W) * 2 - 1) * (W / H) * tan
            uy = ((1 - y / H) * 2 - 1) * tan
            tt = -1.0 / uy                      # floor y = -1
            px, pz = ux * tt, tt
            F = rect_form_factor(px, pz, -0.5, 1.0, 3.5, 5.0, 2.0)
            want = a * Le * F                   # a/pi * (pi * Le * F)
            assert np.allclose(img[y, x], want, rtol=0.04, atol=1e-4), (x, y, img[y, x], want)
            checked += 1
    assert checked >= 30
    floor = np.isin(idx, [0, 1])
    ux = ((np.arange(W)[None, :] / W) * 2 - 1) * (W / H) * tan
    uy = ((1 - np.arange(H)[:, None] / H) * 2 - 1) * tan
    Fs = np.array([[rect_form_factor(ux[0, x] * (-1 / uy[y, 0]), -1 / uy[y, 0], -0.5, 1.0, 3.5, 5.0, 2.0) if floor[y, x] else 0.0
                    for x in range(W)] for y in range(H)])
    ratio = img[floor].sum(axis=0) / (a * Le * Fs[floor].sum())
    assert np.allclose(ratio, 1.0, rtol=0.006), ratio   # all floor pixels together: 0.6 %


@pytest.mark.parametrize("impl", IMPLS)
def test_fully_occluded_emitter_gives_exactly_zero(impl, request, scenes):
    mats = np.array([scenes.material((0.8, 0.8, 0.8)), scenes.material((0, 0, 0), emission=(10, 6, 2)), scenes.material((0, 0, 0))],
                    scenes.MATERIAL_DTYPE)
    img, idx = render(impl, request, light_scene(blocker=True), mats, 12, 8, 64, cam=dict(tan_half_fov=0.5), max_bounces=1, rr_start=1000, nee=1,
                      jitter=1, seed=5, sky_scale=0.0)
    floor = np.isin(idx, [0, 1])
    assert floor.sum() >= 30 and np.all(img[floor] == 0.0)


# ---- dielectric slab at normal incidence -----------------------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_dielectric_slab_normal_incidence(impl, request, scenes):
    ior = 1.5
    r0 = ((1 - ior) / (1 + ior)) ** 2
    b = Builder()
    b.add(quad([-50, -50, 3], [50, -50, 3], [50, 50, 3], [-50, 50, 3]), 0)      # slab, front face (normal -z: towards the camera)
    b.add(quad([-50, -50, 4], [-50, 50, 4], [50, 50, 4], [50, -50, 4]), 0)      # slab, back face (normal +z: out of the glass)
    b.add(quad([-90, -90, 9], [90, -90, 9], [90, 90, 9], [-90, 90, 9]), 1)      # red emitter behind the slab, facing the camera
    b.add(quad([-90, -90, -2], [-90, 90, -2], [90, 90, -2], [90, -90, -2]), 2)  # green emitter behind the camera, facing +z
    v, t, m = b.mesh()
    n = np.cross(v[t[:, 1]] - v[t[:, 0]], v[t[:, 2]] - v[t[:, 0]])
    assert n[0, 2] < 0 and n[2, 2] > 0 and n[4, 2] < 0 and n[6, 2] > 0
    mats = np.array([scenes.material((1, 1, 1), scenes.DIELECTRIC, param=ior), scenes.material((0, 0, 0), emission=(1, 0, 0)),
                     scenes.material((0, 0, 0), emission=(0, 1, 0))], scenes.MATERIAL_DTYPE)
    N = 65536
    img, idx = render(impl, request, (v, t, m), mats, 2, 2, N, max_bounces=64, rr_start=1000, nee=0, jitter=0, seed=6, sky_scale=0.0)
    T, R = img[1, 1, 0], img[1, 1, 1]  # pixel (1, 1) of a 2x2 image without jitter: u = v = 0.5, the ray is (0, 0, 1) exactly
    want_T, want_R = (1 - r0) / (1 + r0), 2 * r0 / (1 + r0)
    sT = math.sqrt(want_T * (1 - want_T) / N)
    assert abs(T - want_T) <= 5 * sT and abs(R - want_R) <= 5 * sT, (T, want_T, R, want_R)
    assert abs(T + R - 1.0) < 1e-9  # nothing is absorbed, every path ends on one of the two emitters


# ---- GGX-coated diffuse: directional albedo by quadrature ------------------------------------------------------------
def coated_albedo_quadrature(cos_o, alpha, base, f0=0.04, n_theta=1024, n_phi=1024):
    """int f cos_i dw_i over the hemisphere for f = F(o.h) D(h) G1(o) G1(i) / (4 cos_o cos_i) + base (1 - F(cos_o)) / pi,
    midpoint rule in (theta_i, phi_i)."""
    th = (np.arange(n_theta) + 0.5) * (0.5 * np.pi / n_theta)
    ph = (np.arange(n_phi) + 0.5) * (2 * np.pi / n_phi)
    TH, PH = np.meshgrid(th, ph, indexing="ij")
    wi = np.stack([np.sin(TH) * np.cos(PH), np.sin(TH) * np.sin(PH), np.cos(TH)], -1)
    wo = np.array([math.sqrt(1 - cos_o ** 2), 0.0, cos_o])
    h = wi + wo
    h /= np.linalg.norm(h, axis=-1, keepdims=True)
    noh, voh, noi = h[..., 2], (h * wo).sum(-1), wi[..., 2]
    a2 = alpha * alpha
    D = a2 / (np.pi * ((noh * noh) * (a2 - 1) + 1) ** 2)
    g1 = lambda x: 2 * x / (x + np.sqrt(a2 + (1 - a2) * x * x))
    F = f0 + (1 - f0) * (1 - voh) ** 5
    spec = F * D * g1(cos_o) * g1(noi) / (4 * cos_o * noi)
    dw = np.sin(TH) * (0.5 * np.pi / n_theta) * (2 * np.pi / n_phi)
    e_spec = float((spec * noi * dw).sum())
    e_diff = base * (1 - (f0 + (1 - f0) * (1 - cos_o) ** 5))
    return e_spec + e_diff


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("alpha", [0.5, 0.25])
def test_ggx_coated_directional_albedo(impl, request, scenes, alpha):
    base = 1.0
    b = Builder()
    b.add(quad([-400, -1, -400], [400, -1, -400], [400, -1, 400], [-400, -1, 400]), 0)  # plane y = -1, normal +y
    v, t, m = b.mesh()
    assert np.cross(v[t[0, 1]] - v[t[0, 0]], v[t[0, 2]] - v[t[0, 0]])[1] > 0
    mats = np.array([scenes.material((base, base, base), scenes.GGX_COATED, param=alpha)], scenes.MATERIAL_DTYPE)
    env = np.ones((1, 1, 3), np.float32)  # uniform environment of radiance 1
    W, H, N, tan = 4, 16, 16384, 1.0
    img, idx = render(impl, request, (v, t, m), mats, W, H, N, cam=dict(tan_half_fov=tan), env=env, max_bounces=2, rr_start=1000, nee=0, jitter=0,
                      seed=7, sky_scale=1.0)
    rows = 0
    for y in range(H // 2 + 2, H):
        assert (idx[y] == 0).all() or (idx[y] == 1).all() or np.isin(idx[y], [0, 1]).all()
        got, want = [], []
        for x in range(W):
            ux = ((x / W) * 2 - 1) * (W / H) * tan
            uy = ((1 - y / H) * 2 - 1) * tan
            cos_o = -uy / math.sqrt(ux * ux + uy * uy + 1)
            want.append(coated_albedo_quadrature(cos_o, alpha, base))
            got.append(img[y, x].mean())
            assert np.allclose(img[y, x], img[y, x, 0])  # grey material, white light
        got, want = np.array(got), np.array(want)
        assert np.all(want <= 1.0 + 1e-6)
        assert np.allclose(got, want, rtol=0.05), (y, got, want)          # one pixel: 16 384 samples of a heavy-ish tailed weight
        assert abs(got.mean() / want.mean() - 1) < 0.025, (y, got.mean(), want.mean())
        rows += 1
    assert rows >= 5
    # energy: a white base under a white furnace reflects at most 1 (what is missing is the single-scattering loss of G1 G1)
    hit = np.isin(idx, [0, 1])
    assert img[hit].mean() <= 1.0 and img[hit].mean() > 0.75

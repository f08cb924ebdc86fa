"""north_star's image bar: "converged images agree within a stated per-pixel RMSE / relative-MSE tolerance".

Two statements, both through the C ABI:
 1. Same RNG streams (the reference's PCG seeding, CPT.cpp:192-195): the GPU image IS the reference image --
    RMSE <= 1e-6 and relMSE <= 1e-10 of the float image (accum / frames); measured 0 (bit-identical accumulation,
    tests/test_gpu_spheres.py).  The same bar for the mesh integrator against the CPU twin.
 2. INDEPENDENT RNG streams (GPU: counter-based Philox; CPU: the reference's PCG): both estimators converge to the same
    image -- relMSE(GPU Philox, reference PCG) <= 4 / N of the pixel-mean-normalised variance bound, falling ~1/N as
    the sample count grows (N = 64 -> 512 must shrink the error at least 4x of the ideal 8x).  A biased kernel
    (wrong cosine lobe, wrong roulette weight, missing sky term) fails this however many samples it takes.
relMSE = mean over pixels and channels of (a - b)^2 / (b^2 + 1e-2), b = the reference image."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RMSE_TOL = 1e-6      # same-stream bar, linear radiance units (image values are O(1))
RELMSE_TOL = 1e-10


def rmse(a, b):
    return float(np.sqrt(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)))


def relmse(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.mean((a - b) ** 2 / (b * b + 1e-2)))


def reference_image(oracle, sph, W, H, N):
    """The reference's own code (Tier 1) when its library travelled to this box, else the twin (bit-identical to it)."""
    if oracle.ref_available():
        s = oracle.RefSession(sph, W, H)
        s.render(N)
        acc, _, fc = s.read()
        assert fc == N
    else:
        acc, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H), sph, N)
    return acc[..., :3] / np.float32(N)


def gpu_image(pt, sph, W, H, N, **cfg):
    with pt.PathTracer(pt.default_config(**cfg)) as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(sph)
        tr.render(N)
        acc, fc = tr.read_accum()
    assert fc == N
    return acc[..., :3] / np.float32(N)


def test_same_streams_rmse_against_the_reference(pt, oracle, scenes):
    W, H, N = 320, 180, 64
    sph = scenes.app_spheres()
    ref = reference_image(oracle, sph, W, H, N)
    gpu = gpu_image(pt, sph, W, H, N)
    e, r = rmse(gpu, ref), relmse(gpu, ref)
    print(f"spheres, same streams, {N} spp: RMSE {e:.3e} relMSE {r:.3e}")
    assert e <= RMSE_TOL and r <= RELMSE_TOL


def test_independent_streams_converge_to_the_reference_image(pt, oracle, scenes):
    W, H = 96, 54
    sph = scenes.app_spheres()
    ref = reference_image(oracle, sph, W, H, 2048)           # the reference estimator, well converged
    err = {}
    for N in (64, 512):
        gpu = gpu_image(pt, sph, W, H, N, rng_mode=pt.RNG_PHILOX, seed=11)   # other RNG, other streams, no jitter
        err[N] = relmse(gpu, ref)
    print(f"spheres, Philox vs the reference's PCG: relMSE {err[64]:.3e} at 64 spp, {err[512]:.3e} at 512 spp")
    assert err[512] <= 2e-4                      # absolute bar at 512 spp (measured 4.8e-5: Monte-Carlo noise only)
    assert err[64] / err[512] >= 4.0             # ~1/N: the two estimators have the same expectation


def test_mesh_same_streams_rmse_against_the_twin(pt, oracle, scenes):
    W, H, N = 96, 54, 32
    v, t, m, mats = scenes.cornell_box(sphere_res=8)
    kw = dict(rng_mode=1, max_bounces=6, sky_scale=0.0, jitter=1, seed=5, nee=1)
    with pt.PathTracer(pt.default_config(samples_per_pass=32, **kw)) as tr:
        tr.set_camera(pt.default_camera(tan_half_fov=0.36))
        tr.set_resolution(W, H)
        tr.upload_mesh(v, t, m, mats)
        tr.render(N)
        acc, _ = tr.read_accum()
    want, _, _, _ = oracle.TwinMesh(v, t, m, mats).render(oracle.twin_params(W, H, cam_tan_half_fov=0.36, **kw), N)
    a, b = acc[..., :3] / np.float32(N), want[..., :3] / np.float32(N)
    e, r = rmse(a, b), relmse(a, b)
    print(f"cornell, same streams, {N} spp: RMSE {e:.3e} relMSE {r:.3e}")
    assert e <= RMSE_TOL and r <= RELMSE_TOL


def test_mesh_nee_and_plain_path_tracing_converge_to_the_same_image(pt, scenes):
    """Two different estimators of the same integral on the GPU (next-event estimation on / off) agree within their
    noise: the light-sampling weights and the emission bookkeeping (no double counting) are right."""
    W, H = 64, 36
    v, t, m, mats = scenes.cornell_box(sphere_res=8)
    img = {}
    for nee, N in ((1, 512), (0, 4096)):
        kw = dict(rng_mode=1, max_bounces=6, sky_scale=0.0, jitter=1, seed=3 + nee, nee=nee)
        with pt.PathTracer(pt.default_config(samples_per_pass=128, **kw)) as tr:
            tr.set_camera(pt.default_camera(tan_half_fov=0.36))
            tr.set_resolution(W, H)
            tr.upload_mesh(v, t, m, mats)
            tr.render(N)
            acc, _ = tr.read_accum()
        img[nee] = acc[..., :3].astype(np.float64) / N
    # compare 4x4 block means (the plain estimator is noisy per pixel)
    blk = lambda a: a.reshape(H // 4, 4, W // 4, 4, 3).mean(axis=(1, 3))
    a, b = blk(img[1]), blk(img[0])
    r = float(np.mean((a - b) ** 2 / (b * b + 1e-2)))
    print(f"cornell, NEE vs plain path tracing, block means: relMSE {r:.3e}; mean radiance {a.mean():.4f} vs {b.mean():.4f}")
    assert abs(a.mean() - b.mean()) <= 0.03 * b.mean()
    assert r <= 5e-3                             # measured 4.3e-4

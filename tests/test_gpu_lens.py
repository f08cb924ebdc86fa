"""Thin-lens camera (cudapt_camera.lens_radius / focus_distance): origin and direction of every camera
ray equal the CPU twin's bit for bit (lens point by rejection from the Philox stream: no transcendental functions), so
the sphere and mesh images do too, within the bars of the pinhole tests.  The reference has no camera object at all
(CPT.cpp:62-73 hard-codes a pinhole); lens_radius = 0 must leave every pinhole result unchanged."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_frac(a, b):
    return float(np.all(bits(a) == bits(b), axis=-1).mean())


def test_lens_spheres_equal_twin(pt, oracle, scenes):
    sph = scenes.app_spheres()
    W, H, N = 160, 90, 8
    cam = dict(lens_radius=0.15, focus_distance=5.0)
    kw = dict(max_bounces=4, jitter=1, seed=21)
    with pt.PathTracer(pt.default_config(rng_mode=1, **kw)) as tr:
        tr.set_camera(pt.default_camera(**cam))
        tr.set_resolution(W, H)
        tr.upload_spheres(sph)
        tr.render(N)
        acc, _ = tr.read_accum()
        tr.set_camera(pt.default_camera())
        tr.render(N)
        pin, _ = tr.read_accum()
    want = oracle.twin_render_spheres(oracle.twin_params(W, H, rng_mode=1, cam_lens_radius=0.15, cam_focus_distance=5.0, **kw), sph, N)[0]
    f = same_frac(acc, want)
    print(f"thin lens, spheres: {f:.5f} of pixels bit-identical")
    assert f >= 0.9999  # sphere path: the double-precision sincos allowance of tests/test_gpu_spheres.py
    assert not np.array_equal(bits(acc), bits(pin))  # the lens does something


@pytest.mark.parametrize("spp_pass", [2, 32])
def test_lens_mesh_equal_twin(pt, oracle, scenes, spp_pass):
    """spp_pass 32 would use camera-ray packets with a pinhole; with a lens it must fall back to per-ray tracing."""
    scene = scenes.cornell_box(sphere_res=8)
    W, H, N = 96, 54, 32
    kw = dict(max_bounces=5, jitter=1, nee=1, sky_scale=0.0, seed=22)
    with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=spp_pass, **kw)) as tr:
        tr.set_camera(pt.default_camera(tan_half_fov=0.36, lens_radius=0.08, focus_distance=4.0))
        tr.set_resolution(W, H)
        tr.upload_mesh(*scene)
        tr.render(N)
        acc, _ = tr.read_accum()
        prim, t = tr.read_primary_hits()
        st = tr.stats()
    assert st.packet_rays == 0
    want, widx, wt, _ = oracle.TwinMesh(*scene).render(
        oracle.twin_params(W, H, rng_mode=1, cam_tan_half_fov=0.36, cam_lens_radius=0.08, cam_focus_distance=4.0, **kw), N, want_primary=True)
    got = np.where(prim == 0xFFFFFFFF, -1, prim.astype(np.int64))
    ok = (got == widx) & (bits(t) == bits(wt))
    f = same_frac(acc, want)
    print(f"thin lens, Cornell mesh: primary {ok.mean():.5f}, accumulation {f:.5f}")
    assert ok.all() and f == 1.0  # mesh path: every pixel (tests/test_gpu_mesh.py)


def test_lens_in_focus_plane_is_sharp(pt, scenes):
    """A quad AT the focus distance images exactly as with the pinhole (every lens ray of a pixel sample passes through
    the same point of the focal plane), a quad off that plane blurs: its edge pixels take intermediate values."""
    def quad(z, half):
        v = np.array([[-half, -half, z], [half, -half, z], [half, half, z], [-half, half, z]], np.float32)
        return v, np.array([[0, 2, 1], [0, 3, 2]], np.uint32)  # facing the camera (emission is one-sided)

    def coverage(z, lens):
        v, t = quad(z, 0.4931 * z)  # same projected size at every depth; edges between pixel corners
        mats = np.array([scenes.material((0.0, 0.0, 0.0), scenes.LAMBERT, emission=(1.0, 1.0, 1.0))], scenes.MATERIAL_DTYPE)
        cfg = pt.default_config(rng_mode=1, max_bounces=1, jitter=0, sky_scale=0.0, seed=5, samples_per_pass=16)
        with pt.PathTracer(cfg) as tr:
            tr.set_camera(pt.default_camera(lens_radius=lens, focus_distance=4.0))
            tr.set_resolution(64, 64)
            tr.upload_mesh(v, t, np.zeros(2, np.uint32), mats)
            tr.render(64)
            acc, _ = tr.read_accum()
        return acc[..., 0] / 64.0  # emission 1 where the quad is seen, 0 elsewhere: the coverage of the pixel

    sharp, pin = coverage(4.0, 0.3), coverage(4.0, 0.0)
    far = coverage(12.0, 0.3)
    frac = lambda img: float(((img > img.min() + 1e-3) & (img < img.max() - 1e-3)).mean())
    print(f"partially covered pixels: in focus {frac(sharp):.4f}, pinhole {frac(pin):.4f}, out of focus {frac(far):.4f}")
    assert frac(pin) == 0.0 and frac(sharp) <= 0.02 and frac(far) > 0.05

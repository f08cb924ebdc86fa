"""GPU parity of the octahedral environment map (cudapt_upload_environment) against the CPU twin.

The lookup uses only + - * / and comparisons, so the bar is the same as for the integrators themselves: sphere
path bit-identical on >= 99.99 % of pixels, mesh path on >= 99.9 %.  A miss-only pixel is a closed-form check:
after N frames its value is N times the bilinear texel blend, summed in float.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_frac(a, b):
    return float(np.all(bits(a) == bits(b), axis=-1).mean())


def test_spheres_with_environment_match_twin(pt, oracle, scenes):
    W, H, N = 160, 90, 6
    env = scenes.octahedral_environment(48)
    sph = scenes.app_spheres()
    with pt.PathTracer() as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(sph)
        tr.upload_environment(env)
        tr.render(N)
        acc, fc = tr.read_accum()
        want, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H, env=env), sph, N)
        f = same_frac(acc, want)
        print(f"spheres + environment: bit-identical pixels {f:.5f}")
        assert fc == N and f >= 0.9999
        # removing the map restores the reference sky (and restarts accumulation)
        tr.upload_environment(None)
        tr.render(N)
        acc2, fc2 = tr.read_accum()
        want2, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H), sph, N)
        assert fc2 == N and same_frac(acc2, want2) >= 0.9999
        assert same_frac(acc, acc2) < 0.5


def test_mesh_with_environment_matches_twin(pt, oracle, scenes):
    W, H, N = 128, 72, 4
    env = scenes.octahedral_environment(32, sun_dir=(-0.3, 0.5, 0.4))
    v, t, m, mats = scenes.heightfield(n_quads=40)
    kw = dict(max_bounces=6, jitter=1, seed=11, sky_scale=0.5)
    with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=2, **kw)) as tr:
        tr.set_resolution(W, H)
        tr.upload_mesh(v, t, m, mats)
        tr.upload_environment(env)
        tr.render(N)
        acc, fc = tr.read_accum()
    mesh = oracle.TwinMesh(v, t, m, mats)
    want, _, _, _ = mesh.render(oracle.twin_params(W, H, rng_mode=1, env=env, **kw), N)
    f = same_frac(acc, want)
    print(f"mesh + environment: bit-identical pixels {f:.5f}")
    assert fc == N and f == 1.0  # mesh path + octahedral lookup: + - * / only, every pixel


def test_environment_edge_cases(pt, oracle, scenes):
    # 1x1 map: every direction reads the single texel; miss-only image = N * texel exactly
    W, H, N = 64, 36, 3
    env = np.array([[[0.25, 0.5, 2.0]]], np.float32)
    far = np.array([[0.0, 0.0, -50.0, 1.0]], np.float32)  # behind the camera: every primary ray misses
    with pt.PathTracer() as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(far)
        tr.upload_environment(env)
        tr.render(N)
        acc, _ = tr.read_accum()
        px = np.float32(0)
        want = np.zeros(4, np.float32)
        for _ in range(N):
            want = (want + np.array([0.25, 0.5, 2.0, 1.0], np.float32)).astype(np.float32)
        assert np.all(bits(acc) == bits(np.broadcast_to(want, acc.shape)))
        with pytest.raises(pt.CudaptError):
            tr._chk(tr._lib.cudapt_upload_environment(tr._h, env.ctypes.data, 100000))


def test_environment_from_file_matches_twin_and_cli(pt, oracle, scenes, tmp_path):
    """cudapt_load_environment: a lat-long .pfm panorama, resampled on the host, renders exactly like the same texels
    uploaded directly -- and like the CPU twin given those texels; cudapt_cli --env writes the same image."""
    import subprocess
    W, H, N = 96, 54, 4
    yy, xx = np.mgrid[0:32, 0:64].astype(np.float32)
    pano = np.stack([0.3 + xx / 64, 1.2 - yy / 32, 0.5 + 0.0 * xx], axis=-1).astype(np.float32)
    pano[6:9, 40:44] = 40.0                                               # a small sun
    path = str(tmp_path / "sky.pfm")
    with open(path, "wb") as f:
        f.write(b"PF\n64 32\n-1.0\n" + np.ascontiguousarray(pano[::-1], "<f4").tobytes())
    sph = scenes.app_spheres()
    with pt.PathTracer() as tr:
        tr.set_resolution(W, H)
        tr.upload_spheres(sph)
        octa = tr.load_environment(path, 64)
        assert np.array_equal(octa, pt.convert_environment(path, 64))
        tr.render(N)
        a_file, _ = tr.read_accum()
        tr.upload_environment(octa)
        tr.render(N)
        a_direct, _ = tr.read_accum()
        with pytest.raises(pt.CudaptError):
            tr.load_environment(str(tmp_path / "missing.hdr"), 64)
        tr.render(1)                                                      # the context survives the failed load
    assert np.array_equal(a_file.view(np.uint32), a_direct.view(np.uint32))
    want, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H, env=octa), sph, N)
    assert same_frac(a_file, want) >= 0.9999
    cli = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "cudapt_cli")
    if os.path.exists(cli):
        out = str(tmp_path / "cli.pfm")
        subprocess.run([cli, "--size", f"{W}x{H}", "--spp", str(N), "--env", path, "--env-size", "64", "--out", out], check=True,
                       capture_output=True)
        raw = open(out, "rb").read()
        head = f"PF\n{W} {H}\n-1.0\n".encode()
        img = np.frombuffer(raw[len(head):], "<f4").reshape(H, W, 3)[::-1]
        assert np.array_equal(img.view(np.uint32), (a_file[..., :3] / np.float32(N)).view(np.uint32))

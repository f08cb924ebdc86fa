"""BVH build on the device (csrc/lbvh_build.cu, cudapt_build_accel_device): Morton codes + radix sort + binary radix
tree + collapse to 8-wide nodes + quantisation by the refit kernels.  The closest hit does not depend on the tree, so
every image must equal the host-built (SAH) tree's BIT FOR BIT and the CPU twin's within the mesh parity bar; refit,
re-upload and staging work on top of a device-built tree."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def render(pt, scene, W, H, N, device_build, cam=None, spp_pass=2, **kw):
    tr = pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=spp_pass, **kw))
    if cam:
        tr.set_camera(pt.default_camera(**cam))
    tr.set_resolution(W, H)
    tr.upload_mesh(*scene, device_build=device_build)
    tr.render(N)
    acc, _ = tr.read_accum()
    prim, t = tr.read_primary_hits()
    return tr, acc, prim, t


def soup(n, seed, duplicates=True):
    """Random small triangles in front of the camera, with exact duplicates (equal Morton codes), a few degenerate
    ones and a cluster of many triangles in one spot."""
    rng = np.random.default_rng(seed)
    c = rng.uniform([-3, -2, 3], [3, 2, 9], (n, 3)).astype(np.float32)
    v = (c[:, None, :] + rng.uniform(-0.25, 0.25, (n, 3, 3))).astype(np.float32)
    if duplicates:
        v[n // 2:n // 2 + n // 8] = v[:n // 8]                    # exact copies
        v[-n // 8:] = v[0] + rng.uniform(-1e-4, 1e-4, (n // 8, 3, 3)).astype(np.float32)  # one dense cluster
        v[5, 1] = v[5, 0]                                          # degenerate
        v[6, 2] = v[6, 1] = v[6, 0]
    return v.reshape(-1, 3), np.arange(3 * n, dtype=np.uint32).reshape(-1, 3), None, None


@pytest.mark.parametrize("name", ["heightfield", "cornell", "soup", "two_triangles", "soup_small"])
def test_device_built_tree_gives_identical_images(pt, oracle, scenes, name):
    kw = dict(max_bounces=4, jitter=1, seed=31)
    cam = None
    if name == "heightfield":
        scene = scenes.heightfield(n_quads=96)
    elif name == "cornell":
        scene = scenes.cornell_box(sphere_res=8)
        kw.update(nee=1, sky_scale=0.0)
        cam = dict(tan_half_fov=0.36)
    elif name == "soup":
        scene = soup(4000, 3)
    elif name == "soup_small":
        scene = soup(9, 4, duplicates=False)
    else:
        v = np.array([[-2, -2, 5], [2, -2, 5], [2, 2, 5], [-2, 2, 5]], np.float32)
        scene = (v, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), None, None)
    W, H, N = 128, 72, 4
    trd, accd, primd, td = render(pt, scene, W, H, N, True, cam=cam, **kw)
    trh, acch, primh, th = render(pt, scene, W, H, N, False, cam=cam, **kw)
    sd, sh = trd.stats(), trh.stats()
    assert sd.n_triangles == sh.n_triangles == len(scene[1]) and sd.n_nodes >= 1
    assert np.array_equal(primd, primh) and np.array_equal(bits(td), bits(th))
    assert np.array_equal(bits(accd), bits(acch))
    assert sd.rays == sh.rays
    if name in ("heightfield", "cornell"):
        tw = {("cam_" + k): v for k, v in (cam or {}).items()}
        want = oracle.TwinMesh(*scene).render(oracle.twin_params(W, H, rng_mode=1, **kw, **tw), N)[0]
        assert float(np.all(bits(accd) == bits(want), axis=-1).mean()) == 1.0
    print(f"{name}: device build {sd.accel_build_ms:.2f} ms, {sd.n_nodes} nodes | host build {sh.accel_build_ms:.2f} ms, {sh.n_nodes} nodes")
    trd.close(), trh.close()


def test_refit_packets_and_staging_on_a_device_built_tree(pt, scenes):
    scene = scenes.heightfield(n_quads=64)
    v = scene[0]
    kw = dict(max_bounces=4, jitter=1, seed=33)
    tr, acc, _, _ = render(pt, scene, 160, 90, 32, True, spp_pass=32, **kw)   # packets on a device-built tree
    trh, acch, _, _ = render(pt, scene, 160, 90, 32, False, spp_pass=32, **kw)
    assert tr.stats().packet_rays > 0 and np.array_equal(bits(acc), bits(acch))
    w = v.copy()
    w[:, 1] += (0.07 * np.sin(2.5 * v[:, 0])).astype(np.float32)
    tr.update_vertices(w)
    trh.update_vertices(w)
    tr.render(32), trh.render(32)
    moved = tr.read_accum()[0]
    assert np.array_equal(bits(moved), bits(trh.read_accum()[0])) and not np.array_equal(bits(moved), bits(acc))
    chk = tr.accel_checksum()
    tr.reupload_scene()
    assert tr.accel_checksum() == chk
    tr.stage_scene()
    tr.render(32)
    assert np.array_equal(bits(tr.read_accum()[0]), bits(moved))
    tr.close(), trh.close()


def test_device_build_of_one_million_triangles(pt, scenes):
    scene = scenes.heightfield()
    assert len(scene[1]) == 1_002_528
    kw = dict(max_bounces=3, jitter=1, seed=35)
    trd, accd, primd, td = render(pt, scene, 480, 270, 2, True, **kw)
    trh, acch, primh, th = render(pt, scene, 480, 270, 2, False, **kw)
    assert np.array_equal(primd, primh) and np.array_equal(bits(accd), bits(acch))
    t0 = time.perf_counter()
    trd.build_accel_device()
    ms = (time.perf_counter() - t0) * 1e3
    sd, sh = trd.stats(), trh.stats()
    print(f"1,002,528 triangles: device build {ms:.1f} ms (incl. mesh upload and read-back of the tree; {sd.n_nodes} nodes), "
          f"host SAH build {sh.accel_build_ms:.0f} ms ({sh.n_nodes} nodes)")
    assert ms < sh.accel_build_ms
    trd.close(), trh.close()


@pytest.mark.parametrize("name", ["heightfield", "cornell"])
def test_rebuild_in_place_reuses_buffers_and_matches_a_fresh_build(pt, scenes, name):
    """A second device build of a mesh of the SAME size keeps every allocation (mesh_host.cu build_on_device); the
    pinned staging and the host copy of the tree are fetched lazily.  Moved vertices + rebuild must equal a fresh
    tracer's build bit for bit, including through staging, re-upload and with emissive triangles (light list)."""
    kw = dict(max_bounces=4, jitter=1, seed=5)
    cam = None
    if name == "heightfield":
        scene = scenes.heightfield(n_quads=64)
    else:
        scene = scenes.cornell_box(sphere_res=8)
        kw.update(nee=1, sky_scale=0.0)
        cam = dict(tan_half_fov=0.36)
    v0, idx, mid, mats = scene
    rng = np.random.default_rng(2)
    v1 = (v0 + rng.uniform(-0.02, 0.02, v0.shape)).astype(np.float32)
    W, H, N = 96, 54, 4
    _, want, want_prim, _ = render(pt, (v1, idx, mid, mats), W, H, N, True, cam, **kw)

    tr, _, _, _ = render(pt, scene, W, H, N, True, cam, **kw)
    for step in range(3):
        tr.upload_mesh(v1 if step != 1 else v0, idx, mid, mats, device_build=True)
        if step == 2:
            tr.stage_scene()                                   # shadow copy of the rebuilt tree, swapped at render
        tr.reset()
        tr.render(N)
        if step != 1:
            acc, _ = tr.read_accum()
            prim, _ = tr.read_primary_hits()
            assert np.array_equal(prim, want_prim)
            assert np.array_equal(bits(acc), bits(want))
    # after the swap the current node buffer is the former shadow: rebuild into it, then re-upload from staging
    tr.upload_mesh(v1, idx, mid, mats, device_build=True)
    tr.reupload_scene()
    tr.reset()
    tr.render(N)
    acc, _ = tr.read_accum()
    assert np.array_equal(bits(acc), bits(want))

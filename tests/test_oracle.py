"""CPU tests of the oracles: Tier 1 (reference TUs + Embree stand-in) vs Tier 2 (twin) vs golden fixtures
and closed-form facts of the reference source (SURVEY.md 8c "known-answer checks")."""
import os

import numpy as np
import pytest


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_twin_matches_golden_bit_for_bit(oracle, golden):
    """The fixture was produced by the reference's own CPUPathTracer.cpp (Tier 1)."""
    W, H = int(golden["width"]), int(golden["height"])
    p = oracle.twin_params(W, H)
    acc, idx, t, c = oracle.twin_render_spheres(p, golden["spheres"], 4, want_primary=True)
    assert np.array_equal(bits(acc), bits(golden["accum4"]))
    assert np.array_equal(idx, golden["primary_idx"])
    assert np.array_equal(bits(t), bits(golden["primary_t"]))
    acc, _, _, c2 = oracle.twin_render_spheres(p, golden["spheres"], 12, first_frame=5, accum=acc)
    assert np.array_equal(bits(acc), bits(golden["accum16"]))
    assert c.rays + c2.rays == int(golden["rays16"])
    assert np.array_equal(oracle.twin_resolve_rgba8(acc, 16), golden["rgba16"])
    assert np.array_equal(oracle.twin_resolve_rgba8(golden["accum4"], 4), golden["rgba4"])


MESH_KW = dict(rng_mode=1, max_bounces=6, jitter=1, nee=1, sky_scale=0.0, seed=17, cam_tan_half_fov=0.36)


def test_twin_mesh_matches_committed_vectors(oracle, golden_mesh):
    """The twin's mesh integrator (triangles, materials, NEE, Philox, thin lens) reproduces the committed fixture bit
    for bit: a change of the mesh specification cannot slip in unnoticed.  (Twin-generated: parity unpinned by the
    reference, which has no triangle path.)"""
    g = golden_mesh
    W, H = int(g["width"]), int(g["height"])
    mesh = oracle.TwinMesh(g["vertices"], g["triangles"], g["material_id"], g["materials"])
    acc, idx, t, c = mesh.render(oracle.twin_params(W, H, **MESH_KW), 4, want_primary=True)
    assert np.array_equal(bits(acc), bits(g["accum4"])) and c.rays == int(g["rays4"])
    assert np.array_equal(idx, g["primary_idx"]) and np.array_equal(bits(t), bits(g["primary_t"]))
    acc, _, _, c = mesh.render(oracle.twin_params(W, H, **MESH_KW), 32)
    assert np.array_equal(bits(acc), bits(g["accum32"])) and c.rays == int(g["rays32"])
    acc, _, _, _ = mesh.render(oracle.twin_params(W, H, cam_lens_radius=0.05, cam_focus_distance=4.0, **MESH_KW), 32)
    assert np.array_equal(bits(acc), bits(g["accum32_lens"]))


def test_reference_matches_golden(oracle, golden):
    """Tier 1 rebuilt now reproduces the committed fixture (guards compiler/flag drift)."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    W, H = int(golden["width"]), int(golden["height"])
    s = oracle.RefSession(golden["spheres"], W, H)
    assert s.backend_name == "CPU Path Tracer (Embree)"
    s.render(4)
    acc, rgba, fc = s.read()
    assert fc == 4
    assert np.array_equal(bits(acc), bits(golden["accum4"]))
    assert np.array_equal(rgba, golden["rgba4"])


@pytest.mark.parametrize("size", [(64, 36), (40, 40), (33, 17)])
def test_twin_vs_reference_other_sizes(oracle, scenes, size):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built")
    W, H = size
    sph = scenes.app_spheres()
    s = oracle.RefSession(sph, W, H)
    s.render(6)
    acc, rgba, fc = s.read()
    ridx, rt = s.primary()
    p = oracle.twin_params(W, H)
    acc2, idx, t, c = oracle.twin_render_spheres(p, sph, 6, want_primary=True)
    assert np.array_equal(bits(acc), bits(acc2))
    assert np.array_equal(ridx, idx) and np.array_equal(bits(rt), bits(t))
    assert c.rays == s.rays
    assert np.array_equal(rgba, oracle.twin_resolve_rgba8(acc2, 6))


def test_pcg_known_answers(oracle, golden):
    for key, words, floats in zip(golden["pcg_keys"], golden["pcg_words"], golden["pcg_floats"]):
        w, f = oracle.pcg_stream(*[int(k) for k in key], 10)
        assert np.array_equal(w, words)
        assert np.array_equal(bits(f), bits(floats))
    # closed form of CPT.cpp:192-195 and the [0, 1] inclusive range of CPT.cpp:300
    x, y, W, frame = 17, 5, 640, 3
    assert oracle._twin().twin_pcg_seed(x, y, W, frame) == (x + y * W + frame * 982451653) % 2**32
    assert np.float32(0xFFFFFF80) / np.float32(4294967295.0) == np.float32(1.0)


def test_philox_known_answers(oracle):
    """Random123 known-answer vectors for philox4x32-10."""
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
        ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
        ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
         [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
    ]
    for ctr, key, want in kat:
        assert [int(x) for x in oracle.philox(ctr, key)] == want


def test_sincos2pi_accuracy(oracle):
    us = np.concatenate([np.linspace(0, 1, 4001), [0.25, 0.5, 0.75, 1.0, 1e-7, 0.9999999]])
    err = 0.0
    for u in us:
        c, s = oracle.sincos2pi(u)
        uf = float(np.float32(u))
        err = max(err, abs(c - np.cos(2 * np.pi * uf)), abs(s - np.sin(2 * np.pi * uf)))
    assert err < 3e-7


def test_miss_only_pixels_closed_form(oracle, golden):
    """a pixel whose primary ray misses accumulates N*sky(dir) with alpha N, identical every
    frame (no jitter); (2) alpha byte is 255 everywhere."""
    W, H = int(golden["width"]), int(golden["height"])
    miss = golden["primary_idx"] < 0
    assert miss.any()
    a4, a16 = golden["accum4"], golden["accum16"]
    assert np.all(a4[..., 3] == 4) and np.all(a16[..., 3] == 16)
    assert np.all((golden["rgba16"] & 0xFF) == 255)
    # sky formula CPT.cpp:286-292 from the primary direction CPT.cpp:62-73
    ys, xs = np.nonzero(miss)
    f = np.float32
    u = xs.astype(f) * (f(1) / f(W))
    v = f(1) - ys.astype(f) * (f(1) / f(H))
    uvx = (u * f(2) - f(1)) * (f(W) / f(H))
    uvy = v * f(2) - f(1)
    ln = np.sqrt(uvx * uvx + uvy * uvy + f(1))
    dy = uvy / ln
    t = f(0.5) * (dy + f(1))
    sky = np.stack([(f(1) - t) + f(0.5) * t, (f(1) - t) + f(0.7) * t, (f(1) - t) + f(1.0) * t], -1).astype(f)
    acc = np.zeros_like(sky)
    for k in range(16):  # CPT.cpp:77-80: one float add per frame, in frame order
        acc = acc + sky
        if k == 3:
            assert np.array_equal(bits(a4[miss][:, :3]), bits(acc))
    assert np.array_equal(bits(a16[miss][:, :3]), bits(acc))


def test_four_hit_paths_contribute_zero_and_throughput_facts():
    """after RR at bounce 3 the throughput is exactly 1.0f."""
    f = np.float32
    thr = f(0.7) * f(0.7) * f(0.7)
    assert thr / thr == f(1.0)
    assert (f(1.0) * f(0.7)) / f(0.7) == f(1.0)


def test_twin_tile_partition_sums_to_full_image(oracle, scenes):
    W, H = 80, 48
    sph = scenes.app_spheres()
    full, _, _, _ = oracle.twin_render_spheres(oracle.twin_params(W, H), sph, 3)
    parts = []
    for r in range(3):
        p = oracle.twin_params(W, H, part_rank=r, part_world=3, tile_w=16, tile_h=8)
        parts.append(oracle.twin_render_spheres(p, sph, 3)[0])
    owned = sum((a[..., 3] > 0).astype(int) for a in parts)
    assert np.all(owned == 1)
    assert np.array_equal(bits(sum(parts)), bits(full))


def test_twin_frame_ranges_compose(oracle, scenes):
    """Frames are independent streams: rendering 1..6 equals 1..3 then 4..6 (checkpoint/resume)."""
    W, H = 48, 27
    sph = scenes.app_spheres()
    p = oracle.twin_params(W, H)
    a, _, _, _ = oracle.twin_render_spheres(p, sph, 6)
    b, _, _, _ = oracle.twin_render_spheres(p, sph, 3)
    b, _, _, _ = oracle.twin_render_spheres(p, sph, 3, first_frame=4, accum=b)
    assert np.array_equal(bits(a), bits(b))


def test_twin_mesh_bvh_matches_bruteforce(oracle, scenes):
    v, t, m, mats = scenes.cornell_box(sphere_res=12)
    mesh = oracle.TwinMesh(v, t, m, mats)
    rng = np.random.default_rng(7)
    o = np.array([0, 0, 0], np.float32)
    n_hit = 0
    for _ in range(400):
        d = np.array([rng.uniform(-0.3, 0.3), rng.uniform(-0.3, 0.3), 1.0])
        d = (d / np.linalg.norm(d)).astype(np.float32)
        a = mesh.closest(o, d)
        b = mesh.closest(o, d, bruteforce=True)
        assert a[0] == b[0] and (a[1] == b[1] or (np.isinf(a[1]) and np.isinf(b[1])))
        n_hit += a[0] >= 0
    assert n_hit > 300
    # rays from inside the box
    for _ in range(300):
        o = np.array([rng.uniform(-0.9, 0.9), rng.uniform(-0.9, 0.9), rng.uniform(3.1, 4.9)], np.float32)
        d = rng.normal(size=3)
        d = (d / np.linalg.norm(d)).astype(np.float32)
        a = mesh.closest(o, d)
        b = mesh.closest(o, d, bruteforce=True)
        assert a[0] == b[0] and a[1] == b[1]


def test_twin_mesh_cornell_energy(oracle, scenes):
    """Sanity of the mesh integrator: NEE and BSDF-only estimators agree in the mean (same expectation)."""
    v, t, m, mats = scenes.cornell_box(sphere_res=10)
    mesh = oracle.TwinMesh(v, t, m, mats)
    W, H, N = 48, 27, 96
    kw = dict(rng_mode=1, max_bounces=6, sky_scale=0.0, jitter=1, seed=11, cam_tan_half_fov=0.36)
    a, _, _, _ = mesh.render(oracle.twin_params(W, H, nee=1, **kw), N)
    b, _, _, _ = mesh.render(oracle.twin_params(W, H, nee=0, **kw), N * 4)
    ma, mb = a[..., :3].mean() / N, b[..., :3].mean() / (N * 4)
    assert ma > 0.02
    assert abs(ma - mb) / mb < 0.12


def test_post_resolve_defaults_equal_reference_resolve(oracle):
    """The numpy restatement of the resolve/post-process stage degenerates to the reference's resolve
    (CPT.cpp:87-117, here through the twin, which is pinned bit-for-bit to the reference TUs above)."""
    rng = np.random.default_rng(7)
    acc = (rng.random((37, 53, 4), dtype=np.float32) * 9.0).astype(np.float32)
    acc[..., 3] = 6.0
    acc[0, 0, :3] = [-1.0, 0.0, 1e9]
    assert np.array_equal(oracle.post_resolve_rgba8(acc, 6)[0], oracle.twin_resolve_rgba8(acc, 6))
    # Reinhard and ACES known answers at c = 1: 1/2 -> 127, (2.54)/(3.16) -> (uint8)(0.8037975 * 255) = 204
    one = np.full((1, 1, 4), 1.0, np.float32)
    assert (oracle.post_resolve_rgba8(one, 1, tonemap=1)[0][0, 0] >> 24) == 127
    assert (oracle.post_resolve_rgba8(one, 1, tonemap=2)[0][0, 0] >> 24) == 204

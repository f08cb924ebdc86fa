"""Incremental scene update: cudapt_update_vertices refits the BVH8 on the device (csrc/refit.cu).
The device kernels and the host refit execute the same __host__ __device__ code (csrc/bvh8_refit.h), so the bytes of
the node and packed-triangle arrays must be IDENTICAL -- checked through FNV-1a checksums of the device copies -- and
the images too.  The reference rebuilds everything on any change (CPT.cpp:122-131,156-160,362-403)."""
import os
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def wobble(v, k, amp=0.05):
    w = v.copy()
    w[:, 1] += (amp * np.sin(3.0 * v[:, 0] + 0.7 * k) * np.cos(2.0 * v[:, 2] - 0.3 * k)).astype(np.float32)
    return w


def tracer(pt, scene, W=160, H=90):
    tr = pt.PathTracer(pt.default_config(rng_mode=1, max_bounces=5, jitter=1, seed=9, samples_per_pass=2))
    tr.set_resolution(W, H)
    tr.upload_mesh(*scene)
    return tr


@pytest.mark.parametrize("n_quads", [1, 7, 96])
def test_device_refit_bytes_equal_host_refit(pt, scenes, n_quads, monkeypatch):
    scene = scenes.heightfield(n_quads=n_quads)
    v = scene[0]
    dev, host = tracer(pt, scene), tracer(pt, scene)
    assert dev.accel_checksum() == host.accel_checksum()
    for k in range(3):  # repeated refits: each starts from the previous refit's tree
        v2 = wobble(v, k)
        monkeypatch.delenv("CUDAPT_REFIT_HOST", raising=False)
        n0 = dev.stats().kernel_launches
        dev.update_vertices(v2)
        assert dev.stats().kernel_launches > n0  # the device kernels ran
        monkeypatch.setenv("CUDAPT_REFIT_HOST", "1")
        n0 = host.stats().kernel_launches
        host.update_vertices(v2)
        assert host.stats().kernel_launches == n0
        assert dev.accel_checksum() == host.accel_checksum()
        dev.render(2), host.render(2)
        assert np.array_equal(bits(dev.read_accum()[0]), bits(host.read_accum()[0]))
    dev.close(), host.close()


def test_device_refit_then_reupload_and_mixed_paths(pt, scenes, monkeypatch):
    """The pinned staging copy lags behind a device refit: reupload_scene and a following HOST refit must both start
    from the refreshed copy."""
    scene = scenes.heightfield(n_quads=24)
    v = scene[0]
    monkeypatch.delenv("CUDAPT_REFIT_HOST", raising=False)
    tr = tracer(pt, scene)
    tr.update_vertices(wobble(v, 1))
    want = tr.accel_checksum()
    tr.render(2)
    img = tr.read_accum()[0]
    tr.reupload_scene()
    assert tr.accel_checksum() == want
    tr.reset()
    tr.render(2)
    assert np.array_equal(bits(tr.read_accum()[0]), bits(img))
    monkeypatch.setenv("CUDAPT_REFIT_HOST", "1")
    tr.update_vertices(wobble(v, 1))
    assert tr.accel_checksum() == want
    monkeypatch.delenv("CUDAPT_REFIT_HOST")
    tr.update_vertices(v)  # back to the original vertices: the triangle records equal the fresh build's
    fresh = tracer(pt, scene)
    assert tr.accel_checksum()[1] == fresh.accel_checksum()[1]
    tr.render(2), fresh.render(2)
    a, b = tr.read_accum()[0], fresh.read_accum()[0]
    assert np.array_equal(bits(a), bits(b))
    tr.close(), fresh.close()


def test_refit_image_equals_twin_on_moved_vertices(pt, oracle, scenes, monkeypatch):
    """The refitted tree is a different (looser) BVH over the same triangles: closest-hit results do not depend on
    the tree, so the image equals the CPU twin's render of the moved mesh within the mesh parity bar."""
    monkeypatch.delenv("CUDAPT_REFIT_HOST", raising=False)
    scene = scenes.heightfield(n_quads=40)
    v, t, m, mats = scene
    v2 = wobble(v, 2, amp=0.15)
    tr = tracer(pt, scene, 128, 72)
    tr.update_vertices(v2)
    tr.render(2)
    acc = tr.read_accum()[0]
    tr.close()
    want, _, _, _ = oracle.TwinMesh(v2, t, m, mats).render(
        oracle.twin_params(128, 72, rng_mode=1, max_bounces=5, jitter=1, seed=9), 2)
    same = float(np.all(bits(acc) == bits(want), axis=-1).mean())
    print(f"refit vs twin: {same:.5f} of pixels identical")
    assert same == 1.0


def test_refit_time_at_one_million_triangles(pt, scenes, monkeypatch):
    scene = scenes.heightfield()
    v = scene[0]
    assert len(scene[1]) == 1_002_528
    monkeypatch.delenv("CUDAPT_REFIT_HOST", raising=False)
    tr = tracer(pt, scene, 320, 180)
    build_ms = tr.stats().accel_build_ms
    tr.update_vertices(wobble(v, 0))  # first call allocates the refit buffers
    t0 = time.perf_counter()
    tr.update_vertices(wobble(v, 1))
    dev_ms = (time.perf_counter() - t0) * 1e3
    want = tr.accel_checksum()
    monkeypatch.setenv("CUDAPT_REFIT_HOST", "1")
    t0 = time.perf_counter()
    tr.update_vertices(wobble(v, 1))
    host_ms = (time.perf_counter() - t0) * 1e3
    assert tr.accel_checksum() == want
    print(f"1,002,528 triangles: full build {build_ms:.0f} ms, host refit + upload {host_ms:.1f} ms, device refit {dev_ms:.1f} ms "
          f"(includes the numpy->pinned copy and the 6 MB vertex upload)")
    assert dev_ms < host_ms
    tr.close()


def test_stage_scene_double_buffered_upload(pt, scenes, monkeypatch):
    """cudapt_stage_scene copies the scene into the shadow buffers on the copy stream while a render is in flight; the
    render in flight is not disturbed, the next render runs from the new copy with accumulation restarted, several
    rounds alternate the two buffers, and a staged scene after a device refit carries the refitted tree."""
    monkeypatch.delenv("CUDAPT_REFIT_HOST", raising=False)
    scene = scenes.heightfield(n_quads=64)
    v = scene[0]
    tr = tracer(pt, scene, 192, 108)
    tr.render(4)
    ref = tr.read_accum()[0]
    chk = tr.accel_checksum()
    for k in range(3):
        tr.reset()
        tr.render(4)           # in flight ...
        tr.stage_scene()       # ... while the next copy of the scene is uploaded
        acc, fc = tr.read_accum()
        assert fc == 4 and np.array_equal(bits(acc), bits(ref))
        tr.render(4)           # swaps the buffers and restarts accumulation
        acc, fc = tr.read_accum()
        assert fc == 4 and np.array_equal(bits(acc), bits(ref))
        assert tr.accel_checksum() == chk
    tr.update_vertices(wobble(v, 1))
    moved_chk = tr.accel_checksum()
    tr.render(4)
    moved = tr.read_accum()[0]
    tr.stage_scene()           # staging is stale after the device refit: refreshed first
    tr.render(4)
    assert tr.accel_checksum() == moved_chk
    assert np.array_equal(bits(tr.read_accum()[0]), bits(moved)) and not np.array_equal(bits(moved), bits(ref))
    tr.stage_scene()
    tr.update_vertices(v)      # a scene call commits the staged copy before it refits
    tr.render(4)
    assert np.array_equal(bits(tr.read_accum()[0]), bits(ref))
    tr.close()


def test_stage_before_first_stage_renders_are_tracked(pt, scenes, monkeypatch):
    """The write-after-read hazard of the double-buffered upload: renders issued BEFORE the first cudapt_stage_scene
    must be tracked too (the copy stream and the last-use events exist from the build on), so that the second stage
    cannot overwrite buffer 0 while the first render still reads it.  Nothing here synchronises with the host between
    the long render, the two stage calls and the renders in between; then a host refit rewrites the pinned staging
    copy and is staged again.  Every image equals a serial run of the same scene."""
    scene = scenes.heightfield(n_quads=192)
    v = scene[0]
    W, H = 480, 270

    def serial(verts, n):
        with pt.PathTracer(pt.default_config(rng_mode=1, max_bounces=5, jitter=1, seed=21, samples_per_pass=32)) as t2:
            t2.set_resolution(W, H)
            t2.upload_mesh(verts, *scene[1:])
            t2.render(n)
            return t2.read_accum()[0]

    want_a, want_b = serial(v, 32), serial(wobble(v, 2), 32)
    tr = pt.PathTracer(pt.default_config(rng_mode=1, max_bounces=5, jitter=1, seed=21, samples_per_pass=32))
    tr.set_resolution(W, H)
    tr.upload_mesh(*scene)
    tr.render(256)             # long render from buffer 0, issued before any stage: 8 passes in flight
    tr.stage_scene()           # copy #1 -> buffer 1
    tr.render(32)              # swap: buffer 1, accumulation restarted
    tr.stage_scene()           # copy #2 -> buffer 0: has to wait for the 256-frame render
    tr.render(32)              # swap: buffer 0
    acc, fc = tr.read_accum()
    assert fc == 32 and np.array_equal(bits(acc), bits(want_a))
    monkeypatch.setenv("CUDAPT_REFIT_HOST", "1")
    tr.update_vertices(wobble(v, 2))   # host refit: rewrites the pinned staging copy, uploads in line
    tr.render(32)
    tr.stage_scene()                   # the refitted tree again, double-buffered, while the render is in flight
    assert np.array_equal(bits(tr.read_accum()[0]), bits(want_b))
    tr.render(32)
    assert np.array_equal(bits(tr.read_accum()[0]), bits(want_b))
    tr.close()

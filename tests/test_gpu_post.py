"""GPU parity of the resolve/post-process stage and the image writer (8f-4) through the C ABI.

Oracle: oracle.post_resolve_rgba8 (numpy float32 restatement of CPT.cpp:87-117 + the exposure hook at :101-104
and the operators declared in utils/Math.h:29-34).  Bars: defaults, exposure, Reinhard and ACES are bit-exact
(only IEEE + - * /); sRGB (powf) and auto-exposure (logf/expf) within 1 LSB of the 8-bit channel, and identical
on >= 99 % of the channels.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def channels(words):
    w = np.asarray(words, np.uint32)
    return np.stack([(w >> 24) & 0xFF, (w >> 16) & 0xFF, (w >> 8) & 0xFF, w & 0xFF], -1).astype(np.int32)


@pytest.fixture(scope="module")
def rendered(pt, scenes):
    W, H, N = 192, 108, 5
    tr = pt.PathTracer()
    tr.set_resolution(W, H)
    tr.upload_spheres(scenes.app_spheres())
    tr.render(N)
    acc, fc = tr.read_accum()
    assert fc == N
    yield tr, acc, N
    tr.close()


def test_defaults_equal_reference_resolve(pt, oracle, rendered):
    tr, acc, N = rendered
    tr.set_postprocess(pt.default_postprocess())
    got = tr.resolve_rgba8()
    assert np.array_equal(got, oracle.twin_resolve_rgba8(acc, N))  # the reference's own resolve arithmetic
    assert np.array_equal(got, oracle.post_resolve_rgba8(acc, N)[0])
    assert tr.exposure() == 1.0


@pytest.mark.parametrize("exposure,tonemap", [(0.5, 0), (2.5, 0), (1.0, 1), (3.0, 1), (1.0, 2), (0.7, 2)])
def test_exposure_and_tonemap_bit_exact(pt, oracle, rendered, exposure, tonemap):
    tr, acc, N = rendered
    tr.set_postprocess(pt.default_postprocess(exposure=exposure, tonemap=tonemap))
    got = tr.resolve_rgba8()
    want, _ = oracle.post_resolve_rgba8(acc, N, exposure=exposure, tonemap=tonemap)
    assert np.array_equal(got, want)
    # post-processing is a display setting: the accumulation is untouched
    acc2, fc2 = tr.read_accum()
    assert fc2 == N and np.array_equal(acc2.view(np.uint32), acc.view(np.uint32))


@pytest.mark.parametrize("kw", [dict(srgb=1), dict(srgb=1, tonemap=2, exposure=1.3), dict(auto_exposure=1),
                                dict(auto_exposure=1, tonemap=1, srgb=1, target_luminance=0.25)])
def test_srgb_and_auto_exposure_within_one_lsb(pt, oracle, rendered, kw):
    tr, acc, N = rendered
    tr.set_postprocess(pt.default_postprocess(**kw))
    got = channels(tr.resolve_rgba8())
    want_w, want_e = oracle.post_resolve_rgba8(acc, N, **kw)
    want = channels(want_w)
    d = np.abs(got - want)
    print(f"{kw}: max diff {d.max()} LSB, identical channels {np.mean(d == 0):.5f}, exposure {tr.exposure():.6f} vs {want_e:.6f}")
    assert d.max() <= 1  # tolerance: 1 LSB of the 8-bit channel (powf / logf / expf are library functions)
    assert np.mean(d == 0) >= 0.99
    assert np.all(got[..., 3] == 255)
    if kw.get("auto_exposure"):
        assert abs(tr.exposure() - want_e) <= 1e-5 * want_e


def test_image_writer(pt, oracle, rendered, tmp_path):
    tr, acc, N = rendered
    tr.set_postprocess(pt.default_postprocess(tonemap=2))
    H, W = acc.shape[:2]
    ppm, pfm = str(tmp_path / "out.ppm"), str(tmp_path / "out.pfm")
    tr.write_image(ppm)
    tr.write_image(pfm)
    raw = open(ppm, "rb").read()
    head = f"P6\n{W} {H}\n255\n".encode()
    assert raw.startswith(head) and len(raw) == len(head) + 3 * W * H
    rgb = np.frombuffer(raw[len(head):], np.uint8).reshape(H, W, 3)
    want = channels(oracle.post_resolve_rgba8(acc, N, tonemap=2)[0])[..., :3]
    assert np.array_equal(rgb.astype(np.int32), want)
    raw = open(pfm, "rb").read()
    head = f"PF\n{W} {H}\n-1.0\n".encode()
    assert raw.startswith(head) and len(raw) == len(head) + 12 * W * H
    img = np.frombuffer(raw[len(head):], "<f4").reshape(H, W, 3)[::-1]  # bottom row first
    assert np.array_equal(img.view(np.uint32), (acc[..., :3] / np.float32(N)).view(np.uint32))
    # PNG: decoded with nothing but zlib + struct (stored deflate blocks, CRC-32 per chunk, filter 0 scanlines)
    import struct, zlib
    png = str(tmp_path / "out.png")
    tr.write_image(png)
    raw = open(png, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        data = raw[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])
        assert crc == (zlib.crc32(typ + data) & 0xFFFFFFFF)
        chunks.append((typ, data))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (W, H, 8, 6, 0, 0, 0)
    lines = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(H, 1 + 4 * W)
    assert not lines[:, 0].any()
    rgba = lines[:, 1:].reshape(H, W, 4)
    assert np.array_equal(rgba[..., :3].astype(np.int32), want) and np.all(rgba[..., 3] == 255)
    # EXR: parsed with struct alone (OpenEXR 2 single-part scan-line layout), and by OpenCV when its codec is there
    exr = str(tmp_path / "out.exr")
    tr.write_image(exr)
    raw = open(exr, "rb").read()
    assert raw[:4] == bytes([0x76, 0x2F, 0x31, 0x01]) and struct.unpack("<i", raw[4:8]) == (2,)
    pos, attrs = 8, {}
    while raw[pos] != 0:
        e = raw.index(b"\0", pos)
        name = raw[pos:e].decode()
        e2 = raw.index(b"\0", e + 1)
        typ = raw[e + 1:e2].decode()
        (size,) = struct.unpack("<i", raw[e2 + 1:e2 + 5])
        attrs[name] = (typ, raw[e2 + 5:e2 + 5 + size])
        pos = e2 + 5 + size
    pos += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    assert struct.unpack("<4i", attrs["dataWindow"][1]) == (0, 0, W - 1, H - 1) and attrs["displayWindow"] == attrs["dataWindow"]
    ch, names = attrs["channels"][1], []
    q = 0
    while ch[q] != 0:
        e = ch.index(b"\0", q)
        names.append(ch[q:e].decode())
        assert struct.unpack("<i4xii", ch[e + 1:e + 17]) == (2, 1, 1)   # FLOAT, sampling 1 x 1
        q = e + 17
    assert names == ["B", "G", "R"]
    offsets = struct.unpack(f"<{H}Q", raw[pos:pos + 8 * H])
    assert offsets[0] == pos + 8 * H and len(raw) == offsets[-1] + 8 + 12 * W
    img = np.empty((H, W, 3), np.float32)
    for y, off in enumerate(offsets):
        assert struct.unpack("<ii", raw[off:off + 8]) == (y, 12 * W)
        planes = np.frombuffer(raw[off + 8:off + 8 + 12 * W], "<f4").reshape(3, W)
        img[y] = planes[::-1].T                                        # B G R planes -> RGB
    assert np.array_equal(img.view(np.uint32), (acc[..., :3] / np.float32(N)).view(np.uint32))
    try:
        os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
        import cv2
        bgr = cv2.imread(exr, cv2.IMREAD_UNCHANGED)
    except Exception:
        bgr = None
    if bgr is not None:
        assert bgr.dtype == np.float32 and np.array_equal(bgr[..., ::-1].view(np.uint32), img.view(np.uint32))
    with pytest.raises(pt.CudaptError):
        tr.write_image(str(tmp_path / "out.tiff"))
    tr.set_postprocess(pt.default_postprocess())


def test_invalid_postprocess_rejected(pt, rendered):
    tr, _, _ = rendered
    with pytest.raises(pt.CudaptError):
        tr.set_postprocess(pt.default_postprocess(tonemap=7))
    with pytest.raises(pt.CudaptError):
        tr.set_postprocess(pt.default_postprocess(exposure=-1.0))


def test_pinned_result_buffer(pt, scenes):
    """cudapt_host_register: a page-locked caller buffer receives the read-back directly (no staging copy); the
    bytes equal those of an ordinary buffer, and an unregistered pointer cannot be unregistered."""
    with pt.PathTracer() as tr:
        tr.set_resolution(96, 54)
        tr.upload_spheres(scenes.app_spheres())
        tr.render(3)
        plain = tr.resolve_rgba8()
        pinned = np.full((54, 96), 0xDEADBEEF, np.uint32)
        tr.host_register(pinned)
        try:
            out = tr.resolve_rgba8(pinned)
            assert out is pinned and np.array_equal(pinned, plain)
            pinned[...] = 0
            tr.download_rgba8(pinned)
            assert np.array_equal(pinned, plain)
        finally:
            tr.host_unregister(pinned)
        with pytest.raises(pt.CudaptError):
            tr.host_unregister(plain)
        assert np.array_equal(tr.resolve_rgba8(), plain)   # the context still works after the failed call


def test_async_readback_double_buffered(pt, scenes):
    """cudapt_resolve_rgba8_async / cudapt_wait_download: the image lands in a page-locked buffer while later renders
    are already queued; bytes equal the synchronous read-back of the same frame count; ordinary memory is refused."""
    with pt.PathTracer() as tr:
        tr.set_resolution(160, 90)
        tr.upload_spheres(scenes.app_spheres())
        bufs = [np.zeros((90, 160), np.uint32) for _ in range(2)]
        for b in bufs:
            tr.host_register(b)
        try:
            want = []
            for k in range(4):      # reference: synchronous read-back after 1, 2, 3, 4 frames
                tr.render(1)
                want.append(tr.resolve_rgba8().copy())
            tr.reset()
            got = []
            for k in range(4):      # the same, double-buffered: image k is collected after render k+1 was queued
                tr.render(1)
                tr.wait_download()
                if k:
                    got.append(bufs[(k - 1) & 1].copy())
                tr.resolve_rgba8_async(bufs[k & 1])
            tr.render(3)            # queued behind the pending copy; must not disturb it
            tr.wait_download()
            got.append(bufs[3 & 1].copy())
            for k in range(4):
                assert np.array_equal(got[k], want[k]), k
            assert tr.stats().frame_count == 7
            tr.wait_download()      # nothing pending: returns at once
            with pytest.raises(pt.CudaptError):
                tr.resolve_rgba8_async(np.zeros((90, 160), np.uint32))
            # a synchronous read-back right behind an asynchronous one waits for it on the device
            tr.resolve_rgba8_async(bufs[0])
            sync = tr.resolve_rgba8()
            tr.wait_download()
            assert np.array_equal(sync, bufs[0])
            # resizing with a copy in flight waits for it
            tr.resolve_rgba8_async(bufs[1])
            tr.set_resolution(64, 36)
            assert np.array_equal(bufs[1], sync)
        finally:
            tr.wait_download()
            for b in bufs:
                tr.host_unregister(b)


def test_renders_queue_without_blocking_and_stats_add_up(pt, scenes):
    """cudapt_render no longer waits for the previous render's timing events: many renders can be queued back to back
    (more than the ring of event pairs), and render_ms still covers every one of them."""
    v, t, m, mats = scenes.heightfield(n_quads=48)
    with pt.PathTracer(pt.default_config(rng_mode=1, samples_per_pass=32, max_bounces=3, jitter=1, seed=1)) as tr:
        tr.set_resolution(128, 72)
        tr.upload_mesh(v, t, m, mats)
        tr.render(32)
        a1, _ = tr.read_accum()
        tr.reset()
        for _ in range(9):
            tr.render(32)
        tr.synchronize()
        st = tr.stats()
        # every one of the nine renders is in render_ms (each takes about as long as the last)
        assert st.frame_count == 9 * 32 and st.last_render_ms > 0.0 and st.last_trace_ms > 0.0
        assert st.render_ms > 9 * 0.6 * st.last_render_ms
        tr.reset()
        tr.render(32)
        a2, _ = tr.read_accum()
        assert np.array_equal(a1.view(np.uint32), a2.view(np.uint32))

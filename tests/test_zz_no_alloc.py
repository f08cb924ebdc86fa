"""cudapt_auto_exposure_static -- the allocation-free entry point (include/cudapt.h, csrc/static_scratch.cu; no counterpart
in the reference): its scratch is a file-scope `static __device__` array,
it calls no cudaMalloc / cudaFree.  Checked through the C ABI against the EXISTING entry point for the same quantity --
cudapt_resolve_rgba8 with cudapt_postprocess.auto_exposure, read back with cudapt_get_exposure -- on calls in a row that
share the static array: same image twice, then a SMALLER image (stale partials of the larger call stay in the array and
must not be read), then the first size again with other contents.  The two paths run the same operations in the same
order, so the bar is bit equality (tighter than the 1-LSB tolerance tests/test_gpu_post.py allows against numpy).

This file is named to be collected LAST; it passes alone in a fresh process and after
the other tests (it creates its own contexts and assumes nothing about the array's contents).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TARGET = 0.18


def f32_bits(x):
    return int(np.float32(x).view(np.uint32))


def render(pt, scenes, W, H, frames):
    tr = pt.PathTracer()
    tr.set_resolution(W, H)
    tr.upload_spheres(scenes.app_spheres())
    tr.render(frames)
    return tr


def existing_path(pt, tr):
    """The exposure the existing entry point computes for the context's accumulation buffer (scratch from cudaMalloc)."""
    tr.set_postprocess(pt.default_postprocess(auto_exposure=1, target_luminance=TARGET))
    tr.resolve_rgba8()
    return tr.exposure()


def static_path(pt, torch, tr, out):
    st = tr.stats()
    out.fill_(-1.0)
    torch.cuda.synchronize()
    pt.auto_exposure_static(0, tr.accum_device_ptr, tr.width * tr.height, st.frame_count, TARGET, out.data_ptr(), tr.stream_handle)
    tr.synchronize()  # the two kernels were enqueued on the context's stream
    return float(out.cpu()[0])


def test_static_scratch_entry_point_matches_existing_entry_point(pt, scenes):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    out = torch.empty(1, dtype=torch.float32, device="cuda:0")  # the CALLER's output buffer
    big = render(pt, scenes, 320, 180, 8)       # 57 600 pixels -> 57 partials
    small = render(pt, scenes, 96, 54, 3)       # 5 184 pixels  -> 6 partials
    want_big, want_small = existing_path(pt, big), existing_path(pt, small)
    assert 0.0 < want_big < 10.0 and want_big != want_small
    # two calls in a row on the same input
    got1 = static_path(pt, torch, big, out)
    got2 = static_path(pt, torch, big, out)
    assert f32_bits(got1) == f32_bits(want_big) and f32_bits(got2) == f32_bits(want_big), (got1, got2, want_big)
    # a smaller image next: partials 6..56 of the previous call are still in the static array and must be ignored
    got3 = static_path(pt, torch, small, out)
    assert f32_bits(got3) == f32_bits(want_small), (got3, want_small)
    # the first size again, other contents (more frames accumulated)
    big.set_postprocess(pt.default_postprocess())
    big.render(5)
    want_big2 = existing_path(pt, big)
    got4 = static_path(pt, torch, big, out)
    assert f32_bits(got4) == f32_bits(want_big2) and f32_bits(want_big2) != f32_bits(want_big), (got4, want_big2)
    # no device memory was taken by further calls (the first call above resolved the symbol / loaded the module)
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(20):
        static_path(pt, torch, small, out)
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 >= free0, (free0, free1)
    big.close()
    small.close()


def test_static_scratch_limits(pt):
    """Sizes beyond the stated maximum are rejected with a clear error before anything touches the GPU; the maximum itself
    uses the whole array (closed form: a constant image of luminance ~1 gives exposure = target)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    out = torch.zeros(1, dtype=torch.float32, device="cuda:0")
    n = pt.STATIC_MAX_PIXELS
    with pytest.raises(pt.CudaptError) as e:
        pt.auto_exposure_static(0, out.data_ptr(), n + 1, 1, TARGET, out.data_ptr(), 0)
    assert e.value.status == 1 and "CUDAPT_STATIC_MAX_PIXELS" in str(e.value) and str(n) in str(e.value)
    with pytest.raises(pt.CudaptError):
        pt.auto_exposure_static(0, 0, 16, 1, TARGET, out.data_ptr(), 0)          # NULL input
    with pytest.raises(pt.CudaptError) as e:
        pt.auto_exposure_static(0, out.data_ptr(), 16, 0, TARGET, out.data_ptr(), 0)  # no frames
    assert e.value.status == 4
    accum = torch.full((n, 4), 2.0, dtype=torch.float32, device="cuda:0")  # 256 MB, the caller's; frame_count 2 -> rgb = 1
    torch.cuda.synchronize()
    pt.auto_exposure_static(0, accum.data_ptr(), n, 2, TARGET, out.data_ptr(), 0)
    torch.cuda.synchronize()
    got = float(out.cpu()[0])
    assert abs(got / TARGET - 1.0) < 1e-5, got
    # and a small call right after the largest one
    pt.auto_exposure_static(0, accum.data_ptr(), 1000, 4, TARGET, out.data_ptr(), 0)   # rgb = 0.5 -> exposure = 2 * target
    torch.cuda.synchronize()
    assert abs(float(out.cpu()[0]) / (2.0 * TARGET) - 1.0) < 1e-5

"""Fused multi-GPU read-back tail (cudapt_reduce_resolve_peers) on ONE device: three contexts stand in
for three ranks (sample ranges g+1, g+1+3, ...), every "rank" runs the fused kernel for its pixel slice with the
other contexts' accumulation buffers as peers.  Oracle: numpy float32 sums in rank order + oracle.post_resolve_rgba8.
Bar: bit-exact (float sums and RGBA8 words).  The cross-process CUDA-IPC path is exercised by bench.py --gpus N.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size,post", [((160, 90), {}), ((67, 41), dict(tonemap=2, exposure=0.8))])
def test_fused_tail_equals_rank_ordered_host_merge(pt, oracle, scenes, size, post):
    W, H = size
    G, N = 3, 4
    sph = scenes.app_spheres()
    ranks = []
    for g in range(G):
        tr = pt.PathTracer()
        tr.set_resolution(W, H)
        tr.upload_spheres(sph)
        tr.set_frame_range(g + 1, G)
        tr.set_postprocess(pt.default_postprocess(**post))
        tr.render(N)
        ranks.append(tr)
    parts = [tr.read_accum()[0] for tr in ranks]  # also synchronises every context's stream
    want_sum = parts[0].copy()
    for g in range(1, G):
        want_sum = (want_sum + parts[g]).astype(np.float32)  # ((a0 + a1) + a2): rank order
    ptrs = [tr.accum_device_ptr for tr in ranks]
    root = ranks[0]
    for g, tr in enumerate(ranks):  # slices are disjoint; in-place accumulation into the root's buffer
        tr.reduce_resolve_peers(ptrs, g, G * N, root.rgba_device_ptr, root.accum_device_ptr)
        tr.synchronize()
    root.set_frame_count(G * N)
    got_sum, fc = root.read_accum()
    got_rgba = root.download_rgba8()
    assert fc == G * N
    assert np.array_equal(got_sum.view(np.uint32), want_sum.view(np.uint32))
    want_rgba, _ = oracle.post_resolve_rgba8(want_sum, G * N, **post)
    assert np.array_equal(got_rgba, want_rgba)
    # and it is the image a single context renders for the same 12 global frames, up to float re-association
    one = pt.PathTracer()
    one.set_resolution(W, H)
    one.upload_spheres(sph)
    one.render(G * N)
    single, _ = one.read_accum()
    assert np.allclose(single, got_sum, rtol=2e-6, atol=1e-6)
    one.close()
    # argument checks
    with pytest.raises(pt.CudaptError):
        root.reduce_resolve_peers(ptrs, G, G * N, root.rgba_device_ptr)
    root.set_postprocess(pt.default_postprocess(auto_exposure=1))
    with pytest.raises(pt.CudaptError):
        root.reduce_resolve_peers(ptrs, 0, G * N, root.rgba_device_ptr)
    for tr in ranks:
        tr.close()

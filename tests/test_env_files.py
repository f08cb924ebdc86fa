"""Environment-map files (csrc/env_load.cpp, cudapt_convert_environment / cudapt_load_environment): .pfm and Radiance
.hdr latitude-longitude panoramas, read without an image library and resampled to the octahedral map the kernels use.
Host-only part here (no GPU): the file readers against files written by this test, the resampling against a numpy
restatement, the error paths.  The GPU half (render with a loaded map == the twin with the same texels) is in
tests/test_gpu_env.py."""
import importlib
import os
import struct

import numpy as np
import pytest

pt = importlib.import_module("software-path-tracer_b200")
needs_lib = pytest.mark.skipif(not os.path.exists(pt.LIB_PATH), reason="libcudapt.so not built")


def panorama(W=64, H=32, seed=0):
    """Smooth HDR lat-long test image: gradient + a bright 'sun', values 0.01 .. ~60."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    img = np.stack([0.2 + 0.8 * x / W, 0.1 + 0.9 * (1 - y / H), 0.3 + 0.2 * np.cos(6.28 * x / W)], axis=-1)
    img += 60.0 * np.exp(-(((x - 0.7 * W) / 2.0) ** 2 + ((y - 0.25 * H) / 2.0) ** 2))[..., None]
    img += 0.01 * rng.random((H, W, 3))
    return img.astype(np.float32)


def write_pfm(path, img, little=True, grey=False):
    H, W = img.shape[:2]
    data = img[::-1, :, 0] if grey else img[::-1]                      # bottom row first
    with open(path, "wb") as f:
        f.write(f"{'Pf' if grey else 'PF'}\n{W} {H}\n{'-1.0' if little else '1.0'}\n".encode())
        f.write(np.ascontiguousarray(data, "<f4" if little else ">f4").tobytes())


def to_rgbe(img):
    m = img.max(axis=-1)
    e = np.where(m > 1e-32, np.floor(np.log2(np.maximum(m, 1e-38))) + 1, 0)   # m < 2^e
    scale = np.where(m > 1e-32, 256.0 / np.exp2(e), 0.0)
    rgb = np.clip(np.floor(img * scale[..., None]), 0, 255).astype(np.uint8)
    return np.concatenate([rgb, np.where(m > 1e-32, e + 128, 0).astype(np.uint8)[..., None]], axis=-1)


def from_rgbe(rgbe):
    e = rgbe[..., 3].astype(np.int32)
    s = np.where(e > 0, np.exp2((e - 136).astype(np.float64)), 0.0)
    return (rgbe[..., :3].astype(np.float64) * s[..., None]).astype(np.float32)


def write_hdr(path, rgbe, rle):
    H, W = rgbe.shape[:2]
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\n# written by tests/test_env_files.py\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n")
        f.write(f"-Y {H} +X {W}\n".encode())
        for y in range(H):
            if not rle:
                f.write(rgbe[y].tobytes())
                continue
            f.write(bytes([2, 2, W >> 8, W & 255]))
            for c in range(4):
                row, x = rgbe[y, :, c], 0
                while x < W:
                    run = 1
                    while x + run < W and run < 127 and row[x + run] == row[x]:
                        run += 1
                    if run >= 4:
                        f.write(bytes([128 + run, int(row[x])]))
                        x += run
                    else:                                        # literal of up to 128 bytes, stopped before a long run
                        n = 1
                        while x + n < W and n < 128 and not (x + n + 3 < W and row[x + n] == row[x + n + 1] == row[x + n + 2] == row[x + n + 3]):
                            n += 1
                        f.write(bytes([n]) + row[x:x + n].tobytes())
                        x += n


def octahedral_from_latlong(img, size):
    """numpy restatement of latlong_to_octahedral (env_load.cpp)."""
    H, W = img.shape[:2]
    j, i = np.mgrid[0:size, 0:size].astype(np.float64)
    px, pz = (i + 0.5) / size * 2 - 1, (j + 0.5) / size * 2 - 1
    y = 1 - np.abs(px) - np.abs(pz)
    x = np.where(y < 0, np.where(px < 0, -1.0, 1.0) * (1 - np.abs(pz)), px)
    z = np.where(y < 0, np.where(pz < 0, -1.0, 1.0) * (1 - np.abs(px)), pz)
    ln = np.sqrt(x * x + y * y + z * z)
    dx, dy, dz = x / ln, y / ln, z / ln
    u = 0.5 + np.arctan2(dx, dz) / (2 * np.pi)
    v = np.arccos(np.clip(dy, -1, 1)) / np.pi
    fx, fy = u * W - 0.5, v * H - 0.5
    x0, y0 = np.floor(fx), np.floor(fy)
    ax, ay = (fx - x0)[..., None], (fy - y0)[..., None]
    x0, y0 = x0.astype(np.int64), y0.astype(np.int64)
    x1, y1 = (x0 + 1) % W, np.clip(y0 + 1, 0, H - 1)
    x0, y0 = x0 % W, np.clip(y0, 0, H - 1)
    im = img.astype(np.float64)
    return ((im[y0, x0] * (1 - ax) + im[y0, x1] * ax) * (1 - ay) + (im[y1, x0] * (1 - ax) + im[y1, x1] * ax) * ay).astype(np.float32)


@needs_lib
@pytest.mark.parametrize("little", [True, False])
def test_pfm_panorama_to_octahedral(tmp_path, little):
    img = panorama()
    path = str(tmp_path / "sky.pfm")
    write_pfm(path, img, little=little)
    got = pt.convert_environment(path, 48)
    want = octahedral_from_latlong(img, 48)
    assert got.shape == (48, 48, 3) and np.allclose(got, want, rtol=1e-6, atol=1e-7)
    # the centre texel of the map is the pole +Y = the top row of the panorama; the sun (u = 0.7: to the right and
    # behind ... of +Z) is the brightest texel and sits in the upper hemisphere (inner diamond)
    assert np.allclose(got[23:25, 23:25].mean(axis=(0, 1)), img[0].mean(axis=0), rtol=0.05)
    jj, ii = np.unravel_index(np.argmax(got[..., 0]), got.shape[:2])
    assert abs((ii + 0.5) / 48 * 2 - 1) + abs((jj + 0.5) / 48 * 2 - 1) < 1.0


@needs_lib
def test_grey_pfm(tmp_path):
    img = panorama()
    path = str(tmp_path / "grey.pfm")
    write_pfm(path, img, grey=True)
    got = pt.convert_environment(path, 16)
    want = octahedral_from_latlong(np.repeat(img[..., :1], 3, axis=-1), 16)
    assert np.allclose(got, want, rtol=1e-6, atol=1e-7)


@needs_lib
@pytest.mark.parametrize("rle", [False, True])
def test_radiance_hdr_panorama(tmp_path, rle):
    img = panorama()
    img[5:9, 10:40] = (0.5, 0.25, 0.125)      # constant stretches, so that the run-length coder emits runs
    img[20:, :] = 0.0                          # and black (exponent byte 0)
    rgbe = to_rgbe(img)
    path = str(tmp_path / ("rle.hdr" if rle else "flat.hdr"))
    write_hdr(path, rgbe, rle)
    got = pt.convert_environment(path, 32)
    want = octahedral_from_latlong(from_rgbe(rgbe), 32)
    assert np.allclose(got, want, rtol=1e-6, atol=1e-7)
    assert np.abs(from_rgbe(rgbe) - img).max() <= img.max() / 128      # the test's own encoder is sane


@needs_lib
def test_bad_files_are_rejected(tmp_path):
    bad = tmp_path / "x.pfm"
    bad.write_bytes(b"P6\n2 2\n255\n" + bytes(12))
    with pytest.raises(pt.CudaptError, match="not a PFM"):
        pt.convert_environment(str(bad), 8)
    trunc = tmp_path / "t.pfm"
    trunc.write_bytes(b"PF\n4 4\n-1.0\n" + bytes(20))
    with pytest.raises(pt.CudaptError, match="truncated"):
        pt.convert_environment(str(trunc), 8)
    hdr = tmp_path / "y.hdr"
    hdr.write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_xyze\n\n-Y 2 +X 2\n" + bytes(16))
    with pytest.raises(pt.CudaptError, match="rgbe"):
        pt.convert_environment(str(hdr), 8)
    with pytest.raises(pt.CudaptError, match="cannot open"):
        pt.convert_environment(str(tmp_path / "missing.hdr"), 8)
    with pytest.raises(pt.CudaptError, match="neither"):
        pt.convert_environment(str(tmp_path / "sky.png"), 8)
    with pytest.raises(pt.CudaptError, match="size"):
        pt.convert_environment(str(bad), 1)

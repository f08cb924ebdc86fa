"""TEST INFRASTRUCTURE -- ctypes bindings of the two CPU oracles.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this package; the
product (software-path-tracer_b200/) never does.

  Tier 1  oracle/_ref/libref_render.so : the reference's own libs/render translation units, compiled
          unmodified against an Embree stand-in (oracle/ref/), driven through render::PathTracer.
  Tier 2  oracle/libtwin.so            : CPU restatement ("twin") of CPUPathTracer.cpp + the mesh
          integrator spec (oracle/twin/).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libref_render.so")
TWIN_SO = os.path.join(_HERE, "libtwin.so")
REFERENCE_ROOT = "/root/reference"


def build(verbose: bool = False) -> None:
    """Compile libtwin.so and, where /root/reference is mounted, _ref/libref_render.so."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout, out.stderr)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


# --------------------------------------------------------------------------------------------------
# Tier 1
# --------------------------------------------------------------------------------------------------
_ref_lib = None


def ref_available() -> bool:
    return os.path.exists(REF_SO)


def _ref():
    global _ref_lib
    if _ref_lib is None:
        lib = ctypes.CDLL(REF_SO)
        lib.ref_create.restype = ctypes.c_void_p
        lib.ref_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        lib.ref_render.restype = ctypes.c_double
        lib.ref_render.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.ref_read.argtypes = [ctypes.c_void_p] * 4
        lib.ref_primary.argtypes = [ctypes.c_void_p] * 3
        lib.ref_ray_count.restype = ctypes.c_uint64
        lib.ref_ray_count.argtypes = [ctypes.c_void_p]
        lib.ref_destroy.argtypes = [ctypes.c_void_p]
        lib.ref_backend_name.restype = ctypes.c_char_p
        lib.ref_backend_name.argtypes = [ctypes.c_void_p]
        _ref_lib = lib
    return _ref_lib


class RefSession:
    """The unmodified reference CPUPathTracer behind render::PathTracer, driven as src/App.cpp does."""

    def __init__(self, spheres: np.ndarray, width: int, height: int):
        self.spheres = np.ascontiguousarray(spheres, dtype=np.float32).reshape(-1, 4)
        self.w, self.h = int(width), int(height)
        self._h = _ref().ref_create(_ptr(self.spheres), len(self.spheres), self.w, self.h)

    def render(self, frames: int, resolve_each_frame: bool = True) -> float:
        """Returns wall seconds spent in render() (+ get_render_result() when resolve_each_frame)."""
        return _ref().ref_render(self._h, int(frames), int(resolve_each_frame))

    def read(self):
        acc = np.zeros((self.h, self.w, 4), np.float32)
        rgba = np.zeros((self.h, self.w), np.uint32)
        fc = ctypes.c_uint32()
        _ref().ref_read(self._h, _ptr(acc), _ptr(rgba), ctypes.byref(fc))
        return acc, rgba, fc.value

    def primary(self):
        idx = np.zeros((self.h, self.w), np.int32)
        t = np.zeros((self.h, self.w), np.float32)
        _ref().ref_primary(self._h, _ptr(idx), _ptr(t))
        return idx, t

    @property
    def rays(self) -> int:
        return _ref().ref_ray_count(self._h)

    @property
    def backend_name(self) -> str:
        return _ref().ref_backend_name(self._h).decode()

    def close(self):
        if self._h:
            _ref().ref_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# --------------------------------------------------------------------------------------------------
# Tier 2
# --------------------------------------------------------------------------------------------------
class TwinParams(ctypes.Structure):
    _fields_ = [
        ("width", ctypes.c_int32), ("height", ctypes.c_int32), ("rng_mode", ctypes.c_int32),
        ("max_bounces", ctypes.c_int32), ("rr_start", ctypes.c_int32), ("tnear", ctypes.c_float),
        ("eps", ctypes.c_float), ("albedo", ctypes.c_float), ("sky_scale", ctypes.c_float),
        ("seed", ctypes.c_uint32), ("jitter", ctypes.c_int32), ("nee", ctypes.c_int32),
        ("cam_origin", ctypes.c_float * 3), ("cam_right", ctypes.c_float * 3), ("cam_up", ctypes.c_float * 3),
        ("cam_forward", ctypes.c_float * 3), ("cam_tan_half_fov", ctypes.c_float), ("threads", ctypes.c_int32),
        ("part_rank", ctypes.c_int32), ("part_world", ctypes.c_int32), ("tile_w", ctypes.c_int32),
        ("tile_h", ctypes.c_int32), ("env_rgb", ctypes.c_void_p), ("env_size", ctypes.c_int32),
        ("cam_lens_radius", ctypes.c_float), ("cam_focus_distance", ctypes.c_float),
    ]


class TwinCounters(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in ("paths", "rays", "shade_invocations", "nodes_visited", "tris_tested")]


MATERIAL_DTYPE = np.dtype([("albedo", np.float32, 3), ("type", np.uint32), ("emission", np.float32, 3),
                           ("param", np.float32)])

_twin_lib = None


def twin_available() -> bool:
    return os.path.exists(TWIN_SO)


def _twin():
    global _twin_lib
    if _twin_lib is None:
        lib = ctypes.CDLL(TWIN_SO)
        lib.twin_mesh_create.restype = ctypes.c_void_p
        lib.twin_mesh_create.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32]
        lib.twin_mesh_destroy.argtypes = [ctypes.c_void_p]
        lib.twin_render_mesh.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32,
                                         ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p]
        lib.twin_render_spheres.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_uint32,
                                            ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p]
        lib.twin_mesh_closest.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_float,
                                          ctypes.c_void_p, ctypes.c_void_p]
        lib.twin_mesh_closest_bruteforce.argtypes = lib.twin_mesh_closest.argtypes
        lib.twin_resolve_rgba8.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
        lib.twin_pcg_seed.restype = ctypes.c_uint32
        lib.twin_pcg_seed.argtypes = [ctypes.c_uint32] * 4
        lib.twin_pcg_next.restype = ctypes.c_float
        lib.twin_pcg_next.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.twin_philox4x32_10.argtypes = [ctypes.c_void_p] * 3
        lib.twin_sincos2pi.argtypes = [ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p]
        _twin_lib = lib
    return _twin_lib


def twin_params(width: int, height: int, **kw) -> TwinParams:
    p = TwinParams()
    _twin().twin_default_params(ctypes.byref(p))
    p.width, p.height = int(width), int(height)
    for k, v in kw.items():
        if k in ("cam_origin", "cam_right", "cam_up", "cam_forward"):
            setattr(p, k, (ctypes.c_float * 3)(*[float(x) for x in v]))
        elif k == "env":  # octahedral environment map, (n, n, 3) float32; kept alive on the params object
            env = np.ascontiguousarray(v, np.float32)
            assert env.ndim == 3 and env.shape[0] == env.shape[1] and env.shape[2] == 3
            p._env_keepalive = env
            p.env_rgb = env.ctypes.data
            p.env_size = env.shape[0]
        else:
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
    return p


def twin_render_spheres(p: TwinParams, spheres, n_frames, first_frame=1, frame_stride=1, accum=None,
                        want_primary=False):
    spheres = np.ascontiguousarray(spheres, dtype=np.float32).reshape(-1, 4)
    if accum is None:
        accum = np.zeros((p.height, p.width, 4), np.float32)
    idx = np.full((p.height, p.width), -1, np.int32) if want_primary else None
    t = np.full((p.height, p.width), np.inf, np.float32) if want_primary else None
    c = TwinCounters()
    rc = _twin().twin_render_spheres(ctypes.byref(p), _ptr(spheres), len(spheres), first_frame, n_frames,
                                     frame_stride, _ptr(accum), _ptr(idx), _ptr(t), ctypes.byref(c))
    assert rc == 0
    return accum, idx, t, c


class TwinMesh:
    def __init__(self, vertices, indices, material_id=None, materials=None):
        self.v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 3)
        self.i = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        self.mid = None if material_id is None else np.ascontiguousarray(material_id, dtype=np.uint32)
        self.mats = None if materials is None else np.ascontiguousarray(materials, dtype=MATERIAL_DTYPE)
        self._h = _twin().twin_mesh_create(_ptr(self.v), len(self.v), _ptr(self.i), len(self.i), _ptr(self.mid),
                                           _ptr(self.mats), 0 if self.mats is None else len(self.mats))

    def render(self, p: TwinParams, n_frames, first_frame=1, frame_stride=1, accum=None, want_primary=False):
        if accum is None:
            accum = np.zeros((p.height, p.width, 4), np.float32)
        idx = np.full((p.height, p.width), -1, np.int32) if want_primary else None
        t = np.full((p.height, p.width), np.inf, np.float32) if want_primary else None
        c = TwinCounters()
        rc = _twin().twin_render_mesh(ctypes.byref(p), self._h, first_frame, n_frames, frame_stride, _ptr(accum),
                                      _ptr(idx), _ptr(t), ctypes.byref(c))
        assert rc == 0
        return accum, idx, t, c

    def closest(self, org, dir, tnear=1e-3, bruteforce=False):
        o = np.ascontiguousarray(org, np.float32)
        d = np.ascontiguousarray(dir, np.float32)
        prim = ctypes.c_int32()
        t = ctypes.c_float()
        fn = _twin().twin_mesh_closest_bruteforce if bruteforce else _twin().twin_mesh_closest
        fn(self._h, _ptr(o), _ptr(d), tnear, ctypes.byref(prim), ctypes.byref(t))
        return prim.value, t.value

    def close(self):
        if self._h:
            _twin().twin_mesh_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def twin_resolve_rgba8(accum: np.ndarray, frame_count: int) -> np.ndarray:
    accum = np.ascontiguousarray(accum, np.float32)
    n = accum.size // 4
    out = np.zeros(accum.shape[:-1], np.uint32)
    _twin().twin_resolve_rgba8(_ptr(accum), n, frame_count, _ptr(out))
    return out


def post_resolve_rgba8(accum: np.ndarray, frame_count: int, exposure: float = 1.0, tonemap: int = 0, srgb: int = 0,
                       auto_exposure: int = 0, target_luminance: float = 0.18):
    """Resolve with the post-process stage restated in numpy float32: every binary operation below is
    one IEEE-754 binary32 operation (numpy does not contract a*b+c), in the order of csrc/post.cu.  Reference
    anchors: CPT.cpp:87-117 (resolve), :101-104 (exposure), Types.h:87-89 (defaults), utils/Math.h:29-34 (declared
    operators).  Returns (rgba8 words, exposure used).  With the defaults it equals twin_resolve_rgba8."""
    f32 = np.float32
    a = np.ascontiguousarray(accum, f32).reshape(-1, 4)
    fc = f32(frame_count)
    c = a[:, :3] / fc
    e = f32(exposure)
    if auto_exposure:
        lum = (f32(0.2126) * c[:, 0] + f32(0.7152) * c[:, 1]) + f32(0.0722) * c[:, 2]
        mean_log = f32(np.log(np.maximum(lum, f32(1e-4))).astype(np.float64).sum() / lum.size)
        e = f32(target_luminance) / np.exp(mean_log, dtype=f32)
    v = c * e
    if tonemap == 1:
        v = v / (f32(1.0) + v)
    elif tonemap == 2:
        num = v * (f32(2.51) * v + f32(0.03))
        den = v * (f32(2.43) * v + f32(0.59)) + f32(0.14)
        v = num / den
    v = np.where(v < 0, f32(0), np.where(v > 1, f32(1), v)).astype(f32)
    if srgb:
        enc = f32(1.055) * np.power(v, f32(1.0) / f32(2.4), dtype=f32) - f32(0.055)
        v = np.where(v <= f32(0.0031308), f32(12.92) * v, enc).astype(f32)
        v = np.where(v < 0, f32(0), np.where(v > 1, f32(1), v)).astype(f32)
    al = a[:, 3] / fc
    al = np.where(al < 0, f32(0), np.where(al > 1, f32(1), al)).astype(f32)
    b = (v * f32(255.0)).astype(np.uint32) & 0xFF  # truncation, CPT.cpp:112
    ab = (al * f32(255.0)).astype(np.uint32) & 0xFF
    out = (b[:, 0] << 24) | (b[:, 1] << 16) | (b[:, 2] << 8) | ab  # Color.h:9
    return out.astype(np.uint32).reshape(np.asarray(accum).shape[:-1]), float(e)


def pcg_stream(x, y, width, frame, n):
    """(raw u32 words, floats) of the reference RNG for one pixel/frame."""
    st = ctypes.c_uint32(_twin().twin_pcg_seed(x, y, width, frame))
    raw = ctypes.c_uint32()
    words, floats = [], []
    for _ in range(n):
        f = _twin().twin_pcg_next(ctypes.byref(st), ctypes.byref(raw))
        words.append(raw.value)
        floats.append(f)
    return np.array(words, np.uint32), np.array(floats, np.float32)


def philox(ctr, key):
    c = np.ascontiguousarray(ctr, np.uint32)
    k = np.ascontiguousarray(key, np.uint32)
    out = np.zeros(4, np.uint32)
    _twin().twin_philox4x32_10(_ptr(c), _ptr(k), _ptr(out))
    return out


def sincos2pi(u):
    c = ctypes.c_float()
    s = ctypes.c_float()
    _twin().twin_sincos2pi(float(u), ctypes.byref(c), ctypes.byref(s))
    return c.value, s.value

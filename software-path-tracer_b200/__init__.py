"""Thin ctypes binding of libcudapt.so -- the C ABI declared in include/cudapt.h.

The product is the CUDA library; this module only exists so that tests/ and bench.py can drive the C ABI
from Python.  There is no CPU fallback: importing works everywhere (so the symbol table can be checked
without a GPU), but creating a context raises when the library or an sm_100 device is missing.

Because the directory name contains '-', import it with
    importlib.import_module("software-path-tracer_b200")
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcudapt.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "cudapt.h")

RNG_REFERENCE_PCG = 0
RNG_PHILOX = 1
MAT_LAMBERT, MAT_MIRROR, MAT_DIELECTRIC, MAT_GGX_COATED = 0, 1, 2, 3


class CudaptError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"cudapt status {status}: {message}")
        self.status = status


class Config(ctypes.Structure):
    _fields_ = [
        ("device", ctypes.c_int32), ("rng_mode", ctypes.c_int32), ("max_bounces", ctypes.c_int32),
        ("rr_start", ctypes.c_int32), ("tnear", ctypes.c_float), ("eps", ctypes.c_float),
        ("albedo", ctypes.c_float), ("sky_scale", ctypes.c_float), ("seed", ctypes.c_uint32),
        ("jitter", ctypes.c_int32), ("nee", ctypes.c_int32), ("samples_per_pass", ctypes.c_int32),
    ]


class Camera(ctypes.Structure):
    _fields_ = [("origin", ctypes.c_float * 3), ("right", ctypes.c_float * 3), ("up", ctypes.c_float * 3),
                ("forward", ctypes.c_float * 3), ("tan_half_fov", ctypes.c_float),
                ("lens_radius", ctypes.c_float), ("focus_distance", ctypes.c_float)]


class Stats(ctypes.Structure):
    _fields_ = [
        ("paths", ctypes.c_uint64), ("rays", ctypes.c_uint64), ("nodes_visited", ctypes.c_uint64),
        ("tris_tested", ctypes.c_uint64), ("shade_invocations", ctypes.c_uint64),
        ("kernel_launches", ctypes.c_uint64), ("render_ms", ctypes.c_double), ("last_render_ms", ctypes.c_double),
        ("accel_build_ms", ctypes.c_double), ("accel_bytes", ctypes.c_uint64), ("frame_count", ctypes.c_uint32),
        ("n_triangles", ctypes.c_uint32), ("n_nodes", ctypes.c_uint32), ("n_spheres", ctypes.c_uint32),
        ("last_trace_ms", ctypes.c_double),
        ("packet_nodes", ctypes.c_uint64), ("packet_tris", ctypes.c_uint64), ("packet_bails", ctypes.c_uint64),
        ("packet_rays", ctypes.c_uint64), ("shadow_rays", ctypes.c_uint64), ("last_packet_ms", ctypes.c_double),
        ("packets", ctypes.c_uint64),
    ]


class PostProcess(ctypes.Structure):
    _fields_ = [("exposure", ctypes.c_float), ("tonemap", ctypes.c_int32), ("srgb", ctypes.c_int32),
                ("auto_exposure", ctypes.c_int32), ("target_luminance", ctypes.c_float)]


TONEMAP_NONE, TONEMAP_REINHARD, TONEMAP_ACES = 0, 1, 2

MATERIAL_DTYPE = np.dtype([("albedo", np.float32, 3), ("type", np.uint32), ("emission", np.float32, 3),
                           ("param", np.float32)])

# every symbol include/cudapt.h declares (tests/test_abi.py checks the header and the .so against this)
EXPORTS = [
    "cudapt_default_config", "cudapt_default_camera", "cudapt_camera_look_at", "cudapt_create", "cudapt_destroy", "cudapt_last_error",
    "cudapt_abi_version", "cudapt_set_resolution", "cudapt_set_config", "cudapt_set_camera",
    "cudapt_upload_spheres", "cudapt_upload_mesh", "cudapt_upload_materials", "cudapt_build_accel",
    "cudapt_build_accel_device", "cudapt_reupload_scene", "cudapt_stage_scene", "cudapt_update_vertices", "cudapt_debug_accel_checksum",
    "cudapt_reset", "cudapt_render", "cudapt_synchronize", "cudapt_resolve_rgba8", "cudapt_host_register", "cudapt_host_unregister",
    "cudapt_resolve_rgba8_async", "cudapt_download_rgba8_async", "cudapt_wait_download", "cudapt_stream_wait_download", "cudapt_read_accum",
    "cudapt_write_accum", "cudapt_read_primary_hits", "cudapt_get_stats", "cudapt_set_counting",
    "cudapt_set_tile_partition", "cudapt_set_frame_range", "cudapt_accum_device_ptr", "cudapt_stream",
    "cudapt_set_frame_count",
    "cudapt_reduce_resolve_peers", "cudapt_rgba_device_ptr", "cudapt_download_rgba8", "cudapt_ipc_export",
    "cudapt_ipc_open", "cudapt_ipc_close",
    "cudapt_upload_environment", "cudapt_load_environment", "cudapt_convert_environment", "cudapt_default_postprocess", "cudapt_set_postprocess", "cudapt_get_exposure", "cudapt_write_image",
    "cudapt_auto_exposure_static",
]
STATIC_MAX_PIXELS = 16777216  # CUDAPT_STATIC_MAX_PIXELS

_lib = None


def load_library() -> ctypes.CDLL:
    """dlopen libcudapt.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CudaptError(-1, f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()); "
                              "there is no CPU fallback")
    lib = ctypes.CDLL(LIB_PATH)
    vp, u32, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int
    lib.cudapt_default_config.argtypes = [ctypes.POINTER(Config)]
    lib.cudapt_default_config.restype = None
    lib.cudapt_default_camera.argtypes = [ctypes.POINTER(Camera)]
    lib.cudapt_default_camera.restype = None
    f3p = ctypes.POINTER(ctypes.c_float)
    lib.cudapt_camera_look_at.argtypes = [ctypes.POINTER(Camera), f3p, f3p, f3p, ctypes.c_float]
    lib.cudapt_camera_look_at.restype = i32
    lib.cudapt_create.argtypes = [ctypes.POINTER(Config), ctypes.POINTER(vp)]
    lib.cudapt_destroy.argtypes = [vp]
    lib.cudapt_destroy.restype = None
    lib.cudapt_last_error.argtypes = [vp]
    lib.cudapt_last_error.restype = ctypes.c_char_p
    lib.cudapt_abi_version.restype = i32
    lib.cudapt_set_resolution.argtypes = [vp, u32, u32]
    lib.cudapt_set_config.argtypes = [vp, ctypes.POINTER(Config)]
    lib.cudapt_set_camera.argtypes = [vp, ctypes.POINTER(Camera)]
    lib.cudapt_upload_spheres.argtypes = [vp, vp, vp, u32]
    lib.cudapt_upload_mesh.argtypes = [vp, vp, u32, vp, u32, vp]
    lib.cudapt_upload_materials.argtypes = [vp, vp, u32]
    lib.cudapt_build_accel.argtypes = [vp]
    lib.cudapt_build_accel_device.argtypes = [vp]
    lib.cudapt_reupload_scene.argtypes = [vp]
    lib.cudapt_stage_scene.argtypes = [vp]
    lib.cudapt_update_vertices.argtypes = [vp, vp, u32]
    lib.cudapt_debug_accel_checksum.argtypes = [vp, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
    lib.cudapt_reset.argtypes = [vp]
    lib.cudapt_render.argtypes = [vp, u32]
    lib.cudapt_synchronize.argtypes = [vp]
    lib.cudapt_resolve_rgba8.argtypes = [vp, vp]
    lib.cudapt_host_register.argtypes = [vp, vp, ctypes.c_size_t]
    lib.cudapt_host_unregister.argtypes = [vp, vp]
    lib.cudapt_resolve_rgba8_async.argtypes = [vp, vp]
    lib.cudapt_download_rgba8_async.argtypes = [vp, vp]
    lib.cudapt_wait_download.argtypes = [vp]
    lib.cudapt_stream_wait_download.argtypes = [vp]
    lib.cudapt_read_accum.argtypes = [vp, vp, ctypes.POINTER(u32)]
    lib.cudapt_write_accum.argtypes = [vp, vp, u32]
    lib.cudapt_read_primary_hits.argtypes = [vp, vp, vp]
    lib.cudapt_get_stats.argtypes = [vp, ctypes.POINTER(Stats)]
    lib.cudapt_set_counting.argtypes = [vp, i32]
    lib.cudapt_set_tile_partition.argtypes = [vp, u32, u32, u32, u32]
    lib.cudapt_set_frame_range.argtypes = [vp, u32, u32]
    lib.cudapt_accum_device_ptr.argtypes = [vp]
    lib.cudapt_accum_device_ptr.restype = vp
    lib.cudapt_stream.argtypes = [vp]
    lib.cudapt_stream.restype = vp
    lib.cudapt_set_frame_count.argtypes = [vp, u32]
    lib.cudapt_upload_environment.argtypes = [vp, vp, u32]
    lib.cudapt_load_environment.argtypes = [vp, ctypes.c_char_p, u32, vp]
    lib.cudapt_convert_environment.argtypes = [ctypes.c_char_p, u32, vp]
    lib.cudapt_reduce_resolve_peers.argtypes = [vp, ctypes.POINTER(vp), u32, u32, u32, vp, vp]
    lib.cudapt_rgba_device_ptr.argtypes = [vp]
    lib.cudapt_rgba_device_ptr.restype = vp
    lib.cudapt_download_rgba8.argtypes = [vp, vp]
    lib.cudapt_ipc_export.argtypes = [vp, vp, vp]
    lib.cudapt_ipc_open.argtypes = [vp, vp, ctypes.POINTER(vp)]
    lib.cudapt_ipc_close.argtypes = [vp, vp]
    lib.cudapt_default_postprocess.argtypes = [ctypes.POINTER(PostProcess)]
    lib.cudapt_default_postprocess.restype = None
    lib.cudapt_set_postprocess.argtypes = [vp, ctypes.POINTER(PostProcess)]
    lib.cudapt_get_exposure.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    lib.cudapt_write_image.argtypes = [vp, ctypes.c_char_p]
    lib.cudapt_auto_exposure_static.argtypes = [i32, vp, u32, u32, ctypes.c_float, vp, vp]
    lib.cudapt_auto_exposure_static.restype = i32
    _lib = lib
    return lib


def default_config(**kw) -> Config:
    c = Config()
    load_library().cudapt_default_config(ctypes.byref(c))
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        setattr(c, k, v)
    return c


def default_camera(**kw) -> Camera:
    c = Camera()
    load_library().cudapt_default_camera(ctypes.byref(c))
    for k, v in kw.items():
        if k in ("tan_half_fov", "lens_radius", "focus_distance"):
            setattr(c, k, float(v))
        else:
            setattr(c, k, (ctypes.c_float * 3)(*[float(x) for x in v]))
    return c


def look_at(position, target, up=(0.0, 1.0, 0.0), vfov_degrees=90.0) -> Camera:
    """The reference's placeholder Camera (setPosition / setTarget / setUp / setPerspective) as a cudapt_camera."""
    c = Camera()
    v3 = lambda x: (ctypes.c_float * 3)(*[float(t) for t in x])
    rc = load_library().cudapt_camera_look_at(ctypes.byref(c), v3(position), v3(target), v3(up), float(vfov_degrees))
    if rc != 0:
        raise ValueError("look_at: degenerate camera (zero view vector, up parallel to it, or fov outside (0, 180))")
    return c


def default_postprocess(**kw) -> PostProcess:
    p = PostProcess()
    load_library().cudapt_default_postprocess(ctypes.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def convert_environment(path: str, size: int = 256) -> np.ndarray:
    """Host only (no GPU): lat-long .pfm / .hdr panorama -> octahedral texels (size, size, 3)."""
    lib = load_library()
    out = np.empty((size, size, 3), np.float32)
    rc = lib.cudapt_convert_environment(os.fsencode(path), size, out.ctypes.data)
    if rc != 0:
        raise CudaptError(rc, (lib.cudapt_last_error(None) or b"").decode())
    return out


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def auto_exposure_static(device: int, accum_device_ptr: int, n_pixels: int, frame_count: int, target_luminance: float,
                         exposure_device_ptr: int, stream: int = 0):
    """cudapt_auto_exposure_static: the allocation-free auto-exposure reduction (device pointers in and out, enqueued on
    `stream`).  Raises CudaptError with the library's text on a non-zero status."""
    lib = load_library()
    rc = lib.cudapt_auto_exposure_static(int(device), ctypes.c_void_p(int(accum_device_ptr)), int(n_pixels), int(frame_count),
                                         float(target_luminance), ctypes.c_void_p(int(exposure_device_ptr)),
                                         ctypes.c_void_p(int(stream)) if stream else None)
    if rc != 0:
        raise CudaptError(rc, (lib.cudapt_last_error(None) or b"").decode())


class PathTracer:
    """One cudapt context.  Methods map 1:1 onto the C ABI; see include/cudapt.h for the reference
    interface each call replaces."""

    def __init__(self, config: Config | None = None, **kw):
        self._lib = load_library()
        self.config = config if config is not None else default_config(**kw)
        h = ctypes.c_void_p()
        rc = self._lib.cudapt_create(ctypes.byref(self.config), ctypes.byref(h))
        if rc != 0:
            raise CudaptError(rc, self._lib.cudapt_last_error(None).decode())
        self._h = h
        self.width = self.height = 0

    def _chk(self, rc: int):
        if rc != 0:
            raise CudaptError(rc, self._lib.cudapt_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self._lib.cudapt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # settings
    def set_resolution(self, width: int, height: int):
        self._chk(self._lib.cudapt_set_resolution(self._h, width, height))
        self.width, self.height = int(width), int(height)

    def set_config(self, config: Config):
        self._chk(self._lib.cudapt_set_config(self._h, ctypes.byref(config)))
        self.config = config

    def set_camera(self, camera: Camera):
        self._chk(self._lib.cudapt_set_camera(self._h, ctypes.byref(camera)))

    # scene
    def upload_spheres(self, center_radius, node_id=None):
        s = np.ascontiguousarray(center_radius, np.float32).reshape(-1, 4)
        ids = None if node_id is None else np.ascontiguousarray(node_id, np.uint32)
        self._chk(self._lib.cudapt_upload_spheres(self._h, _p(s), _p(ids), len(s)))

    def upload_mesh(self, vertices, indices, material_id=None, materials=None, build=True, device_build=False):
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        i = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3)
        m = None if material_id is None else np.ascontiguousarray(material_id, np.uint32)
        self._chk(self._lib.cudapt_upload_mesh(self._h, _p(v), len(v), _p(i), len(i), _p(m)))
        if materials is not None:
            mats = np.ascontiguousarray(materials, MATERIAL_DTYPE)
            self._chk(self._lib.cudapt_upload_materials(self._h, _p(mats), len(mats)))
        if build:
            self.build_accel_device() if device_build else self.build_accel()

    def build_accel(self):
        self._chk(self._lib.cudapt_build_accel(self._h))

    def build_accel_device(self):
        """BVH build on the GPU (linear BVH): fast, lower tree quality, identical images."""
        self._chk(self._lib.cudapt_build_accel_device(self._h))

    def update_vertices(self, vertices):
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        self._chk(self._lib.cudapt_update_vertices(self._h, _p(v), len(v)))

    def accel_checksum(self):
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        self._chk(self._lib.cudapt_debug_accel_checksum(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def reupload_scene(self):
        self._chk(self._lib.cudapt_reupload_scene(self._h))

    def stage_scene(self):
        """Double-buffered re-upload: returns at once, takes effect (and restarts accumulation) at the next render."""
        self._chk(self._lib.cudapt_stage_scene(self._h))

    def upload_environment(self, rgb):
        """Octahedral environment map, (n, n, 3) float32; None restores the procedural sky."""
        if rgb is None:
            self._chk(self._lib.cudapt_upload_environment(self._h, None, 0))
            return
        e = np.ascontiguousarray(rgb, np.float32)
        assert e.ndim == 3 and e.shape[0] == e.shape[1] and e.shape[2] == 3
        self._chk(self._lib.cudapt_upload_environment(self._h, _p(e), e.shape[0]))

    # rendering
    def load_environment(self, path: str, size: int = 256) -> np.ndarray:
        """Lat-long .pfm / .hdr panorama -> octahedral map of `size`, uploaded; returns the texels (size, size, 3)."""
        out = np.empty((size, size, 3), np.float32)
        self._chk(self._lib.cudapt_load_environment(self._h, os.fsencode(path), size, _p(out)))
        return out

    def reset(self):
        self._chk(self._lib.cudapt_reset(self._h))

    def render(self, n_frames: int = 1):
        self._chk(self._lib.cudapt_render(self._h, n_frames))

    def synchronize(self):
        self._chk(self._lib.cudapt_synchronize(self._h))

    # read-back
    def resolve_rgba8(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width), np.uint32)
        self._chk(self._lib.cudapt_resolve_rgba8(self._h, _p(out)))
        return out

    def host_register(self, a: np.ndarray):
        """Page-lock a caller-owned array: read-backs into it skip the staging copy (cudapt_host_register)."""
        self._chk(self._lib.cudapt_host_register(self._h, _p(a), ctypes.c_size_t(a.nbytes)))

    def resolve_rgba8_async(self, out: np.ndarray):
        """Resolve + D2H into the page-locked `out` without waiting (cudapt_resolve_rgba8_async); wait_download() next."""
        self._chk(self._lib.cudapt_resolve_rgba8_async(self._h, _p(out)))

    def download_rgba8_async(self, out: np.ndarray):
        self._chk(self._lib.cudapt_download_rgba8_async(self._h, _p(out)))

    def wait_download(self):
        self._chk(self._lib.cudapt_wait_download(self._h))

    def stream_wait_download(self):
        self._chk(self._lib.cudapt_stream_wait_download(self._h))

    def host_unregister(self, a: np.ndarray):
        self._chk(self._lib.cudapt_host_unregister(self._h, _p(a)))

    def set_postprocess(self, post: PostProcess):
        self._chk(self._lib.cudapt_set_postprocess(self._h, ctypes.byref(post)))

    def exposure(self) -> float:
        e = ctypes.c_float()
        self._chk(self._lib.cudapt_get_exposure(self._h, ctypes.byref(e)))
        return e.value

    def write_image(self, path: str):
        self._chk(self._lib.cudapt_write_image(self._h, os.fsencode(path)))

    def read_accum(self):
        out = np.empty((self.height, self.width, 4), np.float32)
        fc = ctypes.c_uint32()
        self._chk(self._lib.cudapt_read_accum(self._h, _p(out), ctypes.byref(fc)))
        return out, fc.value

    def write_accum(self, accum: np.ndarray, frame_count: int):
        a = np.ascontiguousarray(accum, np.float32)
        assert a.size == self.width * self.height * 4
        self._chk(self._lib.cudapt_write_accum(self._h, _p(a), frame_count))

    def read_primary_hits(self):
        prim = np.empty((self.height, self.width), np.uint32)
        t = np.empty((self.height, self.width), np.float32)
        self._chk(self._lib.cudapt_read_primary_hits(self._h, _p(prim), _p(t)))
        return prim, t

    def stats(self) -> Stats:
        s = Stats()
        self._chk(self._lib.cudapt_get_stats(self._h, ctypes.byref(s)))
        return s

    def set_counting(self, enabled: bool):
        self._chk(self._lib.cudapt_set_counting(self._h, int(enabled)))

    # multi-GPU
    def set_tile_partition(self, rank: int, world: int, tile_w: int = 32, tile_h: int = 32):
        self._chk(self._lib.cudapt_set_tile_partition(self._h, rank, world, tile_w, tile_h))

    def set_frame_range(self, first: int, stride: int):
        self._chk(self._lib.cudapt_set_frame_range(self._h, first, stride))

    def set_frame_count(self, n: int):
        self._chk(self._lib.cudapt_set_frame_count(self._h, n))

    # fused multi-GPU read-back tail
    def reduce_resolve_peers(self, accum_by_rank, rank: int, total_frames: int, root_rgba: int, root_accum: int | None = None):
        arr = (ctypes.c_void_p * len(accum_by_rank))(*[ctypes.c_void_p(int(a)) for a in accum_by_rank])
        self._chk(self._lib.cudapt_reduce_resolve_peers(self._h, arr, len(accum_by_rank), rank, total_frames,
                                                        ctypes.c_void_p(int(root_rgba)),
                                                        ctypes.c_void_p(int(root_accum)) if root_accum else None))

    def download_rgba8(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width), np.uint32)
        self._chk(self._lib.cudapt_download_rgba8(self._h, _p(out)))
        return out

    def ipc_export(self):
        a, r = ctypes.create_string_buffer(64), ctypes.create_string_buffer(64)
        self._chk(self._lib.cudapt_ipc_export(self._h, a, r))
        return a.raw, r.raw

    def ipc_open(self, handle: bytes) -> int:
        p = ctypes.c_void_p()
        self._chk(self._lib.cudapt_ipc_open(self._h, ctypes.c_char_p(handle), ctypes.byref(p)))
        return int(p.value)

    def ipc_close(self, ptr: int):
        self._chk(self._lib.cudapt_ipc_close(self._h, ctypes.c_void_p(int(ptr))))

    @property
    def rgba_device_ptr(self) -> int:
        return int(self._lib.cudapt_rgba_device_ptr(self._h) or 0)

    @property
    def accum_device_ptr(self) -> int:
        return int(self._lib.cudapt_accum_device_ptr(self._h) or 0)

    @property
    def stream_handle(self) -> int:
        return int(self._lib.cudapt_stream(self._h) or 0)

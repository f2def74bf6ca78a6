"""simd-ray-tracer_b200: B200-native backend for SIMD-Ray-Tracer's path-tracing hot path.

The product is the C-ABI shared library libsrt_cuda.so (include/srt_backend.h) plus the
C++ host mirror of the reference's OnInit/OnRender (host/).  This Python module is only a
thin ctypes binding for tests and bench.py; it never computes pixels itself and raises if
the CUDA library is missing (there is no CPU fallback).

Import with importlib (the directory name has a hyphen):
    srt = importlib.import_module("simd-ray-tracer_b200")
"""
import ctypes as C
import os

import numpy as np

from . import build as _build
from . import sharding  # noqa: F401

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = _build.LIB

# ---- POD mirrors of include/srt_backend.h -------------------------------------------------
sphere_dtype = np.dtype([
    ("position", np.float32, 4), ("radius", np.float32), ("_pad", np.float32, 3),
    ("color", np.float32, 4), ("emissive", np.float32, 4),
    ("specular", np.float32), ("ior", np.float32), ("_pad2", np.float32, 2)])
camera_dtype = np.dtype([
    ("position", np.float32, 4), ("z", np.float32, 4), ("x", np.float32, 4), ("y", np.float32, 4),
    ("film_center", np.float32, 4), ("film_w", np.float32), ("film_h", np.float32), ("_pad", np.float32, 2)])
assert sphere_dtype.itemsize == 80 and camera_dtype.itemsize == 96

ACCUM_MEAN, ACCUM_SUM = 0, 1

EXPORTS = [
    "srt_create", "srt_destroy", "srt_last_error", "srt_set_scene", "srt_resize",
    "srt_reset_accumulation", "srt_bind_device_buffers", "srt_render", "srt_poll", "srt_wait",
    "srt_resolve_sum", "srt_read_rgba8", "srt_read_accum", "srt_read_primary",
    "srt_accum_device_ptr", "srt_rgba_device_ptr", "srt_cuda_stream", "srt_device_info",
    "srt_filter_audit", "srt_write_accum", "srt_device_kat",
    "srt_comm_unique_id", "srt_comm_init_rank", "srt_comm_destroy", "srt_comm_info", "srt_set_partition",
    "srt_reduce_accum", "srt_gather_rgba", "srt_gather_accum", "srt_comm_group_begin", "srt_comm_group_end",
    "srt_render_static", "srt_last_kernel",
]
STATIC_MAX_PIXELS = 2 << 20
COMM_ID_BYTES = 128
KAT_PCG, KAT_RANDOM_FLOAT, KAT_STREAM_SEED, KAT_NORMALIZE, KAT_REFLECTANCE, KAT_LINEAR_TO_SRGB, KAT_PACK = range(7)
_KAT_IO = {  # op: (input dtype, values per record in, output dtype, values per record out)
    KAT_PCG: (np.uint64, 1, np.uint32, 16), KAT_RANDOM_FLOAT: (np.uint64, 2, np.float32, 16),
    KAT_STREAM_SEED: (np.uint64, 1, np.uint64, 1), KAT_NORMALIZE: (np.float32, 3, np.float32, 3),
    KAT_REFLECTANCE: (np.float32, 2, np.float32, 1), KAT_LINEAR_TO_SRGB: (np.float32, 1, np.float32, 1),
    KAT_PACK: (np.float32, 4, np.uint32, 1)}


class SceneOptions(C.Structure):
    _fields_ = [("use_sky_color", C.c_int32), ("lane_width", C.c_int32), ("fma_dot", C.c_int32), ("_reserved", C.c_int32)]


class RenderArgs(C.Structure):
    _fields_ = [("first_sample", C.c_uint32), ("n_samples", C.c_uint32), ("sample_stride", C.c_uint32),
                ("accum_mode", C.c_uint32), ("stream_base", C.c_uint64), ("record_primary", C.c_uint32),
                ("_reserved", C.c_uint32)]


class SrtError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen libsrt_cuda.so (no compute). Raises if it has not been built.
    SRT_CUDA_LIB selects another build of the same ABI (the -DSRT_SPIN_LIMIT debug library)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("SRT_CUDA_LIB", LIB_PATH)
    if not os.path.exists(path):
        raise SrtError(f"{path} is missing: run __graft_entry__.build() (nvcc). There is no CPU fallback.")
    L = C.CDLL(path)
    vp = C.c_void_p
    L.srt_create.argtypes = [C.POINTER(vp), C.c_int]
    L.srt_destroy.argtypes = [vp]
    L.srt_destroy.restype = None
    L.srt_last_error.argtypes = [vp]
    L.srt_last_error.restype = C.c_char_p
    L.srt_set_scene.argtypes = [vp, vp, C.c_uint32, C.POINTER(SceneOptions)]
    L.srt_resize.argtypes = [vp, C.c_uint32, C.c_uint32]
    L.srt_reset_accumulation.argtypes = [vp]
    L.srt_bind_device_buffers.argtypes = [vp, vp, vp]
    L.srt_render.argtypes = [vp, vp, C.POINTER(RenderArgs)]
    L.srt_render_static.argtypes = [vp, vp, C.POINTER(RenderArgs)]
    L.srt_poll.argtypes = [vp]
    L.srt_wait.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    L.srt_resolve_sum.argtypes = [vp, C.c_uint32]
    L.srt_filter_audit.argtypes = [vp, vp, C.POINTER(RenderArgs), C.POINTER(C.c_uint64)]
    L.srt_read_rgba8.argtypes = [vp, vp]
    L.srt_read_accum.argtypes = [vp, vp]
    L.srt_read_primary.argtypes = [vp, vp, vp]
    L.srt_write_accum.argtypes = [vp, vp]
    for f in ("srt_accum_device_ptr", "srt_rgba_device_ptr", "srt_cuda_stream"):
        getattr(L, f).argtypes = [vp]
        getattr(L, f).restype = vp
    L.srt_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_uint64)]
    L.srt_last_kernel.argtypes = [vp, C.c_char_p, C.c_uint32]
    L.srt_device_kat.argtypes = [vp, C.c_uint32, vp, C.c_uint32, vp]
    L.srt_comm_unique_id.argtypes = [vp]
    L.srt_comm_init_rank.argtypes = [vp, C.c_int, C.c_int, vp]
    L.srt_comm_destroy.argtypes = [vp]
    L.srt_comm_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.srt_set_partition.argtypes = [vp, C.c_uint32, C.c_uint32]
    L.srt_reduce_accum.argtypes = [vp, C.c_int]
    L.srt_gather_rgba.argtypes = [vp, C.c_int]
    L.srt_gather_accum.argtypes = [vp, C.c_int]
    _lib = L
    return L


class Backend:
    """One handle of the C ABI. Mirrors the call order a reference maintainer would use."""

    def __init__(self, device=0):
        self.L = load_library()
        self.h = C.c_void_p()
        rc = self.L.srt_create(C.byref(self.h), device)
        if rc != 0:
            raise SrtError(f"srt_create failed ({rc}): {self.L.srt_last_error(None).decode()}")
        self.width = self.height = 0

    def _check(self, rc):
        if rc < 0:
            raise SrtError(f"srt error {rc}: {self.L.srt_last_error(self.h).decode()}")
        return rc

    def close(self):
        if self.h:
            self.L.srt_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_scene(self, spheres, use_sky, lane_width=8, fma_dot=True):
        spheres = np.ascontiguousarray(spheres)
        assert spheres.dtype.itemsize == 80
        opt = SceneOptions(int(use_sky), lane_width, int(fma_dot), 0)
        self._check(self.L.srt_set_scene(self.h, spheres.ctypes.data, len(spheres), C.byref(opt)))

    def resize(self, w, h):
        self._check(self.L.srt_resize(self.h, w, h))
        self.width, self.height = w, h

    def reset(self):
        self._check(self.L.srt_reset_accumulation(self.h))

    def bind_device_buffers(self, accum_ptr, rgba_ptr):
        self._check(self.L.srt_bind_device_buffers(self.h, accum_ptr, rgba_ptr))

    def render(self, camera, first_sample=0, n_samples=1, stride=1, mode=ACCUM_MEAN, stream_base=0, record_primary=False):
        cam = np.ascontiguousarray(camera)
        args = RenderArgs(first_sample, n_samples, stride, mode, stream_base, int(record_primary), 0)
        self._check(self.L.srt_render(self.h, cam.ctypes.data, C.byref(args)))

    def render_static(self, camera, first_sample=0, n_samples=1, stride=1, mode=ACCUM_MEAN, stream_base=0):
        """srt_render_static: the render that allocates nothing (scratch in a file-scope device array)."""
        cam = np.ascontiguousarray(camera)
        args = RenderArgs(first_sample, n_samples, stride, mode, stream_base, 0, 0)
        self._check(self.L.srt_render_static(self.h, cam.ctypes.data, C.byref(args)))

    def poll(self):
        return self._check(self.L.srt_poll(self.h)) == 1

    def wait(self):
        ms, rays = C.c_double(), C.c_uint64()
        self._check(self.L.srt_wait(self.h, C.byref(ms), C.byref(rays)))
        return ms.value, rays.value

    def filter_audit(self, camera, first_sample=0, n_samples=1, stream_base=0):
        """Soundness audit of the sphere filters: dict(tests, hits, candidates, violations, slab_survivors,
        bundle_tests, bundle_candidates, bundle_violations)."""
        cam = np.ascontiguousarray(camera)
        args = RenderArgs(first_sample, n_samples, 1, ACCUM_MEAN, stream_base, 0, 0)
        counts = (C.c_uint64 * 8)()
        self._check(self.L.srt_filter_audit(self.h, cam.ctypes.data, C.byref(args), counts))
        return dict(tests=counts[0], hits=counts[1], candidates=counts[2], violations=counts[3], slab_survivors=counts[4],
                    bundle_tests=counts[5], bundle_candidates=counts[6], bundle_violations=counts[7])

    def resolve_sum(self, total_samples):
        self._check(self.L.srt_resolve_sum(self.h, total_samples))

    def read_rgba8(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width), np.uint32)
        self._check(self.L.srt_read_rgba8(self.h, out.ctypes.data))
        return out

    def read_accum(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._check(self.L.srt_read_accum(self.h, out.ctypes.data))
        return out

    def write_accum(self, accum):
        """Upload a host accumulator (H x W x 4 float32) to resume an accumulation."""
        a = np.ascontiguousarray(accum, np.float32)
        assert a.shape == (self.height, self.width, 4)
        self._check(self.L.srt_write_accum(self.h, a.ctypes.data))

    def read_primary(self):
        prim = np.empty((self.height, self.width), np.uint32)
        t = np.empty((self.height, self.width), np.float32)
        self._check(self.L.srt_read_primary(self.h, prim.ctypes.data, t.ctypes.data))
        return prim, t

    def device_kat(self, op, records):
        """Run device function `op` (KAT_*) on `records` (one row per record); returns one row of outputs per record."""
        it, ni, ot, no = _KAT_IO[op]
        a = np.ascontiguousarray(records, it).reshape(-1, ni)
        out = np.zeros((len(a), no), ot)
        self._check(self.L.srt_device_kat(self.h, op, a.ctypes.data, len(a), out.ctypes.data))
        return out

    # ---- several devices (one handle per device / process) --------------------------------
    def comm_init_rank(self, n_ranks, rank, unique_id):
        _preload_nccl()
        buf = np.frombuffer(bytes(unique_id), np.uint8).copy()
        assert buf.nbytes == COMM_ID_BYTES
        self._check(self.L.srt_comm_init_rank(self.h, n_ranks, rank, buf.ctypes.data))

    def comm_destroy(self):
        self._check(self.L.srt_comm_destroy(self.h))

    def comm_info(self):
        _preload_nccl()
        r, n, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self.L.srt_comm_info(self.h, C.byref(r), C.byref(n), C.byref(v)))
        return dict(rank=r.value, n_ranks=n.value, nccl_version=v.value)

    def set_partition(self, part, n_parts):
        self._check(self.L.srt_set_partition(self.h, part, n_parts))

    def reduce_accum(self, root=0):
        self._check(self.L.srt_reduce_accum(self.h, root))

    def gather_rgba(self, root=0):
        self._check(self.L.srt_gather_rgba(self.h, root))

    def gather_accum(self, root=0):
        self._check(self.L.srt_gather_accum(self.h, root))

    def last_kernel(self):
        """The render kernel the last render launched, as ncu prints it."""
        buf = C.create_string_buffer(128)
        self._check(self.L.srt_last_kernel(self.h, buf, 128))
        return buf.value.decode()

    def device_info(self):
        sm, khz, n = C.c_int(), C.c_int(), C.c_uint64()
        self._check(self.L.srt_device_info(self.h, C.byref(sm), C.byref(khz), C.byref(n)))
        return dict(sm_count=sm.value, clock_khz=khz.value, launches=n.value)

    @property
    def accum_ptr(self):
        return self.L.srt_accum_device_ptr(self.h)

    @property
    def rgba_ptr(self):
        return self.L.srt_rgba_device_ptr(self.h)

    @property
    def stream_ptr(self):
        return self.L.srt_cuda_stream(self.h)


_nccl_preloaded = False


def _preload_nccl():
    """libsrt_cuda.so binds NCCL at run time: the libnccl.so.2 already mapped into the process, else the loader's.  In a
    Python process PyTorch may be imported LATER and needs its own (newer) bundled libnccl.so.2 -- glibc would hand it the
    one already loaded under that soname.  So before the library's first NCCL call, map the bundled one if there is one."""
    global _nccl_preloaded
    if _nccl_preloaded:
        return
    _nccl_preloaded = True
    import importlib.util
    import sys
    if "torch" in sys.modules:
        return                                   # torch's NCCL is already in the process
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            path = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return
    except (ImportError, OSError, ValueError):
        pass


def comm_unique_id():
    """128-byte NCCL unique id (made by one rank, handed to every rank's Backend.comm_init_rank)."""
    _preload_nccl()
    L = load_library()
    buf = np.zeros(COMM_ID_BYTES, np.uint8)
    rc = L.srt_comm_unique_id(buf.ctypes.data)
    if rc != 0:
        raise SrtError(f"srt_comm_unique_id failed ({rc}): {L.srt_last_error(None).decode()}")
    return buf.tobytes()


# ---- host mirror (C++): scene tables and camera, no GPU needed ------------------------------
HOSTLIB_PATH = _build.HOSTLIB
DRIVER_PATH = _build.DRIVER
SCENE_SYNTH_1K, SCENE_SYNTH_64K = 3, 4
_hostlib = None


def load_host_library():
    global _hostlib
    if _hostlib is None:
        if not os.path.exists(HOSTLIB_PATH):
            raise SrtError(f"{HOSTLIB_PATH} is missing: run __graft_entry__.build()")
        H = C.CDLL(HOSTLIB_PATH)
        H.srt_host_scene.argtypes = [C.c_uint, C.c_void_p, C.c_uint, C.POINTER(C.c_int), C.c_void_p]
        H.srt_host_camera.argtypes = [C.c_uint, C.c_float, C.c_float, C.c_float, C.c_uint, C.c_uint, C.c_void_p]
        H.srt_host_save_scene.argtypes = [C.c_uint, C.c_char_p]
        H.srt_host_load_scene.argtypes = [C.c_uint, C.c_char_p]
        H.srt_host_reset_scene.argtypes = [C.c_uint]
        H.srt_host_reset_scene.restype = None
        H.srt_host_scene_hash.argtypes = [C.c_uint]
        H.srt_host_scene_hash.restype = C.c_ulonglong
        H.srt_host_save_checkpoint.argtypes = [C.c_char_p, C.c_void_p, C.c_ulonglong, C.c_void_p]
        H.srt_host_load_checkpoint.argtypes = [C.c_char_p, C.c_void_p, C.POINTER(C.c_ulonglong), C.c_void_p, C.c_ulonglong]
        H.srt_host_write_image.argtypes = [C.c_char_p, C.c_void_p, C.c_uint, C.c_uint, C.c_int]
        H.srt_host_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_uint, C.c_uint]
        _hostlib = H
    return _hostlib


def save_scene_file(index, path):
    """Write scene slot `index` in the text format of host/srt_files.cpp."""
    if load_host_library().srt_host_save_scene(index, os.fsencode(path)) != 0:
        raise SrtError(f"cannot write scene file {path}")


def load_scene_file(index, path):
    """Load a scene text file into slot `index` of the host mirror (then host_scene(index))."""
    if load_host_library().srt_host_load_scene(index, os.fsencode(path)) != 0:
        raise SrtError(f"cannot load scene file {path}")


def reset_scene(index):
    """Undo load_scene_file: slot `index` is rebuilt as the built-in / synthetic scene on next use."""
    load_host_library().srt_host_reset_scene(index)


def write_image(path, rgba, kind=0):
    """Write a presented image (H x W uint32 RGBA8, row 0 = bottom) as PPM or PNG (kind 0: by extension, 1 PPM, 2 PNG)."""
    a = np.ascontiguousarray(rgba, np.uint32)
    if a.ndim != 2 or load_host_library().srt_host_write_image(os.fsencode(path), a.ctypes.data, a.shape[1], a.shape[0], int(kind)) != 0:
        raise SrtError(f"cannot write image {path}")


def write_pfm(path, accum):
    """Write a linear accumulator (H x W x 4 float32) as an RGB PFM image."""
    a = np.ascontiguousarray(accum, np.float32)
    if a.ndim != 3 or a.shape[2] != 4 or load_host_library().srt_host_write_pfm(os.fsencode(path), a.ctypes.data, a.shape[1], a.shape[0]) != 0:
        raise SrtError(f"cannot write image {path}")


def host_scene_hash(index):
    """FNV-1a of scene slot `index` (sphere records + sky flag): what a checkpoint pins its scene with."""
    return int(load_host_library().srt_host_scene_hash(index))


def save_checkpoint(path, accum, samples, scene_index, stream_base=0, lane_width=8, fma_dot=1):
    """Write an accumulator checkpoint (host/srt_files.cpp): accum is H x W x 4 float32."""
    a = np.ascontiguousarray(accum, np.float32)
    hdr = np.array([a.shape[1], a.shape[0], samples, scene_index, lane_width, fma_dot], np.uint32)
    if load_host_library().srt_host_save_checkpoint(os.fsencode(path), hdr.ctypes.data, stream_base, a.ctypes.data) != 0:
        raise SrtError(f"cannot write checkpoint {path}")


def load_checkpoint(path):
    """-> (accum H x W x 4 float32, dict(width, height, samples, scene_index, lane_width, fma_dot, stream_base))."""
    H = load_host_library()
    hdr = np.zeros(6, np.uint32)
    base = C.c_ulonglong()
    if H.srt_host_load_checkpoint(os.fsencode(path), hdr.ctypes.data, C.byref(base), None, 0) != 0:
        raise SrtError(f"{path} is not a checkpoint")
    a = np.zeros((int(hdr[1]), int(hdr[0]), 4), np.float32)
    if H.srt_host_load_checkpoint(os.fsencode(path), hdr.ctypes.data, C.byref(base), a.ctypes.data, a.size) != 0:
        raise SrtError(f"{path} is not a checkpoint")
    keys = ("width", "height", "samples", "scene_index", "lane_width", "fma_dot")
    return a, {**{k: int(v) for k, v in zip(keys, hdr)}, "stream_base": int(base.value)}


def host_scene(index):
    """(spheres, use_sky, header) of scene `index` as built by host/srt_scenes.cpp."""
    H = load_host_library()
    sky = C.c_int()
    hdr = np.zeros(6, np.float32)
    n = H.srt_host_scene(index, None, 0, C.byref(sky), hdr.ctypes.data)
    if n <= 0:
        raise SrtError(f"bad scene index {index}")
    sp = np.zeros(n, sphere_dtype)
    H.srt_host_scene(index, sp.ctypes.data, n, None, None)
    return sp, bool(sky.value), hdr


def host_camera(index, width, height, header=None):
    H = load_host_library()
    if header is None:
        header = host_scene(index)[2]
    cam = np.zeros(1, camera_dtype)
    H.srt_host_camera(index, float(header[3]), float(header[4]), float(header[5]), width, height, cam.ctypes.data)
    return cam

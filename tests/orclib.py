"""ctypes loaders for the CPU checkers (TEST INFRASTRUCTURE ONLY).

  Oracle  -> oracle/liborc.so            the C restatement (oracle/srt_oracle.c)
  Ref     -> oracle/_ref/libsrt_ref_w{4,8}.so   the reference's own code, compiled by
             oracle/build_ref.py (present only where /root/reference was available
             at build time, or shipped prebuilt to the GPU box)
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
SPHERE_BYTES = 80

sphere_dtype = np.dtype([
    ("position", np.float32, 4), ("radius", np.float32), ("_pad", np.float32, 3),
    ("color", np.float32, 4), ("emissive", np.float32, 4),
    ("specular", np.float32), ("ior", np.float32), ("_pad2", np.float32, 2)])
assert sphere_dtype.itemsize == SPHERE_BYTES
camera_dtype = np.dtype([
    ("position", np.float32, 4), ("z", np.float32, 4), ("x", np.float32, 4), ("y", np.float32, 4),
    ("film_center", np.float32, 4), ("film_w", np.float32), ("film_h", np.float32), ("_pad", np.float32, 2)])
assert camera_dtype.itemsize == 96


def build_oracle(force=False):
    so = os.path.join(ORACLE_DIR, "liborc.so")
    src = os.path.join(ORACLE_DIR, "srt_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-mfma", "-fPIC", "-shared", "-pthread",
                               src, "-o", so, "-lm"])
    return so


class Oracle:
    def __init__(self):
        L = C.CDLL(build_oracle())
        self.L = L
        L.orc_pcg.restype = C.c_uint32
        L.orc_pcg.argtypes = [C.POINTER(C.c_uint64)]
        L.orc_random_float.restype = C.c_float
        L.orc_random_float.argtypes = [C.POINTER(C.c_uint64), C.c_float, C.c_float]
        L.orc_stream_seed.restype = C.c_uint64
        L.orc_stream_seed.argtypes = [C.c_uint64]
        L.orc_reflectance.restype = C.c_float
        L.orc_reflectance.argtypes = [C.c_float, C.c_float]
        L.orc_linear_to_srgb.restype = C.c_float
        L.orc_linear_to_srgb.argtypes = [C.c_float]
        L.orc_pack.restype = C.c_uint32
        L.orc_pack.argtypes = [C.c_void_p]
        L.orc_normalize.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_builtin_scene.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.POINTER(C.c_int), C.c_void_p]
        L.orc_camera_from_orbit.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_uint32, C.POINTER(C.c_uint64), C.c_int]

    def render_sum(self, spheres, sky, lane_width, fma, cam, w, h, first, n, stride, accum=None, stream_base=0, threads=0):
        spheres = np.ascontiguousarray(spheres)
        if accum is None:
            accum = np.zeros((h, w, 4), np.float32)
        rays = C.c_uint64()
        self.L.orc_render_sum.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32,
                                          C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64), C.c_int]
        r = self.L.orc_render_sum(spheres.ctypes.data, len(spheres), int(sky), lane_width, int(fma), cam.ctypes.data, w, h,
                                  first, n, stride, stream_base, accum.ctypes.data, C.byref(rays), threads or (os.cpu_count() or 1))
        assert r == 0, r
        return accum, rays.value

    def resolve_sum(self, accum, total):
        h, w = accum.shape[:2]
        rgba = np.zeros((h, w), np.uint32)
        self.L.orc_resolve_sum.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        self.L.orc_resolve_sum.restype = None
        self.L.orc_resolve_sum(accum.ctypes.data, rgba.ctypes.data, h * w, total)
        return rgba

    def scene(self, idx):
        sp = np.zeros(488, sphere_dtype)
        sky = C.c_int()
        hdr = np.zeros(6, np.float32)
        n = self.L.orc_builtin_scene(idx, sp.ctypes.data, 488, C.byref(sky), hdr.ctypes.data)
        assert n > 0
        return sp[:n].copy(), bool(sky.value), hdr

    def camera(self, hdr, w, h):
        cam = np.zeros(1, camera_dtype)
        la = np.ascontiguousarray(hdr[:3], np.float32)
        self.L.orc_camera_from_orbit(la.ctypes.data, float(hdr[3]), float(hdr[4]), float(hdr[5]), w, h, cam.ctypes.data)
        return cam

    def render(self, spheres, sky, lane_width, fma, cam, w, h, first=0, n=1, stream_base=0,
               accum=None, threads=0, tap_sample=0):
        spheres = np.ascontiguousarray(spheres)
        if accum is None:
            accum = np.zeros((h, w, 4), np.float32)
        rgba = np.zeros((h, w), np.uint32)
        prim = np.full((h, w), 0xFFFFFFFF, np.uint32)
        t = np.full((h, w), 1e30, np.float32)
        rays = C.c_uint64()
        r = self.L.orc_render(spheres.ctypes.data, len(spheres), int(sky), lane_width, int(fma), cam.ctypes.data, w, h,
                              first, n, stream_base, accum.ctypes.data, rgba.ctypes.data, prim.ctypes.data,
                              t.ctypes.data, tap_sample, C.byref(rays), threads or (os.cpu_count() or 1))
        assert r == 0, r
        return dict(accum=accum, rgba=rgba, prim=prim, t=t, rays=rays.value)


def ref_path(tag):
    return os.path.join(ORACLE_DIR, "_ref", f"libsrt_ref_{tag}.so")


def have_ref():
    return os.path.exists(ref_path("w4")) and os.path.exists(ref_path("w8"))


class Ref:
    """The reference's own compiled code. One instance per lane width (separate .so)."""

    def __init__(self, tag, max_threads=0):
        L = C.CDLL(ref_path(tag))
        self.L = L
        self.tag = tag
        L.srt_ref_render.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_int,
                                     C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        L.srt_ref_scene.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_int), C.c_void_p]
        L.srt_ref_camera.argtypes = [C.c_void_p]
        L.srt_ref_rng.argtypes = [C.c_uint64, C.c_uint32, C.c_float, C.c_float, C.c_int, C.c_void_p]
        L.srt_ref_thread_seed.restype = C.c_uint64
        L.srt_ref_thread_seed.argtypes = [C.c_uint32]
        L.srt_ref_stream_seed.restype = C.c_uint64
        L.srt_ref_stream_seed.argtypes = [C.c_uint64]
        L.srt_ref_linear_to_srgb.restype = C.c_float
        L.srt_ref_linear_to_srgb.argtypes = [C.c_float]
        L.srt_ref_pack.restype = C.c_uint32
        L.srt_ref_pack.argtypes = [C.c_void_p]
        L.srt_ref_reflectance.restype = C.c_float
        L.srt_ref_reflectance.argtypes = [C.c_float, C.c_float]
        L.srt_ref_normalize.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.srt_ref_set_scene.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        self.threads = L.srt_ref_init(max_threads)
        self.width = L.srt_ref_simd_width()

    def scene(self, idx):
        sp = np.zeros(488, sphere_dtype)
        cnt = C.c_uint32()
        sky = C.c_int()
        hdr = np.zeros(6, np.float32)
        sz = self.L.srt_ref_scene(idx, sp.ctypes.data, 488, C.byref(cnt), C.byref(sky), hdr.ctypes.data)
        assert sz == SPHERE_BYTES, sz
        return sp[:cnt.value].copy(), bool(sky.value), hdr

    def set_scene(self, slot, spheres, sky, header):
        spheres = np.ascontiguousarray(spheres)
        header = np.ascontiguousarray(header, np.float32)
        r = self.L.srt_ref_set_scene(slot, spheres.ctypes.data, len(spheres), int(sky), header.ctypes.data)
        assert r == 0, r

    def render(self, scene, w, h, spp, simd=1, threads=0, flags=3, stream_base=0, tap_sample=0):
        rgba = np.zeros((h, w), np.uint32)
        accum = np.zeros((h, w, 4), np.float32)
        prim = np.zeros((h, w), np.uint32)
        t = np.zeros((h, w), np.float32)
        rays = C.c_uint64()
        ms = C.c_double()
        r = self.L.srt_ref_render(scene, w, h, spp, simd, threads, flags, stream_base, rgba.ctypes.data,
                                  accum.ctypes.data, prim.ctypes.data, t.ctypes.data, tap_sample,
                                  C.byref(rays), C.byref(ms))
        assert r == 0, r
        cam = np.zeros(24, np.float32)
        self.L.srt_ref_camera(cam.ctypes.data)
        return dict(accum=accum, rgba=rgba, prim=prim, t=t, rays=rays.value, ms=ms.value, camera=cam)

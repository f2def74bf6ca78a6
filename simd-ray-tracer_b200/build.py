"""In-tree build of the CUDA backend (libsrt_cuda.so) and the headless host driver.

nvcc cross-compiles for sm_100a without a GPU.  -fmad=false is REQUIRED for parity: the
reference's IEEE profile has no fused multiply-adds, and the FMA profile has them only
where the device code spells __fmaf_rn (see csrc/srt_device.cuh).
"""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")
LIB = os.path.join(PKG, "libsrt_cuda.so")
DRIVER = os.path.join(PKG, "srt_headless")
HOSTLIB = os.path.join(PKG, "libsrt_host.so")

DBGLIB = os.path.join(PKG, "libsrt_cuda_dbg.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false",
              "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-Xptxas", "-v", "-ldl"]
# The debug library: same ABI, every wait loop of the pool kernel traps after SRT_SPIN_LIMIT iterations and
# reports its site (csrc/srt_device.cuh); only the lane_width 8 + FMA kernels (SRT_FAST_BUILD) so that it builds fast.
DEBUG_FLAGS = ["-DSRT_SPIN_LIMIT=200000000", "-DSRT_FAST_BUILD"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_library(force=False, verbose=False, extra=(), out=None):
    out = out or LIB
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    srcs.append(os.path.join(ROOT, "include", "srt_backend.h"))
    if force or _newer(out, srcs):
        cmd = ["nvcc", *NVCC_FLAGS, *extra, os.path.join(CSRC, "srt_backend.cu"), "-o", out]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or res.returncode != 0:
            print(" ".join(cmd))
            print(res.stdout + res.stderr)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed building {os.path.basename(out)}")
    return out


def build_debug_library(force=False, verbose=False):
    return build_library(force, verbose, extra=DEBUG_FLAGS, out=DBGLIB)


def build_driver(force=False, verbose=False):
    if not os.path.isdir(HOST):
        return None
    srcs = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST)) if f.endswith((".cpp", ".h"))]
    if not srcs:
        return None
    srcs_all = srcs + [os.path.join(ROOT, "include", "srt_backend.h")]
    if force or _newer(DRIVER, srcs_all) or _newer(DRIVER, [LIB]):
        common = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-I", os.path.join(ROOT, "include")]
        link = ["-L", PKG, "-lsrt_cuda", "-Wl,-rpath,$ORIGIN", "-lm", "-pthread"]
        tables = [os.path.join(HOST, f) for f in ("srt_scenes.cpp", "srt_files.cpp")]
        cmds = [
            # the headless driver: OnInit / OnRender over the C ABI
            [*common, os.path.join(HOST, "srt_app.cpp"), *tables, os.path.join(HOST, "srt_headless.cpp"), "-o", DRIVER, *link],
            # scene tables, camera and file formats for CPU-side users (tests, the reference arm of bench.py):
            # no backend call in it, hence no CUDA library behind it
            [*common, "-fPIC", "-shared", *tables, os.path.join(HOST, "srt_host_capi.cpp"), "-o", HOSTLIB, "-lm"],
        ]
        for cmd in cmds:
            res = subprocess.run(cmd, capture_output=True, text=True)
            if verbose or res.returncode != 0:
                print(" ".join(cmd))
                print(res.stdout + res.stderr)
            if res.returncode != 0:
                raise RuntimeError("g++ failed building the host mirror")
    return DRIVER


def build_all(force=False, verbose=False):
    build_library(force, verbose)
    build_debug_library(force, verbose)
    build_driver(force, verbose)

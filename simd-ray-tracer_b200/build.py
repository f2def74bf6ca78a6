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

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false",
              "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_library(force=False, verbose=False):
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    srcs.append(os.path.join(ROOT, "include", "srt_backend.h"))
    if force or _newer(LIB, srcs):
        cmd = ["nvcc", *NVCC_FLAGS, os.path.join(CSRC, "srt_backend.cu"), "-o", LIB]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or res.returncode != 0:
            print(" ".join(cmd))
            print(res.stdout + res.stderr)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed building libsrt_cuda.so")
    return LIB


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
        app = [os.path.join(HOST, f) for f in ("srt_app.cpp", "srt_scenes.cpp", "srt_files.cpp")]
        cmds = [
            [*common, *app, os.path.join(HOST, "srt_headless.cpp"), "-o", DRIVER, *link],
            [*common, "-fPIC", "-shared", *app, os.path.join(HOST, "srt_host_capi.cpp"), "-o", HOSTLIB, *link],
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
    build_driver(force, verbose)

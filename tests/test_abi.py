"""CPU: the C-ABI library loads and exports every symbol include/srt_backend.h declares,
and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "srt_backend.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(srt):
    L = srt.load_library()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/srt_backend.h but not exported"
    assert sorted(srt.EXPORTS) == names


def test_struct_sizes(srt):
    assert srt.sphere_dtype.itemsize == 80      # scalar_sphere, main.cpp:17-21
    assert srt.camera_dtype.itemsize == 96
    assert C.sizeof(srt.SceneOptions) == 16
    assert C.sizeof(srt.RenderArgs) == 32


def test_no_cpu_fallback(srt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(srt.SrtError) as e:
        srt.Backend(0)
    assert "no CPU fallback" in str(e.value) or "no CUDA device" in str(e.value)


def test_product_never_touches_oracle():
    """The product tree must not reference the oracle (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "simd-ray-tracer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liborc" not in txt and "srt_oracle" not in txt and "orclib" not in txt and "_ref/" not in txt, f


def test_comm_calls_fail_loudly_without_nccl():
    """The multi-device entry points bind NCCL at run time; when it cannot be loaded they return SRT_ERR_COMM (-5) with a
    message -- no host-copy fallback.  (A fresh process: the binding caches the first dlopen.)"""
    import subprocess
    import sys
    code = (
        "import ctypes as C, os, sys\n"
        "os.environ['SRT_NCCL_LIB'] = '/nonexistent/libnccl.so.2'\n"
        f"L = C.CDLL({os.path.join(ROOT, 'simd-ray-tracer_b200', 'libsrt_cuda.so')!r})\n"
        "L.srt_last_error.restype = C.c_char_p\n"
        "L.srt_last_error.argtypes = [C.c_void_p]\n"
        "buf = (C.c_ubyte * 128)()\n"
        "rc = L.srt_comm_unique_id(buf)\n"
        "print(rc, L.srt_last_error(None).decode())\n"
        "rc2 = L.srt_comm_group_begin()\n"
        "print(rc2)\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[0].startswith("-5 ") and "SRT_NCCL_LIB" in lines[0] and "cannot be loaded" in lines[0]
    assert lines[1] == "-5"

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

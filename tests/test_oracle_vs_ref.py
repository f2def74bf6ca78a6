"""CPU: restatement vs the reference's own compiled code (oracle/_ref), live.
Skipped where oracle/_ref was never built (it needs /root/reference at build time)."""
import numpy as np
import pytest

import orclib

pytestmark = pytest.mark.skipif(not orclib.have_ref(), reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def refs():
    return {"w4": orclib.Ref("w4"), "w8": orclib.Ref("w8")}


CASES = [(8, 1, "w8", 1), (4, 0, "w4", 1), (1, 0, "w4", 0), (1, 0, "w8", 0)]


@pytest.mark.parametrize("scene", [0, 1, 2])
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("size", [(200, 112, 2), (33, 65, 3)])
def test_bit_exact(oracle, refs, scene, case, size):
    W, fma, lib, simd = case
    w, h, spp = size
    sp, sky, hdr = oracle.scene(scene)
    cam = oracle.camera(hdr, w, h)
    r = refs[lib].render(scene, w, h, spp, simd=simd, flags=3, threads=0)
    o = oracle.render(sp, sky, W, fma, cam, w, h, 0, spp)
    for k in ("accum", "rgba", "prim", "t"):
        assert np.array_equal(o[k].view(np.uint32), r[k].view(np.uint32)), k
    assert o["rays"] == r["rays"]


def test_ref_parity_mode_is_thread_invariant(refs):
    a = refs["w8"].render(2, 96, 54, 2, simd=1, threads=1, flags=3)
    b = refs["w8"].render(2, 96, 54, 2, simd=1, threads=0, flags=3)
    assert np.array_equal(a["accum"].view(np.uint32), b["accum"].view(np.uint32))


def test_stream_base(oracle, refs):
    sp, sky, hdr = oracle.scene(1)
    cam = oracle.camera(hdr, 64, 36)
    r = refs["w8"].render(1, 64, 36, 2, simd=1, flags=3, stream_base=987654321012)
    o = oracle.render(sp, sky, 8, 1, cam, 64, 36, 0, 2, stream_base=987654321012)
    assert np.array_equal(o["accum"].view(np.uint32), r["accum"].view(np.uint32))


def test_synthetic_scene_through_reference_code(oracle, refs):
    """The reference's own RenderTile on the synthetic 1,024-sphere scene (injected through
    srt_ref_set_scene) == the restatement, bit for bit (cameras from the host mirror)."""
    import importlib
    srt = importlib.import_module("simd-ray-tracer_b200")
    sp, sky, hdr = srt.host_scene(srt.SCENE_SYNTH_1K)
    cam = srt.host_camera(srt.SCENE_SYNTH_1K, 64, 36)
    for lib, W, fma, simd in (("w8", 8, 1, 1), ("w4", 4, 0, 1), ("w4", 1, 0, 0)):
        refs[lib].set_scene(1, sp, sky, hdr)
        r = refs[lib].render(1, 64, 36, 2, simd=simd, flags=3)
        o = oracle.render(sp, sky, W, fma, cam, 64, 36, 0, 2)
        for k in ("accum", "rgba", "prim", "t"):
            assert np.array_equal(o[k].view(np.uint32), r[k].view(np.uint32)), (lib, k)
        assert o["rays"] == r["rays"]
        b_sp, b_sky, b_hdr = oracle.scene(1)
        refs[lib].set_scene(1, b_sp, b_sky, b_hdr)      # restore the built-in scene for later tests

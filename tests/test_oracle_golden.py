"""CPU: the C restatement (oracle/srt_oracle.c) against the committed golden vectors that
tests/golden/make_goldens.py produced from the reference's own compiled code."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import orclib

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODES = {"w8f": (8, 1), "w4i": (4, 0), "scalar": (1, 0)}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(G, "ref_goldens.npz"))


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(G, "ref_kat.json")) as f:
        return json.load(f)


def f32(bits):
    return np.array(bits, np.uint32).view(np.float32)


@pytest.mark.parametrize("scene", [0, 1, 2])
def test_scene_tables(oracle, gold, scene):
    sp, sky, hdr = oracle.scene(scene)
    assert sp.view(np.uint8).reshape(len(sp), 80).tobytes() == gold[f"scene{scene}_spheres"].tobytes()
    assert int(sky) == int(gold[f"scene{scene}_sky"][0])
    assert hdr.tobytes() == gold[f"scene{scene}_header"].tobytes()


@pytest.mark.parametrize("scene", [0, 1, 2])
@pytest.mark.parametrize("mode", list(MODES))
def test_images(oracle, gold, scene, mode):
    W, fma = MODES[mode]
    sp, sky, hdr = oracle.scene(scene)
    h, w = gold[f"img_s{scene}_{mode}_rgba"].shape
    cam = oracle.camera(hdr, w, h)
    ref_cam = gold[f"img_s{scene}_{mode}_camera"]
    mine = cam.view(np.float32).reshape(-1)
    for off in (0, 4, 8, 12, 16):
        assert mine[off:off + 3].tobytes() == ref_cam[off:off + 3].tobytes()
    assert mine[20:22].tobytes() == ref_cam[20:22].tobytes()
    o = oracle.render(sp, sky, W, fma, cam, w, h, 0, 2, threads=2)
    for k in ("accum", "rgba", "prim", "t"):
        assert np.array_equal(o[k].view(np.uint32), gold[f"img_s{scene}_{mode}_{k}"].view(np.uint32)), k
    assert o["rays"] == int(gold[f"img_s{scene}_{mode}_rays"][0])


def test_rng_kat(oracle, kat):
    L = oracle.L
    assert [L.orc_stream_seed(i) for i in range(4)] == kat["thread_seeds"]
    # stream contract: the reference's expression except where it zeroes the seed (i % 64 == 41, 57)
    assert [L.orc_stream_seed(i) for i in kat["stream_seed_index"]] == kat["stream_seeds"]
    assert len({L.orc_stream_seed(i) for i in range(1 << 14)}) == 1 << 14          # no two indices share a stream
    for seed, e in kat["rng"].items():
        st = C.c_uint64(int(seed))
        assert [L.orc_pcg(C.byref(st)) for _ in range(16)] == e["pcg"]
        for key, bits in e.items():
            if not key.startswith("float_"):
                continue
            _, mn, mx = key.split("_")
            st = C.c_uint64(int(seed))
            got = np.array([L.orc_random_float(C.byref(st), float(mn), float(mx)) for _ in range(16)], np.float32)
            assert got.view(np.uint32).tolist() == bits


def test_tonemap_pack_kat(oracle, kat):
    L = oracle.L
    xs = f32(kat["srgb_in"])
    for x, want, pk in zip(xs, kat["srgb_out"], kat["pack_out"]):
        got = np.float32(L.orc_linear_to_srgb(C.c_float(float(x))))
        if np.isnan(x):
            assert np.isnan(got)
        else:
            assert int(got.view(np.uint32)) == want
        v = np.array([x, 0.5, 0.001, 1.0], np.float32)
        assert L.orc_pack(v.ctypes.data) == pk


def test_reflectance_normalize_kat(oracle, kat):
    L = oracle.L
    for c, eta, bits in kat["reflectance"]:
        assert int(np.float32(L.orc_reflectance(c, eta)).view(np.uint32)) == bits
    for vin, vout in kat["normalize"]:
        v = f32(vin).copy()
        o = np.zeros(3, np.float32)
        L.orc_normalize(v.ctypes.data, o.ctypes.data)
        assert o.view(np.uint32).tolist() == vout


def test_inverse_max_int_constants():
    """The device code hard-codes InverseMaxInt (base.h:985) as 2^-32 / 2^-31."""
    assert np.float32((np.float32(1.0) - np.float32(0.0)) / np.float64(4294967295.0)) == np.float32(2.0 ** -32)
    assert np.float32((np.float32(0.5) - np.float32(-0.5)) / np.float64(4294967295.0)) == np.float32(2.0 ** -32)
    assert np.float32((np.float32(1.0) - np.float32(-1.0)) / np.float64(4294967295.0)) == np.float32(2.0 ** -31)


def test_sample_split_is_exact(oracle):
    """accum in/out: rendering samples [0,3) then [3,5) equals [0,5) bit for bit."""
    sp, sky, hdr = oracle.scene(0)
    cam = oracle.camera(hdr, 40, 24)
    full = oracle.render(sp, sky, 8, 1, cam, 40, 24, 0, 5)
    a = oracle.render(sp, sky, 8, 1, cam, 40, 24, 0, 3)
    b = oracle.render(sp, sky, 8, 1, cam, 40, 24, 3, 2, accum=a["accum"].copy())
    assert np.array_equal(full["accum"].view(np.uint32), b["accum"].view(np.uint32))
    assert full["rays"] == a["rays"] + b["rays"]

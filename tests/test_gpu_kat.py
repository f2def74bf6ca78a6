"""GPU: function-level known answers ON THE DEVICE.  tests/golden/ref_kat.json holds outputs of
the reference's own compiled PCG / RandomFloat (base.h:954-989), thread-seed scrambler (main.cpp:667-676),
LinearToSRGB / ColorFromV4 (main.cpp:312-346) at 0, 0.0031308 +- 1 ulp, > 1, +-inf and NaN, Reflectance
(main.cpp:292-300) and Normalize (x64_math.h:234-245).  srt_device_kat runs the CUDA functions the render kernels
are built from on the same inputs."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(G, "ref_kat.json")) as f:
        return json.load(f)


def test_device_pcg_and_random_float(backend, srt, kat):
    seeds = [int(s) for s in kat["rng"]]
    got = backend.device_kat(srt.KAT_PCG, np.array(seeds, np.uint64))
    for s, row in zip(seeds, got):
        assert [int(x) for x in row] == kat["rng"][str(s)]["pcg"]
    for (mn, mx) in [(-1.0, 1.0), (-0.5, 0.5), (0.0, 1.0), (2.0, 5.0), (0.15, 1.0)]:
        inv = np.float32(np.float64(np.float32(mx) - np.float32(mn)) / 4294967295.0)      # InverseMaxInt, base.h:985
        rec = np.zeros((len(seeds), 2), np.uint64)
        rec[:, 0] = seeds
        rec[:, 1] = int(np.float32(inv).view(np.uint32)) | (int(np.float32(mn).view(np.uint32)) << 32)
        got = backend.device_kat(srt.KAT_RANDOM_FLOAT, rec)
        for s, row in zip(seeds, got):
            assert [int(x) for x in row.view(np.uint32)] == kat["rng"][str(s)][f"float_{mn}_{mx}"], (s, mn, mx)


def test_device_stream_seed(backend, srt, kat):
    got = backend.device_kat(srt.KAT_STREAM_SEED, np.arange(4, dtype=np.uint64))
    assert [int(x) for x in got[:, 0]] == kat["thread_seeds"]


def test_device_tonemap_and_pack_edge_values(backend, srt, kat):
    xs = np.array(kat["srgb_in"], np.uint32).view(np.float32)
    got = backend.device_kat(srt.KAT_LINEAR_TO_SRGB, xs)[:, 0]
    for x, g, want in zip(xs, got.view(np.uint32), kat["srgb_out"]):
        if np.isnan(x):
            assert np.isnan(np.uint32(g).view(np.float32)) and np.isnan(np.uint32(want).view(np.float32))
        else:
            assert int(g) == want, float(x)
    rec = np.zeros((len(xs), 4), np.float32)
    rec[:, 0] = xs; rec[:, 1] = 0.5; rec[:, 2] = 0.001; rec[:, 3] = 1.0
    got = backend.device_kat(srt.KAT_PACK, rec)[:, 0]
    assert [int(v) for v in got] == kat["pack_out"]
    # every channel position takes the edge values, not only red
    for ch in (1, 2):
        rec2 = np.full((len(xs), 4), 0.25, np.float32)
        rec2[:, ch] = xs
        g2 = backend.device_kat(srt.KAT_PACK, rec2)[:, 0]
        red = np.array(kat["pack_out"], np.uint32) & 0xFF
        assert np.array_equal((g2 >> (8 * ch)) & 0xFF, red)


def test_device_reflectance_and_normalize(backend, srt, kat):
    rec = np.array([[c, eta] for c, eta, _ in kat["reflectance"]], np.float32)
    got = backend.device_kat(srt.KAT_REFLECTANCE, rec)[:, 0]
    assert [int(v) for v in got.view(np.uint32)] == [b for _, _, b in kat["reflectance"]]
    vin = np.array([v for v, _ in kat["normalize"]], np.uint32).view(np.float32)
    got = backend.device_kat(srt.KAT_NORMALIZE, vin)
    assert np.array_equal(got.view(np.uint32), np.array([o for _, o in kat["normalize"]], np.uint32))

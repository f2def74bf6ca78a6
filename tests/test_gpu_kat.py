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
    # the stream contract around the reference's zero-shift indices (i % 64 == 41, 57) and on 64-bit indices
    got = backend.device_kat(srt.KAT_STREAM_SEED, np.array(kat["stream_seed_index"], dtype=np.uint64))
    assert [int(x) for x in got[:, 0]] == kat["stream_seeds"]
    got = backend.device_kat(srt.KAT_STREAM_SEED, np.arange(1 << 14, dtype=np.uint64))
    assert len(set(int(x) for x in got[:, 0])) == 1 << 14


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


def test_last_kernel_names_the_launched_kernel(srt):
    """srt_last_kernel: the instantiation srt_render picked, spelled as ncu prints it (bench.py matches its ncu profile
    against this name instead of guessing)."""
    b = srt.Backend(0)
    try:
        assert b.last_kernel() == ""
        cases = [(3, 8, True, 64, 36, 16, "render_pool<8, 1, 896, 1, 0, 1>"),      # 1,024 spheres, >= 8 samples: bundle mode
                 (3, 8, True, 64, 36, 2, "render_pool<8, 1, 896, 1, 0, 0>"),
                 (3, 1, False, 64, 36, 2, "render_pool<1, 0, 896, 1, 0, 0>"),       # RenderTileScalar semantics
                 (0, 8, True, 64, 36, 2, "render_paths<8, 1, 0, 4, 4>"),            # 5 spheres: one path per thread
                 (1, 4, False, 64, 36, 2, "render_pool<4, 0, 896, 1, 0, 0>"),       # 256 spheres, SSE profile
                 (4, 8, True, 32, 18, 8, "render_pool<8, 1, 256, 2, 1, 1>")]        # 65,536 spheres: streamed tables
        for scene, lanes, fma, w, h, spp, want in cases:
            sp, sky, hdr = srt.host_scene(scene)
            b.set_scene(sp, sky, lanes, fma)
            b.resize(w, h)
            b.render(srt.host_camera(scene, w, h, hdr), 0, spp)
            b.wait()
            assert b.last_kernel() == want, (scene, lanes, fma, spp)
        b.render_static(srt.host_camera(4, 32, 18, hdr), 8, 8)                       # no bundle mode without allocations
        b.wait()
        assert b.last_kernel() == "render_pool<8, 1, 256, 2, 1, 0>"
    finally:
        b.close()


def test_one_kernel_scenes_of_different_sizes_in_any_order(srt, oracle):
    """The dynamic shared memory limit of a kernel is a per-device function attribute shared by every handle and every
    scene size: it may only be raised.  Small -> large -> medium -> large on one kernel instantiation (and a second
    handle rendering a small scene in between) used to leave the limit at the medium size: `invalid argument` at the next
    launch of the large scene."""
    spheres, sky, hdr = oracle.scene(2)
    w, h = 48, 27
    cam = oracle.camera(hdr, w, h)
    a, b = srt.Backend(0), srt.Backend(0)
    try:
        for n, handle in ((104, a), (482, a), (296, a), (482, a), (80, b), (482, a), (482, b)):
            sub = spheres[:n].copy()
            handle.set_scene(sub, sky, 4, False)
            handle.resize(w, h)
            handle.render(cam, 0, 2)
            _, rays = handle.wait()
            assert handle.last_kernel() == "render_pool<4, 0, 896, 1, 0, 0>"
            o = oracle.render(sub, sky, 4, 0, cam, w, h, 0, 2)
            assert rays == o["rays"] and np.array_equal(handle.read_accum().view(np.uint32), o["accum"].view(np.uint32)), n
    finally:
        a.close()
        b.close()

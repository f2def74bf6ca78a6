"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle, bit for bit.

Tolerance required by BASELINE.json: RGBA8 max error <= 2, PSNR >= 50 dB on the linear
accumulator, primary hit mask / primitive id bit-exact.  What these tests assert is
stronger: every buffer identical to the last bit (tolerance 0)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MODES = [(8, 1), (4, 0), (1, 0), (8, 0), (4, 1)]   # (lane_width, fma_dot)


def _cmp(o, g):
    bad = {}
    for k in ("accum", "rgba", "prim", "t"):
        a, b = o[k].view(np.uint32), g[k].view(np.uint32)
        if not np.array_equal(a, b):
            bad[k] = int((a != b).sum())
    if o["rays"] != g["rays"]:
        bad["rays"] = (o["rays"], g["rays"])
    return bad


def _gpu_render(backend, spheres, sky, W, fma, cam, w, h, first, n, stream_base=0):
    backend.set_scene(spheres, sky, W, bool(fma))
    if (backend.width, backend.height) != (w, h):
        backend.resize(w, h)
    if first == 0:
        backend.reset()
    backend.render(cam, first, n, record_primary=True, stream_base=stream_base)
    ms, rays = backend.wait()
    prim, t = backend.read_primary()
    return dict(accum=backend.read_accum(), rgba=backend.read_rgba8(), prim=prim, t=t, rays=rays, ms=ms)


@pytest.mark.parametrize("scene", [0, 1, 2])
@pytest.mark.parametrize("mode", MODES)
def test_builtin_scenes_bit_exact(backend, oracle, scene, mode):
    W, fma = mode
    spheres, sky, hdr = oracle.scene(scene)
    w, h, spp = 160, 90, 3
    cam = oracle.camera(hdr, w, h)
    o = oracle.render(spheres, sky, W, fma, cam, w, h, 0, spp)
    g = _gpu_render(backend, spheres, sky, W, fma, cam, w, h, 0, spp)
    assert _cmp(o, g) == {}


@pytest.mark.parametrize("size", [(1, 1), (7, 5), (33, 17), (64, 4), (9, 130)])
def test_ragged_sizes(backend, oracle, size):
    w, h = size
    spheres, sky, hdr = oracle.scene(2)
    cam = oracle.camera(hdr, w, h)
    o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, 2)
    g = _gpu_render(backend, spheres, sky, 8, 1, cam, w, h, 0, 2)
    assert _cmp(o, g) == {}


def test_progressive_equals_batched(backend, oracle):
    """1 spp/frame x N frames (config 4 protocol) == one N-spp launch == oracle."""
    spheres, sky, hdr = oracle.scene(0)
    w, h, n = 96, 54, 5
    cam = oracle.camera(hdr, w, h)
    o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, n)
    backend.set_scene(spheres, sky, 8, True)
    backend.resize(w, h)
    rays = 0
    for k in range(n):
        backend.render(cam, k, 1)
        rays += backend.wait()[1]
    assert np.array_equal(backend.read_accum().view(np.uint32), o["accum"].view(np.uint32))
    assert np.array_equal(backend.read_rgba8(), o["rgba"])
    assert rays == o["rays"]


def test_stream_base_and_first_sample(backend, oracle):
    spheres, sky, hdr = oracle.scene(1)
    w, h = 80, 45
    cam = oracle.camera(hdr, w, h)
    o1 = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, 2, stream_base=123456789)
    o2 = oracle.render(spheres, sky, 8, 1, cam, w, h, 2, 2, stream_base=123456789, accum=o1["accum"].copy())
    g1 = _gpu_render(backend, spheres, sky, 8, 1, cam, w, h, 0, 2, stream_base=123456789)
    assert np.array_equal(g1["accum"].view(np.uint32), o1["accum"].view(np.uint32))
    backend.render(cam, 2, 2, stream_base=123456789)
    backend.wait()
    assert np.array_equal(backend.read_accum().view(np.uint32), o2["accum"].view(np.uint32))
    assert np.array_equal(backend.read_rgba8(), o2["rgba"])


def test_sum_mode_sample_sharding(backend, oracle):
    """Sample-sharded SUM accumulation (the multi-GPU path, emulated on one device):
    2 shards of strided samples summed, resolved -> within tolerance of the exact mean."""
    spheres, sky, hdr = oracle.scene(2)
    w, h, n = 80, 45, 8
    cam = oracle.camera(hdr, w, h)
    o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, n)
    backend.set_scene(spheres, sky, 8, True)
    backend.resize(w, h)
    total_rays = 0
    for shard in range(2):
        backend.render(cam, shard, n // 2, stride=2, mode=1)
        total_rays += backend.wait()[1]
    backend.resolve_sum(n)
    acc = backend.read_accum()
    rgba = backend.read_rgba8()
    assert total_rays == o["rays"]
    mse = float(np.mean((acc[..., :3].astype(np.float64) - o["accum"][..., :3].astype(np.float64)) ** 2))
    peak = float(max(o["accum"][..., :3].max(), 1.0))
    psnr = 10 * np.log10(peak * peak / max(mse, 1e-30))
    assert psnr >= 50.0, psnr
    d = np.abs(rgba.view(np.uint8).astype(int) - o["rgba"].view(np.uint8).astype(int)).max()
    assert d <= 2, d


def test_sum_mode_bit_exact_vs_oracle(backend, oracle):
    """Each shard's f32 sum, the two-shard total and the resolved image equal the oracle's
    restatement of the multi-GPU path bit for bit."""
    spheres, sky, hdr = oracle.scene(1)
    w, h, n = 64, 36, 6
    cam = oracle.camera(hdr, w, h)
    backend.set_scene(spheres, sky, 8, True)
    backend.resize(w, h)
    o_acc = np.zeros((h, w, 4), np.float32)
    for shard in range(2):
        backend.render(cam, shard, n // 2, stride=2, mode=1)
        backend.wait()
        o_acc, _ = oracle.render_sum(spheres, sky, 8, 1, cam, w, h, shard, n // 2, 2, accum=o_acc)
        assert np.array_equal(backend.read_accum()[..., :3].view(np.uint32), o_acc[..., :3].view(np.uint32))
    backend.resolve_sum(n)
    o_rgba = oracle.resolve_sum(o_acc, n)
    assert np.array_equal(backend.read_accum().view(np.uint32), o_acc.view(np.uint32))
    assert np.array_equal(backend.read_rgba8(), o_rgba)

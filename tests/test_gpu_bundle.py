"""GPU: bundle mode of the pool kernel (csrc/srt_bundle.cuh) -- launches of >= 8 samples on scenes of >= 384
spheres (forced here with SRT_BUNDLE=k on smaller scenes the oracle finishes quickly) hand out single
samples as work items, so that a primary warp pass holds the samples of one pixel, runs the level-1 slab
test once per (pixel, sphere) for the whole bundle, and folds the finished samples into the running mean
in sample order from staging rows.

Evidence that none of this changes a bit:
  1. bundle launches equal the CPU oracle (every buffer, ray count) for chunk sizes 4..64, sample counts that
     are not multiples of the chunk, resumed accumulations, 1x1 to ragged images, and bands so small that
     staging rows are reused thousands of times within a launch;
  2. bundle mode on == bundle mode off at a size the oracle would take minutes for, resident and streamed;
  3. the strided SUM shards of the multi-GPU path;
  4. srt_filter_audit: 0 level-0 violations over every (primary ray, sphere) pair, also for cameras that look
     along the slab axis (pixels that must fall back to the per-ray scan) and for translated scenes.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bufs(b, rays):
    prim, t = b.read_primary()
    return dict(accum=b.read_accum(), rgba=b.read_rgba8(), prim=prim, t=t, rays=rays)


def _same(a, b, keys=("accum", "rgba", "prim", "t")):
    bad = {k: int((a[k].view(np.uint32) != b[k].view(np.uint32)).sum()) for k in keys
           if not np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32))}
    if a["rays"] != b["rays"]:
        bad["rays"] = (a["rays"], b["rays"])
    return bad


def _run(srt, env, monkeypatch, spheres, sky, W, fma, cam, w, h, launches):
    """launches: [(first, n)]; a fresh backend so that the SRT_* knobs are read."""
    for k in ("SRT_BUNDLE", "SRT_BUNDLE_BAND_ROWS", "SRT_BUNDLE_STAGE_LIMIT", "SRT_KERNEL_VARIANT", "SRT_POOL_CHUNK"):
        monkeypatch.delenv(k, raising=False)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    b = srt.Backend(0)
    try:
        b.set_scene(spheres, sky, W, bool(fma))
        b.resize(w, h)
        rays = 0
        for first, n in launches:
            b.render(cam, first, n, record_primary=(first == 0))
            rays += b.wait()[1]
        return _bufs(b, rays)
    finally:
        b.close()


@pytest.mark.parametrize("env", [{"SRT_BUNDLE": "6"}, {"SRT_BUNDLE": "2"}, {"SRT_BUNDLE": "3", "SRT_BUNDLE_BAND_ROWS": "1"}, {"SRT_BUNDLE": "5", "SRT_BUNDLE_BAND_ROWS": "2"}],
                         ids=["S64", "S4", "S8_band1row", "S32_band2rows"])
@pytest.mark.parametrize("mode", [(8, 1), (4, 0), (1, 0)])
def test_bundle_equals_oracle(srt, oracle, monkeypatch, env, mode):
    W, fma = mode
    spheres, sky, hdr = srt.host_scene(2)
    for (w, h, launches) in (((1, 1, [(0, 70)])), (5, 3, [(0, 24)]), (33, 17, [(0, 3), (3, 9)]), (64, 36, [(0, 13)]), (9, 50, [(0, 8), (8, 8)])):
        cam = srt.host_camera(2, w, h, hdr)
        total = sum(n for _, n in launches)
        o = oracle.render(spheres, sky, W, fma, cam, w, h, 0, total)
        g = _run(srt, env, monkeypatch, spheres, sky, W, fma, cam, w, h, launches)
        assert _same(o, g) == {}, (w, h, launches)


@pytest.mark.parametrize("scene,variant,size", [(3, "0", (240, 135, 16)), (1, "0", (160, 90, 40)), (2, "12", (96, 54, 20)), (4, "0", (48, 27, 8))],
                         ids=["synth1k", "scene1", "scene2_streamed", "synth64k"])
def test_bundle_on_equals_bundle_off(srt, monkeypatch, scene, variant, size):
    w, h, spp = size
    spheres, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h, hdr)
    on = _run(srt, {"SRT_KERNEL_VARIANT": variant, "SRT_BUNDLE": "6"}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, spp)])
    off = _run(srt, {"SRT_KERNEL_VARIANT": variant, "SRT_BUNDLE": "0"}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, spp)])
    assert _same(off, on) == {}
    small = _run(srt, {"SRT_KERNEL_VARIANT": variant, "SRT_BUNDLE": "3", "SRT_BUNDLE_BAND_ROWS": "3"}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, spp)])
    assert _same(off, small) == {}


@pytest.mark.parametrize("limit", ["200000", "40000", "100"], ids=["half_bands", "one_block_row", "no_bundle"])
def test_bundle_short_of_memory(srt, oracle, monkeypatch, limit):
    """The staging rows do not fit (simulated with SRT_BUNDLE_STAGE_LIMIT): srt_render halves the band until they do
    and, when not even one row of blocks fits, runs the launch without bundle mode.  Same bits every time."""
    spheres, sky, hdr = srt.host_scene(2)
    w, h, spp = 64, 36, 12
    cam = srt.host_camera(2, w, h, hdr)
    o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, spp)
    g = _run(srt, {"SRT_BUNDLE": "6", "SRT_BUNDLE_STAGE_LIMIT": limit}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, spp)])
    monkeypatch.delenv("SRT_BUNDLE_STAGE_LIMIT", raising=False)
    assert _same(o, g) == {}


def test_bundle_sum_shards(srt, oracle, monkeypatch):
    """The multi-GPU path: rank g adds samples k = g (mod N) to a plain sum, in k order."""
    for k in ("SRT_BUNDLE_BAND_ROWS", "SRT_KERNEL_VARIANT", "SRT_POOL_CHUNK"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("SRT_BUNDLE", "6")
    spheres, sky, hdr = srt.host_scene(2)
    w, h, n, shards = 40, 30, 36, 3
    cam = srt.host_camera(2, w, h, hdr)
    b = srt.Backend(0)
    try:
        b.set_scene(spheres, sky, 8, True)
        b.resize(w, h)
        o_acc = np.zeros((h, w, 4), np.float32)
        for shard in range(shards):
            b.render(cam, shard, n // shards, stride=shards, mode=1)
            b.wait()
            o_acc, _ = oracle.render_sum(spheres, sky, 8, 1, cam, w, h, shard, n // shards, shards, accum=o_acc)
            assert np.array_equal(b.read_accum()[..., :3].view(np.uint32), o_acc[..., :3].view(np.uint32)), shard
    finally:
        b.close()


def _look_camera(srt, pos, target, w, h, film_w=1.0):
    cam = np.zeros(1, srt.camera_dtype)
    pos = np.asarray(pos, np.float64); target = np.asarray(target, np.float64)
    z = pos - target; z /= np.linalg.norm(z)
    up = np.array([0.0, 1.0, 0.0]) if abs(z[1]) < 0.99 else np.array([1.0, 0.0, 0.0])
    x = np.cross(up, z); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    cam["position"][0, :3] = pos; cam["z"][0, :3] = z; cam["x"][0, :3] = x; cam["y"][0, :3] = y
    cam["film_center"][0, :3] = (pos - z).astype(np.float32)
    cam["film_w"][0] = film_w; cam["film_h"][0] = film_w * h / w
    return cam


@pytest.mark.parametrize("axis", ["0", "1", "2"])
def test_bundle_camera_along_every_axis(srt, oracle, monkeypatch, axis):
    """A camera that looks straight down one coordinate axis: for the launch whose slab axis is that axis the
    central pixels cannot be bundled (their rays are parallel to the slab planes' axis) and must take the
    per-ray scan; the image equals the oracle's for every forced slab axis, and the audit stays clean."""
    spheres, sky, hdr = srt.host_scene(2)
    w, h, spp = 48, 27, 9
    for pos in ((0.3, 9.0, 0.2), (9.0, 0.4, 0.1), (0.2, 0.5, 9.0)):
        cam = _look_camera(srt, pos, (0.0, 0.3, 0.0), w, h)
        o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, spp)
        g = _run(srt, {"SRT_SLAB_AXIS": axis, "SRT_BUNDLE": "6"}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, spp)])
        assert _same(o, g) == {}, pos
    monkeypatch.delenv("SRT_SLAB_AXIS", raising=False)


@pytest.mark.parametrize("scene", [1, 2, 3])
def test_bundle_audit(backend, srt, scene):
    spheres, sky, hdr = srt.host_scene(scene)
    w, h = (96, 54) if scene == 3 else (160, 90)
    cam = srt.host_camera(scene, w, h, hdr)
    backend.set_scene(spheres, sky, 8, True)
    backend.resize(w, h)
    a = backend.filter_audit(cam, 0, 4)
    assert a["bundle_tests"] > 0 and a["bundle_violations"] == 0 and a["violations"] == 0, a
    # level 0 is about as selective as the per-ray level 1 it replaces
    assert a["bundle_candidates"] * 20 < a["bundle_tests"], a
    # translated far from the origin (rounding of the film point grows with |film centre|)
    s2, c2 = spheres.copy(), cam.copy()
    for off in (700.0, -30000.0):
        s2["position"][:, :3] = spheres["position"][:, :3] + np.float32(off)
        c2["position"][0, :3] = cam["position"][0, :3] + np.float32(off)
        c2["film_center"][0, :3] = cam["film_center"][0, :3] + np.float32(off)
        backend.set_scene(s2, sky, 8, True)
        a = backend.filter_audit(c2, 0, 2)
        assert a["bundle_violations"] == 0 and a["violations"] == 0, (off, a)


def test_bundle_fuzz_against_non_bundle(srt, monkeypatch):
    """Random image sizes, sample counts, resumed launches, chunk sizes and band sizes on the 1,024-sphere scene:
    every buffer of a bundle-mode accumulation equals the non-bundle kernel's (which the other tests pin to the
    oracle).  Exercises ragged block edges, short last bands, short last chunks and staging-row reuse together."""
    rng = np.random.default_rng(20261105)
    spheres, sky, hdr = srt.host_scene(3)
    for case in range(24):
        w, h = int(rng.integers(1, 150)), int(rng.integers(1, 90))
        n1, n2 = int(rng.integers(8, 70)), int(rng.integers(0, 20))
        launches = [(0, n1)] + ([(n1, n2)] if n2 else [])
        env = {"SRT_BUNDLE": str(int(rng.integers(2, 7)))}
        if rng.integers(0, 2):
            env["SRT_BUNDLE_BAND_ROWS"] = str(int(rng.integers(1, 6)))
        cam = srt.host_camera(3, w, h, hdr)
        on = _run(srt, env, monkeypatch, spheres, sky, 8, 1, cam, w, h, launches)
        off = _run(srt, {"SRT_BUNDLE": "0"}, monkeypatch, spheres, sky, 8, 1, cam, w, h, launches)
        assert _same(off, on) == {}, (case, w, h, launches, env)


def test_bundle_lanes_past_the_table_end(srt, oracle, monkeypatch):
    """n_pad % 128 != 0 (scene 2: 482 spheres -> 488): in the last 128-sphere round of the level-0 scan the lanes past
    the table end used to test a dummy {0, 0, r' = -1} record, and |nou| <= wl * eps - 1 HOLDS for a far camera that
    looks at the origin through a coarse image (eps ~ 0.05) -- queueing sphere indices >= n_pad.  Such lanes now
    contribute nothing and padding radii are -1e30.  Far cameras, tiny images, every slab axis: bundle mode equals the
    CPU oracle and the non-bundle kernel."""
    spheres, sky, hdr = srt.host_scene(2)
    assert len(spheres) % 128 != 0 and ((len(spheres) + 7) // 8 * 8) % 128 != 0
    for axis in ("0", "1", "2"):
        for (pos, w, h) in (((60.0, 25.0, 40.0), 16, 16), ((0.0, 100.0, 3.0), 24, 12), ((-35.0, 2.0, -80.0), 64, 36), ((20.0, 20.0, 20.0), 8, 8)):
            cam = _look_camera(srt, pos, (0.0, 0.0, 0.0), w, h, film_w=1.0)
            on = _run(srt, {"SRT_BUNDLE": "6", "SRT_SLAB_AXIS": axis}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, 64)])
            off = _run(srt, {"SRT_BUNDLE": "0", "SRT_SLAB_AXIS": axis}, monkeypatch, spheres, sky, 8, 1, cam, w, h, [(0, 64)])
            assert _same(off, on) == {}, (axis, pos, w, h)
        o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, 64)
        assert _same(o, on) == {}, axis
    monkeypatch.delenv("SRT_SLAB_AXIS", raising=False)

"""GPU: the ray-pair kernel (csrc/srt_pairs.cuh) -- a conservative fused-arithmetic sphere
filter followed by the reference's exact arithmetic on the survivors.

Three kinds of evidence that the filter never changes a result:
  1. srt_filter_audit: for every (path segment, sphere) pair of whole renders, the reference's
     hit predicate implies the filter predicate (violations == 0), on the built-in scenes, the
     synthetic 1,024-sphere scene and on adversarial scenes (large coordinates, tiny spheres,
     far camera).
  2. The ray-pair kernel and the sphere-pair kernel (literal reference arithmetic for every
     sphere) give bit-identical buffers at sizes the CPU oracle would take too long for.
  3. Scenes built to trigger the rare paths (sticky InsideSphere replay, exact t ties, rays
     starting inside several spheres) match the CPU oracle bit for bit in every mode.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MODES = [(8, 1), (4, 0), (1, 0)]   # (lane_width, fma_dot)


def _render(backend, spheres, sky, W, fma, cam, w, h, n, first=0):
    backend.set_scene(spheres, sky, W, bool(fma))
    backend.resize(w, h)
    backend.render(cam, first, n, record_primary=True)
    ms, rays = backend.wait()
    prim, t = backend.read_primary()
    return dict(accum=backend.read_accum(), rgba=backend.read_rgba8(), prim=prim, t=t, rays=rays)


def _same(a, b):
    bad = {k: int((a[k].view(np.uint32) != b[k].view(np.uint32)).sum()) for k in ("accum", "rgba", "prim", "t")
           if not np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32))}
    if a["rays"] != b["rays"]:
        bad["rays"] = (a["rays"], b["rays"])
    return bad


def _mk(srt, rows):
    """rows: (x, y, z, radius, (r,g,b), (er,eg,eb), specular, ior)"""
    sp = np.zeros(len(rows), srt.sphere_dtype)
    for i, (x, y, z, r, col, em, spec, ior) in enumerate(rows):
        sp["position"][i, :3] = (x, y, z)
        sp["radius"][i] = r
        sp["color"][i, :3] = col
        sp["emissive"][i, :3] = em
        sp["specular"][i] = spec
        sp["ior"][i] = ior
    return sp


@pytest.mark.parametrize("scene", [1, 2, 3])
@pytest.mark.parametrize("mode", MODES)
def test_filter_audit_scenes(backend, srt, scene, mode):
    W, fma = mode
    spheres, sky, hdr = srt.host_scene(scene)
    w, h = (96, 54) if scene == 3 else (160, 90)
    cam = srt.host_camera(scene, w, h, hdr)
    backend.set_scene(spheres, sky, W, bool(fma))
    backend.resize(w, h)
    a = backend.filter_audit(cam, 0, 2)
    assert a["tests"] > 0 and a["hits"] > 0
    assert a["violations"] == 0 and a["bundle_violations"] == 0, a
    assert a["candidates"] >= a["hits"]
    assert a["candidates"] < 4 * a["hits"] + a["tests"] // 500, a   # the filter stays selective


def _shifted(spheres, cam, off):
    s, c = spheres.copy(), cam.copy()
    s["position"][:, :3] += np.float32(off)
    c["position"][0, :3] += np.float32(off)
    c["film_center"][0, :3] += np.float32(off)
    return s, c


@pytest.mark.parametrize("off", [300.0, 5000.0, -40000.0])
def test_filter_audit_large_coordinates(backend, srt, oracle, off):
    """The filter's error terms grow with |C|^2 + |O|^2: translate a scene far from the origin.
    It must stay sound (only less selective) and the image must still equal the oracle's."""
    spheres, sky, hdr = srt.host_scene(1)
    w, h = 64, 36
    s2, c2 = _shifted(spheres, srt.host_camera(1, w, h, hdr), off)
    backend.set_scene(s2, sky, 8, True)
    backend.resize(w, h)
    a = backend.filter_audit(c2, 0, 2)
    assert a["violations"] == 0 and a["bundle_violations"] == 0, a
    o = oracle.render(s2, sky, 8, 1, c2, w, h, 0, 2)
    g = _render(backend, s2, sky, 8, 1, c2, w, h, 2)
    assert _same(o, g) == {}


def test_filter_audit_tiny_far_spheres(backend, srt, oracle):
    """Small spheres seen from far away: r^2 is orders of magnitude below the rounding noise of
    |C|^2 and |O|^2, the regime where a badly bounded filter would drop grazing hits."""
    rng = np.random.default_rng(7)
    rows = [(0.0, -1000.0, 0.0, 1000.0, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 0.0)]
    for i in range(199):
        x, z = rng.uniform(-40, 40, 2)
        r = float(rng.uniform(0.01, 0.3))
        kind = i % 4
        rows.append((float(x), r, float(z), r, tuple(rng.uniform(0.2, 1, 3)), (2, 2, 2) if kind == 0 else (0, 0, 0),
                     1.0 if kind == 1 else 0.0, 1.5 if kind == 2 else 0.0))
    sp = _mk(srt, rows)
    w, h = 96, 54
    cam = np.zeros(1, srt.camera_dtype)
    # camera 900 units away looking at the cluster through a narrow film
    pos = np.array([0.0, 300.0, 850.0], np.float32)
    zdir = pos / np.linalg.norm(pos)
    xdir = np.cross([0, 1, 0], zdir); xdir /= np.linalg.norm(xdir)
    ydir = np.cross(zdir, xdir)
    cam["position"][0, :3] = pos; cam["z"][0, :3] = zdir; cam["x"][0, :3] = xdir; cam["y"][0, :3] = ydir
    cam["film_center"][0, :3] = pos - zdir.astype(np.float32)
    cam["film_w"][0] = 0.1; cam["film_h"][0] = 0.1 * h / w
    for W, fma in MODES:
        backend.set_scene(sp, True, W, bool(fma))
        backend.resize(w, h)
        a = backend.filter_audit(cam, 0, 2)
        assert a["hits"] > 0 and a["violations"] == 0 and a["bundle_violations"] == 0, (W, fma, a)
        o = oracle.render(sp, True, W, fma, cam, w, h, 0, 2)
        g = _render(backend, sp, True, W, fma, cam, w, h, 2)
        assert _same(o, g) == {}, (W, fma)


@pytest.mark.parametrize("mode", MODES)
def test_pair_kernel_equals_sphere_pair_kernel(backend, srt, monkeypatch, mode):
    """1,024-sphere scene, 480x270, 4 spp, all bounces: the production pool kernel (filter +
    select, CTA-owned slots, primary / secondary passes), resident and streamed, and the kernel
    that runs the reference's arithmetic on every sphere agree on every bit."""
    W, fma = mode
    spheres, sky, hdr = srt.host_scene(3)
    w, h = 480, 270
    cam = srt.host_camera(3, w, h, hdr)
    monkeypatch.setenv("SRT_KERNEL_VARIANT", "7")     # sphere-pair / scalar kernel, one path per thread
    ref = _render(backend, spheres, sky, W, fma, cam, w, h, 4)
    monkeypatch.setenv("SRT_KERNEL_VARIANT", "9")     # pool kernel (forced)
    pool = _render(backend, spheres, sky, W, fma, cam, w, h, 4)
    monkeypatch.setenv("SRT_KERNEL_VARIANT", "12")    # pool kernel, tables streamed through per-warp TMA stages
    streamed = _render(backend, spheres, sky, W, fma, cam, w, h, 4)
    monkeypatch.delenv("SRT_KERNEL_VARIANT")
    auto = _render(backend, spheres, sky, W, fma, cam, w, h, 4)
    assert _same(ref, pool) == {}
    assert _same(ref, streamed) == {}
    assert _same(ref, auto) == {}
    assert ref["rays"] > w * h * 4


def test_pool_kernel_progressive_and_tiny_images(backend, srt, oracle):
    """Pool kernel corner cases: fewer pixels than path slots (most slots retire at once),
    progressive 1-spp launches that resume the accumulator, ragged sizes."""
    spheres, sky, hdr = srt.host_scene(2)
    for (w, h) in ((1, 1), (5, 3), (33, 17), (70, 9)):
        cam = srt.host_camera(2, w, h, hdr)
        o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, 3)
        backend.set_scene(spheres, sky, 8, True)
        backend.resize(w, h)
        rays = 0
        for k in range(3):
            backend.render(cam, k, 1, record_primary=(k == 0))
            rays += backend.wait()[1]
        assert np.array_equal(backend.read_accum().view(np.uint32), o["accum"].view(np.uint32)), (w, h)
        assert np.array_equal(backend.read_rgba8(), o["rgba"]), (w, h)
        assert rays == o["rays"]


@pytest.mark.parametrize("mode", MODES)
def test_streamed_64k_scene_vs_oracle(backend, srt, oracle, mode):
    """BASELINE configs[4]: 65,536 spheres (3 x 1 MiB of tables, not shared-memory resident):
    the streamed pool kernel equals the CPU oracle bit for bit, and the filter audit is clean."""
    W, fma = mode
    spheres, sky, hdr = srt.host_scene(4)
    assert len(spheres) > 65000
    w, h = 64, 36
    cam = srt.host_camera(4, w, h, hdr)
    o = oracle.render(spheres, sky, W, fma, cam, w, h, 0, 1)
    g = _render(backend, spheres, sky, W, fma, cam, w, h, 1)
    assert _same(o, g) == {}
    backend.set_scene(spheres, sky, W, bool(fma))
    backend.resize(16, 9)
    a = backend.filter_audit(srt.host_camera(4, 16, 9, hdr), 0, 1)
    assert a["hits"] > 0 and a["violations"] == 0 and a["bundle_violations"] == 0, a


def test_streamed_ragged_scene_sizes(backend, srt, oracle, monkeypatch):
    """Streaming corner cases: table sizes that are not multiples of the 64-sphere stage or of
    the 32-sphere mask word (last chunk of 8, 16, 24, 40 ... spheres)."""
    spheres, sky, hdr = srt.host_scene(2)
    w, h = 48, 27
    cam = srt.host_camera(2, w, h, hdr)
    monkeypatch.setenv("SRT_KERNEL_VARIANT", "12")
    for n in (65, 72, 100, 129, 200, 257, 482):
        sp = spheres[:n]
        o = oracle.render(sp, sky, 8, 1, cam, w, h, 0, 2)
        g = _render(backend, sp, sky, 8, 1, cam, w, h, 2)
        assert _same(o, g) == {}, n


@pytest.mark.parametrize("mode", [(8, 1), (4, 0), (1, 0), (8, 0), (4, 1)])
def test_nested_and_tied_spheres(backend, srt, oracle, mode):
    """Rare paths of the select step: rays that start inside several nested glass spheres whose
    indices share a SIMD lane (sticky InsideSphere, main.cpp:425 -> replay_inside), duplicated
    spheres (exact t ties across lanes and groups), and a sphere enclosing the camera."""
    W, fma = mode
    glass = ((1, 1, 1), (0, 0, 0), 0.0, 1.5)
    rows = [(0.0, -100.5, -1.0, 100.0, (0.8, 0.8, 0.0), (0, 0, 0), 0.0, 0.0)]          # 0 ground
    rows += [(0.0, 0.0, -1.0, 0.9, *glass)]                                             # 1 outer glass
    rows += [(50.0 + i, 50.0, 50.0, 0.1, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 0.0) for i in range(7)]   # 2..8 filler
    rows += [(0.0, 0.0, -1.0, 0.5, *glass)]                                             # 9: same lane as 1 (W=8 and 4)
    rows += [(0.2, 0.1, -1.0, 0.2, (0.9, 0.2, 0.2), (1, 1, 1), 0.0, 0.0)]               # 10 emissive core
    rows += [(60.0 + i, 50.0, 50.0, 0.1, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 0.0) for i in range(6)]   # 11..16 filler
    rows += [(0.0, 0.0, -1.0, 0.3, *glass)]                                             # 17: lane of 1 and 9 again
    rows += [(-1.4, 0.0, -1.0, 0.5, (0.8, 0.6, 0.2), (0, 0, 0), 1.0, 0.0)] * 3          # 18..20 exact duplicates (ties)
    rows += [(1.4, 0.0, -1.0, 0.5, (0.1, 0.2, 0.5), (0, 0, 0), 0.0, 0.0)] * 2           # 21..22 duplicates
    rows += [(0.0, 0.0, 0.0, 30.0, *glass)]                                             # 23 encloses the camera
    rows += [(70.0 + i, 50.0, 50.0, 0.1, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 0.0) for i in range(48)]  # pad to >= 64 spheres
    sp = _mk(srt, rows)
    w, h = 96, 54
    cam = np.zeros(1, srt.camera_dtype)
    cam["position"][0, :3] = (0, 0, 1.5); cam["z"][0, :3] = (0, 0, 1); cam["x"][0, :3] = (1, 0, 0); cam["y"][0, :3] = (0, 1, 0)
    cam["film_center"][0, :3] = (0, 0, 0.5); cam["film_w"][0] = 1.0; cam["film_h"][0] = h / w
    o = oracle.render(sp, True, W, fma, cam, w, h, 0, 3)
    g = _render(backend, sp, True, W, fma, cam, w, h, 3)
    assert _same(o, g) == {}
    backend.set_scene(sp, True, W, bool(fma))
    a = backend.filter_audit(cam, 0, 2)
    assert a["violations"] == 0 and a["bundle_violations"] == 0, a


@pytest.mark.parametrize("chunk", ["", "1", "3", "1000"])
def test_pool_sample_chunks_fold_in_order(srt, oracle, monkeypatch, chunk):
    """Pool work items are (sample chunk, pixel); a pixel's chunks must fold into the running mean
    in sample order even when far more path slots than pixels exist (every slot of the grid then
    holds a later chunk of some pixel and waits for the preceding one)."""
    if chunk:
        monkeypatch.setenv("SRT_POOL_CHUNK", chunk)
    else:
        monkeypatch.delenv("SRT_POOL_CHUNK", raising=False)
    spheres, sky, hdr = srt.host_scene(2)
    b = srt.Backend(0)
    try:
        for (w, h, spp, first) in ((1, 1, 40, 0), (5, 3, 24, 0), (33, 17, 9, 2), (160, 90, 7, 0)):
            cam = srt.host_camera(2, w, h, hdr)
            o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, first + spp)
            b.set_scene(spheres, sky, 8, True)
            b.resize(w, h)
            rays = 0
            if first:
                b.render(cam, 0, first)
                rays += b.wait()[1]
            b.render(cam, first, spp)
            rays += b.wait()[1]
            assert np.array_equal(b.read_accum().view(np.uint32), o["accum"].view(np.uint32)), (w, h)
            assert np.array_equal(b.read_rgba8(), o["rgba"]) and rays == o["rays"], (w, h)
        # strided SUM shards (the multi-GPU path) chunked the same way
        w, h, n = 40, 30, 10
        cam = srt.host_camera(2, w, h, hdr)
        b.set_scene(spheres, sky, 8, True)
        b.resize(w, h)
        o_acc = np.zeros((h, w, 4), np.float32)
        for shard in range(2):
            b.render(cam, shard, n // 2, stride=2, mode=1)
            b.wait()
            o_acc, _ = oracle.render_sum(spheres, sky, 8, 1, cam, w, h, shard, n // 2, 2, accum=o_acc)
            assert np.array_equal(b.read_accum()[..., :3].view(np.uint32), o_acc[..., :3].view(np.uint32))
    finally:
        b.close()


@pytest.mark.parametrize("axis", ["0", "1", "2"])
def test_slab_axis_never_changes_results(srt, oracle, monkeypatch, axis):
    """Level 1 (slab test): the slab axis affects selectivity only.  Every axis gives the oracle's
    bits on a resident and on a streamed (variant 12) run, and a clean audit."""
    monkeypatch.setenv("SRT_SLAB_AXIS", axis)
    spheres, sky, hdr = srt.host_scene(2)
    w, h = 96, 54
    cam = srt.host_camera(2, w, h, hdr)
    o = oracle.render(spheres, sky, 8, 1, cam, w, h, 0, 2)
    for variant in ("0", "12"):
        monkeypatch.setenv("SRT_KERNEL_VARIANT", variant)
        b = srt.Backend(0)
        try:
            assert _same(o, _render(b, spheres, sky, 8, 1, cam, w, h, 2)) == {}, (axis, variant)
            a = b.filter_audit(cam, 0, 1)
            assert a["violations"] == 0 and a["slab_survivors"] >= a["candidates"] >= a["hits"], a
        finally:
            b.close()


@pytest.mark.parametrize("mode", MODES)
def test_slab_degenerate_rays_and_far_camera(backend, srt, oracle, mode):
    """Rays parallel to the slab axis (U undefined -> every sphere is a candidate), a column of
    spheres along that axis, axis-aligned view directions, and a camera far outside the scene's
    origin bound (the slab radii are inflated for the camera distance)."""
    W, fma = mode
    diff = ((0.7, 0.3, 0.3), (0, 0, 0), 0.0, 0.0)
    glass = ((1, 1, 1), (0, 0, 0), 0.0, 1.5)
    lamp = ((0.9, 0.9, 0.9), (2, 2, 2), 0.0, 0.0)
    # a column along y (centres spread along y only -> the slab axis is x or z), mirrors and glass in it
    rows = [(0.0, -1000.5, 0.0, 1000.0, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 0.0)]
    rows += [(0.0, 0.6 * i, 0.0, 0.25, *(glass if i % 3 == 0 else (lamp if i % 3 == 1 else diff))) for i in range(70)]
    sp = _mk(srt, rows)
    w, h = 64, 48
    views = []
    for pos, z in (((0.0, 60.0, 0.0), (0, 1, 0)),         # straight down the column: primary rays ~ parallel to y
                   ((6.0, 20.0, 0.0), (1, 0, 0)),          # along -x
                   ((0.0, 20.0, 6.0), (0, 0, 1)),          # along -z
                   ((0.0, 20.0, 9000.0), (0, 0, 1))):      # far camera: |O| >> scene bound
        cam = np.zeros(1, srt.camera_dtype)
        zz = np.array(z, np.float32)
        xx = np.array((1, 0, 0) if z[0] == 0 else (0, 0, -1), np.float32)
        yy = np.cross(zz, xx).astype(np.float32)
        cam["position"][0, :3] = pos; cam["z"][0, :3] = zz; cam["x"][0, :3] = xx; cam["y"][0, :3] = yy
        cam["film_center"][0, :3] = np.array(pos, np.float32) - zz
        cam["film_w"][0] = 1.0; cam["film_h"][0] = h / w
        views.append(cam)
    for cam in views:
        o = oracle.render(sp, True, W, fma, cam, w, h, 0, 2)
        g = _render(backend, sp, True, W, fma, cam, w, h, 2)
        assert _same(o, g) == {}
        a = backend.filter_audit(cam, 0, 1)
        assert a["violations"] == 0 and a["bundle_violations"] == 0, a


def test_full_size_4k_properties(srt, monkeypatch):
    """BASELINE configs[2] at its full image size (3840x2160, 1,024 spheres), too large for the CPU
    oracle: size-independent properties instead.  (1) the production kernel equals the kernel that
    runs the reference's arithmetic on every sphere, bit for bit, on all 8.3 M pixels; (2) two
    launches of 32 spp equal one launch of 64 spp; (3) a resumed accumulator (srt_write_accum) continues
    identically; (4) 4 strided SUM shards resolve to within 1 LSB / 60 dB of the exact running mean.
    64 spp: the production launch is render_pool in BUNDLE mode (64-sample chunks, two staging bands at 4K) -- the
    kernel bench.py times; tests/test_gpu_fullsize.py pins the 256-spp launch itself to the reference."""
    w, h, spp = 3840, 2160, 64
    spheres, sky, hdr = srt.host_scene(3)
    cam = srt.host_camera(3, w, h, hdr)
    b = srt.Backend(0)
    try:
        monkeypatch.setenv("SRT_KERNEL_VARIANT", "7")
        ref = _render(b, spheres, sky, 8, 1, cam, w, h, spp)
        monkeypatch.delenv("SRT_KERNEL_VARIANT")
        auto = _render(b, spheres, sky, 8, 1, cam, w, h, spp)
        assert _same(ref, auto) == {}
        assert ref["rays"] > w * h * spp
        b.reset()
        b.render(cam, 0, 32); r1 = b.wait()[1]
        half = b.read_accum()
        b.render(cam, 32, 32); r2 = b.wait()[1]
        assert np.array_equal(b.read_accum().view(np.uint32), ref["accum"].view(np.uint32)) and r1 + r2 == ref["rays"]
        b.reset()
        b.write_accum(half)
        b.render(cam, 32, 32); b.wait()
        assert np.array_equal(b.read_accum().view(np.uint32), ref["accum"].view(np.uint32))
        assert np.array_equal(b.read_rgba8(), ref["rgba"])
        b.reset()
        rays = 0
        for shard in range(4):
            b.render(cam, shard, 16, stride=4, mode=1)
            rays += b.wait()[1]
        b.resolve_sum(spp)
        assert rays == ref["rays"]
        acc = b.read_accum()[..., :3].astype(np.float64)
        want = ref["accum"][..., :3].astype(np.float64)
        mse = float(np.mean((acc - want) ** 2))
        peak = float(max(want.max(), 1.0))
        assert 10 * np.log10(peak * peak / max(mse, 1e-30)) >= 60.0
        d = np.abs(b.read_rgba8().view(np.uint8).astype(np.int16) - ref["rgba"].view(np.uint8).astype(np.int16)).max()
        assert d <= 1, d
    finally:
        b.close()

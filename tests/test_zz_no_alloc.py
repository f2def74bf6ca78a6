"""GPU: srt_render_static -- the render entry point that allocates nothing (its scratch, the pool kernel's per-pixel chunk
counters, is a file-scope device array) -- against srt_render, through the C ABI, bit for bit (the tolerance of every
image test in this repository), on two calls in a row.  Collected last on purpose (file name),
and this entry point was added outside SURVEY.md section 8."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bits(b):
    return b.read_accum().view(np.uint32).copy(), b.read_rgba8().copy()


@pytest.mark.parametrize("scene,w,h,n1,n2", [(3, 320, 180, 6, 5),      # pool kernel, several chunks: the static counters are read
                                             (3, 200, 120, 12, 9),     # srt_render takes bundle mode here, the static variant the plain pool kernel
                                             (0, 256, 144, 3, 2),      # 5 spheres: one path per thread, no scratch at all
                                             (4, 96, 54, 2, 2)])       # 65,536 spheres: streamed tables
def test_static_render_equals_render_on_two_calls_in_a_row(srt, scene, w, h, n1, n2):
    spheres, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h, hdr)
    ref = srt.Backend(0)
    sta = srt.Backend(0)
    try:
        for b in (ref, sta):
            b.set_scene(spheres, sky, 8, True)
            b.resize(w, h)
        # call 1: a fresh accumulation
        ref.render(cam, 0, n1)
        _, rays_ref = ref.wait()
        sta.render_static(cam, 0, n1)
        _, rays_sta = sta.wait()
        a0, r0 = _bits(ref)
        a1, r1 = _bits(sta)
        assert rays_ref == rays_sta and np.array_equal(a0, a1) and np.array_equal(r0, r1)
        # call 2, straight after: continues the accumulation; the shared scratch still holds call 1's counters
        ref.render(cam, n1, n2)
        _, rays_ref = ref.wait()
        sta.render_static(cam, n1, n2)
        _, rays_sta = sta.wait()
        a0, r0 = _bits(ref)
        a1, r1 = _bits(sta)
        assert rays_ref == rays_sta and np.array_equal(a0, a1) and np.array_equal(r0, r1)
    finally:
        ref.close()
        sta.close()


def test_static_render_shares_its_scratch_between_handles(srt):
    """Two handles take turns on the one array (serialised: each call is waited for); stale counters of the other handle,
    written with the same launch tags, must not leak into a launch."""
    spheres, sky, hdr = srt.host_scene(3)
    w, h = 160, 90
    cam = srt.host_camera(3, w, h, hdr)
    ref = srt.Backend(0)
    ref.set_scene(spheres, sky, 8, True)
    ref.resize(w, h)
    ref.render(cam, 0, 7)
    ref.wait()
    want_a, want_r = _bits(ref)
    ref.close()
    hs = [srt.Backend(0) for _ in range(2)]
    try:
        for b in hs:
            b.set_scene(spheres, sky, 8, True)
            b.resize(w, h)
        for k in (0, 3):                       # both handles issue launch tags 0, 1 on the shared array
            for b in hs:
                b.render_static(cam, k, 3 if k == 0 else 4)
                b.wait()
        for b in hs:
            a, r = _bits(b)
            assert np.array_equal(a, want_a) and np.array_equal(r, want_r)
    finally:
        for b in hs:
            b.close()


def test_static_render_limits_and_no_device_allocation(srt):
    import torch
    spheres, sky, hdr = srt.host_scene(3)
    b = srt.Backend(0)
    try:
        b.set_scene(spheres, sky, 8, True)
        b.resize(2048, 1025)                                           # one row more than SRT_STATIC_MAX_PIXELS
        assert 2048 * 1025 > srt.STATIC_MAX_PIXELS
        with pytest.raises(srt.SrtError, match="exceeds the static scratch"):
            b.render_static(srt.host_camera(3, 2048, 1025, hdr), 0, 2)
        w, h = 1920, 1080                                               # the largest BASELINE image that fits
        assert w * h <= srt.STATIC_MAX_PIXELS
        b.resize(w, h)
        cam = srt.host_camera(3, w, h, hdr)
        b.render_static(cam, 0, 2)                                      # first use loads the kernel module
        b.wait()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        b.render_static(cam, 2, 3)
        _, rays = b.wait()
        assert torch.cuda.mem_get_info(0)[0] == free0 and rays > w * h * 3   # the device's free memory did not move
        acc, rgba = _bits(b)
    finally:
        b.close()
    ref = srt.Backend(0)
    try:
        ref.set_scene(spheres, sky, 8, True)
        ref.resize(w, h)
        ref.render(cam, 0, 5)
        ref.wait()
        a, r = _bits(ref)
        assert np.array_equal(a, acc) and np.array_equal(r, rgba)
    finally:
        ref.close()

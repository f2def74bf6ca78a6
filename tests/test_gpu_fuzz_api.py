"""GPU: random call sequences through the C ABI on two handles of one device -- scene uploads of different sizes and
arithmetic profiles, resizes, resets, renders of 1..24 samples (one-path kernel, pool kernel with and without bundle mode),
srt_render_static, accumulator read-back / write-back -- every render compared with the CPU oracle bit for bit.
State that leaks between calls, scenes or handles (launch caches, done-tags, fold weights, counters re-armed by the last
CTA, function attributes shared by the handles of a device) shows up here."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MODES = [(8, 1), (4, 0), (1, 0)]


class Model:
    """What one handle should hold, replayed on the oracle."""

    def __init__(self, srt, oracle, rng):
        self.srt, self.oracle, self.rng = srt, oracle, rng
        self.b = srt.Backend(0)
        self.k = 0
        self.rays_total = 0
        self.scene = None
        self.size = None

    def close(self):
        self.b.close()

    def new_scene(self, pools):
        src, sky = pools[int(self.rng.integers(len(pools)))]
        n = int(self.rng.choice([3, 5, 17, 40, 64, 100, 256, 300, 482, 700, 1024, 1024, 8200]))   # 8,200: tables streamed through TMA stages
        n = min(n, len(src))
        self.scene = (src[:n].copy(), sky)
        self.mode = MODES[int(self.rng.integers(3))]
        self.b.set_scene(self.scene[0], sky, self.mode[0], bool(self.mode[1]))
        if self.size:
            if n > 2000:
                self.resize()                 # keep the oracle's cost bounded for the 8,200-sphere table
            else:
                self.b.reset()
        self.k, self.rays_total = 0, 0

    def resize(self):
        big = len(self.scene[0]) > 2000
        self.size = (int(self.rng.integers(1, 24 if big else 70)), int(self.rng.integers(1, 16 if big else 40)))
        self.b.resize(*self.size)
        self.k, self.rays_total = 0, 0

    def reset(self):
        self.b.reset()
        self.k, self.rays_total = 0, 0

    def render(self, cam_of, static):
        n = int(self.rng.choice([1, 1, 2, 3, 7, 8, 9, 16, 24]))
        w, h = self.size
        cam = cam_of(w, h)
        if static:
            self.b.render_static(cam, self.k, n)
        else:
            self.b.render(cam, self.k, n)
        _, rays = self.b.wait()
        W, fma = self.mode
        o = self.oracle.render(self.scene[0], self.scene[1], W, fma, cam, w, h, 0, self.k + n)
        assert rays == o["rays"] - self.rays_total, ("rays", self.describe(n, static))
        assert np.array_equal(self.b.read_accum().view(np.uint32), o["accum"].view(np.uint32)), ("accum", self.describe(n, static))
        assert np.array_equal(self.b.read_rgba8(), o["rgba"]), ("rgba", self.describe(n, static))
        self.k += n
        self.rays_total = o["rays"]

    def roundtrip(self):
        acc = self.b.read_accum()
        self.b.reset()
        self.b.write_accum(acc)

    def describe(self, n, static):
        return dict(spheres=len(self.scene[0]), mode=self.mode, size=self.size, k=self.k, n=n, static=static, kernel=self.b.last_kernel())


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_random_call_sequences_match_the_oracle(srt, oracle, seed):
    rng = np.random.default_rng(1000 + seed)
    pools = []
    for idx in (1, 2, 3, 4):
        sp, sky, hdr = srt.host_scene(idx)
        pools.append((sp[:8200].copy(), sky))
    _, _, hdr3 = srt.host_scene(3)

    def cam_of(w, h):
        return srt.host_camera(3, w, h, hdr3)           # one camera for every scene: only the sphere tables change

    ms = [Model(srt, oracle, rng), Model(srt, oracle, rng)]
    try:
        for m in ms:
            m.new_scene(pools)
            m.resize()
        for step in range(70):
            m = ms[int(rng.integers(2))]
            op = rng.random()
            if op < 0.12:
                m.new_scene(pools)
            elif op < 0.20:
                m.resize()
            elif op < 0.26:
                m.reset()
            elif op < 0.32 and m.k:
                m.roundtrip()
            else:
                m.render(cam_of, static=bool(rng.random() < 0.3))
    finally:
        for m in ms:
            m.close()

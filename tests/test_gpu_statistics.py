"""Statistical check of the RNG stream contract against the reference AS SHIPPED.

The bit-exact tests compare with the reference's code under the parity pins; this one compares with what the
reference renders without them (its own per-thread sequential RNG and NormalizeFast, one worker thread;
tests/golden/make_stat_goldens.py): at 512 samples per pixel the two images must differ by Monte-Carlo noise only.
The yardstick for "noise" is the difference of two GPU renders with unrelated stream bases.  Stream contract v1
(the reference's seed scrambler taken literally: every path index = 41 or 57 mod 64 on one stream) fails every
check: measured on B200 with a library built from the v1 scrambler, scene 0 / scene 2: overall rms ratio 2.9 / 3.5
(v2: 1.01 / 1.00), column structure 5.2 / 6.8 (v2: 0.97 / 1.02), collapsed columns against the rest 14.9 / 18.3
(v2: 1.07 / 0.98) -- and two v1 renders with different stream bases were IDENTICAL in those columns."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_asshipped_stats.npz")


def _render(backend, sp, sky, mode, cam, w, h, spp, base):
    backend.set_scene(sp, sky, mode[0], bool(mode[1]))
    if (backend.width, backend.height) != (w, h):
        backend.resize(w, h)
    backend.reset()
    backend.render(cam, 0, spp, stream_base=base)
    backend.wait()
    return np.clip(backend.read_accum()[..., :3].astype(np.float64), 0.0, 2.0)


def _rms(a):
    return float(np.sqrt(np.mean(np.square(a))))


@pytest.mark.parametrize("scene,mode", [(0, (8, 1)), (2, (8, 1)), (1, (1, 0))])   # the profiles the goldens were rendered in
def test_contract_render_differs_from_the_shipped_reference_by_noise_only(backend, oracle, scene, mode):
    gold = np.load(GOLD)
    w, h, spp = (int(x) for x in gold["size"])
    sp, sky, hdr = oracle.scene(scene)
    cam = oracle.camera(hdr, w, h)
    mine, ref_cam = cam.view(np.float32).reshape(-1), gold[f"s{scene}_camera"].astype(np.float32)
    for off, n in ((0, 3), (4, 3), (8, 3), (12, 3), (16, 3), (20, 2)):        # v3 fields are padded to 4 floats
        assert mine[off:off + n].tobytes() == ref_cam[off:off + n].tobytes()
    ref = np.clip(gold[f"s{scene}_accum"].astype(np.float64), 0.0, 2.0)
    a = _render(backend, sp, sky, mode, cam, w, h, spp, 0)
    b = _render(backend, sp, sky, mode, cam, w, h, spp, 1 << 40)
    d_ref, d_self = a - ref, a - b
    # same noise level overall, no offset
    assert 0.85 < _rms(d_ref) / _rms(d_self) < 1.15
    assert abs(d_ref.mean()) < 5.0 * _rms(d_self) / np.sqrt(d_ref.size)
    # no structure along columns or rows (means over the other axis and the colour channels)
    for axis in (0, 1):
        assert _rms(d_ref.mean(axis=(axis, 2))) < 1.5 * _rms(d_self.mean(axis=(axis, 2))), axis
    # the residue classes the reference's scrambler collapses are as converged as every other column
    x = np.arange(w)
    bad = (x % 64 == 41) | (x % 64 == 57)
    assert bad.any()
    assert _rms(d_ref[:, bad]) < 1.3 * _rms(d_ref[:, ~bad])
    assert _rms(d_self[:, bad]) < 1.3 * _rms(d_self[:, ~bad])

"""GPU: the headless C++ driver (OnInit/OnRender mirror) end to end vs the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run_driver(srt, tmp_path, *args):
    acc = tmp_path / "accum.f32"
    ppm = tmp_path / "out.ppm"
    cmd = [srt.DRIVER_PATH, *map(str, args), "--accum", str(acc), "--out", str(ppm)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    stats = json.loads(res.stdout.strip().splitlines()[-1])
    return stats, np.fromfile(acc, np.float32), ppm.read_bytes()


@pytest.mark.parametrize("scene,simd,lanes,fma", [(0, 1, 8, 1), (1, 1, 4, 0), (2, 0, 8, 1), (2, 1, 8, 1)])
def test_driver_progressive_matches_oracle(srt, oracle, tmp_path, scene, simd, lanes, fma):
    w, h, spp = 96, 54, 3
    stats, acc, ppm = run_driver(srt, tmp_path, "--scene", scene, "--width", w, "--height", h, "--spp", spp,
                                 "--simd", simd, "--lanes", lanes, "--fma", fma)
    sp, sky, hdr = oracle.scene(scene)
    cam = oracle.camera(hdr, w, h)
    W = lanes if simd else 1
    o = oracle.render(sp, sky, W, fma if simd else 0, cam, w, h, 0, spp)
    assert np.array_equal(acc.view(np.uint32), o["accum"].reshape(-1).view(np.uint32))
    assert stats["rays"] == o["rays"] and stats["samples_accumulated"] == spp and stats["frames"] == spp
    # PPM: rows flipped, RGB bytes of the RGBA8 image
    head = f"P6\n{w} {h}\n255\n".encode()
    assert ppm.startswith(head)
    rgb = np.frombuffer(ppm[len(head):], np.uint8).reshape(h, w, 3)
    want = o["rgba"].view(np.uint8).reshape(h, w, 4)[::-1, :, :3]
    assert np.array_equal(rgb, want)


def test_driver_samples_per_frame_and_orbit(srt, oracle, tmp_path):
    w, h = 64, 36
    s1, a1, _ = run_driver(srt, tmp_path, "--scene", 0, "--width", w, "--height", h, "--spp", 4, "--spf", 4)
    s2, a2, _ = run_driver(srt, tmp_path, "--scene", 0, "--width", w, "--height", h, "--spp", 4, "--spf", 1, "--orbit", 0)
    assert np.array_equal(a1.view(np.uint32), a2.view(np.uint32)) and s1["rays"] == s2["rays"]
    s3, a3, _ = run_driver(srt, tmp_path, "--scene", 0, "--width", w, "--height", h, "--spp", 2, "--orbit", 3)
    assert s3["samples_accumulated"] == 2 and np.isfinite(a3).all()
    assert not np.array_equal(a3[: a1.size], a1)            # camera moved -> different image

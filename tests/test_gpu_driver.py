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


def test_write_accum_resume_is_bit_exact(srt, oracle):
    """N2: 4 spp, read the accumulator, upload it into a NEW backend, 4 more spp == 8 spp straight."""
    w, h = 160, 90
    sp, sky, hdr = srt.host_scene(2)
    cam = srt.host_camera(2, w, h)
    for lanes, fma in ((8, 1), (1, 0)):
        full = srt.Backend()
        full.set_scene(sp, sky, lanes, fma); full.resize(w, h)
        full.render(cam, 0, 8); _, rays_full = full.wait()
        a = srt.Backend()
        a.set_scene(sp, sky, lanes, fma); a.resize(w, h)
        a.render(cam, 0, 4); _, r1 = a.wait()
        half = a.read_accum()
        b = srt.Backend()
        b.set_scene(sp, sky, lanes, fma); b.resize(w, h)
        b.write_accum(half)
        b.render(cam, 4, 4); _, r2 = b.wait()
        assert np.array_equal(b.read_accum().view(np.uint32), full.read_accum().view(np.uint32))
        assert np.array_equal(b.read_rgba8(), full.read_rgba8()) and r1 + r2 == rays_full
        for x in (full, a, b):
            x.close()


def test_driver_checkpoint_resume_and_scene_file(srt, oracle, tmp_path):
    w, h = 96, 54
    ck = tmp_path / "half.ckpt"
    sf = tmp_path / "two.scene"
    s_full, a_full, ppm_full = run_driver(srt, tmp_path, "--scene", 1, "--width", w, "--height", h, "--spp", 6)
    s1, _, _ = run_driver(srt, tmp_path, "--scene", 1, "--width", w, "--height", h, "--spp", 2, "--checkpoint", ck)
    acc, hd = srt.load_checkpoint(ck)
    assert hd["samples"] == 2 and hd["width"] == w and hd["height"] == h and hd["scene_index"] == 1
    s2, a2, ppm2 = run_driver(srt, tmp_path, "--scene", 1, "--spp", 4, "--resume", ck)
    assert s2["resumed_from"] == 2 and s2["samples_accumulated"] == 6
    assert np.array_equal(a2.view(np.uint32), a_full.view(np.uint32)) and ppm2 == ppm_full
    assert s1["rays"] + s2["rays"] == s_full["rays"]
    # a checkpoint for another scene is refused
    res = subprocess.run([srt.DRIVER_PATH, "--scene", "0", "--spp", "1", "--resume", str(ck)], capture_output=True, text=True, timeout=120)
    assert res.returncode != 0 and "checkpoint" in res.stderr
    # scene file: dump scene 2, load it into slot 0, render == scene 2 rendered directly (same default camera)
    res = subprocess.run([srt.DRIVER_PATH, "--scene", "2", "--dump-scene", str(sf)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    s_a, a_a, _ = run_driver(srt, tmp_path, "--scene", 2, "--width", w, "--height", h, "--spp", 2)
    s_b, a_b, _ = run_driver(srt, tmp_path, "--scene", 0, "--scene-file", sf, "--width", w, "--height", h, "--spp", 2)
    assert np.array_equal(a_a.view(np.uint32), a_b.view(np.uint32)) and s_a["rays"] == s_b["rays"] and s_b["spheres"] == s_a["spheres"]


def test_driver_pfm_and_thread_count(srt, tmp_path):
    """N2/N3: the PFM dump is the accumulator (RGB, bottom-up rows); --threads is accepted and changes nothing."""
    w, h = 48, 27
    pfm = tmp_path / "a.pfm"
    s1, a1, _ = run_driver(srt, tmp_path, "--scene", 1, "--width", w, "--height", h, "--spp", 2, "--pfm", pfm, "--threads", 7)
    s2, a2, _ = run_driver(srt, tmp_path, "--scene", 1, "--width", w, "--height", h, "--spp", 2)
    assert np.array_equal(a1.view(np.uint32), a2.view(np.uint32)) and s1["rays"] == s2["rays"]
    raw = pfm.read_bytes()
    head = f"PF\n{w} {h}\n-1.0\n".encode()
    assert raw.startswith(head) and len(raw) == len(head) + w * h * 12
    rgb = np.frombuffer(raw[len(head):], np.float32).reshape(h, w, 3)
    assert np.array_equal(rgb.view(np.uint32), a1.reshape(h, w, 4)[..., :3].view(np.uint32))

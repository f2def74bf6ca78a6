"""GPU: the multi-device path behind the C ABI (include/srt_backend.h, "several devices").

  * image partition (srt_set_partition): handles that render interleaved rows of 8x4 blocks with the exact running
    mean; the sum of their buffers is BIT-IDENTICAL to a whole-image render -- checked on one GPU with two and three
    handles (no NCCL needed: the rows a handle does not own are zero), with bundle mode, resumed accumulations and
    srt_write_accum;
  * with >= 2 GPUs: the library's own NCCL collectives (srt_comm_init_rank, srt_gather_rgba / srt_gather_accum,
    srt_reduce_accum), one Python thread per handle, and `srt_headless --gpus 2` == `--gpus 1` byte for byte.
"""
import json
import subprocess
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


def _whole(srt, spheres, sky, cam, w, h, launches):
    b = srt.Backend(0)
    try:
        b.set_scene(spheres, sky, 8, True)
        b.resize(w, h)
        rays = 0
        for first, n in launches:
            b.render(cam, first, n)
            rays += b.wait()[1]
        return b.read_accum(), b.read_rgba8(), rays
    finally:
        b.close()


@pytest.mark.parametrize("scene,w,h,launches,parts", [(2, 97, 61, [(0, 3)], 2), (3, 160, 90, [(0, 12), (12, 9)], 3), (0, 64, 5, [(0, 2)], 4),
                                                       (3, 130, 70, [(0, 70)], 2)])
def test_partitions_sum_to_the_whole_image(srt, scene, w, h, launches, parts):
    spheres, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h, hdr)
    acc0, rgba0, rays0 = _whole(srt, spheres, sky, cam, w, h, launches)
    acc = np.zeros_like(acc0)
    rgba = np.zeros_like(rgba0)
    rays = 0
    owned = np.zeros((h, w), np.int32)
    for p in range(parts):
        b = srt.Backend(0)
        try:
            b.set_scene(spheres, sky, 8, True)
            b.resize(w, h)
            b.set_partition(p, parts)
            for first, n in launches:
                b.render(cam, first, n)
                rays += b.wait()[1]
            a, r = b.read_accum(), b.read_rgba8()
        finally:
            b.close()
        rows = (np.arange(h) // 4) % parts == p
        assert not a[~rows].any() and not r[~rows].any()          # foreign rows stay zero: the combine step is a sum
        owned += (r != 0)
        acc += a                                                    # x + 0 = x
        rgba += r
    assert (owned == 1).all()
    assert np.array_equal(acc.view(np.uint32), acc0.view(np.uint32)) and np.array_equal(rgba, rgba0) and rays == rays0


def test_partitioned_write_accum_keeps_own_rows(srt):
    spheres, sky, hdr = srt.host_scene(2)
    w, h = 80, 45
    cam = srt.host_camera(2, w, h, hdr)
    acc4, _, _ = _whole(srt, spheres, sky, cam, w, h, [(0, 4)])
    acc9, rgba9, _ = _whole(srt, spheres, sky, cam, w, h, [(0, 9)])
    tot_a, tot_r = np.zeros_like(acc9), np.zeros_like(rgba9)
    for p in range(2):
        b = srt.Backend(0)
        try:
            b.set_scene(spheres, sky, 8, True)
            b.resize(w, h)
            b.set_partition(p, 2)
            b.write_accum(acc4)                                     # the WHOLE checkpointed accumulator: the handle keeps its rows
            b.render(cam, 4, 5)
            b.wait()
            tot_a += b.read_accum(); tot_r += b.read_rgba8()
        finally:
            b.close()
    assert np.array_equal(tot_a.view(np.uint32), acc9.view(np.uint32)) and np.array_equal(tot_r, rgba9)


def test_partition_argument_checks(srt):
    b = srt.Backend(0)
    try:
        for part, n in ((1, 1), (0, 0), (5, 4)):
            with pytest.raises(srt.SrtError):
                b.set_partition(part, n)
        with pytest.raises(srt.SrtError):
            b.gather_rgba(0)                                        # no communicator
        with pytest.raises(srt.SrtError):
            b.resize(70000, 8)                                      # pixel coordinates are 16-bit in the path slots
        assert b.comm_info()["n_ranks"] == 0
    finally:
        b.close()


def _rank_thread(srt, rank, n, uid, scene, w, h, spp, out, mode):
    try:
        spheres, sky, hdr = srt.host_scene(scene)
        cam = srt.host_camera(scene, w, h, hdr)
        b = srt.Backend(rank)
        b.comm_init_rank(n, rank, uid)
        b.set_scene(spheres, sky, 8, True)
        b.resize(w, h)
        if mode == "bands":
            b.set_partition(rank, n)
            b.render(cam, 0, spp)
            b.gather_rgba(0)
            b.gather_accum(-1)                                      # all-gather flavour: every rank gets the accumulator
            _, rays = b.wait()
            out[rank] = dict(rays=rays, accum=b.read_accum(), rgba=b.read_rgba8() if rank == 0 else None, info=b.comm_info())
        else:
            b.render(cam, rank, spp // n, stride=n, mode=srt.ACCUM_SUM)
            b.reduce_accum(0)
            if rank == 0:
                b.resolve_sum(spp)
            _, rays = b.wait()
            out[rank] = dict(rays=rays, accum=b.read_accum() if rank == 0 else None, rgba=b.read_rgba8() if rank == 0 else None)
        b.comm_destroy()
        b.close()
    except Exception as e:   # noqa: BLE001 - reported by the main thread
        out[rank] = e


@pytest.mark.parametrize("mode", ["bands", "samples"])
def test_two_gpus_through_the_library_collectives(srt, mode):
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    n, scene, w, h, spp = 2, 3, 320, 180, 16
    uid = srt.comm_unique_id()
    out = {}
    ts = [threading.Thread(target=_rank_thread, args=(srt, r, n, uid, scene, w, h, spp, out, mode)) for r in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(300)
    for r in range(n):
        assert not isinstance(out.get(r), Exception), out.get(r)
    spheres, sky, hdr = srt.host_scene(scene)
    acc0, rgba0, rays0 = _whole(srt, spheres, sky, srt.host_camera(scene, w, h, hdr), w, h, [(0, spp)])
    assert out[0]["rays"] + out[1]["rays"] == rays0
    if mode == "bands":                                             # bit-identical to one device
        assert out[0]["info"]["n_ranks"] == 2 and out[0]["info"]["nccl_version"] > 0
        assert np.array_equal(out[0]["rgba"], rgba0)
        for r in range(n):
            assert np.array_equal(out[r]["accum"].view(np.uint32), acc0.view(np.uint32))
    else:                                                           # sum-then-divide: the stated tolerance
        d = np.abs(out[0]["rgba"].view(np.uint8).astype(np.int16) - rgba0.view(np.uint8).astype(np.int16)).max()
        assert d <= 2
        a, want = out[0]["accum"][..., :3].astype(np.float64), acc0[..., :3].astype(np.float64)
        peak = max(want.max(), 1.0)
        assert 10 * np.log10(peak * peak / max(np.mean((a - want) ** 2), 1e-30)) >= 50.0


def test_headless_driver_on_two_gpus_is_byte_identical(srt, tmp_path):
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    outs = []
    for g in (1, 2):
        ppm, acc = tmp_path / f"g{g}.ppm", tmp_path / f"g{g}.f32"
        res = subprocess.run([srt.DRIVER_PATH, "--scene", "3", "--width", "256", "--height", "144", "--spp", "12", "--spf", "4", "--gpus", str(g),
                              "--out", str(ppm), "--accum", str(acc)], capture_output=True, text=True, timeout=300)
        assert res.returncode == 0, res.stderr
        stats = json.loads(res.stdout.strip().splitlines()[-1])
        assert stats["gpus"] == g
        outs.append((ppm.read_bytes(), acc.read_bytes(), stats["rays"]))
    assert outs[0] == outs[1]


def test_primary_taps_under_an_image_partition(srt, oracle):
    """record_primary with srt_set_partition on one device: own rows equal the unpartitioned taps, foreign rows read as all
    bits set, and a part that owns no row at all (image lower than n_parts block rows) still answers (found by scripts/soak.py)."""
    sp, sky, hdr = oracle.scene(2)
    b = srt.Backend(0)
    try:
        b.set_scene(sp, sky, 8, True)
        for (w, h, n_parts) in ((96, 40, 3), (50, 3, 2)):
            cam = oracle.camera(hdr, w, h)
            b.resize(w, h)
            b.set_partition(0, 1)
            b.reset()
            b.render(cam, 0, 2, record_primary=True)
            b.wait()
            prim_all, t_all = b.read_primary()
            rows = np.arange(h)
            for part in range(n_parts):
                b.set_partition(part, n_parts)
                b.reset()
                b.render(cam, 0, 2, record_primary=True)
                b.wait()
                prim, t = b.read_primary()
                own = ((rows // 4) % n_parts) == part
                assert np.array_equal(prim[own], prim_all[own]) and np.array_equal(t[own].view(np.uint32), t_all[own].view(np.uint32))
                assert (prim[~own] == 0xFFFFFFFF).all() and (t[~own].view(np.uint32) == 0xFFFFFFFF).all()
    finally:
        b.set_partition(0, 1)
        b.close()

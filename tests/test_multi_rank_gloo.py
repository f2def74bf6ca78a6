"""CPU, world_size 2 over gloo: the N>1 host logic -- both partitions of a frame (row bands: sum-reduce of disjoint
rows == gather, bit-identical; samples: sum-reduce + resolve, within tolerance).  The CPU oracle stands in for the
device renderer; on the GPUs the same plan runs inside the library (srt_set_partition / srt_gather_rgba /
srt_reduce_accum over NCCL, tests/test_gpu_multi.py)."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP = 40, 24, 8


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import importlib
    import torch
    import orclib
    srt = importlib.import_module("simd-ray-tracer_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    O = orclib.Oracle()
    sp, sky, hdr = O.scene(2)
    cam = O.camera(hdr, W, H)
    shard = srt.sharding.plan(rank, world, SPP)
    acc, rays = O.render_sum(sp, sky, 8, 1, cam, W, H, shard["first"], shard["count"], shard["stride"], threads=1)
    t = torch.from_numpy(acc)
    srt.sharding.reduce_sum(dist, t, dst=0)
    r = torch.tensor([rays], dtype=torch.int64)
    dist.reduce(r, dst=0)
    # image partition: this rank's row bands of the exact image, zero elsewhere; integer sum of the RGBA8 words
    exact = O.render(sp, sky, 8, 1, cam, W, H, 0, SPP, threads=1)
    rows = srt.sharding.band_rows(rank, world, H)
    mine_rgba = np.zeros((H, W), np.int64)
    mine_rgba[rows] = exact["rgba"][rows]
    mine_acc = np.zeros((H, W, 4), np.float32)
    mine_acc[rows] = exact["accum"][rows]
    tr, ta = torch.from_numpy(mine_rgba), torch.from_numpy(mine_acc)
    srt.sharding.reduce_sum(dist, tr, dst=0)
    srt.sharding.reduce_sum(dist, ta, dst=0)
    if rank == 0:
        np.save(os.path.join(out_dir, "sum.npy"), t.numpy())
        np.save(os.path.join(out_dir, "rays.npy"), r.numpy())
        np.save(os.path.join(out_dir, "band_rgba.npy"), tr.numpy())
        np.save(os.path.join(out_dir, "band_acc.npy"), ta.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sample_sharding(tmp_path, oracle):
    import importlib
    srt = importlib.import_module("simd-ray-tracer_b200")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    total = np.load(tmp_path / "sum.npy")
    rays = int(np.load(tmp_path / "rays.npy")[0])
    sp, sky, hdr = oracle.scene(2)
    cam = oracle.camera(hdr, W, H)
    # the same shards summed in one process: bitwise equal (f32 add of two partial sums)
    a0, r0 = oracle.render_sum(sp, sky, 8, 1, cam, W, H, 0, SPP // 2, 2, threads=1)
    a1, r1 = oracle.render_sum(sp, sky, 8, 1, cam, W, H, 1, SPP // 2, 2, threads=1)
    assert np.array_equal((a0 + a1)[..., :3].view(np.uint32), total[..., :3].view(np.uint32))
    assert rays == r0 + r1
    # resolved image vs the exact running mean of the reference: within the stated tolerance
    rgba = oracle.resolve_sum(total, SPP)
    exact = oracle.render(sp, sky, 8, 1, cam, W, H, 0, SPP)
    assert exact["rays"] == rays                                        # identical sample set
    d = np.abs(rgba.view(np.uint8).astype(int) - exact["rgba"].view(np.uint8).astype(int)).max()
    assert d <= 2
    mse = float(np.mean((total[..., :3].astype(np.float64) - exact["accum"][..., :3].astype(np.float64)) ** 2))
    peak = float(max(exact["accum"][..., :3].max(), 1.0))
    assert 10 * np.log10(peak * peak / max(mse, 1e-30)) >= 50.0
    # row bands: the reduced buffers ARE the exact image
    assert np.array_equal(np.load(tmp_path / "band_rgba.npy").astype(np.uint32), exact["rgba"])
    assert np.array_equal(np.load(tmp_path / "band_acc.npy").view(np.uint32), exact["accum"].view(np.uint32))


def test_band_plan_covers_every_row_once():
    import importlib
    srt = importlib.import_module("simd-ray-tracer_b200")
    for world in (1, 2, 3, 4, 8):
        for height in (1, 4, 5, 135, 720, 2160):
            seen = sorted(y for r in range(world) for y in srt.sharding.band_rows(r, world, height))
            assert seen == list(range(height))
    # 4K over 8 ranks: 540 block rows -> 67 or 68 each (0.7 % imbalance at most)
    sizes = [len(srt.sharding.band_rows(r, 8, 2160)) for r in range(8)]
    assert max(sizes) - min(sizes) <= 4 and sum(sizes) == 2160


def test_plan_covers_every_sample_once():
    import importlib
    srt = importlib.import_module("simd-ray-tracer_b200")
    for world in (1, 2, 4, 8):
        seen = []
        for r in range(world):
            p = srt.sharding.plan(r, world, 256)
            seen += [p["first"] + i * p["stride"] for i in range(p["count"])]
        assert sorted(seen) == list(range(256))
    with pytest.raises(ValueError):
        srt.sharding.plan(0, 3, 256)

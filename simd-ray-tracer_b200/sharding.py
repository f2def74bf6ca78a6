"""How a frame is split over ranks -- pure host logic, the mirror of what the C ABI does on the
devices (srt_set_partition / srt_gather_rgba, srt_reduce_accum), so that the plan can be tested without a GPU
(gloo, world_size 2) and bench.py can report it.

Two partitions of a frame, as north_star allows ("by sample index (or image tiles)"):

  bands    rank g renders ALL samples of every n-th row of 8x4-pixel blocks (block rows g, g + n, ...) with the exact
           running mean; rows of other ranks stay zero in its buffers, so ONE sum-reduce of the RGBA8 words (33 MB at
           4K) is the gather, and the image is bit-identical to a single-device render.  The reference partitions a
           frame the same way in spirit: 32x32 tiles pulled by its workers (main.cpp:9, 362-368, 824-827).
  samples  rank g renders samples k = g (mod n) of every pixel into an f32 SUM, one sum-reduce of the accumulator
           (132.7 MB at 4K) to rank 0, which divides and tonemaps: within the stated tolerance of the running mean.
"""
BLOCK_ROWS = 4          # pixel rows per row of 8x4 blocks (csrc/srt_path.cuh item_to_pixel)


def plan(rank, world, spp):
    """Sample partition: rank `rank` of `world` renders samples first, first+stride, ... (count of them).
    Every (sample, pixel) path owns its RNG stream, so the union over ranks is exactly the
    sample set a single device renders."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    if spp % world:
        raise ValueError(f"spp {spp} is not divisible by world size {world}")
    return dict(first=rank, count=spp // world, stride=world)


def band_rows(rank, world, height):
    """Image partition: the pixel rows rank `rank` owns -- block row by = y // 4 with by % world == rank."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return [y for y in range(height) if (y // BLOCK_ROWS) % world == rank]


def reduce_sum(dist, tensor, dst=0):
    """One sum-reduce to `dst` over torch.distributed (the CPU tests' stand-in for the library's NCCL call)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(tensor, dst=dst, op=dist.ReduceOp.SUM)
    return tensor

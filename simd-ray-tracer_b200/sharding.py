"""Sample sharding across ranks: which samples a rank renders, and the
reduce + resolve step.  Pure host logic: works with any torch.distributed backend (NCCL on the
GPU box, gloo in the CPU tests)."""


def plan(rank, world, spp):
    """Rank `rank` of `world` renders samples first, first+stride, ... (count of them).
    Every (sample, pixel) path owns its RNG stream, so the union over ranks is exactly the
    sample set a single device renders."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    if spp % world:
        raise ValueError(f"spp {spp} is not divisible by world size {world}")
    return dict(first=rank, count=spp // world, stride=world)


def reduce_sum(dist, tensor, dst=0):
    """One sum-reduce of the f32 accumulator to `dst` (the only data-path collective)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(tensor, dst=dst, op=dist.ReduceOp.SUM)
    return tensor

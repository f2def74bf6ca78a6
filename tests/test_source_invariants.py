"""CPU: invariants of the CUDA sources that no CPU test can execute and that a refactoring could drop silently.

Each one was a real bug; the GPU tests cover the behaviour, these keep the reason in the code."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "simd-ray-tracer_b200", "csrc")


def _read(name):
    with open(os.path.join(CSRC, name)) as f:
        return f.read()


def test_warp_reconverges_after_the_ring_pops():
    """render_pool: the lanes leave pool_pop's spin loop one by one; without a __syncwarp() right after the pops the early
    lanes may run a whole scan while the others have not read their ring entries (lost / duplicated slot ids)."""
    src = _read("srt_backend.cu")
    pops = [m.end() for m in re.finditer(r"pool_pop\(ring, RCAP - 1u, pos \+ 32u \+ lane", src)]
    assert len(pops) == 1
    tail = src[pops[0]:pops[0] + 2500]
    sync, fence, use = tail.find("__syncwarp();"), tail.find("__threadfence_block();"), tail.find("SRT_SU(")
    assert 0 < sync < fence < use, "the pops must be followed by __syncwarp() and a block fence before any slot state is read"


def test_ring_push_publishes_after_every_lane_wrote():
    src = _read("srt_pool.cuh")
    body = src[src.index("__device__ __forceinline__ void pool_push("):]
    write, fence, sync, publish = (body.find(s) for s in ("*e = (ring_t)id;", "__threadfence_block();", "__syncwarp();", "atomicAdd(&ctl->avail[q]"))
    assert 0 < write < fence < sync < publish


def test_stream_seed_has_no_zero_shift():
    """A shift count of zero in the reference's scrambler zeroes the seed: every index = 41, 57 (mod 64) shared one stream."""
    for name in ("srt_device.cuh",):
        src = _read(name)
        body = src[src.index("uint64_t stream_seed(uint64_t i)"):][:700]
        assert "k1 ? k1 : 32u" in body and "k2 ? k2 : 32u" in body
    with open(os.path.join(ROOT, "oracle", "srt_oracle.c")) as f:
        c = f.read()
    body = c[c.index("uint64_t orc_stream_seed(uint64_t i)"):][:700]
    assert "k1 ? k1 : 32" in body and "k2 ? k2 : 32" in body

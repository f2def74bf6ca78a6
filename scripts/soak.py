#!/usr/bin/env python3
"""Soak run on a GPU box: many random (scene size, image size, sample count, arithmetic profile) launches at sizes
that fill the machine, on two handles, hashes written to a JSON file.

    python scripts/soak.py --cases 300 --out gpurun_out/soak_release.json
    python scripts/soak.py --cases 300 --modes 8:1 --out gpurun_out/soak_release81.json
    SRT_CUDA_LIB=simd-ray-tracer_b200/libsrt_cuda_dbg.so python scripts/soak.py --cases 300 --modes 8:1 --out gpurun_out/soak_debug81.json
    python scripts/soak.py --compare gpurun_out/soak_release81.json gpurun_out/soak_debug81.json

What it is for: the pool kernel hands path slots between warps through rings guarded by spin waits, and the last
CTA of a launch re-arms the counters of the next one.  A lost wake-up or a stale counter would show up as a hang
(release library; run it under `timeout`) or as a trap with a site code (the -DSRT_SPIN_LIMIT debug library), and a
race as a hash that differs between the two runs or between two launches of the same job (every case renders twice).
Parity with the oracle is the tests' business (tests/test_gpu_fuzz_api.py); this only needs results to be reproducible.
"""
import argparse
import hashlib
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:24]


def compare(a, b):
    A, B = json.load(open(a)), json.load(open(b))
    bad = [k for k in A["cases"] if A["cases"][k] != B["cases"].get(k)]
    print(f"{len(A['cases'])} cases, {len(bad)} differ", bad[:5])
    return 1 if bad or len(A["cases"]) != len(B["cases"]) else 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "soak.json"))
    ap.add_argument("--compare", nargs=2)
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--fixed", default="", help="n,w,h,spp,frames: every case is this job (scene pool and handle still vary)")
    ap.add_argument("--first", type=int, default=0, help="skip the launches of the cases before this one (parameters are still drawn)")
    ap.add_argument("--modes", default="8:1,4:0,1:0,8:0,4:1", help="lane_width:fma_dot list (the debug library only holds 8:1)")
    args = ap.parse_args()
    if args.compare:
        sys.exit(compare(*args.compare))
    srt = importlib.import_module("simd-ray-tracer_b200")
    rng = np.random.default_rng(args.seed)
    modes = [tuple(int(x) for x in m.split(":")) for m in args.modes.split(",")]
    pools = [srt.host_scene(i) for i in (1, 2, 3, 4)]
    hdr3 = pools[2][2]
    handles = [srt.Backend(0), srt.Backend(0)]
    sizes = [3, 40, 64, 100, 256, 383, 384, 482, 700, 1024, 1024, 1024, 2000, 3000, 8200, 8200, 20000, 65536]
    out, t0, rays_total = {}, time.time(), 0
    for case in range(args.cases):
        b = handles[int(rng.integers(2))]
        sp, sky, _ = pools[int(rng.integers(len(pools)))]
        n = min(int(rng.choice(sizes)), len(sp))
        W, fma = modes[int(rng.integers(5)) % len(modes)]
        big = n > 4000
        w = int(rng.integers(1, 500 if big else 1400))
        h = int(rng.integers(1, 300 if big else 800))
        spp = int(rng.choice([1, 1, 2, 7, 8, 9, 16, 33] if big else [1, 1, 2, 7, 8, 9, 16, 33, 64]))
        frames = int(rng.integers(1, 4))
        # the other launch parameters of the C ABI: sample stride / offset (sample partition), sum mode, image partition
        # (this handle renders part `part` of `n_parts`), the no-allocation entry point, the primary-ray taps
        stride = int(rng.choice([1, 1, 1, 2, 3]))
        offset = int(rng.integers(stride))
        mode = srt.ACCUM_SUM if (rng.random() < 0.2 or stride > 1) else srt.ACCUM_MEAN      # the exact running mean needs stride 1
        n_parts = int(rng.choice([1, 1, 2, 3, 4]))
        part = int(rng.integers(n_parts))
        static = bool(rng.random() < 0.2)
        taps = bool(rng.random() < 0.15)
        if args.fixed:
            n, w, h, spp, frames = (int(x) for x in args.fixed.split(","))
            n = min(n, len(sp))
            stride, offset, mode, n_parts, part, static, taps = 1, 0, srt.ACCUM_MEAN, 1, 0, False, False
        cam = srt.host_camera(3, w, h, hdr3)
        if case < args.first:
            continue
        b.set_scene(sp[:n].copy(), sky, W, bool(fma))
        b.resize(w, h)
        b.set_partition(part, n_parts)
        static = static and w * h <= srt.STATIC_MAX_PIXELS
        got = []
        key = f"{case}:n{n}_W{W}f{fma}_{w}x{h}_{spp}spp_x{frames}_s{stride}+{offset}_m{mode}_p{part}/{n_parts}{'_static' if static else ''}{'_taps' if taps else ''}"
        if args.verbose:
            print(key, flush=True)
        try:
            for rep in range(2):                        # the same job twice: must be reproducible
                b.reset()
                rays = 0
                for f in range(frames):
                    first = offset + f * spp * stride
                    if static:
                        b.render_static(cam, first, spp, stride=stride, mode=mode)
                    else:
                        b.render(cam, first, spp, stride=stride, mode=mode, record_primary=taps and f == 0)
                    rays += b.wait()[1]
                extra = sha(np.concatenate([x.view(np.uint32).ravel() for x in b.read_primary()])) if taps and not static else ""
                got.append((sha(b.read_accum()), sha(b.read_rgba8()), int(rays), extra))
        except Exception as e:
            print("FAILED", key, b.last_kernel(), e, flush=True)
            sys.exit(3)
        if got[0] != got[1]:
            print("NOT REPRODUCIBLE", key, got, b.last_kernel(), flush=True)
            sys.exit(2)
        out[key] = list(got[0]) + [b.last_kernel()]
        rays_total += got[0][2] * 2
        b.set_partition(0, 1)
    for b in handles:
        b.close()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump({"lib": os.environ.get("SRT_CUDA_LIB", "release"), "seconds": round(time.time() - t0, 1), "rays": rays_total, "cases": out}, f, indent=0)
    kernels = sorted({v[-1] for v in out.values()})
    print(f"{len(out)} cases, {rays_total / 1e9:.2f} G rays, {time.time() - t0:.1f} s, {len(kernels)} kernels", flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""As-shipped reference renders for the STATISTICAL check of the RNG stream contract (tests/test_gpu_statistics.py).

    python tests/golden/make_stat_goldens.py

The bit-exact goldens (make_goldens.py, make_full_goldens.py) run the reference's compiled code under the parity
pins of DESIGN.md section 3, so they cannot tell whether the pins themselves bias the image.  This file holds what
the reference renders WITHOUT them -- flags=0: its own per-thread sequential RNG seeded as OnInit does
(main.cpp:667-678) and its own NormalizeFast; one worker thread, so the render is reproducible -- at a sample count
where the Monte-Carlo noise is small.  The GPU test renders the same frames under the stream contract and checks
that the two images differ by noise only, with no structure along pixel columns or rows (the never-converging
columns x % 64 = 41, 57 of contract v1 fail it).  Build container only; the .npz is committed.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import orclib  # noqa: E402

W, H, SPP = 128, 72, 512
CASES = [(0, "w8", 1), (2, "w8", 1), (1, "w4", 0)]     # (scene, lib, EnableSIMD): AVX2 RenderTile twice, RenderTileScalar once


def main():
    out = {"size": np.array([W, H, SPP])}
    for scene, lib, simd in CASES:
        R = orclib.Ref(lib, 1)
        t0 = time.time()
        r = R.render(scene, W, H, SPP, simd=simd, threads=1, flags=0)
        out[f"s{scene}_accum"] = np.ascontiguousarray(r["accum"][..., :3]).astype(np.float32)
        out[f"s{scene}_camera"] = r["camera"]
        print(f"scene {scene}: {int(r['rays'])} rays, {time.time() - t0:.1f} s", flush=True)
    np.savez_compressed(os.path.join(HERE, "ref_asshipped_stats.npz"), **out)
    print("wrote", os.path.getsize(os.path.join(HERE, "ref_asshipped_stats.npz")), "bytes")


if __name__ == "__main__":
    main()

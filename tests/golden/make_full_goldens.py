#!/usr/bin/env python3
"""Full-size golden HASHES from the reference's OWN compiled code (oracle/_ref), SURVEY 8c items 3-4.

    python tests/golden/make_full_goldens.py [--only KEY_SUBSTRING] [--threads N] [--bench]

For the BASELINE configs at their stated sizes -- scenes 0-2 at 1280x720 x 1 spp (configs[0], the
reference's start-up window, main.cpp:649-650) and 1920x1080 x 1 / 4 / 64 spp (configs[1]) in the three
arithmetic profiles -- and, with --bench, for the benchmarked launch itself (configs[2]: synthetic
1,024-sphere scene, 3840x2160, 256 spp, AVX2 profile; ~45 minutes on 8 cores), this writes the SHA-256 of
every buffer the path produces: f32 accumulator, RGBA8 image, primary-ray primitive id and t, plus the ray
count.  Hashes keep the repository small; the GPU tests (tests/test_gpu_fullsize.py) and bench.py's
`parity_check` compare whole device buffers against them.  Runs in the build container only
(/root/reference must have been compiled by oracle/build_ref.py); the JSON is committed.
Resumable: cases already in the JSON are skipped.
"""
import argparse
import hashlib
import importlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import orclib  # noqa: E402

OUT = os.path.join(HERE, "ref_full_hashes.json")
MODES = {"w8f": ("w8", 1), "w4i": ("w4", 1), "scalar": ("w4", 0)}   # name -> (lib, EnableSIMD)
CASES = [(1280, 720, 1), (1920, 1080, 1), (1920, 1080, 4), (1920, 1080, 64)]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def digest(r):
    return {"accum": sha(r["accum"]), "rgba": sha(r["rgba"]), "prim": sha(r["prim"]), "t": sha(r["t"]), "rays": int(r["rays"]),
            "camera": [int(x) for x in np.asarray(r["camera"], np.float32).view(np.uint32)]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--bench", action="store_true", help="also the 3840x2160 x 256 spp bench launch (scene 3, w8f)")
    args = ap.parse_args()
    out = {}
    if os.path.exists(OUT):
        with open(OUT) as f:
            out = json.load(f)
    refs = {"w4": orclib.Ref("w4", args.threads), "w8": orclib.Ref("w8", args.threads)}

    def save():
        with open(OUT + ".tmp", "w") as f:
            json.dump(out, f, indent=0, sort_keys=True)
        os.replace(OUT + ".tmp", OUT)

    todo = []
    for s in range(3):
        for name, (lib, simd) in MODES.items():
            for (w, h, spp) in CASES:
                todo.append((f"s{s}_{name}_{w}x{h}_{spp}spp", lib, simd, s, w, h, spp, None))
    if args.bench:
        srt = importlib.import_module("simd-ray-tracer_b200")
        for (idx, w, h, spp, key) in [(3, 3840, 2160, 256, "bench_synth1k_4k_256spp"), (3, 480, 270, 16, "s3_w8f_480x270_16spp"),
                                      (4, 240, 135, 1, "s4_w8f_240x135_1spp")]:
            todo.append((key, "w8", 1, 1, w, h, spp, srt.host_scene(idx)))
    for key, lib, simd, scene, w, h, spp, inject in todo:
        if key in out or (args.only and args.only not in key):
            continue
        R = refs[lib]
        if inject is not None:       # synthetic scene -> the reference's own Scenes[1] slot
            sp, sky, hdr = inject
            R.set_scene(1, sp, sky, hdr)
        t0 = time.time()
        # flags=3: the parity pins of DESIGN.md section 3 (per-path RNG streams, IEEE NormalizeFast)
        r = R.render(scene, w, h, spp, simd=simd, threads=args.threads, flags=3)
        out[key] = digest(r)
        out[key]["seconds"] = round(time.time() - t0, 1)
        print(key, out[key]["rays"], "rays", out[key]["seconds"], "s", flush=True)
        save()
    save()


if __name__ == "__main__":
    main()

"""Scratch perf probe (not the bench): kernel variants / regen thresholds on the config scenes."""
import importlib, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
srt = importlib.import_module("simd-ray-tracer_b200")
b = srt.Backend(0)
print(b.device_info())

def run(name, scene, w, h, spp, W=8, fma=True, reps=3, variant=0, thr=None):
    sp, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h)
    os.environ["SRT_KERNEL_VARIANT"] = str(variant)
    if thr is None: os.environ.pop("SRT_REGEN_THRESHOLD", None)
    else: os.environ["SRT_REGEN_THRESHOLD"] = str(thr)
    try:
        b.set_scene(sp, sky, W, fma)
    except srt.SrtError as e:
        print(name, "skipped:", e); return
    b.resize(w, h)
    best = None
    for r in range(reps):
        b.reset(); b.render(cam, 0, spp); ms, rays = b.wait()
        if best is None or ms < best[0]: best = (ms, rays)
    ms, rays = best
    npad = (len(sp) + 7) // 8 * 8
    tests = rays * npad
    print(json.dumps(dict(name=name, n=len(sp), w=w, h=h, spp=spp, W=W, fma=fma, variant=variant, thr=thr, ms=round(ms, 3), rays=rays,
                          mrays=round(rays / ms / 1e3, 2), gtests=round(tests / ms / 1e6, 2),
                          tflops=round(20 * tests / ms / 1e9, 2))), flush=True)

for variant in (1, 0):
    run("synth1k", 3, 1920, 1080, 4, 8, True, variant=variant)
    run("synth1k", 3, 1920, 1080, 4, 4, False, variant=variant)
    run("scene2", 2, 1920, 1080, 16, variant=variant)
    run("scene1", 1, 1920, 1080, 16, variant=variant)
for thr in (1, 2, 4, 8, 12, 16, 24, 32):
    run("scene0", 0, 1920, 1080, 16, thr=thr)
for thr in (1, 2, 4, 8):
    run("scene1", 1, 1920, 1080, 16, thr=thr)
for thr in (1, 2, 4):
    run("synth1k", 3, 1920, 1080, 4, thr=thr)
run("scene0", 0, 1920, 1080, 1)
run("scene0", 0, 1920, 1080, 64)

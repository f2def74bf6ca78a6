"""Scratch perf probe (not the bench): times render launches on built-in and synthetic scenes."""
import importlib, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orclib
srt = importlib.import_module("simd-ray-tracer_b200")
O = orclib.Oracle()
b = srt.Backend(0)
print(b.device_info())

def synth(n, seed=1):
    rng = np.random.default_rng(seed)
    sp = np.zeros(n + 1, srt.sphere_dtype)
    sp["position"][:n, 0] = rng.uniform(-12, 12, n); sp["position"][:n, 1] = rng.uniform(0, 6, n); sp["position"][:n, 2] = rng.uniform(-12, 12, n)
    sp["radius"][:n] = rng.uniform(0.05, 0.35, n)
    sp["color"][:n, :3] = rng.uniform(0.2, 1.0, (n, 3))
    sp["specular"][:n] = (rng.uniform(size=n) < 0.5)
    sp["position"][n] = (0, -62.5, 0, 0); sp["radius"][n] = 62.5; sp["color"][n, :3] = 0.5
    return sp

def run(name, spheres, sky, hdr, w, h, spp, W=8, fma=True, reps=3):
    cam = O.camera(hdr, w, h)
    b.set_scene(spheres, sky, W, fma); b.resize(w, h)
    best = None
    for r in range(reps):
        b.reset(); b.render(cam, 0, spp); ms, rays = b.wait()
        if best is None or ms < best[0]: best = (ms, rays)
    ms, rays = best
    npad = (len(spheres) + 7) // 8 * 8
    tests = rays * npad
    print(json.dumps(dict(name=name, n=len(spheres), w=w, h=h, spp=spp, W=W, fma=fma, ms=round(ms, 3), rays=rays,
                          mrays=round(rays / ms / 1e3, 2), gtests=round(tests / ms / 1e6, 2),
                          tflops=round(20 * tests / ms / 1e9, 2))))

for s in range(3):
    sp, sky, hdr = O.scene(s)
    run(f"scene{s}", sp, sky, hdr, 1920, 1080, 1)
    run(f"scene{s}", sp, sky, hdr, 1920, 1080, 16)
hdr = np.array([0, 1, 0, 14.0, 0.6, 4.0], np.float32)
sp = synth(1024)
for W, fma in [(8, True), (4, False), (1, False)]:
    run("synth1024", sp, True, hdr, 1920, 1080, 4, W, fma)
run("synth8191", synth(8191), True, hdr, 960, 540, 2)

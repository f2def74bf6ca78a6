import importlib, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
srt = importlib.import_module("simd-ray-tracer_b200")
b = srt.Backend(0)
def run(name, scene, w, h, spp, W=8, fma=True, reps=3, variant=0, thr=None, pc=0):
    sp, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h)
    os.environ["SRT_KERNEL_VARIANT"] = str(variant)
    os.environ["SRT_PAIR_CONFIG"] = str(pc)
    if thr is None: os.environ.pop("SRT_REGEN_THRESHOLD", None)
    else: os.environ["SRT_REGEN_THRESHOLD"] = str(thr)
    b.set_scene(sp, sky, W, fma); b.resize(w, h)
    best = None
    for r in range(reps):
        b.reset(); b.render(cam, 0, spp); ms, rays = b.wait()
        if best is None or ms < best[0]: best = (ms, rays)
    ms, rays = best
    tests = rays * ((len(sp) + 7) // 8 * 8)
    print(json.dumps(dict(name=name, n=len(sp), spp=spp, W=W, fma=fma, variant=variant, pc=pc, ms=round(ms, 3),
                          mrays=round(rays / ms / 1e3, 2), gtests=round(tests / ms / 1e6, 2), tflops=round(20 * tests / ms / 1e9, 2))), flush=True)
for a in sys.argv[1:]:
    f = a.split(",")
    name, scene, spp, W, fma, variant = f[:6]
    run(name, int(scene), 1920, 1080, int(spp), int(W), fma == "1", variant=int(variant), pc=int(f[6]) if len(f) > 6 else 0)

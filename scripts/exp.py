"""Ad-hoc A/B runs on the GPU box: python scripts/exp.py 'name,scene,w,h,spp[,ENV=VAL;ENV=VAL...]' ...
The backend reads its SRT_* knobs at srt_create / srt_render, so every spec gets a fresh Backend."""
import importlib, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
srt = importlib.import_module("simd-ray-tracer_b200")
if os.environ.get("SRT_EXP_LIB"):
    srt.LIB_PATH = os.path.join(ROOT, os.environ["SRT_EXP_LIB"])   # an experimental build of the library
def run(name, scene, w, h, spp, env, reps=3):
    saved = {}
    for k, v in env.items():
        saved[k] = os.environ.get(k); os.environ[k] = v
    b = srt.Backend(0)
    sp, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h)
    b.set_scene(sp, sky, 8, True); b.resize(w, h)
    best = None
    for r in range(reps):
        b.reset(); b.render(cam, 0, spp); ms, rays = b.wait()
        if best is None or ms < best[0]: best = (ms, rays)
    ms, rays = best
    print(json.dumps(dict(name=name, n=len(sp), w=w, h=h, spp=spp, env=env, ms=round(ms, 4), mrays=round(rays / ms / 1e3, 2))), flush=True)
    del b
    for k, v in saved.items():
        if v is None: os.environ.pop(k, None)
        else: os.environ[k] = v
for a in sys.argv[1:]:
    f = a.split(",")
    env = dict(kv.split("=") for kv in f[5].split(";")) if len(f) > 5 and f[5] else {}
    run(f[0], int(f[1]), int(f[2]), int(f[3]), int(f[4]), env)

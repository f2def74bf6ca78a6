"""One short render for ncu: python scripts/prof_one.py <scene> <w> <h> <spp> [variant]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
srt = importlib.import_module("simd-ray-tracer_b200")
scene, w, h, spp = (int(x) for x in sys.argv[1:5])
if len(sys.argv) > 5: os.environ["SRT_KERNEL_VARIANT"] = sys.argv[5]
b = srt.Backend(0)
sp, sky, hdr = srt.host_scene(scene)
cam = srt.host_camera(scene, w, h)
b.set_scene(sp, sky, 8, True); b.resize(w, h)
for i in range(2):
    b.reset(); b.render(cam, 0, spp); ms, rays = b.wait()
    print(f"scene {scene} {w}x{h} {spp}spp: {ms:.3f} ms, {rays} rays, {rays*((len(sp)+7)//8*8)/ms/1e6:.1f} Gtests/s")

"""Filter audit statistics per scene: python scripts/audit_stats.py  (GPU)"""
import importlib, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
srt = importlib.import_module("simd-ray-tracer_b200")
b = srt.Backend(0)
for scene, w, h in ((1, 320, 180), (2, 320, 180), (3, 192, 108), (4, 24, 14)):
    sp, sky, hdr = srt.host_scene(scene)
    cam = srt.host_camera(scene, w, h, hdr)
    b.set_scene(sp, sky, 8, True); b.resize(w, h)
    a = b.filter_audit(cam, 0, 2)
    a.update(scene=scene, n=len(sp), cand_per_hit=round(a["candidates"] / max(a["hits"], 1), 3),
             hits_per_ray=round(a["hits"] * ((len(sp) + 7) // 8 * 8) / a["tests"], 3))
    print(json.dumps(a), flush=True)

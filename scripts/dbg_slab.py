import importlib, os, sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
srt = importlib.import_module("simd-ray-tracer_b200")
import orclib
orc = orclib.Oracle()
scene, w, h, spp = 2, 96, 54, 2
sp, sky, hdr = orc.scene(scene)
cam = orc.camera(hdr, w, h)
for W, fma in ((8, 1), (8, 0), (4, 1), (4, 0), (1, 0)):
    o = orc.render(sp, sky, W, fma, cam, w, h, 0, spp)
    for variant in ("0", "12"):
        os.environ["SRT_KERNEL_VARIANT"] = variant
        for rep in range(3):
            b = srt.Backend(0)
            b.set_scene(sp, sky, W, bool(fma)); b.resize(w, h)
            b.render(cam, 0, spp, record_primary=True); ms, rays = b.wait()
            acc = b.read_accum()
            diff = (acc.view(np.uint32) != o["accum"].view(np.uint32)).any(axis=2)
            a = b.filter_audit(cam, 0, spp)
            print("W", W, "fma", fma, "variant", variant, "rep", rep, "diff pixels", int(diff.sum()), "rays", rays, o["rays"], "audit viol", a["violations"], "slab", a["slab_survivors"], "cand", a["candidates"], "hits", a["hits"])
            b.close()

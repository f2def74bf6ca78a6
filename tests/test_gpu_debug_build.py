"""GPU: the -DSRT_SPIN_LIMIT debug build of the library (libsrt_cuda_dbg.so: every wait loop of the pool kernel traps
with a site code after N iterations instead of hanging the GPU) has the same ABI and renders the same bits."""
import hashlib
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import hashlib, importlib, sys
sys.path.insert(0, %r)
srt = importlib.import_module("simd-ray-tracer_b200")
out = []
for scene, w, h, spp in ((3, 320, 180, 20), (4, 96, 54, 8), (0, 200, 100, 3)):
    sp, sky, hdr = srt.host_scene(scene)
    b = srt.Backend(0)
    b.set_scene(sp, sky, 8, True)
    b.resize(w, h)
    b.render(srt.host_camera(scene, w, h, hdr), 0, spp)
    ms, rays = b.wait()
    out.append((rays, hashlib.sha256(b.read_accum().tobytes()).hexdigest(), hashlib.sha256(b.read_rgba8().tobytes()).hexdigest()))
    b.close()
print(out)
"""


def test_debug_library_renders_the_same_bits(srt):
    dbg = os.path.join(srt.PKG_DIR, "libsrt_cuda_dbg.so")
    assert os.path.exists(dbg), "run __graft_entry__.build()"
    res = []
    for lib in (None, dbg):
        env = dict(os.environ)
        env.pop("SRT_CUDA_LIB", None)
        if lib:
            env["SRT_CUDA_LIB"] = lib
        r = subprocess.run([sys.executable, "-c", SCRIPT % ROOT], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr
        res.append(r.stdout.strip().splitlines()[-1])
    assert res[0] == res[1]

"""CPU: the C++ host mirror (host/srt_scenes.cpp, srt_app.cpp) against goldens dumped from
the reference's own compiled code, and the properties of the synthetic stress scenes."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(G, "ref_goldens.npz"))


@pytest.mark.parametrize("scene", [0, 1, 2])
def test_builtin_scene_tables_match_reference(srt, gold, scene):
    sp, sky, hdr = srt.host_scene(scene)
    assert sp.view(np.uint8).reshape(len(sp), 80).tobytes() == gold[f"scene{scene}_spheres"].tobytes()
    assert int(sky) == int(gold[f"scene{scene}_sky"][0])
    assert hdr.tobytes() == gold[f"scene{scene}_header"].tobytes()


@pytest.mark.parametrize("scene", [0, 1, 2])
def test_camera_matches_reference(srt, gold, scene):
    h, w = gold[f"img_s{scene}_w8f_rgba"].shape
    cam = srt.host_camera(scene, w, h).view(np.float32).reshape(-1)
    ref = gold[f"img_s{scene}_w8f_camera"]
    for off in (0, 4, 8, 12, 16):
        assert cam[off:off + 3].tobytes() == ref[off:off + 3].tobytes()
    assert cam[20:22].tobytes() == ref[20:22].tobytes()


@pytest.mark.parametrize("scene,count", [(3, 1024), (4, 65536)])
def test_synthetic_scenes(srt, scene, count):
    sp, sky, hdr = srt.host_scene(scene)
    assert len(sp) == count and sky
    sp2, _, _ = srt.host_scene(scene)
    assert sp.tobytes() == sp2.tobytes()                       # deterministic
    p = sp["position"][1:, :3].astype(np.float64)
    r = sp["radius"][1:].astype(np.float64)
    assert (p[:, 1] - r >= -1e-6).all()                        # resting on / above the ground plane
    # non-overlap (checked on a sample; the generator enforces it for all pairs)
    idx = np.random.default_rng(0).choice(len(p), 400, replace=False)
    d = np.linalg.norm(p[idx, None, :] - p[None, :, :], axis=2)
    rr = r[idx, None] + r[None, :]
    d[np.arange(len(idx)), idx] = np.inf
    assert (d > rr).all()
    assert (sp["ior"] == 0).all()
    frac_em = (sp["emissive"][1:, :3].sum(1) > 0).mean()
    assert 0.08 < frac_em < 0.17

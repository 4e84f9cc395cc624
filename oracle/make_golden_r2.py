"""Generate tests/golden/golden_r2.npz from the UNMODIFIED reference (oracle/_ref/ref_driver).

    python oracle/make_golden_r2.py          (build container only: needs /root/reference)

Second set of reference vectors:
  samples/<scene>/<cam>/ms/<spp>   per-SAMPLE radiance of MultiScatterGaussians, [32, 32, spp, 4] = (r, g, b, bounce index when
                                   the path ended).  ref_driver walks integrator.h:557-702 around the reference's own routines
                                   and checks that the si-ordered sums reproduce the reference's render bit for bit
                                   ("render_mismatch": 0), so these ARE the reference's samples.
  spheres_env0/...                 RayMarchingSpheres (test_integrators.h:23-135) with Scene::env_color = (0,0,0): the
                                   random_device-seeded environment term (:95-117) vanishes, the image is deterministic.
                                   one_sphere at config C0's full 512^2 (every 4th pixel of every 4th row is kept) and both
                                   sphere scenes at 128^2.
  far/... , x50/...                MultiScatter / FreeFlight renders and per-sample dumps of rand300 translated by
                                   (+100, -50, +200) and of rand300 scaled x50 (covariances x2500, densities x2500 so the
                                   optical depths stay put): coordinates far from the origin, where the slack of the
                                   culling boxes has to follow the rounding of the reference's FP32 quadric test.
Scene files written next to the others: tests/golden/scenes/rand300_far.txt, rand300_x50.txt.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_binding as ob  # noqa: E402

_spec = importlib.util.spec_from_file_location("scenegen", os.path.join(ROOT, "3dg-vol-renderer_b200", "scenegen.py"))
scenegen = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(scenegen)

GOLD = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLD, "scenes")
CAMS = {
    "pin": dict(cam="pinhole", pos="0,1,6", look="0,1,0", fov="0.78539816339"),
    "pin2": dict(cam="pinhole", pos="2.5,1.8,3.5", look="0,0.9,0", fov="0.6"),
    "ortho": dict(cam="ortho", pos="0,1,6", look="0,1,0"),
    # the "pin" camera moved / scaled with the scene
    "pin_far": dict(cam="pinhole", pos="100,-49,206", look="100,-49,200", fov="0.78539816339"),
    "pin_x50": dict(cam="pinhole", pos="0,50,300", look="0,50,0", fov="0.78539816339"),
}
SHIFT = np.array([100.0, -50.0, 200.0])
SCALE = 50.0


def write_scene(path, scene):
    """like scenegen.write_gmm but with enough digits for shifted / scaled coordinates (the reader takes any float)"""
    with open(path, "w") as f:
        for l in scene["lights"]:
            f.write("l " + " ".join(f"{v:.6g}" for v in l) + "\n")
        f.write("\n")
        for i in range(len(scene["density"])):
            f.write("g " + " ".join(f"{v:.9g}" for v in scene["mean"][i]) + "  " + " ".join(f"{v:.9g}" for v in scene["cov6"][i])
                    + f"  {scene['density'][i]:.9g} {scene['albedo'][i]:.9g}\n")


def derived_scenes():
    base = scenegen.load_gmm(os.path.join(SCENES, "rand300.txt"))
    far = {k: np.array(v, np.float64) for k, v in base.items()}
    far["mean"] = far["mean"] + SHIFT
    far["lights"][:, 0:3] += SHIFT
    write_scene(os.path.join(SCENES, "rand300_far.txt"), far)
    x50 = {k: np.array(v, np.float64) for k, v in base.items()}
    x50["mean"] *= SCALE; x50["cov6"] *= SCALE * SCALE; x50["density"] *= SCALE * SCALE   # norm ~ s^-3, path ~ s
    x50["lights"][:, 0:3] *= SCALE; x50["lights"][:, 3:6] *= SCALE * SCALE      # the same irradiance at the scaled distances
    write_scene(os.path.join(SCENES, "rand300_x50.txt"), x50)


def main():
    ob.build()
    assert ob.have_ref(), "oracle/_ref/ref_driver missing (needs /root/reference)"
    derived_scenes()
    out = {}
    tmp = tempfile.mkdtemp()
    p = os.path.join(tmp, "o.f32")
    W = H = 32
    for sc, cam, spp in (("two_gauss", "ortho", 4), ("rand50", "pin", 16), ("rand300", "pin", 16), ("rand2k", "pin2", 4),
                         ("rand300_far", "pin_far", 16), ("rand300_x50", "pin_x50", 16)):
        r = ob.run_ref("samples", scene=os.path.join(SCENES, sc + ".txt"), w=W, h=H, spp=spp, out=p, **CAMS[cam])
        assert r["render_mismatch"] == 0
        out[f"samples/{sc}/{cam}/ms/{spp}"] = np.fromfile(p, np.float32).reshape(H, W, spp, 4)
    for sc, cam in (("rand300_far", "pin_far"), ("rand300_x50", "pin_x50")):
        for integ in ("ss", "ms"):
            ob.run_ref("render", scene=os.path.join(SCENES, sc + ".txt"), integrator=integ, w=W, h=H, spp=16, out=p, **CAMS[cam])
            out[f"render/{sc}/{cam}/{integ}/16"] = np.fromfile(p, np.float32).reshape(H, W, 3)
    # ---- spheres with a black environment
    for sc, wh in (("one_sphere", 512), ("one_sphere", 128), ("three_spheres", 128)):
        ob.run_ref("spheres", scene=os.path.join(SCENES, sc + ".txt"), w=wh, h=wh, step=0.01, env=5, envcolor="0,0,0", out=p,
                   **CAMS["ortho"])
        img = np.fromfile(p, np.float32).reshape(wh, wh, 3)
        out[f"spheres_env0/{sc}/ortho/{wh}"] = img[::4, ::4].copy() if wh == 512 else img
    path = os.path.join(GOLD, "golden_r2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB,", len(out), "arrays")


if __name__ == "__main__":
    main()

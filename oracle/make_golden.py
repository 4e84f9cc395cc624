"""Generate tests/golden/* from the UNMODIFIED reference (oracle/_ref/ref_driver).

Run in the build container (needs /root/reference to have built _ref/ref_driver):
    python oracle/make_golden.py
Writes small scene files (tests/golden/scenes/*.txt, produced by the seeded generator
3dg-vol-renderer_b200/scenegen.py) and tests/golden/golden_v1.npz holding, for those scenes,
the reference's own outputs: known-answer vectors (PCG stream, erfinv_approx, intersect_direct,
optical_depth, mu_t, solve_for_t_given_tau, transmittance_up_to, event lists, camera rays),
float32 renders of FreeFlightGaussians / MultiScatterGaussians, recorded per-pixel Gaussian lists,
and two SFD + Adam iterations with injected sign vectors.
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

TWO_GAUSS = """l 0.0 4.0 0.5  12.0 11.0 10.0
g -0.25 1.0 0.0   0.07 0.03 0.0 0.06 0.01 0.15   2.5 0.3
g 0.3 1.1 0.1   0.05 -0.02 0.01 0.08 0.0 0.04   1.5 0.9
"""
ONE_SPHERE = """l 0.0 4.0 0.0 35.0 35.0 35.0
s 0.0 1.0 0.0 1.0 0.1 0.7
"""
THREE_SPHERES = """l 0.0 4.0 0.0 35.0 35.0 35.0
l 3.0 2.0 2.0 10.0 5.0 0.0
s 0.0 1.0 0.0 1.0 0.1 0.7
s 0.8 1.3 0.2 0.6 0.3 1.2
s -1.2 0.6 -0.5 0.5 0.0 2.0
"""


def make_scenes():
    os.makedirs(SCENES, exist_ok=True)
    with open(os.path.join(SCENES, "two_gauss.txt"), "w") as f:
        f.write(TWO_GAUSS)
    with open(os.path.join(SCENES, "one_sphere.txt"), "w") as f:
        f.write(ONE_SPHERE)
    with open(os.path.join(SCENES, "three_spheres.txt"), "w") as f:
        f.write(THREE_SPHERES)
    scenegen.write_gmm(os.path.join(SCENES, "rand50.txt"), scenegen.generate(50, 11, diameter_range=(0.1, 0.5)))
    scenegen.write_gmm(os.path.join(SCENES, "rand300.txt"), scenegen.generate(300, 12, diameter_range=(0.05, 0.25)))
    scenegen.write_gmm(os.path.join(SCENES, "rand2k.txt"), scenegen.generate(2000, 13, nonuniform=True,
                                                                               diameter_range=(0.01, 0.07)))


CAMS = {
    "pin": dict(cam="pinhole", pos="0,1,6", look="0,1,0", fov="0.78539816339"),
    "pin2": dict(cam="pinhole", pos="2.5,1.8,3.5", look="0,0.9,0", fov="0.6"),
    "ortho": dict(cam="ortho", pos="0,1,6", look="0,1,0"),
}


def main():
    ob.build()
    assert ob.have_ref(), "oracle/_ref/ref_driver missing (needs /root/reference)"
    make_scenes()
    out = {}
    tmp = tempfile.mkdtemp()
    # --- known-answer vectors
    for sc in ("two_gauss", "rand50", "rand300"):
        p = os.path.join(tmp, f"kat_{sc}.bin")
        ob.run_ref("kat", scene=os.path.join(SCENES, sc + ".txt"), n=300, seed=5, out=p, **CAMS["pin"])
        for tag, count, payload in ob.read_sections(p):
            dt = {1: np.uint32, 2: np.float64}.get(tag, np.float32)
            out[f"kat/{sc}/{tag}"] = np.frombuffer(payload, dt).copy()
    # --- renders
    W = H = 32
    for sc, camname, spp in (("two_gauss", "pin", 16), ("two_gauss", "ortho", 4), ("rand50", "pin", 16),
                             ("rand50", "pin2", 9), ("rand300", "pin", 16), ("rand2k", "pin", 16),
                             ("rand2k", "pin2", 4)):
        for integ in ("ss", "ms"):
            p = os.path.join(tmp, "img.f32")
            rec = os.path.join(tmp, "rec.bin")
            kw = dict(scene=os.path.join(SCENES, sc + ".txt"), integrator=integ, w=W, h=H, spp=spp, out=p,
                      **CAMS[camname])
            if integ == "ms":
                kw["record"] = rec
            ob.run_ref("render", **kw)
            key = f"render/{sc}/{camname}/{integ}/{spp}"
            out[key] = np.fromfile(p, np.float32).reshape(H, W, 3)
            if integ == "ms":
                raw = np.fromfile(rec, np.int32)
                out[key + "/record"] = raw  # [npix, (count, idx*count)*]
    # min_bounces variant
    p = os.path.join(tmp, "img.f32")
    ob.run_ref("render", scene=os.path.join(SCENES, "rand50.txt"), integrator="ms", w=W, h=H, spp=16, out=p,
               min_bounces=1, **CAMS["pin"])
    out["render/rand50/pin/ms/16/minb1"] = np.fromfile(p, np.float32).reshape(H, W, 3)
    # --- SFD iterations with injected signs (target = a 64 spp render of the unperturbed scene)
    sc = os.path.join(SCENES, "rand50.txt")
    Ws = Hs = 24
    tgt = os.path.join(tmp, "tgt.f32")
    ob.run_ref("render", scene=sc, integrator="ms", w=Ws, h=Hs, spp=64, out=tgt, **CAMS["pin"])
    D = 50 * 11
    S, iters = 2, 2
    rs = np.random.RandomState(99)
    signs = (rs.randint(0, 2, size=(iters, S, D)).astype(np.int8) * 2 - 1).astype(np.int8)
    sp = os.path.join(tmp, "signs.i8")
    signs.tofile(sp)
    pk = os.path.join(tmp, "pack.f32")
    ob.run_ref("pack", scene=sc, out=pk)
    params0 = np.fromfile(pk, np.float32)
    # jitter the start so that there is a non-trivial loss
    params0 = (params0 + rs.normal(size=D).astype(np.float32) * np.float32(0.05)).astype(np.float32)
    params0.tofile(pk)
    ob.run_ref("sfd", scene=sc, target=tgt, signs=sp, S=S, iters=iters, w=Ws, h=Hs, spp=16, lr=0.01, params_in=pk,
               out=os.path.join(tmp, "sfd"), **CAMS["pin"])
    out["sfd/target"] = np.fromfile(tgt, np.float32).reshape(Hs, Ws, 3)
    out["sfd/signs"] = signs
    out["sfd/params0"] = params0
    it = -1
    for tag, count, payload in ob.read_sections(os.path.join(tmp, "sfd.bin")):
        if tag == 10:
            it += 1
        dt = np.float64 if tag in (11, 14) else np.float32
        out[f"sfd/iter{it}/{tag}"] = np.frombuffer(payload, dt).copy()
    # --- the reference's pack_parameters of a file scene (Eigen eigen-solver conventions)
    out["pack/rand50"] = np.fromfile(pk, np.float32) * 0 + np.fromfile(pk, np.float32)  # jittered copy (as used)
    ob.run_ref("pack", scene=sc, out=pk)
    out["pack/rand50/raw"] = np.fromfile(pk, np.float32)
    # --- spheres (statistical only: env directions are random_device-seeded in the reference)
    for sc_name in ("one_sphere", "three_spheres"):
        p = os.path.join(tmp, "sph.f32")
        ob.run_ref("spheres", scene=os.path.join(SCENES, sc_name + ".txt"), w=48, h=48, step=0.01, env=5, out=p,
                   **CAMS["ortho"])
        out[f"spheres/{sc_name}/ortho"] = np.fromfile(p, np.float32).reshape(48, 48, 3)
    np.savez_compressed(os.path.join(GOLD, "golden_v1.npz"), **out)
    sz = os.path.getsize(os.path.join(GOLD, "golden_v1.npz"))
    print(f"wrote {len(out)} arrays, {sz / 1024:.1f} KiB")


if __name__ == "__main__":
    main()

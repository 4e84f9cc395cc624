"""Generate tests/golden/golden_march_v1.npz from the UNMODIFIED reference (oracle/_ref/ref_driver march ...):
float32 images of TestIntegrator, PureRayMarching and RayMarchingGaussians on the bundled small scenes.

    python oracle/make_golden_march.py

The reference draws the marchers' environment directions from a random_device-seeded generator, so two
kinds of vectors are kept: `env0/...` rendered with Scene::env_color = (0,0,0), where that term vanishes and
the image is deterministic (per-pixel parity), and `env/...` with the default sky colour (means only).
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_binding as ob  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLD, "scenes")
CAM = dict(cam="pinhole", pos="0,1,6", look="0,1,0", fov="0.78539816339")
CASES = (("two_gauss", 16, 0.05), ("rand50", 16, 0.05), ("rand300", 12, 0.1))


def main():
    ob.build()
    out = {}
    p = os.path.join(tempfile.mkdtemp(), "m.f32")
    for sc, wh, step in CASES:
        for integ in ("test", "march_pure", "march_gauss"):
            for tag, envcolor in (("env0", "0,0,0"), ("env", "0.53,0.81,0.92")):
                ob.run_ref("march", scene=os.path.join(SCENES, sc + ".txt"), integrator=integ, w=wh, h=wh, step=step, env=4,
                           envcolor=envcolor, out=p, **CAM)
                out[f"{tag}/{sc}/{integ}"] = np.fromfile(p, np.float32).reshape(wh, wh, 3)
        out[f"meta/{sc}"] = np.array([wh, step, 4], np.float64)
    path = os.path.join(GOLD, "golden_march_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()

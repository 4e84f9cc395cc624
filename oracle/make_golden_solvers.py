"""Generate tests/golden/golden_solvers.npz: per-sample radiance of the UNMODIFIED reference with each of its distance solvers.

    make -C oracle ref_solvers && python oracle/make_golden_solvers.py      (build container only: needs /root/reference)

The reference selects its solver with a #define in distance_solvers.h:142-147 (ANALYTIC_PLUS_NEWTON as shipped); its #if / #elif
chain (:157-186) tests BISECTION, NEWTON and ANALYTIC_PLUS_BISECTION first, so `-D<mode>` on the compiler's command line selects
that solver without touching the headers (oracle/Makefile, target ref_solvers -> oracle/_ref/ref_driver_<mode>).
  samples/<mode>/<scene>/<cam>/ms/<spp>   [24, 24, spp, 4] = (r, g, b, bounce index when the path ended); ref_driver checks that the
                                          si-ordered sums reproduce the reference's own render bit for bit ("render_mismatch": 0)
Modes are named by this repository's solver numbers (include/volrender.h): 1 ANALYTIC_PLUS_BISECTION, 2 NEWTON, 3 BISECTION
(0 = ANALYTIC_PLUS_NEWTON is in golden_r2.npz; 4 = UNIFORM draws from std::random_device in the reference: unpinned).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLD, "scenes")
MODES = {1: "ANALYTIC_PLUS_BISECTION", 2: "NEWTON", 3: "BISECTION"}
CAMS = {"pin": dict(cam="pinhole", pos="0,1,6", look="0,1,0", fov="0.78539816339")}
CASES = [("rand50", "pin", 16), ("rand300", "pin", 16)]


def main():
    out = {}
    p = os.path.join(tempfile.mkdtemp(), "o.f32")
    W = H = 24
    for mode, name in MODES.items():
        exe = os.path.join(ROOT, "oracle", "_ref", "ref_driver_" + name)
        assert os.path.exists(exe), f"{exe} missing: make -C oracle ref_solvers"
        for sc, cam, spp in CASES:
            args = [exe, "samples", f"scene={os.path.join(SCENES, sc + '.txt')}", f"w={W}", f"h={H}", f"spp={spp}", f"out={p}"] + \
                   [f"{k}={v}" for k, v in CAMS[cam].items()]
            r = subprocess.run(args, check=True, capture_output=True, text=True)
            assert {1: "Analytic Inverse + Bisection", 2: "Newton-Raphson", 3: "Bisection"}[mode] in r.stdout, r.stdout[:400]
            assert json.loads(r.stdout.strip().splitlines()[-1])["render_mismatch"] == 0
            out[f"samples/{mode}/{sc}/{cam}/ms/{spp}"] = np.fromfile(p, np.float32).reshape(H, W, spp, 4)
    path = os.path.join(GOLD, "golden_solvers.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB,", len(out), "arrays")


if __name__ == "__main__":
    sys.exit(main())

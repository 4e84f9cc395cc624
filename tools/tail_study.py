#!/usr/bin/env python
"""One rank's share of config C2 at N = 8 (rows r, r + 8, ... of the 1024^2 x 256 spp image) on ONE GPU: where the time of a
short pass goes.  Prints ms per render for a few pool sizes / part counts / drain thresholds; VR_WAVE_TIMELINE=1 adds
the per-batch timeline of the library on stderr.  Not a bench: no L2 flush, wall clock around a synchronous render."""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (scene + config table)

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--slots", type=int, default=0)
    ap.add_argument("--parts", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--config", default="c2")
    a = ap.parse_args()
    vr = bench._pkg("volrender")
    cfg = bench.CONFIGS[a.config]
    scene = bench.make_scene(cfg)
    ctx = vr.Context(0)
    if a.slots: ctx.set_option("wave_slots", a.slots)
    if a.parts: ctx.set_option("wave_parts", a.parts)
    cam = vr.camera_basis("pinhole", (0, 1, 6), look=(0, 1, 0), fov=0.78539816339)
    ctx.set_lights(scene["lights"], bench.ENV); ctx.set_camera(cam); ctx.set_gmm(scene)
    W, H, spp = cfg["W"], cfg["H"], cfg["spp"]
    p = vr.Context.params(vr.INTEGRATOR_MS, W, H, spp, part=(0, a.ranks))
    for _ in range(2): ctx.render_device(p)
    ts = []; st = None
    for _ in range(a.reps):
        t = time.perf_counter(); st = ctx.render_device(p); ts.append((time.perf_counter() - t) * 1e3)
    n = W * len(range(0, H, a.ranks)) * spp
    print(f"ranks {a.ranks} slots {a.slots} parts {a.parts} env {os.environ.get('VR_WAVE_DRAIN_DIV','-')}: best {min(ts):.2f} ms median {sorted(ts)[len(ts)//2]:.2f} ms -> "
          f"{n / min(ts) * 1e-3:.1f} Msamples/s on this share, x{a.ranks} = {a.ranks * n / min(ts) * 1e-3:.0f}; kernel ms {st['kernel_ms']:.2f}")

if __name__ == "__main__":
    main()

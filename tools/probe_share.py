"""One rank's share of a C2 step at N = 8 (rows r, r + 8, ... of the 1024^2 image, 256 spp = 33.5 M samples), on one GPU:
where the end-to-end time goes (vr_set_gmm, render call, device kernels) and the wave-part / pool-size sweep.
usage (GPU box): python tools/probe_share.py [world [parts:slots_per_sm,...]]"""
import sys, time, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from conftest import load_pkg_module
vr = load_pkg_module("volrender"); sg = load_pkg_module("scenegen")
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
scene = sg.generate(10000, 2, nonuniform=True)
cam = vr.camera_basis("pinhole", (0, 1, 6), look=(0, 1, 0), fov=0.78539816339)
out = np.zeros((1024, 1024, 3), np.float32)
p = vr.Context.params(vr.INTEGRATOR_MS, 1024, 1024, 256, part=(0, world))
samples = len(range(0, 1024, world)) * 1024 * 256
combos = ((2, 0), (1, 0), (3, 0), (4, 0), (2, 148 * 16384), (2, 148 * 32768), (4, 148 * 32768))
if len(sys.argv) > 2:   # e.g. 2:32768,3:40960  (parts:slots per SM)
    combos = tuple((int(c.split(':')[0]), 148 * int(c.split(':')[1])) for c in sys.argv[2].split(','))
for parts, slots in combos:
    ctx = vr.Context(0)
    ctx.set_lights(scene["lights"], (0.53, 0.81, 0.92)); ctx.set_camera(cam)
    ctx.set_option("wave_parts", parts)
    if slots:
        ctx.set_option("wave_slots", slots)
    best = None
    for it in range(5):
        t0 = time.perf_counter(); ctx.set_gmm(scene); t1 = time.perf_counter()
        img, st = ctx.render(p, out=out); t2 = time.perf_counter()
        row = ((t1 - t0) * 1e3, (t2 - t1) * 1e3, st["kernel_ms"], st["total_ms"])
        if it >= 2 and (best is None or row[0] + row[1] < best[0] + best[1]):
            best = row
    print("parts %d slots %8d: set_gmm %6.2f ms, render call %7.2f ms (kernels %7.2f, device total %7.2f) -> %6.1f Msamples/s device, %6.1f e2e"
          % (parts, slots or 148 * 32768, *best, samples / best[3] * 1e-3, samples / (best[0] + best[1]) * 1e-3), flush=True)
    ctx.close()

import sys, time, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from conftest import load_pkg_module
vr = load_pkg_module("volrender"); sg = load_pkg_module("scenegen")
from oracle import oracle_binding as ob
scene = sg.generate(10000, 2, nonuniform=True)
oc = ob.make_camera("pinhole", (0, 1, 6), look=(0, 1, 0))
cam = dict(kind=int(oc.kind), pos=np.array(oc.pos[:], np.float32), view=np.array(oc.view[:], np.float32), right=np.array(oc.right[:], np.float32), up=np.array(oc.up[:], np.float32), pinhole=np.array(oc.pinhole[:], np.float32))
ctx = vr.Context(0)
ctx.set_lights(scene["lights"], (0.53, 0.81, 0.92)); ctx.set_camera(cam)
out = np.zeros((1024, 1024, 3), np.float32)
p = vr.Context.params(vr.INTEGRATOR_MS, 1024, 1024, 256)
for it in range(3):
    t0 = time.perf_counter(); ctx.set_gmm(scene); t1 = time.perf_counter()
    img, st = ctx.render(p, out=out); t2 = time.perf_counter()
    print("set_gmm %.2f ms, render call %.2f ms (device kernels %.2f, total_ms %.2f)" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, st["kernel_ms"], st["total_ms"]))

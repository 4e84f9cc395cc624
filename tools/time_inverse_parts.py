"""Where does an inverse iteration spend its time?  (GPU box)  python tools/time_inverse_parts.py [spp]"""
import sys, time, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import load_pkg_module
import numpy as np
vr = load_pkg_module("volrender"); sg = load_pkg_module("scenegen")
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
scene = sg.generate(10000, 2, nonuniform=True)
ctx = vr.Context(0)
ctx.set_lights(scene["lights"], (0.53, 0.81, 0.92)); ctx.set_camera(vr.camera_basis("pinhole", (0, 1, 6), look=(0, 1, 0)))
params = sg.pack_parameters(scene)
def t(fn, n=5):
    fn(); t0 = time.perf_counter()
    for _ in range(n): fn()
    return (time.perf_counter() - t0) / n * 1e3
print("set_gmm (H2D + prepare + host SAH build + upload) ms:", round(t(lambda: ctx.set_gmm(scene)), 2))
print("set_gmm_params ms:", round(t(lambda: ctx.set_gmm_params(params)), 2))
for rec in (False, True):
    rp = vr.Context.params(vr.INTEGRATOR_MS, 512, 512, spp, record=rec)
    print(f"render_device 512^2 x {spp} spp record={rec} ms:", round(t(lambda: ctx.render_device(rp)), 2), ctx.render_device(rp)["kernel_launches"], "launches")
eps = np.tile(sg.EPS11, 10000)
ctx.sfd_begin(params, eps); img, _ = ctx.render(vr.Context.params(vr.INTEGRATOR_MS, 512, 512, spp)); ctx.sfd_set_target(img)
signs = np.where(np.random.RandomState(1).randint(0, 2, size=(4, 110000)) > 0, 1, -1).astype(np.int8)
rp = vr.Context.params(vr.INTEGRATOR_MS, 512, 512, spp, record=True)
def acc():
    ctx.sfd_reset(4); ctx.sfd_accumulate(rp, signs)
print("sfd_accumulate (1 view, S=4: 5 scene rebuilds + 5 recorded renders + losses + F_diff) ms:", round(t(acc, 3), 2))
print("sfd_apply ms:", round(t(lambda: ctx.sfd_apply(signs, None), 3), 2))

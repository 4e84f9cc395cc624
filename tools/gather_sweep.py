#!/usr/bin/env python
"""Conservativeness of the culling boxes vs their slack (vr_bvh.h inflate_boxes, option "box_kappa"): overlap sets of camera
and interior rays through the wavefront traversal kernel (vr_kat_gather) against the oracle's gmm.h:457-515 restatement, and
the cost of the slack on a C2 render.  Checker tooling: imports oracle/."""
import importlib.util, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_binding as ob


def _pkg(name):
    spec = importlib.util.spec_from_file_location("volrender_b200_" + name, os.path.join(ROOT, "3dg-vol-renderer_b200", name + ".py"))
    mod = importlib.util.module_from_spec(spec); sys.modules["volrender_b200_" + name] = mod; spec.loader.exec_module(mod); return mod


vr = _pkg("volrender"); sg = _pkg("scenegen")
ENV = (0.53, 0.81, 0.92)
ctx = vr.Context(0)
scene = sg.generate(10000, 2, nonuniform=True)
oc = ob.make_camera("pinhole", (0, 1, 6), look=(0, 1, 0), fov=0.78539816339)
cd = dict(kind=int(oc.kind), pos=np.array(oc.pos[:], np.float32), view=np.array(oc.view[:], np.float32), right=np.array(oc.right[:], np.float32),
          up=np.array(oc.up[:], np.float32), pinhole=np.array(oc.pinhole[:], np.float32))
ctx.set_lights(scene["lights"], ENV); ctx.set_camera(cd)
rs = np.random.RandomState(1)
n = 60000
# camera rays through random film points (the far-origin worst case) and rays between interior points
uv = rs.uniform(0, 1, size=(n // 2, 2)).astype(np.float32)
o1 = np.empty((n // 2, 3), np.float32); d1 = np.empty((n // 2, 3), np.float32)
L = ob.lib()
import ctypes as C
for i in range(n // 2):
    oo = np.empty(3, np.float32); dd = np.empty(3, np.float32)
    L.vo_camera_sample_ray(C.byref(oc), float(uv[i, 0]), float(uv[i, 1]), oo, dd); o1[i] = oo; d1[i] = dd
a = rs.uniform((-1.1, -0.1, -1.1), (1.1, 2.1, 1.1), size=(n // 2, 3)).astype(np.float32)
b = rs.uniform((-1.1, -0.1, -1.1), (1.1, 2.1, 1.1), size=(n // 2, 3)).astype(np.float32)
d2 = b - a; d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(np.float32)
o = np.concatenate([o1, a]); d = np.concatenate([d1, d2])
gm = ob.Gmm(scene)
t0 = time.time()
ref = []
for i in range(n):
    _, _, idx = gm.events(o[i], d[i]); ref.append(frozenset(idx.tolist()))
print(f"oracle sets: {time.time() - t0:.1f} s, mean overlaps per ray {np.mean([len(r) for r in ref]):.2f}", flush=True)
p = vr.Context.params(vr.INTEGRATOR_MS, 1024, 1024, 16)
for kappa in (0.0, 0.5, 1.0, 1.5, 2.0, 4.0):
    ctx.set_option("box_kappa", kappa)
    ctx.set_gmm(scene)
    cnt, idx = ctx.kat_gather(o, d, cap=16)
    miss_cam = miss_int = extra = 0
    for i in range(n):
        got = frozenset(idx[i, :min(cnt[i], 16)].tolist())
        if len(ref[i]) > 16:
            continue
        if ref[i] - got:
            if i < n // 2: miss_cam += 1
            else: miss_int += 1
        if got - ref[i]:
            extra += 1
    ctx.render_device(p); ctx.render_device(p)
    ms = np.mean([ctx.render_device(p)["kernel_ms"] for _ in range(3)])
    print(f"kappa {kappa}: rays with a missed component: camera {miss_cam} / {n // 2}, interior {miss_int} / {n // 2}; rays with a spurious one {extra}; "
          f"C2 16 spp kernel {ms:.2f} ms", flush=True)
ctx.close()

"""ctypes binding of oracle/liboracle.so (the plain-C restatement) and helpers to run
oracle/_ref/ref_driver (the untouched reference compiled here).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never from the product package.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_DRIVER = os.path.join(HERE, "_ref", "ref_driver")

SOLVER_ANALYTIC_NEWTON, SOLVER_ANALYTIC_BISECTION, SOLVER_NEWTON, SOLVER_BISECTION = 0, 1, 2, 3
INTEGRATOR_SS, INTEGRATOR_MS = 0, 1

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")


class Camera(C.Structure):
    _fields_ = [("kind", C.c_int), ("pos", C.c_float * 3), ("view", C.c_float * 3), ("right", C.c_float * 3),
                ("up", C.c_float * 3), ("pinhole", C.c_float * 3)]


def build(force: bool = False) -> None:
    """Compile liboracle.so (and _ref/ref_driver when /root/reference exists)."""
    src = os.path.join(HERE, "volpath_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, LIB_PATH], stdout=subprocess.DEVNULL)
    drv = os.path.join(HERE, "ref_driver.cpp")
    if os.path.isdir("/root/reference/include") and (
            force or not os.path.exists(REF_DRIVER) or os.path.getmtime(REF_DRIVER) < os.path.getmtime(drv)):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(LIB_PATH)
    L.vo_splitmix64.restype = C.c_uint64; L.vo_splitmix64.argtypes = [C.c_uint64]
    L.vo_derive_path_seed.restype = C.c_uint64; L.vo_derive_path_seed.argtypes = [C.c_int] * 3
    L.vo_pcg_stream.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _u32p, _f32p]
    L.vo_erfinv_approx.restype = C.c_double; L.vo_erfinv_approx.argtypes = [C.c_double]
    L.vo_gmm_create.restype = C.c_void_p; L.vo_gmm_create.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p]
    L.vo_gmm_create_from_params.restype = C.c_void_p; L.vo_gmm_create_from_params.argtypes = [C.c_int, _f32p]
    L.vo_gmm_destroy.argtypes = [C.c_void_p]
    L.vo_gmm_num_nodes.restype = C.c_int; L.vo_gmm_num_nodes.argtypes = [C.c_void_p]
    L.vo_gmm_get_derived.argtypes = [C.c_void_p, _f32p, _f32p]
    L.vo_camera_make.argtypes = [C.c_int, _f32p, _f32p, C.c_float, C.POINTER(Camera)]
    L.vo_camera_sample_ray.argtypes = [C.POINTER(Camera), C.c_float, C.c_float, _f32p, _f32p]
    L.vo_intersect_direct.restype = C.c_int
    L.vo_intersect_direct.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.vo_optical_depth.restype = C.c_float
    L.vo_optical_depth.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.c_float, C.c_float]
    L.vo_mu_t.restype = C.c_float; L.vo_mu_t.argtypes = [C.c_void_p, C.c_int, _f32p]
    L.vo_solve_for_t_given_tau.restype = C.c_int
    L.vo_solve_for_t_given_tau.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.c_float, C.c_float, C.c_float,
                                           C.POINTER(C.c_float)]
    L.vo_transmittance_batch.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, _f32p]
    L.vo_intersect_events.restype = C.c_int
    L.vo_intersect_events.argtypes = [C.c_void_p, _f32p, _f32p, C.c_int, _f32p, _i32p, _u32p]
    L.vo_transmittance_brute.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, _f32p]
    L.vo_intersect_all.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.c_int, _i32p, _u32p]
    L.vo_render.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int,
                            C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, C.c_void_p,
                            C.c_void_p, C.c_int]
    L.vo_pixel_samples.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, C.POINTER(Camera), C.c_int, C.c_int,
                                   C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
    L.vo_render_samples.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, C.c_void_p]
    L.vo_libm_batch.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p]
    L.vo_num_threads.restype = C.c_int
    L.vo_counters_get.argtypes = [np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")]
    L.vo_pixel_losses.argtypes = [_f32p, _f32p, C.c_int, _f32p]
    L.vo_sfd_fdiff.argtypes = [C.c_int, C.c_int, _i32p, _u32p, _i32p, _u32p, C.c_int, _f32p, _f32p, _f64p]
    L.vo_sfd_accumulate.argtypes = [C.c_int, _f64p, _i8p, _f32p, _f64p]
    L.vo_default_eps.argtypes = [C.c_int, _f32p]
    L.vo_adam_step.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p, C.POINTER(C.c_int), C.c_float, C.c_float,
                               C.c_float, C.c_float]
    L.vo_unpack_params.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p]
    L.vo_render_spheres.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p, C.c_int, _f32p, _f32p, _f32p,
                                    C.POINTER(Camera), C.c_int, C.c_int, C.c_float, C.c_int, _f32p]
    _lib = L
    return L


def _c(a, dt=np.float32):
    return np.ascontiguousarray(a, dtype=dt)


def make_camera(kind: str, pos, look=None, view_dir=None, fov: float = 0.25 * np.pi) -> Camera:
    """kind 'pinhole'|'ortho'. view_dir = normalize(look - pos) in float32, as tests/main.cpp:32 does."""
    pos = _c(pos)
    if view_dir is None:
        d = _c(look) - pos
        view_dir = d / np.sqrt(np.float32(np.dot(d, d)), dtype=np.float32)
    cam = Camera()
    lib().vo_camera_make(0 if kind == "pinhole" else 1, pos, _c(view_dir), np.float32(fov), C.byref(cam))
    return cam


class Gmm:
    def __init__(self, scene: dict | None = None, params=None):
        L = lib()
        if params is not None:
            p = _c(params)
            self.n = p.size // 11
            self.h = L.vo_gmm_create_from_params(self.n, p)
        else:
            self.n = int(scene["density"].shape[0])
            self.h = L.vo_gmm_create(self.n, _c(scene["mean"]), _c(scene["cov6"]), _c(scene["density"]),
                                     _c(scene["albedo"]))

    def __del__(self):
        if getattr(self, "h", None):
            lib().vo_gmm_destroy(self.h)
            self.h = None

    def derived(self):
        inv = np.empty((self.n, 9), np.float32); norm = np.empty(self.n, np.float32)
        lib().vo_gmm_get_derived(self.h, inv, norm)
        return inv, norm

    def transmittance(self, o, d, tmax):
        o = _c(o); d = _c(d); tmax = _c(tmax)
        out = np.empty(o.shape[0], np.float32)
        lib().vo_transmittance_batch(self.h, o.shape[0], o, d, tmax, out)
        return out

    def transmittance_brute(self, o, d, tmax):
        """transmittance_up_to with every component asked (no hierarchy)"""
        o = _c(o); d = _c(d); tmax = _c(tmax)
        out = np.empty(o.shape[0], np.float32)
        lib().vo_transmittance_brute(self.h, o.shape[0], o, d, tmax, out)
        return out

    def intersect_all(self, o, d, cap=64):
        """brute force: per ray the components the FP32 quadric test accepts, no hierarchy -> (counts[n], idx[n, cap])"""
        o = _c(o); d = _c(d)
        cnt = np.zeros(o.shape[0], np.int32); idx = np.zeros((o.shape[0], cap), np.uint32)
        lib().vo_intersect_all(self.h, o.shape[0], o, d, cap, cnt, idx)
        return cnt, idx

    def events(self, o, d, cap=4096):
        t = np.empty(cap, np.float32); e = np.empty(cap, np.int32); idx = np.empty(cap, np.uint32)
        n = lib().vo_intersect_events(self.h, _c(o), _c(d), cap, t, e, idx)
        return t[:n], e[:n], idx[:n]


def render(gmm: Gmm, lights, env, cam: Camera, integrator: int, W: int, H: int, spp: int, min_bounces: int = 5,
           solver: int = SOLVER_ANALYTIC_NEWTON, window=None, record_cap: int = 0):
    """Integrator::render restated. Returns img[H,W,3] (and (counts, idx) when record_cap > 0)."""
    lights = _c(lights).reshape(-1, 6)
    lpos = _c(lights[:, 0:3]); lint = _c(lights[:, 3:6])
    if lpos.size == 0:
        lpos = np.zeros(3, np.float32); lint = np.zeros(3, np.float32)
    x0, y0, x1, y1 = window if window is not None else (0, 0, W, H)
    out = np.zeros((H, W, 3), np.float32)
    cnt = idx = None
    pc = pi = None
    if record_cap > 0:
        cnt = np.zeros(W * H, np.int32); idx = np.zeros((W * H, record_cap), np.uint32)
        pc = cnt.ctypes.data_as(C.c_void_p); pi = idx.ctypes.data_as(C.c_void_p)
    lib().vo_render(gmm.h, lights.shape[0], lpos, lint, _c(env), C.byref(cam), integrator, solver, W, H, spp,
                    min_bounces, x0, y0, x1, y1, out.reshape(-1), pc, pi, record_cap)
    return (out, cnt, idx) if record_cap > 0 else out


COUNTER_NAMES = ("scatter_rays", "shadow_rays", "hits", "optical_depth", "mu_t", "analytic", "newton", "bisection",
                 "paths")


def counters_reset():
    lib().vo_counters_reset()


def counters() -> dict:
    """algorithmic-work counters accumulated by render() since counters_reset()"""
    c = np.zeros(len(COUNTER_NAMES), np.uint64)
    lib().vo_counters_get(c)
    return dict(zip(COUNTER_NAMES, (int(v) for v in c)))


def pixel_samples(gmm: Gmm, lights, env, cam, integrator, W, H, spp, x, y, min_bounces=5,
                  solver=SOLVER_ANALYTIC_NEWTON):
    lights = _c(lights).reshape(-1, 6)
    out = np.zeros((spp, 3), np.float32)
    lib().vo_pixel_samples(gmm.h, lights.shape[0], _c(lights[:, 0:3]), _c(lights[:, 3:6]), _c(env), C.byref(cam),
                           integrator, solver, W, H, spp, min_bounces, x, y, out.reshape(-1))
    return out


def render_samples(gmm: Gmm, lights, env, cam: Camera, integrator: int, W: int, H: int, spp: int, min_bounces: int = 5,
                   solver: int = SOLVER_ANALYTIC_NEWTON, window=None, perturb: int = 0, margins: bool = False):
    """Per-sample radiance of a pixel window: array [wh, ww, spp, 4] = (r, g, b, bounce index at the end of the path)
    and, with margins=True, [wh, ww, spp] smallest relative margin of each sample's discrete decisions.
    perturb 1/2: free-flight distances +/- 1 ulp, 3/4: sampled optical depths * (1 +/- 2^-21), 5/6: distances * (1 +/- 2^-20)."""
    lights = _c(lights).reshape(-1, 6)
    lpos = _c(lights[:, 0:3]); lint = _c(lights[:, 3:6])
    if lpos.size == 0:
        lpos = np.zeros(3, np.float32); lint = np.zeros(3, np.float32)
    x0, y0, x1, y1 = window if window is not None else (0, 0, W, H)
    out = np.zeros((y1 - y0, x1 - x0, spp, 4), np.float32)
    mg = np.zeros((y1 - y0, x1 - x0, spp), np.float32) if margins else None
    lib().vo_render_samples(gmm.h, lights.shape[0], lpos, lint, _c(env), C.byref(cam), integrator, solver, W, H, spp,
                            min_bounces, x0, y0, x1, y1, perturb, out.reshape(-1),
                            mg.ctypes.data_as(C.c_void_p) if margins else None)
    return (out, mg) if margins else out


def sample_sensitivity(gmm: Gmm, lights, env, cam, integrator, W, H, spp, rtol=2e-5, atol=1e-7, response=False, **kw):
    """Samples whose value moves by more than rtol under a few-ulp perturbation of a free-flight distance (+/- 1 ulp and
    +/- 8 ulp: where coordinates are large the scatter POSITION is quantised much coarser than the distance, so a single ulp
    may not cross a rounding boundary that a few do) or of a sampled optical depth (perturb modes 1..6): returns (samples[..., 4], sensitive[...] bool, margin[...]) and, with
    response=True, also the largest absolute move of any channel under the four perturbations."""
    base, mg = render_samples(gmm, lights, env, cam, integrator, W, H, spp, margins=True, **kw)
    sens = np.zeros(base.shape[:-1], bool)
    resp = np.zeros(base.shape[:-1], np.float32)
    for mode in (1, 2, 3, 4, 5, 6):
        alt = render_samples(gmm, lights, env, cam, integrator, W, H, spp, perturb=mode, **kw)
        dev = np.abs(alt[..., :3] - base[..., :3])
        d = dev > atol + rtol * np.abs(base[..., :3])
        sens |= d.any(axis=-1) | (alt[..., 3] != base[..., 3])
        resp = np.maximum(resp, dev.max(axis=-1))
    return (base, sens, mg, resp) if response else (base, sens, mg)


def libm_batch(x):
    x = _c(x)
    s = np.empty_like(x); c = np.empty_like(x); a = np.empty_like(x)
    lib().vo_libm_batch(x.size, x, s, c, a)
    return s, c, a


def render_spheres(smm: dict, env, cam, W, H, step=0.01, env_samples=5):
    lights = _c(smm["lights"]).reshape(-1, 6)
    out = np.zeros((H, W, 3), np.float32)
    lib().vo_render_spheres(int(smm["radius"].shape[0]), _c(smm["center"]), _c(smm["radius"]), _c(smm["sigma_a"]),
                            _c(smm["sigma_s"]), lights.shape[0], _c(lights[:, 0:3]), _c(lights[:, 3:6]), _c(env),
                            C.byref(cam), W, H, np.float32(step), env_samples, out.reshape(-1))
    return out


MARCH_TEST, MARCH_PURE, MARCH_GAUSS = 0, 1, 2


def render_march(gmm: Gmm, lights, env, cam, W, H, mode, step=0.01, env_samples=20):
    """TestIntegrator / PureRayMarching / RayMarchingGaussians restated (vo_render_march)."""
    lights = _c(lights).reshape(-1, 6)
    lpos = _c(lights[:, 0:3]); lint = _c(lights[:, 3:6])
    if lpos.size == 0:
        lpos = np.zeros(3, np.float32); lint = np.zeros(3, np.float32)
    out = np.zeros((H, W, 3), np.float32)
    f = lib().vo_render_march
    f.restype = None
    f(C.c_void_p(gmm.h) if not isinstance(gmm.h, C.c_void_p) else gmm.h, C.c_int(lights.shape[0]), lpos.ctypes.data_as(C.c_void_p),
      lint.ctypes.data_as(C.c_void_p), _c(env).ctypes.data_as(C.c_void_p), C.byref(cam), C.c_int(W), C.c_int(H), C.c_int(mode),
      C.c_float(step), C.c_int(env_samples), out.ctypes.data_as(C.c_void_p))
    return out


# ------------------------------------------------------------------------------------------
# the compiled reference (oracle/_ref/ref_driver)
# ------------------------------------------------------------------------------------------
def have_ref() -> bool:
    return os.path.exists(REF_DRIVER)


def run_ref(cmd: str, threads: int | None = None, **kw) -> dict:
    """Run `ref_driver <cmd> key=value...`; returns the JSON line it prints last."""
    args = [REF_DRIVER, cmd] + [f"{k}={v}" for k, v in kw.items()]
    env = dict(os.environ)
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    out = subprocess.run(args, check=True, capture_output=True, text=True, env=env).stdout
    return json.loads(out.strip().splitlines()[-1])


def read_sections(path: str) -> list:
    """Parse the [u32 tag][u32 count][u64 bytes][payload] sections ref_driver writes."""
    secs = []
    with open(path, "rb") as f:
        data = f.read()
    off = 0
    while off < len(data):
        tag, count = np.frombuffer(data, np.uint32, 2, off)
        nbytes = int(np.frombuffer(data, np.uint64, 1, off + 8)[0])
        secs.append((int(tag), int(count), data[off + 16: off + 16 + nbytes]))
        off += 16 + nbytes
    return secs

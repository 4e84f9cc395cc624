"""ctypes binding of libvolrender_b200.so (the C-ABI in include/volrender.h).

Used by tests/, bench.py and __graft_entry__.py.  The C++ drop-in API (host/include/*.h) binds the
same entry points.  There is no CPU fallback anywhere: if the library or a CUDA device is missing,
construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# VR_LIB selects a tuning variant built by csrc/Makefile (VARIANT=...); tests and the bench use the default
LIB_PATH = os.environ.get("VR_LIB") or os.path.join(HERE, "libvolrender_b200.so")
CSRC = os.path.join(HERE, "csrc")

INTEGRATOR_SS, INTEGRATOR_MS, INTEGRATOR_MARCH_SPHERES = 0, 1, 2
INTEGRATOR_TEST, INTEGRATOR_PURE_MARCH, INTEGRATOR_MARCH_GAUSSIANS = 3, 4, 5
SOLVER_ANALYTIC_NEWTON, SOLVER_ANALYTIC_BISECTION, SOLVER_NEWTON, SOLVER_BISECTION, SOLVER_UNIFORM = 0, 1, 2, 3, 4
CAMERA_PINHOLE, CAMERA_ORTHO = 0, 1

STATUS = {0: "VR_OK", 1: "VR_ERR_INVALID_ARGUMENT", 2: "VR_ERR_CUDA", 3: "VR_ERR_NO_DEVICE", 4: "VR_ERR_NO_SCENE",
          5: "VR_ERR_HIT_OVERFLOW", 6: "VR_ERR_OUT_OF_MEMORY", 7: "VR_ERR_UNSUPPORTED"}

# every symbol include/volrender.h declares (tests/test_cabi.py checks the header against this and the .so)
SYMBOLS = ["vr_create", "vr_destroy", "vr_last_error", "vr_version", "vr_device_count", "vr_stream", "vr_set_gmm",
           "vr_set_gmm_params", "vr_set_smm", "vr_set_lights", "vr_set_camera", "vr_render", "vr_render_device",
           "vr_image_dev", "vr_get_recorded", "vr_get_samples", "vr_transmittance", "vr_transmittance_noalloc", "vr_transmittance_noalloc_max_rays", "vr_sfd_begin", "vr_sfd_set_target", "vr_sfd_fdiff",
           "vr_sfd_set_target_view", "vr_sfd_select_view", "vr_sfd_apply", "vr_sfd_reset_fdiff", "vr_sfd_accumulate", "vr_sfd_read", "vr_sfd_fdiff_dev", "vr_comm_available", "vr_comm_version",
           "vr_comm_unique_id", "vr_comm_init_rank", "vr_comm_init_all", "vr_comm_destroy", "vr_comm_size", "vr_sfd_allreduce", "vr_sfd_allreduce_all",
           "vr_sfd_get_params", "vr_sfd_get_state", "vr_sfd_set_state", "vr_adam_step", "vr_unpack_params", "vr_set_option", "vr_kat_pcg", "vr_kat_gaussian",
           "vr_kat_camera_rays", "vr_kat_events", "vr_kat_gather", "vr_kat_fdiff", "vr_kat_libm", "vr_kat_erfinv"]


class RenderParams(C.Structure):
    _fields_ = [("integrator", C.c_int32), ("solver", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("spp", C.c_int32), ("min_bounces", C.c_int32), ("x0", C.c_int32), ("y0", C.c_int32),
                ("x1", C.c_int32), ("y1", C.c_int32), ("part_rank", C.c_int32), ("part_count", C.c_int32),
                ("record", C.c_int32), ("march_step", C.c_float), ("env_samples", C.c_int32),
                ("reserved", C.c_int32 * 4)]


class RenderStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("total_ms", C.c_double), ("samples", C.c_uint64),
                ("scatter_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("hits", C.c_uint64),
                ("overflow_samples", C.c_uint64), ("kernel_launches", C.c_uint32), ("reserved", C.c_uint32),
                ("node_visits", C.c_uint64), ("quadric_tests", C.c_uint64), ("erf_evals", C.c_uint64), ("mu_evals", C.c_uint64),
                ("analytic_solves", C.c_uint64), ("iterative_solves", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class VolRenderError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{STATUS.get(code, code)}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile libvolrender_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    args = ["make", "-C", CSRC, "-j8"]
    if force:
        subprocess.check_call(["make", "-C", CSRC, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(args, stdout=subprocess.DEVNULL)
    return LIB_PATH


_libs = {}
STATS_LIB_PATH = os.path.join(HERE, "libvolrender_b200_stats.so")   # instrumented twin: executed-work counters (csrc/Makefile)


def lib(path: str | None = None):
    """Load the shared library (no compute call is made; works without a GPU).  path: another build of the same library
    (the instrumented twin STATS_LIB_PATH); default = the product library."""
    path = path or LIB_PATH
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} is missing: run __graft_entry__.build() (there is no CPU fallback)")
    L = C.CDLL(path)
    vp, ci, fp = C.c_void_p, C.c_int, C.c_void_p
    L.vr_create.argtypes = [C.POINTER(vp), ci]
    L.vr_destroy.argtypes = [vp]; L.vr_destroy.restype = None
    L.vr_last_error.argtypes = [vp]; L.vr_last_error.restype = C.c_char_p
    L.vr_version.restype = C.c_char_p
    L.vr_stream.argtypes = [vp]; L.vr_stream.restype = vp
    L.vr_set_gmm.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_set_gmm_params.argtypes = [vp, ci, fp]
    L.vr_set_smm.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_set_lights.argtypes = [vp, ci, fp, fp, fp]
    L.vr_set_camera.argtypes = [vp, ci, fp, fp, fp, fp, fp]
    L.vr_render.argtypes = [vp, C.POINTER(RenderParams), fp, C.POINTER(RenderStats)]
    L.vr_render_device.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(RenderStats)]
    L.vr_image_dev.argtypes = [vp]; L.vr_image_dev.restype = vp
    L.vr_get_recorded.argtypes = [vp, fp, fp, ci]
    L.vr_get_samples.argtypes = [vp, fp, C.c_uint64]
    L.vr_transmittance.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_transmittance_noalloc.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_transmittance_noalloc_max_rays.argtypes = []
    L.vr_sfd_begin.argtypes = [vp, ci, fp, fp, C.c_float, C.c_float, C.c_float, C.c_float]
    L.vr_sfd_set_target.argtypes = [vp, ci, ci, fp]
    L.vr_sfd_set_target_view.argtypes = [vp, ci, ci, ci, fp]
    L.vr_sfd_select_view.argtypes = [vp, ci]
    L.vr_sfd_fdiff.argtypes = [vp, C.POINTER(RenderParams), ci, fp, fp, fp, fp]
    L.vr_sfd_apply.argtypes = [vp, ci, fp, fp, fp]
    L.vr_sfd_reset_fdiff.argtypes = [vp, ci]
    L.vr_sfd_accumulate.argtypes = [vp, C.POINTER(RenderParams), ci, fp]
    L.vr_sfd_read.argtypes = [vp, fp, fp]
    L.vr_sfd_fdiff_dev.argtypes = [vp]; L.vr_sfd_fdiff_dev.restype = vp
    L.vr_comm_unique_id.argtypes = [fp]
    L.vr_comm_init_rank.argtypes = [vp, fp, ci, ci]
    L.vr_comm_init_all.argtypes = [C.POINTER(vp), ci]
    L.vr_comm_destroy.argtypes = [vp]
    L.vr_comm_size.argtypes = [vp]
    L.vr_sfd_allreduce.argtypes = [vp]
    L.vr_sfd_allreduce_all.argtypes = [C.POINTER(vp), ci]
    L.vr_sfd_get_params.argtypes = [vp, fp]
    L.vr_sfd_get_state.argtypes = [vp, fp, fp, fp, C.POINTER(ci)]
    L.vr_sfd_set_state.argtypes = [vp, fp, fp, fp, ci]
    L.vr_adam_step.argtypes = [vp, ci, fp, fp, fp, fp, C.POINTER(ci), C.c_float, C.c_float, C.c_float, C.c_float]
    L.vr_unpack_params.argtypes = [vp, ci, fp, fp, fp, fp, fp]
    L.vr_set_option.argtypes = [vp, C.c_char_p, C.c_double]
    L.vr_kat_pcg.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_kat_gaussian.argtypes = [vp, ci] + [fp] * 11
    L.vr_kat_camera_rays.argtypes = [vp, ci, fp, fp, fp]
    L.vr_kat_events.argtypes = [vp, fp, fp, ci, fp, fp, fp, C.POINTER(ci)]
    L.vr_kat_gather.argtypes = [vp, ci, fp, fp, ci, fp, fp]
    L.vr_kat_fdiff.argtypes = [vp, ci, ci, fp, fp, fp, fp, fp]
    L.vr_kat_libm.argtypes = [vp, ci, fp, fp, fp, fp]
    L.vr_kat_erfinv.argtypes = [vp, ci, fp, fp]
    _libs[path] = L
    return L


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def camera_basis(kind: str, pos, look=None, view_dir=None, fov: float = 0.25 * np.pi):
    """Camera constructor arithmetic in float32 (camera.h:15-22,38-43).  NOTE the reference's
    constructor bodies use the constructor PARAMETER view_dir (it shadows the member), i.e. the
    caller's vector as passed: callers pass normalize(look - pos) like tests/main.cpp:32."""
    f = np.float32
    pos = _f32(pos)
    if view_dir is None:
        d = _f32(look) - pos
        view_dir = (d / np.sqrt(f(d[0] * d[0]) + (f(d[1] * d[1]) + f(d[2] * d[2])), dtype=f)).astype(f)
    varg = _f32(view_dir)

    def normalized(a):
        z = f(a[0] * a[0]) + (f(a[1] * a[1]) + f(a[2] * a[2]))
        return (a / np.sqrt(z, dtype=f)).astype(f) if z > 0 else a

    def cross(a, b):
        return np.array([f(a[1] * b[2]) - f(a[2] * b[1]), f(a[2] * b[0]) - f(a[0] * b[2]),
                         f(a[0] * b[1]) - f(a[1] * b[0])], f)

    view = normalized(varg)
    right = normalized(cross(varg, np.array([0, 1, 0], f)))
    up = normalized(cross(right, varg))
    focal = f(1.0) / np.tan(f(0.5) * f(fov), dtype=f)
    pinhole = (pos + focal * varg).astype(f)
    return dict(kind=CAMERA_PINHOLE if kind == "pinhole" else CAMERA_ORTHO, pos=pos, view=view, right=right, up=up,
                pinhole=pinhole)


class Context:
    """One device context (vr_create .. vr_destroy)."""

    def __init__(self, device: int = 0, lib_path: str | None = None):
        self.L = lib(lib_path)
        self.h = C.c_void_p()
        rc = self.L.vr_create(C.byref(self.h), device)
        if rc:
            msg = self.L.vr_last_error(self.h).decode() if self.h else "no usable CUDA device"
            if self.h:
                self.L.vr_destroy(self.h)
            self.h = None
            raise VolRenderError(rc, msg)
        self.n = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.vr_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise VolRenderError(rc, self.L.vr_last_error(self.h).decode())

    # ---- scene
    def set_gmm(self, scene: dict):
        m, c, d, a = _f32(scene["mean"]), _f32(scene["cov6"]), _f32(scene["density"]), _f32(scene["albedo"])
        self.n = int(d.shape[0])
        self._ck(self.L.vr_set_gmm(self.h, self.n, _ptr(m), _ptr(c), _ptr(d), _ptr(a)))

    def set_gmm_params(self, params):
        p = _f32(params)
        self.n = p.size // 11
        self._ck(self.L.vr_set_gmm_params(self.h, self.n, _ptr(p)))

    def set_smm(self, smm: dict):
        c, r, sa, ss = _f32(smm["center"]), _f32(smm["radius"]), _f32(smm["sigma_a"]), _f32(smm["sigma_s"])
        self._ck(self.L.vr_set_smm(self.h, int(r.shape[0]), _ptr(c), _ptr(r), _ptr(sa), _ptr(ss)))

    def set_lights(self, lights, env):
        lights = _f32(lights).reshape(-1, 6)
        lp, li, e = _f32(lights[:, 0:3]), _f32(lights[:, 3:6]), _f32(env)
        self._ck(self.L.vr_set_lights(self.h, lights.shape[0], _ptr(lp), _ptr(li), _ptr(e)))

    def set_camera(self, cam: dict):
        self._ck(self.L.vr_set_camera(self.h, cam["kind"], _ptr(_f32(cam["pos"])), _ptr(_f32(cam["view"])),
                                      _ptr(_f32(cam["right"])), _ptr(_f32(cam["up"])), _ptr(_f32(cam["pinhole"]))))

    def set_option(self, name: str, value: float):
        self._ck(self.L.vr_set_option(self.h, name.encode(), float(value)))

    # ---- render
    @staticmethod
    def params(integrator, W, H, spp, min_bounces=5, solver=SOLVER_ANALYTIC_NEWTON, window=None, part=(0, 1),
               record=False, march_step=0.01, env_samples=5) -> RenderParams:
        x0, y0, x1, y1 = window if window is not None else (0, 0, W, H)
        return RenderParams(integrator, solver, W, H, spp, min_bounces, x0, y0, x1, y1, part[0], part[1],
                            int(record), march_step, env_samples)

    def render(self, p: RenderParams, out=None):
        """vr_render: returns (img[H,W,3] float32 on the host, stats dict)."""
        if out is None:
            out = np.zeros((p.height, p.width, 3), np.float32)
        st = RenderStats()
        self._ck(self.L.vr_render(self.h, C.byref(p), _ptr(out), C.byref(st)))
        return out, st.as_dict()

    def render_into(self, p: RenderParams, host_ptr: int):
        st = RenderStats()
        self._ck(self.L.vr_render(self.h, C.byref(p), C.c_void_p(host_ptr), C.byref(st)))
        return st.as_dict()

    def render_device(self, p: RenderParams):
        st = RenderStats()
        self._ck(self.L.vr_render_device(self.h, C.byref(p), C.byref(st)))
        return st.as_dict()

    def image_dev(self) -> int:
        return int(self.L.vr_image_dev(self.h) or 0)

    def recorded(self, W, H, cap):
        cnt = np.zeros(W * H, np.int32); idx = np.zeros((W * H, cap), np.uint32)
        self._ck(self.L.vr_get_recorded(self.h, _ptr(cnt), _ptr(idx), cap))
        return cnt, idx

    def samples(self, n_pixels, spp):
        """per-sample (r, g, b, bounce) of the last single-pass free-flight render: [n_pixels, spp, 4]"""
        out = np.empty((n_pixels, spp, 4), np.float32)
        self._ck(self.L.vr_get_samples(self.h, _ptr(out), n_pixels * spp))
        return out

    def transmittance(self, o, d, tmax):
        o, d, tmax = _f32(o), _f32(d), _f32(tmax)
        out = np.empty(tmax.shape[0], np.float32)
        self._ck(self.L.vr_transmittance(self.h, tmax.shape[0], _ptr(o), _ptr(d), _ptr(tmax), _ptr(out)))
        return out

    def transmittance_noalloc(self, o, d, tmax):
        """vr_transmittance_noalloc: the same kernel with the rays staged in the library's fixed device array (at most
        noalloc_max_rays() rays, no cudaMalloc inside the call; calls that overlap on one GPU race)"""
        o, d, tmax = _f32(o), _f32(d), _f32(tmax)
        out = np.empty(tmax.shape[0], np.float32)
        self._ck(self.L.vr_transmittance_noalloc(self.h, tmax.shape[0], _ptr(o), _ptr(d), _ptr(tmax), _ptr(out)))
        return out

    def noalloc_max_rays(self) -> int:
        return int(self.L.vr_transmittance_noalloc_max_rays())

    # ---- inverse
    def sfd_begin(self, params, eps, lr=1e-2, beta1=0.9, beta2=0.999, adam_eps=1e-8):
        p, e = _f32(params), _f32(eps)
        self._ck(self.L.vr_sfd_begin(self.h, p.size // 11, _ptr(p), _ptr(e), lr, beta1, beta2, adam_eps))
        self._sfd_n = p.size // 11

    def sfd_set_target(self, target):
        t = _f32(target)
        self._ck(self.L.vr_sfd_set_target(self.h, t.shape[1], t.shape[0], _ptr(t)))

    def sfd_set_target_view(self, view, target):
        t = _f32(target)
        self._ck(self.L.vr_sfd_set_target_view(self.h, view, t.shape[1], t.shape[0], _ptr(t)))

    def sfd_select_view(self, view):
        self._ck(self.L.vr_sfd_select_view(self.h, view))

    def sfd_fdiff(self, rp: RenderParams, signs):
        s = np.ascontiguousarray(signs, np.int8)
        S = s.shape[0]
        F = np.zeros((S, self._sfd_n), np.float64); ls = np.zeros(1, np.float64); npx = np.zeros(1, np.int64)
        self._ck(self.L.vr_sfd_fdiff(self.h, C.byref(rp), S, _ptr(s), _ptr(F), _ptr(ls), _ptr(npx)))
        return F, float(ls[0]), int(npx[0])

    def sfd_reset(self, S):
        self._ck(self.L.vr_sfd_reset_fdiff(self.h, S))

    def sfd_accumulate(self, rp: RenderParams, signs):
        s = np.ascontiguousarray(signs, np.int8)
        self._ck(self.L.vr_sfd_accumulate(self.h, C.byref(rp), s.shape[0], _ptr(s)))

    def sfd_read(self, S):
        F = np.zeros((S, self._sfd_n), np.float64); ls = np.zeros(1, np.float64)
        self._ck(self.L.vr_sfd_read(self.h, _ptr(F), _ptr(ls)))
        return F, float(ls[0])

    # ---- the inverse path's collective (csrc/vr_comm.cu): NCCL all-reduce of the fixed-point partial sums on the context's stream
    def comm_init_rank(self, unique_id: bytes, nranks: int, rank: int):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._ck(self.L.vr_comm_init_rank(self.h, C.cast(buf, C.c_void_p), nranks, rank))

    def comm_size(self) -> int:
        return int(self.L.vr_comm_size(self.h))

    def sfd_allreduce(self):
        self._ck(self.L.vr_sfd_allreduce(self.h))

    def sfd_fdiff_dev(self) -> int:
        return int(self.L.vr_sfd_fdiff_dev(self.h) or 0)

    def sfd_apply(self, signs, F=None):
        s = np.ascontiguousarray(signs, np.int8)
        grad = np.zeros(self._sfd_n * 11, np.float32)
        Fp = _ptr(np.ascontiguousarray(F, np.float64)) if F is not None else None
        self._ck(self.L.vr_sfd_apply(self.h, s.shape[0], _ptr(s), Fp, _ptr(grad)))
        return grad

    def sfd_params(self):
        p = np.zeros(self._sfd_n * 11, np.float32)
        self._ck(self.L.vr_sfd_get_params(self.h, _ptr(p)))
        return p

    def sfd_get_state(self):
        D = self._sfd_n * 11
        p, m, v = (np.zeros(D, np.float32) for _ in range(3)); t = C.c_int(0)
        self._ck(self.L.vr_sfd_get_state(self.h, _ptr(p), _ptr(m), _ptr(v), C.byref(t)))
        return p, m, v, t.value

    def sfd_set_state(self, p, m, v, t):
        self._ck(self.L.vr_sfd_set_state(self.h, _ptr(_f32(p)), _ptr(_f32(m)), _ptr(_f32(v)), int(t)))

    def adam_step(self, params, grads, m, v, t, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8):
        tt = C.c_int(t)
        self._ck(self.L.vr_adam_step(self.h, params.size, _ptr(params), _ptr(_f32(grads)), _ptr(m), _ptr(v), C.byref(tt),
                                     lr, beta1, beta2, eps))
        return tt.value

    def unpack_params(self, params):
        p = _f32(params); n = p.size // 11
        mean = np.empty((n, 3), np.float32); cov6 = np.empty((n, 6), np.float32)
        den = np.empty(n, np.float32); alb = np.empty(n, np.float32)
        self._ck(self.L.vr_unpack_params(self.h, n, _ptr(p), _ptr(mean), _ptr(cov6), _ptr(den), _ptr(alb)))
        return mean, cov6, den, alb

    # ---- known-answer hooks
    def kat_pcg(self, xysi):
        a = np.ascontiguousarray(xysi, np.int32).reshape(-1, 3); n = a.shape[0]
        seed = np.empty(n, np.uint64); u = np.empty((n, 4), np.uint32); f = np.empty((n, 4), np.float32)
        self._ck(self.L.vr_kat_pcg(self.h, n, _ptr(a), _ptr(seed), _ptr(u), _ptr(f)))
        return seed, u, f

    def kat_gaussian(self, o, d, gidx, frac):
        o, d, frac = _f32(o), _f32(d), _f32(frac); g = np.ascontiguousarray(gidx, np.int32); n = g.shape[0]
        hit = np.empty(n, np.int32); ok = np.empty(n, np.int32)
        t0, t1, tau, mu, ts = (np.empty(n, np.float32) for _ in range(5))
        self._ck(self.L.vr_kat_gaussian(self.h, n, _ptr(o), _ptr(d), _ptr(g), _ptr(frac), _ptr(hit), _ptr(t0), _ptr(t1),
                                        _ptr(tau), _ptr(mu), _ptr(ok), _ptr(ts)))
        return dict(hit=hit, t0=t0, t1=t1, tau=tau, mu=mu, ok=ok, t_solved=ts)

    def kat_camera_rays(self, uv):
        uv = _f32(uv).reshape(-1, 2); n = uv.shape[0]
        o = np.empty((n, 3), np.float32); d = np.empty((n, 3), np.float32)
        self._ck(self.L.vr_kat_camera_rays(self.h, n, _ptr(uv), _ptr(o), _ptr(d)))
        return o, d

    def kat_events(self, o, d, cap=4096):
        t = np.empty(cap, np.float32); e = np.empty(cap, np.int32); idx = np.empty(cap, np.uint32); n = C.c_int(0)
        self._ck(self.L.vr_kat_events(self.h, _ptr(_f32(o)), _ptr(_f32(d)), cap, _ptr(t), _ptr(e), _ptr(idx), C.byref(n)))
        return t[:n.value], e[:n.value], idx[:n.value]

    def kat_gather(self, o, d, cap=16):
        """overlap sets of a batch of rays through the wavefront traversal kernel: counts[n], idx[n, cap] (original indices)"""
        o, d = _f32(o), _f32(d)
        n = o.shape[0]
        cnt = np.empty(n, np.int32); idx = np.empty((n, cap), np.uint32)
        self._ck(self.L.vr_kat_gather(self.h, n, _ptr(o), _ptr(d), cap, _ptr(cnt), _ptr(idx)))
        return cnt, idx

    def kat_fdiff(self, n, bits_base, bits_plus, loss_base, loss_plus):
        """F_diff[n] in int64 fixed point (2^-32) from per-pixel recorded-set bitmaps [npix, (n+31)//32] and losses [npix]"""
        b0 = np.ascontiguousarray(bits_base, np.uint32); b1 = np.ascontiguousarray(bits_plus, np.uint32)
        l0, l1 = _f32(loss_base), _f32(loss_plus)
        F = np.empty(n, np.int64)
        self._ck(self.L.vr_kat_fdiff(self.h, l0.size, n, _ptr(b0), _ptr(b1), _ptr(l0), _ptr(l1), _ptr(F)))
        return F

    def kat_libm(self, x):
        x = _f32(x); n = x.size
        s, c, a = (np.empty(n, np.float32) for _ in range(3))
        self._ck(self.L.vr_kat_libm(self.h, n, _ptr(x), _ptr(s), _ptr(c), _ptr(a)))
        return s, c, a

    def kat_erfinv(self, x):
        x = np.ascontiguousarray(x, np.float64); y = np.empty_like(x)
        self._ck(self.L.vr_kat_erfinv(self.h, x.size, _ptr(x), _ptr(y)))
        return y


def comm_available() -> bool:
    return bool(lib().vr_comm_available())


def comm_unique_id() -> bytes:
    """ncclGetUniqueId through the library (rank 0 calls it, the 128 bytes go to every rank by the launcher's means)"""
    buf = (C.c_char * 128)()
    rc = lib().vr_comm_unique_id(C.cast(buf, C.c_void_p))
    if rc:
        raise VolRenderError(rc, "vr_comm_unique_id failed (NCCL unavailable?)")
    return bytes(buf)


def _handles(ctxs):
    arr = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    return arr


def comm_init_all(ctxs):
    """one process driving several devices: ncclCommInitAll over the contexts (contexts on one device: peer-copy sums)"""
    rc = lib().vr_comm_init_all(_handles(ctxs), len(ctxs))
    if rc:
        raise VolRenderError(rc, ctxs[0].L.vr_last_error(ctxs[0].h).decode())


def sfd_allreduce_all(ctxs):
    rc = lib().vr_sfd_allreduce_all(_handles(ctxs), len(ctxs))
    if rc:
        raise VolRenderError(rc, ctxs[0].L.vr_last_error(ctxs[0].h).decode())

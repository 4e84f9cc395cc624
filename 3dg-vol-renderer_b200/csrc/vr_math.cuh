// vr_math.cuh -- device arithmetic of the Gaussian volume path, rounding-faithful.
//
// The reference evaluates everything in FP32 with catastrophic cancellation (C - B^2/4A with
// |p| ~ 6 and sigma ~ 0.01 loses 5 digits) and the medium is optically so thick (mu_t ~ 2e4) that a
// 1-ulp shift of a scatter position moves a shadow-ray transmittance by ~1e-3.  Parity at 1e-4 is
// therefore only reachable by executing the SAME sequence of IEEE operations.  Every
// parity-critical expression below is written with __fmul_rn/__fadd_rn/__fmaf_rn/__fdiv_rn/
// __fsqrt_rn (no contraction, correctly rounded) in the reference's evaluation order:
//   Eigen dot()         = fma chain       (extern/eigen-3.4.0 Core/InnerProduct.h, numext::fma)
//   Eigen squaredNorm() = e0 + (e1 + e2)  (Core/Redux.h)
//   Eigen 3x3 * 3x1     = row . v as e0 + (e1 + e2), plain mul/add
// sinf/cosf/acosf are ports of the algorithms glibc 2.39 runs on the reference's host (they are
// not correctly rounded, so CUDA's or a double-rounded version disagrees in 1-8% of arguments; provenance: NOTICE).
// tests/test_gpu_parity_r2.py::test_libm_ports_on_every_argument_the_renderer_can_produce compares them with the
// host libm on all 2^24 arguments of each call site: acosf bit for bit, sinf / cosf within one ulp on every argument
// and bit for bit on > 99.99 % (the host picks an FMA or a non-FMA build of the same polynomials at run time).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace vr {

#define VR_PI_D 3.14159265358979323846

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ f3 scale3(float s, f3 a) { return mk3(__fmul_rn(s, a.x), __fmul_rn(s, a.y), __fmul_rn(s, a.z)); }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, __fmul_rn(a.x, b.x))); }
__device__ __forceinline__ float sqnorm3(f3 a) {
    return __fadd_rn(__fmul_rn(a.x, a.x), __fadd_rn(__fmul_rn(a.y, a.y), __fmul_rn(a.z, a.z)));
}
__device__ __forceinline__ f3 normalized3(f3 a) {
    float z = sqnorm3(a);
    if (z > 0.0f) { float s = __fsqrt_rn(z); return mk3(__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)); }
    return a;
}
__device__ __forceinline__ float row3(float m0, float m1, float m2, f3 v) {
    return __fadd_rn(__fmul_rn(m0, v.x), __fadd_rn(__fmul_rn(m1, v.y), __fmul_rn(m2, v.z)));
}

// ----------------------------------------------------------------------------------------------
// libm ports (glibc 2.39: sysdeps/ieee754/flt-32/{s_sinf.c,s_cosf.c,s_sincosf.h,e_acosf.c})
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ float sincos_poly(double x, double x2, bool neg_table, int n) {
    // cosine polynomial c0..c4 is negated in table[1]; sine coefficients are shared
    const double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5, c3 = -0x1.6c087e89a359dp-10,
                 c4 = 0x1.99343027bf8c3p-16, s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7,
                 s3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        double x3 = __dmul_rn(x, x2);
        double t1 = __dadd_rn(s2, __dmul_rn(x2, s3));
        double x7 = __dmul_rn(x3, x2);
        double s = __dadd_rn(x, __dmul_rn(x3, s1));
        return (float)__dadd_rn(s, __dmul_rn(x7, t1));
    } else {
        double sg = neg_table ? -1.0 : 1.0;
        double x4 = __dmul_rn(x2, x2);
        double q2 = __dadd_rn(sg * c3, __dmul_rn(x2, sg * c4));
        double q1 = __dadd_rn(sg * c1, __dmul_rn(x2, sg * c2));
        double x6 = __dmul_rn(x4, x2);
        double c = __dadd_rn(sg * c0, __dmul_rn(x2, q1));
        return (float)__dadd_rn(c, __dmul_rn(x6, q2));
    }
}
__device__ __forceinline__ uint32_t abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ffu; }
__device__ __forceinline__ double sincos_reduce(double x, int* np) {
    double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    int n = (__double2int_rz(r) + 0x800000) >> 24;
    *np = n;
    return __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
}
// valid for |y| < 120 (the renderer only passes [0, 2pi]); larger arguments fall back to CUDA's
__device__ __forceinline__ float ref_sinf(float y) {
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return sincos_poly(x, __dmul_rn(x, x), false, 0);
    }
    if (abstop12(y) >= abstop12(120.0f)) return sinf(y);
    int n;
    x = sincos_reduce(x, &n);
    double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    return sincos_poly(__dmul_rn(x, s), __dmul_rn(x, x), (n & 2) != 0, n);
}
__device__ __forceinline__ float ref_cosf(float y) {
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return sincos_poly(x, __dmul_rn(x, x), false, 1);
    }
    if (abstop12(y) >= abstop12(120.0f)) return cosf(y);
    int n;
    x = sincos_reduce(x, &n);
    double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    return sincos_poly(__dmul_rn(x, s), __dmul_rn(x, x), (n & 2) != 0, n ^ 1);
}
__device__ __forceinline__ float acos_pq(float z) {
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f,
                pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f, qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f,
                qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
#define VR_MA(a, b, c) __fadd_rn(a, __fmul_rn(b, c))
    float p = __fmul_rn(z, VR_MA(pS0, z, VR_MA(pS1, z, VR_MA(pS2, z, VR_MA(pS3, z, VR_MA(pS4, z, pS5))))));
    float q = VR_MA(1.0f, z, VR_MA(qS1, z, VR_MA(qS2, z, VR_MA(qS3, z, qS4))));
#undef VR_MA
    return __fdiv_rn(p, q);
}
__device__ __forceinline__ float ref_acosf(float x) {
    const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
    int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) return hx > 0 ? 0.0f : __fadd_rn(pi, __fmul_rn(2.0f, pio2_lo));
    if (ix > 0x3f800000) return CUDART_NAN_F;
    if (ix < 0x3f000000) {
        if (ix <= 0x32800000) return __fadd_rn(pio2_hi, pio2_lo);
        float r = acos_pq(__fmul_rn(x, x));
        return __fsub_rn(pio2_hi, __fsub_rn(x, __fsub_rn(pio2_lo, __fmul_rn(x, r))));
    } else if (hx < 0) {
        float z = __fmul_rn(__fadd_rn(1.0f, x), 0.5f);
        float r = acos_pq(z);
        float s = __fsqrt_rn(z);
        float w = __fsub_rn(__fmul_rn(r, s), pio2_lo);
        return __fsub_rn(pi, __fmul_rn(2.0f, __fadd_rn(s, w)));
    } else {
        float z = __fmul_rn(__fsub_rn(1.0f, x), 0.5f);
        float s = __fsqrt_rn(z);
        float df = __int_as_float(__float_as_int(s) & 0xfffff000);
        float c = __fdiv_rn(__fsub_rn(z, __fmul_rn(df, df)), __fadd_rn(s, df));
        float r = acos_pq(z);
        float w = __fadd_rn(__fmul_rn(r, s), c);
        return __fmul_rn(2.0f, __fadd_rn(df, w));
    }
}

// ----------------------------------------------------------------------------------------------
// RNG -- include/rng.h:13-57 (PCG32 with the reference's non-standard shift-OR)
// ----------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint64_t derive_path_seed(int x, int y, int si) {
    uint64_t s = ((uint64_t)(int64_t)si << 32) | ((uint64_t)(int64_t)y << 16) | (uint64_t)(int64_t)x;
    return splitmix64(s);
}
struct Pcg {
    uint64_t state;  // inc is always (1 << 1) | 1 = 3 for the path streams (PCG32(seed, 1))
    uint64_t inc;
    __host__ __device__ __forceinline__ uint32_t next_u32() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + inc;
        uint32_t shifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        uint32_t lsh = ((0u - rot) + 1u) & 31u;
        return (shifted >> rot) | (shifted << lsh);
    }
    __host__ __device__ __forceinline__ void seed(uint64_t seed_state, uint64_t seq) {
        state = 0; inc = (seq << 1) | 1;
        next_u32();
        state += seed_state;
        next_u32();
    }
    __host__ __device__ __forceinline__ float uniform() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
};

// ----------------------------------------------------------------------------------------------
// Gaussian -- include/gaussian.h.  Device record: 3 x float4 (48 B, 128-bit loads)
//   q0 = (mean.x, mean.y, mean.z, m00)  q1 = (m01, m02, m11, m12)  q2 = (m22, density, norm, albedo)
// ----------------------------------------------------------------------------------------------
struct GaussRec {
    f3 mean;
    float m00, m01, m02, m11, m12, m22;
    float density, norm, albedo;
};
__device__ __forceinline__ GaussRec load_gauss(const float4* __restrict__ g, uint32_t i) {
    float4 q0 = __ldg(g + 3 * (size_t)i), q1 = __ldg(g + 3 * (size_t)i + 1), q2 = __ldg(g + 3 * (size_t)i + 2);
    GaussRec r;
    r.mean = mk3(q0.x, q0.y, q0.z);
    r.m00 = q0.w; r.m01 = q1.x; r.m02 = q1.y; r.m11 = q1.z; r.m12 = q1.w; r.m22 = q2.x;
    r.density = q2.y; r.norm = q2.z; r.albedo = q2.w;
    return r;
}
__device__ __forceinline__ f3 mulM(const GaussRec& g, f3 v) {
    return mk3(row3(g.m00, g.m01, g.m02, v), row3(g.m01, g.m11, g.m12, v), row3(g.m02, g.m12, g.m22, v));
}
// quadratic coefficients of the ray against the 3-sigma ellipsoid, gaussian.h:126-140 / :208-219
struct RayQuad { float A, B, C0; };  // C0 = p^T M p (the intersection uses C0 - 9)
__device__ __forceinline__ RayQuad ray_quad(const GaussRec& g, f3 o, f3 d) {
    f3 p = sub3(o, g.mean);
    f3 Md = mulM(g, d);
    f3 Mp = mulM(g, p);
    RayQuad q;
    q.A = dot3(d, Md);
    q.B = __fmul_rn(2.0f, dot3(p, Md));
    q.C0 = dot3(p, Mp);
    return q;
}
// gaussian.h:141-164 (R = 3).  t_enter is clamped to 0.
__device__ __forceinline__ bool quad_intersect(const RayQuad& q, float* t_enter, float* t_exit) {
    float C = __fsub_rn(q.C0, 9.0f);
    float disc = __fsub_rn(__fmul_rn(q.B, q.B), __fmul_rn(__fmul_rn(4.0f, q.A), C));
    if (disc < 0.0f) return false;
    float sq = __fsqrt_rn(disc);
    float twoA = __fmul_rn(2.0f, q.A);
    float t0 = __fdiv_rn(__fsub_rn(-q.B, sq), twoA);
    float t1 = __fdiv_rn(__fadd_rn(-q.B, sq), twoA);
    if (t0 > t1) { float t = t0; t0 = t1; t1 = t; }
    if (t1 < 0.0f) return false;
    *t_enter = (t0 >= 0.0f) ? t0 : 0.0f;
    *t_exit = t1;
    return true;
}
// per-(ray, Gaussian) constants of optical_depth (gaussian.h:208-231): tau(ta,tb) = K * (erf(arg(tb)) - erf(arg(ta)))
struct DepthConst { float A, B, K, den; };
__device__ __forceinline__ DepthConst depth_const(const GaussRec& g, const RayQuad& q) {
    DepthConst c;
    c.A = q.A; c.B = q.B;
    float dn = __fmul_rn(g.density, g.norm);
    float pref = (float)__dmul_rn((double)dn, __dsqrt_rn(__ddiv_rn(VR_PI_D, (double)__fmul_rn(2.0f, q.A))));
    float e = __fmul_rn(-0.5f, __fsub_rn(q.C0, __fdiv_rn(__fmul_rn(q.B, q.B), __fmul_rn(4.0f, q.A))));
    c.K = __fmul_rn(pref, expf(e));
    c.den = __fmul_rn(2.0f, __fsqrt_rn(__fmul_rn(2.0f, q.A)));
    return c;
}
__device__ __forceinline__ float erf_arg(const DepthConst& c, float t) {
    return __fdiv_rn(__fadd_rn(c.B, __fmul_rn(__fmul_rn(2.0f, c.A), t)), c.den);
}
// one out-of-line copy of erff per kernel: it is evaluated at a dozen call sites and inlining each
// (~60 instructions) bloated the path kernels past the instruction caches (profiles/r01_v0)
static __device__ __noinline__ float erff_ni(float x) { return erff(x); }
__device__ __forceinline__ float depth_between(const DepthConst& c, float t0, float t1) {
    float F1 = erff_ni(erf_arg(c, t1));
    float F0 = erff_ni(erf_arg(c, t0));
    return __fmul_rn(c.K, __fsub_rn(F1, F0));
}
// gaussian.h:111-117
__device__ __forceinline__ float gauss_mu_t(const GaussRec& g, f3 x) {
    f3 d = sub3(x, g.mean);
    f3 h = mk3(__fmul_rn(-0.5f, d.x), __fmul_rn(-0.5f, d.y), __fmul_rn(-0.5f, d.z));
    f3 hm = mk3(row3(g.m00, g.m01, g.m02, h), row3(g.m01, g.m11, g.m12, h), row3(g.m02, g.m12, g.m22, h));
    float e = dot3(hm, d);
    return __fmul_rn(g.density, __fmul_rn(g.norm, expf(e)));
}
// gaussian.h:10-25 (Winitzki a = 0.14; the approximation error is part of the reference's result, F6)
__device__ __forceinline__ double erfinv_approx(double x) {
    if (isnan(x)) return CUDART_NAN;
    if (x <= -1.0) return -CUDART_INF;
    if (x >= 1.0) return CUDART_INF;
    const double a = 0.14;
    double sign = (x < 0.0) ? -1.0 : 1.0;
    double ln_term = log(__dsub_rn(1.0, __dmul_rn(x, x)));
    double first = __dadd_rn(2.0 / (VR_PI_D * a), __ddiv_rn(ln_term, 2.0));
    double inside = __dsub_rn(__dmul_rn(first, first), __ddiv_rn(ln_term, a));
    if (inside < 0.0) inside = 0.0;
    return sign * __dsqrt_rn(__dsub_rn(__dsqrt_rn(inside), first));
}
// gaussian.h:235-297, all double
static __device__ __noinline__ bool solve_t_given_tau(const GaussRec& g, const RayQuad& q, float t0, float tb, float target_tau,
                                                      float* t_out) {
    double A = (double)q.A;
    if (!(A > 0.0) || !isfinite(A)) return false;
    double B = (double)q.B;  // 2.0 * (double)dot == (double)(2.0f * dot): scaling by 2 is exact
    double C = (double)q.C0;
    double sqrtA = __dsqrt_rn(A);
    double pref = __dmul_rn(__dmul_rn((double)g.density, (double)g.norm), __dsqrt_rn(__ddiv_rn(VR_PI_D, __dmul_rn(2.0, A))));
    double ef = exp(__dmul_rn(-0.5, __dsub_rn(C, __ddiv_rn(__dmul_rn(B, B), __dmul_rn(4.0, A)))));
    double denom = __dmul_rn(pref, ef);
    if (!(denom > 0.0) || !isfinite(denom)) return false;
    double k = __dmul_rn(__dmul_rn(2.0, __dsqrt_rn(2.0)), sqrtA);
    double erf_t0 = erf(__ddiv_rn(__dadd_rn(B, __dmul_rn(__dmul_rn(2.0, A), (double)t0)), k));
    double target_erf = __dadd_rn(__ddiv_rn((double)target_tau, denom), erf_t0);
    const double one_eps = 1.0 - 1e-14;
    if (target_erf >= one_eps) { *t_out = tb; return true; }
    if (target_erf <= -one_eps) { *t_out = t0; return true; }
    if (!isfinite(target_erf)) return false;
    double arg_t = erfinv_approx(target_erf);
    double tc = __ddiv_rn(__dsub_rn(__dmul_rn(k, arg_t), B), __dmul_rn(2.0, A));
    if (!isfinite(tc)) return false;
    if (tc < (double)t0 - 1e-6) tc = (double)t0;
    if (tc > (double)tb + 1e-6) tc = (double)tb;
    *t_out = (float)tc;
    return true;
}

// ----------------------------------------------------------------------------------------------
// Camera -- include/camera.h:45-73 (+ Ray ctor normalisation, ray.h:11-12)
// ----------------------------------------------------------------------------------------------
struct CameraPod { int kind; float pos[3], view[3], right[3], up[3], pinhole[3]; };
struct RayOD { f3 o, d; };
__device__ __forceinline__ RayOD camera_ray(const CameraPod& c, float uu, float vv) {
    f3 pos = mk3(c.pos[0], c.pos[1], c.pos[2]);
    f3 right = mk3(c.right[0], c.right[1], c.right[2]);
    f3 up = mk3(c.up[0], c.up[1], c.up[2]);
    RayOD r;
    if (c.kind == 0) {
        float u = __fsub_rn(1.0f, __fmul_rn(uu, 2.0f));
        float v = __fsub_rn(__fmul_rn(vv, 2.0f), 1.0f);
        r.o = add3(add3(pos, scale3(u, right)), scale3(v, up));
        f3 d = sub3(mk3(c.pinhole[0], c.pinhole[1], c.pinhole[2]), r.o);
        r.d = normalized3(normalized3(d));
    } else {
        float u = __fsub_rn(__fmul_rn(uu, 2.0f), 1.0f);
        float v = __fsub_rn(1.0f, __fmul_rn(vv, 2.0f));
        r.o = add3(add3(pos, scale3(u, right)), scale3(v, up));
        r.d = normalized3(normalized3(mk3(c.view[0], c.view[1], c.view[2])));
    }
    return r;
}
// integrator.h:32-44
__device__ __forceinline__ f3 sample_uniform_direction(Pcg& rng) {
    float xi1 = rng.uniform();
    float xi2 = rng.uniform();
    float theta = (float)__dmul_rn(6.283185307179586, (double)xi1);  // 2.0f * pi(double) * xi1
    float phi = ref_acosf(__fsub_rn(1.0f, __fmul_rn(2.0f, xi2)));
    float sp = ref_sinf(phi);
    return mk3(__fmul_rn(sp, ref_cosf(theta)), __fmul_rn(sp, ref_sinf(theta)), ref_cosf(phi));
}

}  // namespace vr

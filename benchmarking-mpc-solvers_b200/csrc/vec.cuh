// vec.cuh — the arithmetic vocabulary of the float32 rollout code: one set of functions for a scalar lane (float)
// and for a PACKED PAIR of lanes (f2).  On sm_100a the packed pair maps to FFMA2 / FADD2 / FMUL2 — two FP32
// operations per issued instruction at the scalar pipe throughput (tools/pipebench.cu) — so a thread that rolls out
// two samples spends half the issue slots on its FMA-pipe work.  min/max, MUFU and sign manipulation have no packed
// form and are done per lane.
//
// Transcendentals are written for the bounded arguments of the three models (angles of a few tens of radians at
// most), with no slow path:
//   vsincos<false>  reduction modulo pi by the round-to-nearest magic number (no F2I/I2F on the XU pipe), two-term
//                   Cody-Waite with FMA, minimax polynomials on [-pi/2, pi/2] (sin degree 9: 1.5e-7 abs, cos degree
//                   10: 0.9e-7 abs), sign (-1)^n applied with integer ops
//   vsincos<true>   MUFU.SIN / MUFU.COS (abs err ~4e-7 on [-pi, pi]) — the "fast_math" flavour
//   vrcp            MUFU.RCP (1 ulp); quotients are formed as a * vrcp(b)
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace mpc {

struct f2 {
    float2 v;
    __device__ __forceinline__ f2() {}
    __device__ __forceinline__ f2(float s) : v(make_float2(s, s)) {}  // broadcast
    __device__ __forceinline__ f2(float a, float b) : v(make_float2(a, b)) {}
    __device__ __forceinline__ explicit f2(float2 q) : v(q) {}
};

template <typename V>
struct VT;
template <>
struct VT<float> {
    static constexpr int N = 1;
    typedef float S;
};
template <>
struct VT<f2> {
    static constexpr int N = 2;
    typedef float S;
};
template <>
struct VT<double> {
    static constexpr int N = 1;
    typedef double S;
};

// ---- lane access
__device__ __forceinline__ float lane(float a, int) { return a; }
__device__ __forceinline__ double lane(double a, int) { return a; }
__device__ __forceinline__ float lane(const f2 &a, int i) { return i ? a.v.y : a.v.x; }
template <typename V>
__device__ __forceinline__ typename VT<V>::S lane_of(const V &a, int i) { return lane(a, i); }  // where `lane` is taken

// ---- arithmetic
__device__ __forceinline__ float vfma(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double vfma(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ f2 vfma(const f2 &a, const f2 &b, const f2 &c) { return f2(__ffma2_rn(a.v, b.v, c.v)); }
__device__ __forceinline__ f2 operator+(const f2 &a, const f2 &b) { return f2(__fadd2_rn(a.v, b.v)); }
__device__ __forceinline__ f2 operator-(const f2 &a) { return f2(-a.v.x, -a.v.y); }
__device__ __forceinline__ f2 operator-(const f2 &a, const f2 &b) { return f2(__fadd2_rn(a.v, make_float2(-b.v.x, -b.v.y))); }
__device__ __forceinline__ f2 operator*(const f2 &a, const f2 &b) { return f2(__fmul2_rn(a.v, b.v)); }

__device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ double vmin(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ double vmax(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ f2 vmin(const f2 &a, const f2 &b) { return f2(fminf(a.v.x, b.v.x), fminf(a.v.y, b.v.y)); }
__device__ __forceinline__ f2 vmax(const f2 &a, const f2 &b) { return f2(fmaxf(a.v.x, b.v.x), fmaxf(a.v.y, b.v.y)); }
template <typename V>
__device__ __forceinline__ V vclip(const V &a, const V &lo, const V &hi) {
    return vmin(vmax(a, lo), hi);
}
__device__ __forceinline__ float vabs(float a) { return fabsf(a); }
__device__ __forceinline__ f2 vabs(const f2 &a) { return f2(fabsf(a.v.x), fabsf(a.v.y)); }

__device__ __forceinline__ float rcp_approx(float a) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
__device__ __forceinline__ float vrcp(float a) { return rcp_approx(a); }
__device__ __forceinline__ f2 vrcp(const f2 &a) { return f2(rcp_approx(a.v.x), rcp_approx(a.v.y)); }

// ---- bit tricks
__device__ __forceinline__ float xor_bits(float a, uint32_t m) { return __uint_as_float(__float_as_uint(a) ^ m); }

constexpr float RN_MAGIC = 12582912.0f;  // 1.5 * 2^23: (x + M) - M == rint(x) for |x| < 2^22, n in the low mantissa bits
constexpr float F_PI_HI = 3.14159274f, F_PI_LO = -8.74227766e-08f, F_INV_PI = 0.318309873f;
constexpr float F_2PI_HI = 6.28318548f, F_2PI_LO = -1.74845553e-07f, F_INV_2PI = 0.159154937f;
constexpr float F_HALF_PI = 1.57079637f;

// x - 2*pi*rint(x / 2pi): the angle wrapped to [-pi, pi]
template <typename V>
__device__ __forceinline__ V vwrap_pi(const V &x) {
    const V t = vfma(x, V(F_INV_2PI), V(RN_MAGIC));
    const V nf = t - V(RN_MAGIC);
    V r = vfma(nf, V(-F_2PI_HI), x);
    return vfma(nf, V(-F_2PI_LO), r);
}

// floor(y) for |y| < 2^22 without the XU pipe: round to nearest, step down where that went up
__device__ __forceinline__ float vfloor(float y) {
    const float n = (y + RN_MAGIC) - RN_MAGIC;
    return n - (n > y ? 1.0f : 0.0f);
}
__device__ __forceinline__ f2 vfloor(const f2 &y) {
    const f2 n = (y + f2(RN_MAGIC)) - f2(RN_MAGIC);
    return n - f2(n.v.x > y.v.x ? 1.0f : 0.0f, n.v.y > y.v.y ? 1.0f : 0.0f);
}

// minimax coefficients on [-pi/2, pi/2] (fitted in float64, rounded to float32, errors measured in float32 evaluation)
constexpr float SC_S1 = -0.1666666567325592f, SC_S2 = 0.008333242498338223f, SC_S3 = -0.0001982271351153031f,
                SC_S4 = 2.6346453978476347e-06f;
constexpr float SC_C1 = -0.5f, SC_C2 = 0.0416666604578495f, SC_C3 = -0.0013888678513467312f,
                SC_C4 = 2.477601810824126e-05f, SC_C5 = -2.6323152724216925e-07f;

template <typename V>
__device__ __forceinline__ V sin_poly(const V &r, const V &r2) {
    V p = vfma(V(SC_S4), r2, V(SC_S3));
    p = vfma(p, r2, V(SC_S2));
    p = vfma(p, r2, V(SC_S1));
    return vfma(p, r * r2, r);
}
template <typename V>
__device__ __forceinline__ V cos_poly(const V &r2) {
    V p = vfma(V(SC_C5), r2, V(SC_C4));
    p = vfma(p, r2, V(SC_C3));
    p = vfma(p, r2, V(SC_C2));
    p = vfma(p, r2, V(SC_C1));
    return vfma(p, r2, V(1.0f));
}

__device__ __forceinline__ void flip_by_parity(float t, float *s, float *c) {
    const uint32_t sg = __float_as_uint(t) << 31;  // low mantissa bit of (x/pi + M) = parity of n
    *s = xor_bits(*s, sg);
    *c = xor_bits(*c, sg);
}

template <bool FAST>
__device__ __forceinline__ void vsincos(float x, float *s, float *c) {
    if (FAST) {
        *s = __sinf(x);
        *c = __cosf(x);
        return;
    }
    const float t = fmaf(x, F_INV_PI, RN_MAGIC);
    const float nf = t - RN_MAGIC;
    float r = fmaf(nf, -F_PI_HI, x);
    r = fmaf(nf, -F_PI_LO, r);
    const float r2 = r * r;
    *s = sin_poly(r, r2);
    *c = cos_poly(r2);
    flip_by_parity(t, s, c);
}
template <bool FAST>
__device__ __forceinline__ void vsincos(const f2 &x, f2 *s, f2 *c) {
    if (FAST) {
        *s = f2(__sinf(x.v.x), __sinf(x.v.y));
        *c = f2(__cosf(x.v.x), __cosf(x.v.y));
        return;
    }
    const f2 t = vfma(x, f2(F_INV_PI), f2(RN_MAGIC));
    const f2 nf = t - f2(RN_MAGIC);
    f2 r = vfma(nf, f2(-F_PI_HI), x);
    r = vfma(nf, f2(-F_PI_LO), r);
    const f2 r2 = r * r;
    *s = sin_poly(r, r2);
    *c = cos_poly(r2);
    flip_by_parity(t.v.x, &s->v.x, &c->v.x);
    flip_by_parity(t.v.y, &s->v.y, &c->v.y);
}

// ---- angles in half-turns (p = theta / pi): the reduction p - rint(p) is three exact additions, pi lives in the
// polynomial coefficients.  Minimax fits on |f| <= 1/2 (float64 fit, float32 evaluation): sin 1.5e-7, cos 1.2e-7 abs.
constexpr float SP_0 = 3.141592741e+00f, SP_1 = -5.167710304e+00f, SP_2 = 2.550077677e+00f, SP_3 = -5.982926488e-01f,
                SP_4 = 7.766041905e-02f;
constexpr float CP_1 = -4.934802055e+00f, CP_2 = 4.058711529e+00f, CP_3 = -1.335241079e+00f, CP_4 = 2.350802720e-01f,
                CP_5 = -2.463775314e-02f;

template <typename V>
__device__ __forceinline__ void sincospi_poly(const V &f, V *s, V *c) {
    const V u = f * f;
    V p = vfma(V(SP_4), u, V(SP_3));
    p = vfma(p, u, V(SP_2));
    p = vfma(p, u, V(SP_1));
    p = vfma(p, u, V(SP_0));
    *s = p * f;
    V q = vfma(V(CP_5), u, V(CP_4));
    q = vfma(q, u, V(CP_3));
    q = vfma(q, u, V(CP_2));
    q = vfma(q, u, V(CP_1));
    *c = vfma(q, u, V(1.0f));
}

// sin(pi p), cos(pi p)
template <bool FAST>
__device__ __forceinline__ void vsincospi(float p, float *s, float *c) {
    if (FAST) {
        *s = __sinf(p * F_PI_HI);
        *c = __cosf(p * F_PI_HI);
        return;
    }
    const float t = p + RN_MAGIC;
    const float f = p - (t - RN_MAGIC);
    sincospi_poly(f, s, c);
    flip_by_parity(t, s, c);
}
template <bool FAST>
__device__ __forceinline__ void vsincospi(const f2 &p, f2 *s, f2 *c) {
    if (FAST) {
        *s = f2(__sinf(p.v.x * F_PI_HI), __sinf(p.v.y * F_PI_HI));
        *c = f2(__cosf(p.v.x * F_PI_HI), __cosf(p.v.y * F_PI_HI));
        return;
    }
    const f2 t = p + f2(RN_MAGIC);
    const f2 f = p - (t - f2(RN_MAGIC));
    sincospi_poly(f, s, c);
    flip_by_parity(t.v.x, &s->v.x, &c->v.x);
    flip_by_parity(t.v.y, &s->v.y, &c->v.y);
}

// p - 2 rint(p / 2): half-turns wrapped to [-1, 1]
template <typename V>
__device__ __forceinline__ V vwrap_halfturns(const V &p) {
    const V t = vfma(p, V(0.5f), V(RN_MAGIC));
    return vfma(t - V(RN_MAGIC), V(-2.0f), p);
}

// sin of an angle already in [-pi, pi]: sin(x) = sign(x) * sin(min(|x|, pi - |x|)), argument in [0, pi/2]
template <bool FAST>
__device__ __forceinline__ float vsin_wrapped(float x) {
    if (FAST) return __sinf(x);
    const float a = fabsf(x);
    const float r = fminf(a, (F_PI_HI - a) + F_PI_LO);
    return xor_bits(sin_poly(r, r * r), __float_as_uint(x) & 0x80000000u);
}
template <bool FAST>
__device__ __forceinline__ f2 vsin_wrapped(const f2 &x) {
    if (FAST) return f2(__sinf(x.v.x), __sinf(x.v.y));
    const f2 a = vabs(x);
    const f2 r = vmin(a, (f2(F_PI_HI) - a) + f2(F_PI_LO));
    const f2 s = sin_poly(r, r * r);
    return f2(xor_bits(s.v.x, __float_as_uint(x.v.x) & 0x80000000u), xor_bits(s.v.y, __float_as_uint(x.v.y) & 0x80000000u));
}

}  // namespace mpc

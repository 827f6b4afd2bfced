// envs.cuh — the three MPCBenchmark models as inlined device code: cost of (x_t, u_t) then one dynamics step,
// exactly the order of Model.predict (models/model.py:26-44), state kept in registers by the caller.
//
//   Pendulum   models/pendulum_model.py:28-50,58-59
//   CartPole   models/cartpole_swingup_model.py:63-104
//   Acrobot    models/acrobot_model.py:99-176,185-220,270-282
//
// Each model is a template over the lane type V (vec.cuh):
//  * V = float or f2 (two samples per thread, packed FFMA2/FADD2/FMUL2) — throughput flavour, written for the GPU,
//    not transliterated: constants folded on the host (EnvConst, layout "f" in capi.cu:fold_env_params), the diagonal
//    cost as a sum of squares of sqrt(W_i)-scaled residuals (one FMA per residual, one per square), trig shared
//    (cos(a - pi/2) = sin a, angle addition for theta1 + theta2), x^10 by squaring, all quotients of one `_dsdt` /
//    one CartPole step through ONE reciprocal, angle wraps by round-to-nearest instead of loops, no slow paths.
//    TRIG selects MUFU.SIN/COS instead of the polynomial sincos (0 = polynomial, 1 = MUFU; Acrobot: per-stage splits).
//  * V = double (parity mode, compiled with -fmad=false): the DYNAMICS follow the reference's numpy expression tree
//    operation by operation (same association, true divisions, cos(th1+th2-pi/2) evaluated as written, fmod-based
//    floored remainder, repeated 2*pi wrap), because these systems are chaotic: an ulp of difference per step is
//    amplified ~1e4x over a 100-step Acrobot rollout and would eat the 1e-12 parity budget.  Costs do not feed
//    back into the state, so they keep the cheaper forms (x^10 by squaring).
// Every rewrite is validated against the oracle and the reference's golden vectors (tests/test_gpu_parity.py).
//
// Interface:  Env::NS, Env::NG, Env::CHUNK (steps whose costs are summed in V before they reach the float64 sum),
//   prep_goal(k, g_row[NS+1] float64, gs[NG])   per-solve goal terms in the form step() consumes
//   step(k, x, u, gs) -> cost of (x, u); advances x in place
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <type_traits>

#include "vec.cuh"

namespace mpc {

enum { ENV_PENDULUM = 0, ENV_CARTPOLE = 1, ENV_ACROBOT = 2 };

// Folded model constants, one flat array per env (see fold_env_params in capi.cu for both layouts).
template <typename V>
struct EnvConst {
    V c[24];
    V w[15];  // generic-cost flavour (GEN = 1): upper triangle, row major, of the symmetrised effective cost matrix
              // (w_ii = s W_ii, w_ij = s (W_ij + W_ji), s = -1 for CartPole whose _state_cost returns -costs)
};

// the one-sample view of the constants (lane 0 of every packed pair)
template <typename V>
__device__ __forceinline__ EnvConst<typename VT<V>::S> scalar_consts(const EnvConst<V> &k) {
    EnvConst<typename VT<V>::S> o;
#pragma unroll
    for (int i = 0; i < 24; ++i) o.c[i] = lane(k.c[i], 0);
#pragma unroll
    for (int i = 0; i < 15; ++i) o.w[i] = lane(k.w[i], 0);
    return o;
}

template <typename V>
__device__ __forceinline__ typename VT<V>::S k0(const EnvConst<V> &k, int i) {
    return lane(k.c[i], 0);
}

__device__ __forceinline__ double clipd(double a, double lo, double hi) { return fmin(fmax(a, lo), hi); }

// zd^T W zd for a full (non-diagonal) W, as einsum("bi,ij,bj->b") of the models' _state_cost
// (pendulum_model.py:46-50, cartpole_swingup_model.py:95-104, acrobot_model.py:172-176), through the symmetrised
// upper triangle k.w: sum_i zd_i (sum_{j >= i} w_ij zd_j)
template <int NZ, typename V>
__device__ __forceinline__ V quad_form(const EnvConst<V> &k, const V (&zd)[NZ]) {
    V cost = V((typename VT<V>::S)0);
    int idx = 0;
#pragma unroll
    for (int i = 0; i < NZ; ++i) {
        V row = k.w[idx++] * zd[i];
#pragma unroll
        for (int j = i + 1; j < NZ; ++j) row = vfma(k.w[idx++], zd[j], row);
        cost = vfma(zd[i], row, cost);
    }
    return cost;
}
constexpr double D_PI = 3.14159265358979323846;

// ------------------------------------------------------------------------------------------------ Pendulum
// float layout f: 0 A = -c_sin*dt  1 B = c_u*dt  2 dt  3 max_speed  4 max_torque  5..7 sqrt(W0..W2)
//   with c_sin = -3g/(2l), c_u = 3/(m l^2)
template <typename V, int TRIG, int GEN = 0>  // GEN: generic cost (full W, goal rows in z units; envs.cuh header)
struct Pendulum {
    static constexpr int NS = 2, NG = 3, CHUNK = 4;
    static constexpr int ENV = ENV_PENDULUM;
    static constexpr int GENERIC = GEN;
    typedef typename VT<V>::S S;
    typedef V Vec;
    typedef Pendulum<S, 0, GEN> Scalar;  // one sample per thread, accurate trig: nominal rollouts
    // reference units <-> the units step() keeps the state in (identity here)
    static __device__ __forceinline__ void enter(const EnvConst<V> &, S *) {}
    static __device__ __forceinline__ void leave(const EnvConst<V> &, const S *x, double *out) {
#pragma unroll
        for (int i = 0; i < NS; ++i) out[i] = (double)x[i];
    }
    // cost = sum W_i (z_i - g_i)^2 with z = -[thn, om, u]  ==  sum (sW_i*[thn, om, u]_i + sW_i*g_i)^2
    static __device__ __forceinline__ void prep_goal(const EnvConst<V> &k, const double *g, S *gs) {
#pragma unroll
        for (int i = 0; i < 3; ++i) gs[i] = GEN ? (S)g[i] : k0(k, 5 + i) * (S)g[i];
    }
    static __device__ __forceinline__ V step(const EnvConst<V> &k, V (&x)[NS], const V &u, const S *gs) {
        const V th = x[0], om = x[1];
        // angle_normalize: ((th + pi) mod 2pi) - pi == th - 2pi*floor((th + pi)/2pi).  Round-to-nearest instead of
        // floor differs only when th is an odd multiple of pi exactly (+pi instead of -pi): same cost (thn^2) and
        // |sin| < 1e-7 either way, and it needs no compare/select on the ALU pipe.
        const V thn = vwrap_pi(th);
        V cost;
        if constexpr (GEN) {
            const V zd[3] = {-thn - V(gs[0]), -om - V(gs[1]), -u - V(gs[2])};  // z = -[thn, om, u]; zd = z - g
            cost = quad_form<3>(k, zd);
        } else {
            const V d0 = vfma(k.c[5], thn, V(gs[0])), d1 = vfma(k.c[6], om, V(gs[1])), d2 = vfma(k.c[7], u, V(gs[2]));
            cost = vfma(d0, d0, vfma(d1, d1, d2 * d2));   // u NOT clipped in the cost   pendulum_model.py:40-50
        }
        // dynamics: sin(th + pi) = -sin(th) = -sin(thn)                                      pendulum_model.py:28-38
        const V uc = vclip(u, -k.c[4], k.c[4]);
        const V s = vsin_wrapped<(TRIG != 0)>(thn);
        const V om2 = vfma(k.c[0], s, vfma(k.c[1], uc, om));
        x[0] = vfma(om2, k.c[2], th);  // integrates with the UNclipped new speed
        x[1] = vclip(om2, -k.c[3], k.c[3]);
        return cost;
    }
};

// parity layout c: 0 c_sin=-3g/(2l)  1 c_u=3/(m l^2)  2 dt  3 max_speed  4 max_torque  5..7 W0..W2
template <int TRIG, int GEN>
struct Pendulum<double, TRIG, GEN> {
    static constexpr int NS = 2, NG = 3, CHUNK = 1;
    static constexpr int ENV = ENV_PENDULUM;
    static constexpr int GENERIC = GEN;
    typedef double S;
    typedef double Vec;
    typedef Pendulum<double, TRIG, GEN> Scalar;
    // reference units <-> the units step() keeps the state in (identity here)
    static __device__ __forceinline__ void enter(const EnvConst<double> &, double *) {}
    static __device__ __forceinline__ void leave(const EnvConst<double> &, const double *x, double *out) {
#pragma unroll
        for (int i = 0; i < NS; ++i) out[i] = (double)x[i];
    }
    static __device__ __forceinline__ void prep_goal(const EnvConst<double> &, const double *g, double *gs) {
#pragma unroll
        for (int i = 0; i < 3; ++i) gs[i] = g[i];
    }
    // numpy's expression tree (pendulum_model.py:28-50,58-59)
    static __device__ __forceinline__ double step(const EnvConst<double> &k, double (&x)[NS], double u, const double *g) {
        const double th = x[0], om = x[1];
        const double two_pi = 2 * D_PI;
        double r = fmod(th + D_PI, two_pi);  // np.remainder: fmod, then sign fix-up
        if (r != 0.0) {
            if (r < 0) r += two_pi;
        } else {
            r = 0.0;
        }
        const double thn = r - D_PI;
        const double d0 = -thn - g[0], d1 = -om - g[1], d2 = -u - g[2];
        double cost;
        if constexpr (GEN) {
            const double zd[3] = {d0, d1, d2};
            cost = quad_form<3>(k, zd);
        } else {
            cost = (d0 * k.c[5]) * d0 + (d1 * k.c[6]) * d1 + (d2 * k.c[7]) * d2;
        }
        const double uc = clipd(u, -k.c[4], k.c[4]);
        double om2 = om + (k.c[0] * sin(th + D_PI) + k.c[1] * uc) * k.c[2];
        x[0] = th + om2 * k.c[2];
        x[1] = clipd(om2, -k.c[3], k.c[3]);
        return cost;
    }
};

// ------------------------------------------------------------------------------------------------ CartPole
// float layout f: 0 1/x_threshold  1 4*force_mag  2 dt  3 6*force_mag/l
//    4 A1=-2 m_p l  5 A2=3 m_p g  6 A4=4 b  7 D1=4 M  8 D2=3 m_p
//    9 B1/l=-3 m_p  10 B2/l=6 M g/l  11 6 b/l  12 u clip (1)   13..17 sqrt(-W_i)
//   (theta_acc's denominator 4 l M - 3 m_p l c^2 is l times x_acc's, so one reciprocal serves both)
template <typename V, int TRIG, int GEN = 0>
struct CartPole {
    static constexpr int NS = 4, NG = 5, CHUNK = 4;
    static constexpr int ENV = ENV_CARTPOLE;
    static constexpr int GENERIC = GEN;
    typedef typename VT<V>::S S;
    typedef V Vec;
    typedef CartPole<S, 0, GEN> Scalar;  // one sample per thread, accurate trig: nominal rollouts
    // reference units <-> the units step() keeps the state in (identity here)
    static __device__ __forceinline__ void enter(const EnvConst<V> &, S *) {}
    static __device__ __forceinline__ void leave(const EnvConst<V> &, const S *x, double *out) {
#pragma unroll
        for (int i = 0; i < NS; ++i) out[i] = (double)x[i];
    }
    // z = [xc, pd, 1 - cos th, thd, u]; residual_i = sW_i (z_i - g_i)
    static __device__ __forceinline__ void prep_goal(const EnvConst<V> &k, const double *g, S *gs) {
        if constexpr (GEN) {
#pragma unroll
            for (int i = 0; i < 5; ++i) gs[i] = (S)g[i];
        } else {
            gs[0] = -k0(k, 13) * (S)g[0];
            gs[1] = -k0(k, 14) * (S)g[1];
            gs[2] = k0(k, 15) * (S)(1.0 - g[2]);
            gs[3] = -k0(k, 16) * (S)g[3];
            gs[4] = -k0(k, 17) * (S)g[4];
        }
    }
    static __device__ __forceinline__ V step(const EnvConst<V> &k, V (&x)[NS], const V &u, const S *gs) {
        const V p = x[0], pd = x[1], th = x[2], thd = x[3];
        V s, c;
        vsincos<(TRIG != 0)>(th, &s, &c);
        // _transform: xc = q^2 + q^10, tc = 1 - cos(th)                      cartpole_swingup_model.py:87-93
        const V q = p * k.c[0];
        const V q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        const V xc = vfma(q8, q2, q2);
        V cost;
        if constexpr (GEN) {
            const V zd[5] = {xc - V(gs[0]), pd - V(gs[1]), (V(1.0f) - c) - V(gs[2]), thd - V(gs[3]), u - V(gs[4])};
            cost = quad_form<5>(k, zd);
        } else {
            const V d0 = vfma(k.c[13], xc, V(gs[0])), d1 = vfma(k.c[14], pd, V(gs[1])), d2 = vfma(-k.c[15], c, V(gs[2])),
                    d3 = vfma(k.c[16], thd, V(gs[3])), d4 = vfma(k.c[17], u, V(gs[4]));
            cost = vfma(d0, d0, vfma(d1, d1, vfma(d2, d2, vfma(d3, d3, d4 * d4))));
        }
        // _dynamics (explicit Euler, old velocities)                          cartpole_swingup_model.py:63-85
        const V uc = vclip(u, -k.c[12], k.c[12]);
        const V t1 = (thd * thd) * s, sc = s * c;
        V nx = vfma(k.c[4], t1, k.c[1] * uc);           // A1 thd^2 s + 4 a
        nx = vfma(k.c[5], sc, nx);                      // + A2 s c
        nx = vfma(-k.c[6], pd, nx);                     // - 4 b pd
        V nt = vfma(-k.c[11], pd, k.c[3] * uc) * c;     // 6 (a - b pd) c / l
        nt = vfma(k.c[10], s, nt);                      // + B2 s / l
        nt = vfma(k.c[9], t1 * c, nt);                  // + B1 thd^2 s c / l
        const V r = vrcp(vfma(-k.c[8], c * c, k.c[7])); // 1 / (4 M - 3 m_p c^2)
        x[0] = vfma(pd, k.c[2], p);
        x[2] = vfma(thd, k.c[2], th);
        x[1] = vfma(nx * r, k.c[2], pd);
        x[3] = vfma(nt * r, k.c[2], thd);
        return cost;
    }
};

// parity layout c: 0 1/x_threshold  1 force_mag  2 dt  3 b
//    4 A1=-2 m_p l  5 A2=3 m_p g  6 A4=4 b  7 D1=4 M  8 D2=3 m_p
//    9 B1=-3 m_p l 10 B2=6 M g   11 E1=4 l M 12 E2=3 m_p l
//   13..17 Wc_i = -W_i  (cost = -(zd W zd), W negative diagonal)   18 x_threshold
template <int TRIG, int GEN>
struct CartPole<double, TRIG, GEN> {
    static constexpr int NS = 4, NG = 5, CHUNK = 1;
    static constexpr int ENV = ENV_CARTPOLE;
    static constexpr int GENERIC = GEN;
    typedef double S;
    typedef double Vec;
    typedef CartPole<double, TRIG, GEN> Scalar;
    // reference units <-> the units step() keeps the state in (identity here)
    static __device__ __forceinline__ void enter(const EnvConst<double> &, double *) {}
    static __device__ __forceinline__ void leave(const EnvConst<double> &, const double *x, double *out) {
#pragma unroll
        for (int i = 0; i < NS; ++i) out[i] = (double)x[i];
    }
    static __device__ __forceinline__ void prep_goal(const EnvConst<double> &, const double *g, double *gs) {
#pragma unroll
        for (int i = 0; i < 5; ++i) gs[i] = g[i];
    }
    // numpy's expression tree (cartpole_swingup_model.py:63-104)
    static __device__ __forceinline__ double step(const EnvConst<double> &k, double (&x)[NS], double u, const double *g) {
        const double p = x[0], pd = x[1], th = x[2], thd = x[3];
        const double s = sin(th), c = cos(th);
        const double q = p / k.c[18];
        const double q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        const double z[5] = {q2 + q8 * q2, pd, 1 - c, thd, u};
        double acc = 0.0;
        if constexpr (GEN) {
            double zd[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) zd[i] = z[i] - g[i];
            acc = quad_form<5>(k, zd);
        } else {
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const double d = z[i] - g[i];
                acc += (d * k.c[13 + i]) * d;
            }
        }
        const double a = clipd(u, -1.0, 1.0) * k.c[1];
        const double thd2 = thd * thd, c2 = c * c;
        const double xacc = (((k.c[4] * thd2) * s + (k.c[5] * s) * c + 4 * a) - k.c[6] * pd) / (k.c[7] - k.c[8] * c2);
        const double tacc = ((((k.c[9] * thd2) * s) * c + k.c[10] * s) + (6 * (a - k.c[3] * pd)) * c) /
                            (k.c[11] - k.c[12] * c2);
        x[0] = p + pd * k.c[2];
        x[2] = th + thd * k.c[2];
        x[1] = pd + xacc * k.c[2];
        x[3] = thd + tacc * k.c[2];
        return acc;
    }
};

// ------------------------------------------------------------------------------------------------ Acrobot
// float layout f: 0 h  1 A = m1 lc1^2 + m2 (l1^2 + lc2^2) + I1 + I2   2 B = 2 m2 l1 lc2     (d1 = A + B cos th2)
//    3 C = m2 lc2^2 + I2   4 D = m2 l1 lc2                                                    (d2 = C + D cos th2)
//    5 P1 = (m1 lc1 + m2 l1) g   6 P2 = m2 lc2 g   7 max_vel1  8 max_vel2   9 h/2  10 h/6  11 sqrt(W0)  12 sqrt(W4)
//    13..15 W1..W3 (their residuals are constants of the solve)   16 h/(2 pi)  17 h/pi  18 h/(6 pi)
template <typename V, int TRIG, int GEN = 0>
struct Acrobot {
    static constexpr int NS = 4, NG = GEN ? 5 : 3, CHUNK = 2;
    static constexpr int ENV = ENV_ACROBOT;
    static constexpr int GENERIC = GEN;
    typedef typename VT<V>::S S;
    typedef V Vec;
    typedef Acrobot<S, 0, GEN> Scalar;  // one sample per thread, accurate trig: nominal rollouts
    // step() keeps the two angles in HALF-TURNS (theta / pi): range reductions become exact additions (vec.cuh).
    // The all-MUFU flavour (TRIG == 1) keeps them in RADIANS instead: MUFU.SIN/COS take radians (one FMUL each for their
    // own range scaling), so half-turns would cost a second multiply per stage angle — 8 instructions per step.
    static constexpr bool RAD = TRIG == 1;
    static __device__ __forceinline__ void enter(const EnvConst<V> &, S *x) {
        if constexpr (!RAD) {
            x[0] *= 0.318309886f;
            x[1] *= 0.318309886f;
        }
    }
    static __device__ __forceinline__ void leave(const EnvConst<V> &, const S *x, double *out) {
        out[0] = RAD ? (double)x[0] : (double)x[0] * D_PI;
        out[1] = RAD ? (double)x[1] : (double)x[1] * D_PI;
        out[2] = (double)x[2];
        out[3] = (double)x[3];
    }
    // z = [-cos th1 - cos(th1+th2) - 2, 0, 0, 0, u]                                   acrobot_model.py:162-176
    static __device__ __forceinline__ void prep_goal(const EnvConst<V> &k, const double *g, S *gs) {
        if constexpr (GEN) {
#pragma unroll
            for (int i = 0; i < 5; ++i) gs[i] = (S)g[i];
        } else {
            gs[0] = k0(k, 11) * (S)(2.0 + g[0]);  // residual_0 = -(sW0 (c1 + c12) + gs0)
            gs[1] = -k0(k, 12) * (S)g[4];
            gs[2] = k0(k, 13) * (S)(g[1] * g[1]) + k0(k, 14) * (S)(g[2] * g[2]) + k0(k, 15) * (S)(g[3] * g[3]);
        }
    }

    // sin/cos of RK4 stage STAGE's angle (half-turns).  TRIG picks, per stage, the polynomial sincospi (FMA pipe, 1.5e-7
    // abs) or MUFU.SIN/COS (XU pipe, ~4e-7 abs plus the rounding of the radian argument):
    //   0 polynomial everywhere        1 MUFU everywhere (argument unreduced)
    //   2 stage 1 polynomial, stages 2-4 MUFU       3 stages 2 and 3 (weight 2/6 each in the RK4 sum) polynomial, 1 and 4 MUFU
    //   4 MUFU everywhere, the unwrapped stage 2-4 angles first reduced modulo one turn (exact in half-turns), which
    //     halves the rounding of the radian argument        5 = 2 with the reduction of 4
    template <int STAGE>
    static __device__ __forceinline__ void stage_sincos(const V &p, V *s, V *c) {
        constexpr bool poly = TRIG == 0 || ((TRIG == 2 || TRIG == 5) && STAGE == 1) || (TRIG == 3 && (STAGE == 2 || STAGE == 3));
        constexpr bool reduce = (TRIG == 4 || TRIG == 5) && STAGE != 1;
        if constexpr (poly)
            vsincospi<false>(p, s, c);
        else if constexpr (reduce)
            vsincospi<true>(vwrap_halfturns(p), s, c);
        else if constexpr (RAD)
            vsincos<true>(p, s, c);
        else
            vsincospi<true>(p, s, c);
    }

    // "book" _dsdt (acrobot_model.py:99-132) given sin/cos of both angles; returns the angular accelerations.
    // The reference's two quotients are the solution of the 2x2 mass-matrix system
    //     [d1 d2; d2 C] [acc1; acc2] = [-phi1; n1],   n1 = a - D w1^2 s2 - phi2,
    // (substitute acc1 = -(d2 acc2 + phi1)/d1 into the second row), so with det = C d1 - d2^2 and ONE reciprocal
    //     acc2 = (d1 n1 + d2 phi1) / det,     acc1 = -(C phi1 + d2 n1) / det.
    static __device__ __forceinline__ void accel(const EnvConst<V> &k, const V &s1, const V &c1, const V &s2,
                                                 const V &c2, const V &w1, const V &w2, const V &a, V *a1, V *a2) {
        const V d1 = vfma(k.c[2], c2, k.c[1]);
        const V d2 = vfma(k.c[4], c2, k.c[3]);
        const V s12 = vfma(s1, c2, c1 * s2);  // sin(th1 + th2) = cos(th1 + th2 - pi/2)
        const V phi2 = k.c[6] * s12;
        const V Ds2 = k.c[4] * s2;
        // phi1 = -D w2^2 s2 - 2 D w2 w1 s2 + P1 sin(th1) + phi2
        const V phi1 = vfma(-(Ds2 * w2), vfma(V(2.0f), w1, w2), vfma(k.c[5], s1, phi2));
        const V r = vrcp(vfma(k.c[3], d1, -(d2 * d2)));
        const V n1 = vfma(-Ds2, w1 * w1, a) - phi2;
        *a2 = vfma(d2, phi1, n1 * d1) * r;
        *a1 = vfma(k.c[3], phi1, d2 * n1) * (-r);
    }

    static __device__ __forceinline__ V step(const EnvConst<V> &k, V (&x)[NS], const V &u, const S *gs) {
        const V p1 = x[0], p2 = x[1], w1 = x[2], w2 = x[3];  // p = theta / pi
        // angle increments: h / pi, h / (2 pi), h / (6 pi) in half-turns; h, h / 2, h / 6 in radians
        const V h = k.c[0], h2 = k.c[9], h6 = k.c[10], hp = RAD ? k.c[0] : k.c[17], h2p = RAD ? k.c[9] : k.c[16],
                h6p = RAD ? k.c[10] : k.c[18];
        V s1, c1, s2, c2;
        stage_sincos<1>(p1, &s1, &c1);
        stage_sincos<1>(p2, &s2, &c2);
        const V c12 = vfma(c1, c2, -(s1 * s2));
        V cost;
        if constexpr (GEN) {  // z = [-c1 - c12 - 2, 0, 0, 0, u]
            const V zd[5] = {(-(c1 + c12) - V(2.0f)) - V(gs[0]), -V(gs[1]), -V(gs[2]), -V(gs[3]), u - V(gs[4])};
            cost = quad_form<5>(k, zd);
        } else {
            const V d0 = vfma(k.c[11], c1 + c12, V(gs[0])), d4 = vfma(k.c[12], u, V(gs[1]));
            cost = vfma(d0, d0, vfma(d4, d4, V(gs[2])));
        }
        // one classical RK4 step, torque u (unclipped) held constant            acrobot_model.py:134-160,270-282
        V k1a, k1b, k2a, k2b, k3a, k3b, k4a, k4b;
        accel(k, s1, c1, s2, c2, w1, w2, u, &k1a, &k1b);
        const V k2w1 = vfma(h2, k1a, w1), k2w2 = vfma(h2, k1b, w2);
        stage_sincos<2>(vfma(h2p, w1, p1), &s1, &c1);
        stage_sincos<2>(vfma(h2p, w2, p2), &s2, &c2);
        accel(k, s1, c1, s2, c2, k2w1, k2w2, u, &k2a, &k2b);
        const V k3w1 = vfma(h2, k2a, w1), k3w2 = vfma(h2, k2b, w2);
        stage_sincos<3>(vfma(h2p, k2w1, p1), &s1, &c1);
        stage_sincos<3>(vfma(h2p, k2w2, p2), &s2, &c2);
        accel(k, s1, c1, s2, c2, k3w1, k3w2, u, &k3a, &k3b);
        const V k4w1 = vfma(h, k3a, w1), k4w2 = vfma(h, k3b, w2);
        stage_sincos<4>(vfma(hp, k3w1, p1), &s1, &c1);
        stage_sincos<4>(vfma(hp, k3w2, p2), &s2, &c2);
        accel(k, s1, c1, s2, c2, k4w1, k4w2, u, &k4a, &k4b);
        // y' = y + h/6 (k1 + 2 k2 + 2 k3 + k4): the velocity rows with the shared sum k2 + k3
        const V m1 = k2w1 + k3w1, m2 = k2w2 + k3w2, ma = k2a + k3a, mb = k2b + k3b;
        const V n0 = vfma(h6p, vfma(V(2.0f), m1, w1) + k4w1, p1);
        const V n1 = vfma(h6p, vfma(V(2.0f), m2, w2) + k4w2, p2);
        const V n2 = vfma(h6, vfma(V(2.0f), ma, k1a) + k4a, w1);
        const V n3 = vfma(h6, vfma(V(2.0f), mb, k1b) + k4b, w2);
        // wrap(.., -pi, pi) by repeated -+2pi  ==  nearest multiple of 2pi removed
        x[0] = RAD ? vwrap_pi(n0) : vwrap_halfturns(n0);
        x[1] = RAD ? vwrap_pi(n1) : vwrap_halfturns(n1);
        x[2] = vclip(n2, -k.c[7], k.c[7]);
        x[3] = vclip(n3, -k.c[8], k.c[8]);
        return cost;
    }
};

// parity layout c: 0 dt  1 A  2 B  3 C  4 D  5 P1  6 P2  7 max_vel1  8 max_vel2   9..13 W0..W4
//   14 m1 lc1^2  15 l1^2 + lc2^2  16 2 l1 lc2  17 m2  18 I1  19 I2  20 lc2^2  21 l1 lc2  22 D ("book") or 0 ("nips")
template <int TRIG, int GEN>
struct Acrobot<double, TRIG, GEN> {
    static constexpr int NS = 4, NG = 5, CHUNK = 1;
    static constexpr int ENV = ENV_ACROBOT;
    static constexpr int GENERIC = GEN;
    typedef double S;
    typedef double Vec;
    typedef Acrobot<double, TRIG, GEN> Scalar;
    // reference units <-> the units step() keeps the state in (identity here)
    static __device__ __forceinline__ void enter(const EnvConst<double> &, double *) {}
    static __device__ __forceinline__ void leave(const EnvConst<double> &, const double *x, double *out) {
#pragma unroll
        for (int i = 0; i < NS; ++i) out[i] = (double)x[i];
    }
    static __device__ __forceinline__ void prep_goal(const EnvConst<double> &, const double *g, double *gs) {
#pragma unroll
        for (int i = 0; i < 5; ++i) gs[i] = g[i];
    }

    // _dsdt: the reference's expression tree, term by term (acrobot_model.py:114-132)
    static __device__ __forceinline__ void dsdt_literal(const EnvConst<double> &k, const double (&y)[4], double a,
                                                        double (&d)[4]) {
        const double th1 = y[0], th2 = y[1], w1 = y[2], w2 = y[3];
        const double half_pi = D_PI / 2.;
        const double cth2 = cos(th2), sth2 = sin(th2);
        const double d1 = k.c[14] + k.c[17] * (k.c[15] + k.c[16] * cth2) + k.c[18] + k.c[19];
        const double d2 = k.c[17] * (k.c[20] + k.c[21] * cth2) + k.c[19];
        const double phi2 = k.c[6] * cos(th1 + th2 - half_pi);
        const double phi1 = ((-k.c[4] * (w2 * w2)) * sth2 - ((k.c[2] * w2) * w1) * sth2) + k.c[5] * cos(th1 - half_pi) + phi2;
        // c[22] = D for the "book" equations, 0 for "nips" (acrobot_model.py:124-132: the nips form has no w1^2 term;
        // subtracting its exact zero leaves the reference's value bit for bit)
        const double acc2 = (((a + (d2 / d1) * phi1) - (k.c[22] * (w1 * w1)) * sth2) - phi2) / (k.c[3] - (d2 * d2) / d1);
        const double acc1 = -(d2 * acc2 + phi1) / d1;
        d[0] = w1;
        d[1] = w2;
        d[2] = acc1;
        d[3] = acc2;
    }

    static __device__ __forceinline__ double wrap(double v) {  // acrobot_model.py:185-203 with m = -pi, M = pi
        const double diff = D_PI - (-D_PI);
        while (v > D_PI) v = v - diff;
        while (v < -D_PI) v = v + diff;
        return v;
    }

    // acrobot_model.py:134-176, rk4 at :270-282
    static __device__ __forceinline__ double step(const EnvConst<double> &k, double (&x)[NS], double u, const double *g) {
        const double z0 = -cos(x[0]) - cos(x[1] + x[0]) - 2;
        const double zz[5] = {z0, 0.0, 0.0, 0.0, u};
        double acc = 0.0;
        if constexpr (GEN) {
            double zd[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) zd[i] = zz[i] - g[i];
            acc = quad_form<5>(k, zd);
        } else {
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const double d = zz[i] - g[i];
                acc += (d * k.c[9 + i]) * d;
            }
        }
        const double dt = k.c[0] - 0.0, dt2 = dt / 2.0;
        const double y0[4] = {x[0], x[1], x[2], x[3]};
        double k1[4], k2[4], k3[4], k4[4], y[4];
        dsdt_literal(k, y0, u, k1);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = y0[i] + dt2 * k1[i];
        dsdt_literal(k, y, u, k2);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = y0[i] + dt2 * k2[i];
        dsdt_literal(k, y, u, k3);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = y0[i] + dt * k3[i];
        dsdt_literal(k, y, u, k4);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = y0[i] + dt / 6.0 * (k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i]);
        x[0] = wrap(y[0]);
        x[1] = wrap(y[1]);
        x[2] = fmin(fmax(y[2], -k.c[7]), k.c[7]);
        x[3] = fmin(fmax(y[3], -k.c[8]), k.c[8]);
        return acc;
    }
};

// ------------------------------------------------------------------------------------------------ model noise
// Model.predict of a model with sensor noise adds N(0, std) to every component of the new state (model.py:42-43),
// after clipping and wrapping.  unit(i) converts the reference's units to the ones step() keeps component i in.
template <class Env>
struct NoiseUnit {
    static __device__ __forceinline__ typename Env::S unit(int) { return (typename Env::S)1; }
};
template <int TRIG, int GEN>
struct NoiseUnit<Acrobot<float, TRIG, GEN>> {  // angles in half-turns
    static __device__ __forceinline__ float unit(int i) { return (i < 2 && TRIG != 1) ? 0.318309886f : 1.0f; }
};
// One sample per thread (float or double lanes).  Philox counter = (sample, step, draw, stream).
template <class Env>
__device__ __forceinline__ void add_model_noise(typename Env::S (&x)[Env::NS], double sigma, uint64_t seed,
                                                uint32_t stream, uint32_t draw, uint32_t k, uint32_t t);

}  // namespace mpc

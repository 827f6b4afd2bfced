// envs.cuh — the three MPCBenchmark models as inlined device code: cost of (x_t, u_t) then one dynamics step,
// exactly the order of Model.predict (models/model.py:26-44), state kept in registers by the caller.
//
//   Pendulum   models/pendulum_model.py:28-50,58-59
//   CartPole   models/cartpole_swingup_model.py:63-104
//   Acrobot    models/acrobot_model.py:99-176,185-220,270-282
//
// Two arithmetic flavours, selected by the scalar type:
//  * float  (throughput mode): written for the GPU, not transliterated — constants folded on the host into EnvConst,
//    trig shared (cos(a - pi/2) = sin a, angle addition for theta1+theta2, one sincos per angle), x^10 as three
//    squarings and a multiply, the four divisions of the Acrobot accelerations as two reciprocals, FMA everywhere.
//  * double (parity mode, compiled with -fmad=false): the DYNAMICS follow the reference's numpy expression tree
//    operation by operation (same association, true divisions, cos(th1+th2-pi/2) evaluated as written, fmod-based
//    floored remainder, repeated 2*pi wrap), because these systems are chaotic: an ulp of difference per step is
//    amplified ~1e4x over a 100-step Acrobot rollout and would eat the 1e-12 parity budget.  Costs do not feed
//    back into the state, so they keep the cheaper forms (x^10 by squaring).
// Every rewrite is validated against the oracle and the reference's golden vectors (tests/test_gpu_parity.py).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <type_traits>

namespace mpc {

enum { ENV_PENDULUM = 0, ENV_CARTPOLE = 1, ENV_ACROBOT = 2 };

// Folded model constants, one flat array per env (see fold_env_params in capi.cu for the layout).
template <typename R>
struct EnvConst {
    R c[24];
};

template <typename R>
struct M;
template <>
struct M<float> {
    static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
    static __device__ __forceinline__ float floor(float a) { return floorf(a); }
    static __device__ __forceinline__ float sin(float a) { return sinf(a); }
    static __device__ __forceinline__ float cos(float a) { return cosf(a); }
    static __device__ __forceinline__ void sincos(float a, float *s, float *c) { sincosf(a, s, c); }
    static __device__ __forceinline__ float min(float a, float b) { return fminf(a, b); }
    static __device__ __forceinline__ float max(float a, float b) { return fmaxf(a, b); }
    static __device__ __forceinline__ float rcp(float a) { return 1.0f / a; }
    static constexpr float PI = 3.14159265358979323846f;
    static constexpr float INV_2PI = 0.15915494309189533577f;
    // 2*pi split for Cody-Waite reduction: HI has trailing zero bits so n*HI is exact for small n
    static constexpr float TWO_PI_HI = 6.28318548202514648f;          // float(2*pi)
    static constexpr float TWO_PI_LO = -1.74845553146951715e-07f;     // 2*pi - float(2*pi)
};
template <>
struct M<double> {
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
    static __device__ __forceinline__ double floor(double a) { return ::floor(a); }
    static __device__ __forceinline__ double sin(double a) { return ::sin(a); }
    static __device__ __forceinline__ double cos(double a) { return ::cos(a); }
    static __device__ __forceinline__ void sincos(double a, double *s, double *c) { ::sincos(a, s, c); }
    static __device__ __forceinline__ double min(double a, double b) { return fmin(a, b); }
    static __device__ __forceinline__ double max(double a, double b) { return fmax(a, b); }
    static __device__ __forceinline__ double rcp(double a) { return 1.0 / a; }
    static constexpr double PI = 3.14159265358979323846;
    static constexpr double INV_2PI = 0.15915494309189533577;
    static constexpr double TWO_PI_HI = 6.283185307179586232;         // double(2*pi)
    static constexpr double TWO_PI_LO = 2.4492935982947064e-16;       // 2*pi - double(2*pi)
};

template <typename R>
__device__ __forceinline__ R clip(R a, R lo, R hi) {
    return M<R>::min(M<R>::max(a, lo), hi);
}

// ------------------------------------------------------------------------------------------------ Pendulum
// c: 0 c_sin=-3g/(2l)  1 c_u=3/(m l^2)  2 dt  3 max_speed  4 max_torque  5..7 W0..W2
template <typename R>
struct Pendulum {
    static constexpr int NS = 2;
    static constexpr int ENV = ENV_PENDULUM;
    // goal row g = [g0, g1, g2]; returns cost, advances x in place
    static __device__ __forceinline__ R step(const EnvConst<R> &k, R (&x)[NS], R u, const R *g) {
        const R th = x[0], om = x[1];
        if constexpr (std::is_same<R, double>::value) {
            // ---- parity flavour: numpy's expression tree (pendulum_model.py:28-50,58-59)
            const double two_pi = 2 * M<double>::PI;
            double r = fmod(th + M<double>::PI, two_pi);  // np.remainder: fmod, then sign fix-up
            if (r != 0.0) {
                if (r < 0) r += two_pi;
            } else {
                r = 0.0;
            }
            const double thn = r - M<double>::PI;
            const double d0 = -thn - g[0], d1 = -om - g[1], d2 = -u - g[2];
            const double cost = (d0 * k.c[5]) * d0 + (d1 * k.c[6]) * d1 + (d2 * k.c[7]) * d2;
            const double uc = clip(u, -k.c[4], k.c[4]);
            double om2 = om + (k.c[0] * sin(th + M<double>::PI) + k.c[1] * uc) * k.c[2];
            x[0] = th + om2 * k.c[2];
            x[1] = clip(om2, -k.c[3], k.c[3]);
            return cost;
        }
        // angle_normalize: ((th + pi) mod 2pi) - pi with numpy's floored remainder  ==  th - 2pi*floor((th+pi)/2pi)
        const R n = M<R>::floor((th + M<R>::PI) * M<R>::INV_2PI);
        R thn = M<R>::fma(-n, M<R>::TWO_PI_HI, th);
        thn = M<R>::fma(-n, M<R>::TWO_PI_LO, thn);
        // z = -[thn, om, u] (u NOT clipped); cost = sum W_i (z_i - g_i)^2              pendulum_model.py:40-50
        const R d0 = -thn - g[0], d1 = -om - g[1], d2 = -u - g[2];
        const R cost = k.c[5] * d0 * d0 + k.c[6] * d1 * d1 + k.c[7] * d2 * d2;
        // dynamics: sin(th + pi) = -sin(th) = -sin(thn)                                  pendulum_model.py:28-38
        const R uc = clip(u, -k.c[4], k.c[4]);
        const R s = M<R>::sin(thn);
        R om2 = om + (k.c[1] * uc - k.c[0] * s) * k.c[2];
        x[0] = th + om2 * k.c[2];  // integrates with the UNclipped new speed
        x[1] = clip(om2, -k.c[3], k.c[3]);
        return cost;
    }
};

// ------------------------------------------------------------------------------------------------ CartPole
// c: 0 1/x_threshold  1 force_mag  2 dt  3 b
//    4 A1=-2 m_p l  5 A2=3 m_p g  6 A4=4 b  7 D1=4 M  8 D2=3 m_p
//    9 B1=-3 m_p l 10 B2=6 M g   11 E1=4 l M 12 E2=3 m_p l
//   13..17 Wc_i = -W_i  (cost = -(zd W zd), W negative diagonal)   18 x_threshold
template <typename R>
struct CartPole {
    static constexpr int NS = 4;
    static constexpr int ENV = ENV_CARTPOLE;
    static __device__ __forceinline__ R step(const EnvConst<R> &k, R (&x)[NS], R u, const R *g) {
        const R p = x[0], pd = x[1], th = x[2], thd = x[3];
        if constexpr (std::is_same<R, double>::value) {
            // ---- parity flavour: numpy's expression tree (cartpole_swingup_model.py:63-104)
            const double s = sin(th), c = cos(th);
            const double q = p / k.c[18];
            const double q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
            const double z[5] = {q2 + q8 * q2, pd, 1 - c, thd, u};
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const double d = z[i] - g[i];
                acc += (d * k.c[13 + i]) * d;
            }
            const double a = clip(u, -1.0, 1.0) * k.c[1];
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
        R s, c;
        M<R>::sincos(th, &s, &c);
        // _transform: xc = q^2 + q^10, tc = 1 - cos(th)                      cartpole_swingup_model.py:87-93
        const R q = p * k.c[0];
        const R q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        const R xc = M<R>::fma(q8, q2, q2);
        const R d0 = xc - g[0], d1 = pd - g[1], d2 = (R(1) - c) - g[2], d3 = thd - g[3], d4 = u - g[4];
        const R cost = k.c[13] * d0 * d0 + k.c[14] * d1 * d1 + k.c[15] * d2 * d2 + k.c[16] * d3 * d3 + k.c[17] * d4 * d4;
        // _dynamics (explicit Euler, old velocities)                          cartpole_swingup_model.py:63-85
        const R a = k.c[1] * clip(u, R(-1), R(1));
        const R thd2 = thd * thd, c2 = c * c;
        const R num_x = k.c[4] * thd2 * s + k.c[5] * s * c + R(4) * a - k.c[6] * pd;
        const R num_t = k.c[9] * thd2 * s * c + k.c[10] * s + R(6) * (a - k.c[3] * pd) * c;
        const R xacc = num_x / (k.c[7] - k.c[8] * c2);
        const R tacc = num_t / (k.c[11] - k.c[12] * c2);
        x[0] = p + pd * k.c[2];
        x[2] = th + thd * k.c[2];
        x[1] = pd + xacc * k.c[2];
        x[3] = thd + tacc * k.c[2];
        return cost;
    }
};

// ------------------------------------------------------------------------------------------------ Acrobot
// c: 0 dt  1 A = m1 lc1^2 + m2 (l1^2 + lc2^2) + I1 + I2   2 B = 2 m2 l1 lc2     (d1 = A + B cos th2)
//    3 C = m2 lc2^2 + I2   4 D = m2 l1 lc2                                        (d2 = C + D cos th2)
//    5 P1 = (m1 lc1 + m2 l1) g   6 P2 = m2 lc2 g   7 max_vel1  8 max_vel2   9..13 W0..W4
//   parity flavour only: 14 m1 lc1^2  15 l1^2 + lc2^2  16 2 l1 lc2  17 m2  18 I1  19 I2  20 lc2^2  21 l1 lc2
template <typename R>
struct Acrobot {
    static constexpr int NS = 4;
    static constexpr int ENV = ENV_ACROBOT;

    // "book" _dsdt (acrobot_model.py:99-132) given sin/cos of both angles; returns angular accelerations.
    static __device__ __forceinline__ void accel(const EnvConst<R> &k, R s1, R c1, R s2, R c2, R w1, R w2, R a,
                                                 R *a1, R *a2) {
        const R d1 = M<R>::fma(k.c[2], c2, k.c[1]);
        const R d2 = M<R>::fma(k.c[4], c2, k.c[3]);
        const R s12 = M<R>::fma(s1, c2, c1 * s2);       // sin(th1 + th2) = cos(th1 + th2 - pi/2)
        const R phi2 = k.c[6] * s12;
        const R Ds2 = k.c[4] * s2;
        // phi1 = -D w2^2 s2 - 2 D w2 w1 s2 + P1 sin(th1) + phi2
        const R phi1 = M<R>::fma(k.c[5], s1, phi2) - Ds2 * w2 * (w2 + R(2) * w1);
        const R r1 = M<R>::rcp(d1);
        const R d2r = d2 * r1;
        // ddth2 = (a + d2/d1 phi1 - D w1^2 s2 - phi2) / (C - d2^2/d1)
        const R num = a + d2r * phi1 - Ds2 * w1 * w1 - phi2;
        const R den = M<R>::fma(-d2, d2r, k.c[3]);
        const R acc2 = num / den;
        *a2 = acc2;
        *a1 = -(M<R>::fma(d2, acc2, phi1)) * r1;
    }

    // parity flavour of _dsdt: the reference's expression tree, term by term (acrobot_model.py:114-132)
    static __device__ __forceinline__ void dsdt_literal(const EnvConst<double> &k, const double (&y)[4], double a,
                                                        double (&d)[4]) {
        const double th1 = y[0], th2 = y[1], w1 = y[2], w2 = y[3];
        const double half_pi = M<double>::PI / 2.;
        const double cth2 = cos(th2), sth2 = sin(th2);
        const double d1 = k.c[14] + k.c[17] * (k.c[15] + k.c[16] * cth2) + k.c[18] + k.c[19];
        const double d2 = k.c[17] * (k.c[20] + k.c[21] * cth2) + k.c[19];
        const double phi2 = k.c[6] * cos(th1 + th2 - half_pi);
        const double phi1 = ((-k.c[4] * (w2 * w2)) * sth2 - ((k.c[2] * w2) * w1) * sth2) + k.c[5] * cos(th1 - half_pi) + phi2;
        const double acc2 = (((a + (d2 / d1) * phi1) - (k.c[4] * (w1 * w1)) * sth2) - phi2) / (k.c[3] - (d2 * d2) / d1);
        const double acc1 = -(d2 * acc2 + phi1) / d1;
        d[0] = w1;
        d[1] = w2;
        d[2] = acc1;
        d[3] = acc2;
    }

    static __device__ __forceinline__ R wrap(R v) {  // acrobot_model.py:185-203 with m = -pi, M = pi
        if constexpr (std::is_same<R, double>::value) {
            const double diff = M<double>::PI - (-M<double>::PI);
            while (v > M<double>::PI) v = v - diff;
            while (v < -M<double>::PI) v = v + diff;
            return v;
        }
        while (v > M<R>::PI) v = (v - M<R>::TWO_PI_HI) - M<R>::TWO_PI_LO;
        while (v < -M<R>::PI) v = (v + M<R>::TWO_PI_HI) + M<R>::TWO_PI_LO;
        return v;
    }

    static __device__ __forceinline__ R step(const EnvConst<R> &k, R (&x)[NS], R u, const R *g) {
        if constexpr (std::is_same<R, double>::value) {
            // ---- parity flavour (acrobot_model.py:134-176, rk4 at :270-282)
            const double z0 = -cos(x[0]) - cos(x[1] + x[0]) - 2;
            const double zz[5] = {z0, 0.0, 0.0, 0.0, u};
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const double d = zz[i] - g[i];
                acc += (d * k.c[9 + i]) * d;
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
        const R th1 = x[0], th2 = x[1], w1 = x[2], w2 = x[3];
        const R h = k.c[0], h2 = h * R(0.5);
        R s1, c1, s2, c2;
        M<R>::sincos(th1, &s1, &c1);
        M<R>::sincos(th2, &s2, &c2);
        // cost: z0 = -cos th1 - cos(th1 + th2) - 2, z1..3 = 0, z4 = u                acrobot_model.py:162-176
        const R c12 = M<R>::fma(c1, c2, -s1 * s2);
        const R d0 = (-c1 - c12 - R(2)) - g[0], d4 = u - g[4];
        const R cost = k.c[9] * d0 * d0 + k.c[10] * g[1] * g[1] + k.c[11] * g[2] * g[2] + k.c[12] * g[3] * g[3] +
                       k.c[13] * d4 * d4;
        // one classical RK4 step, torque u (unclipped) held constant            acrobot_model.py:134-160,270-282
        R k1a, k1b, k2a, k2b, k3a, k3b, k4a, k4b;
        accel(k, s1, c1, s2, c2, w1, w2, u, &k1a, &k1b);
        R y0 = M<R>::fma(h2, w1, th1), y1 = M<R>::fma(h2, w2, th2), y2 = M<R>::fma(h2, k1a, w1), y3 = M<R>::fma(h2, k1b, w2);
        const R k2w1 = y2, k2w2 = y3;
        M<R>::sincos(y0, &s1, &c1);
        M<R>::sincos(y1, &s2, &c2);
        accel(k, s1, c1, s2, c2, y2, y3, u, &k2a, &k2b);
        y0 = M<R>::fma(h2, k2w1, th1); y1 = M<R>::fma(h2, k2w2, th2); y2 = M<R>::fma(h2, k2a, w1); y3 = M<R>::fma(h2, k2b, w2);
        const R k3w1 = y2, k3w2 = y3;
        M<R>::sincos(y0, &s1, &c1);
        M<R>::sincos(y1, &s2, &c2);
        accel(k, s1, c1, s2, c2, y2, y3, u, &k3a, &k3b);
        y0 = M<R>::fma(h, k3w1, th1); y1 = M<R>::fma(h, k3w2, th2); y2 = M<R>::fma(h, k3a, w1); y3 = M<R>::fma(h, k3b, w2);
        const R k4w1 = y2, k4w2 = y3;
        M<R>::sincos(y0, &s1, &c1);
        M<R>::sincos(y1, &s2, &c2);
        accel(k, s1, c1, s2, c2, y2, y3, u, &k4a, &k4b);
        const R h6 = h * R(1.0 / 6.0);
        const R n0 = M<R>::fma(h6, w1 + R(2) * (k2w1 + k3w1) + k4w1, th1);
        const R n1 = M<R>::fma(h6, w2 + R(2) * (k2w2 + k3w2) + k4w2, th2);
        const R n2 = M<R>::fma(h6, k1a + R(2) * (k2a + k3a) + k4a, w1);
        const R n3 = M<R>::fma(h6, k1b + R(2) * (k2b + k3b) + k4b, w2);
        x[0] = wrap(n0);
        x[1] = wrap(n1);
        x[2] = clip(n2, -k.c[7], k.c[7]);
        x[3] = clip(n3, -k.c[8], k.c[8]);
        return cost;
    }
};

}  // namespace mpc

// philox.cuh — Philox4x32-10 counter-based generator + Box-Muller, inlined into the rollout kernels.
//
// Counter layout: (global sample index k, t / 4, draw, stream) ; key = 64-bit seed.  One call yields the four
// standard normals of time steps 4*(t/4) .. 4*(t/4)+3 of sample k, so the stream is a pure function of
// (seed, stream, draw, k, t): independent of block size, tiling and of how K is sharded across GPUs.
//
// The normals are defined in float32 (as cuRAND's curand_normal does) with explicitly rounded intrinsics, so every
// kernel that inlines this function — the rollouts and the materialising kernel behind mpc_philox_noise — produces
// bit-identical values regardless of how the surrounding code is contracted.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mpc {

struct Philox4 {
    uint32_t v[4];
};

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += W0;
        k1 += W1;
    }
    Philox4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}

// two standard normals from two 32-bit words (Box-Muller, float32, MUFU log/sincos: |err| ~ 1e-6, irrelevant for
// synthetic noise; what matters is that it is deterministic)
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float *z0, float *z1) {
    const float u1 = __fmaf_rn((float)a, 2.3283064365386963e-10f, 1.1641532182693481e-10f);  // (a + 0.5) / 2^32
    const float u2 = __fmaf_rn((float)b, 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    const float r = __fsqrt_rn(__fmul_rn(-2.0f, __logf(u1)));
    float s, c;
    __sincosf(__fmul_rn(6.2831853071795865f, u2), &s, &c);
    *z0 = __fmul_rn(r, c);
    *z1 = __fmul_rn(r, s);
}

// the four normals of steps 4*tq .. 4*tq+3 of sample k
__device__ __forceinline__ void philox_normal4(uint64_t seed, uint32_t stream, uint32_t draw, uint32_t k, uint32_t tq,
                                               float (&z)[4]) {
    const Philox4 p = philox4x32_10(k, tq, draw, stream, (uint32_t)seed, (uint32_t)(seed >> 32));
    box_muller(p.v[0], p.v[1], &z[0], &z[1]);
    box_muller(p.v[2], p.v[3], &z[2], &z[3]);
}

}  // namespace mpc

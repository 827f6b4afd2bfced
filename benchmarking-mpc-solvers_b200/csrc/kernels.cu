// kernels.cu — sm_100a device code of the sampling-MPC inner loop.
//
//   mppi_solve_kernel     (mppi.cuh) the whole MPPI solve in one persistent launch: MPPI.f for every sample
//                         (mppi.py:35-48), softmin-weighted noise sum (mppi.py:59-67), planned_us += erg (mppi.py:69)
//                         and the clipped plan's nominal rollout (agent.py:64-69,97-108)
//   mppi_rank_merge_kernel  multi-GPU: merges the gathered rank records and applies them
//   cem_*                 CEM._calc_action (cem.py:54-78): rollout of clip(mu + sd*z), radix top-n_elite select,
//                         elite mean / std refit with alpha smoothing and the epsilon test.
#include "kernels.cuh"
#include "envs.cuh"
#include "philox.cuh"
#include "peer.cuh"

#include <cuda.h>  // CUtensorMap (type only; the encoder is resolved at run time in capi.cu)
#include <float.h>
#include <string.h>

namespace mpc {

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// orders generic-proxy shared-memory accesses before later async-proxy (bulk copy) writes to the same buffer
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(__cvta_generic_to_global(src_gmem)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// TMA: box of a rank-3 tensor map -> shared memory, completing on an mbarrier (SASS: UTMALDG)
__device__ __forceinline__ void tma_load_3d(void *dst_smem, const CUtensorMap *map, int c0, int c1, int c2, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

// ------------------------------------------------------------------------------------------------ reductions
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// all threads get the result; red must hold >= blockDim/32 doubles; contains two barriers
template <bool IS_MIN>
__device__ __forceinline__ double block_reduce(double v, double *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = IS_MIN ? warp_min(v) : warp_sum(v);
    __syncthreads();  // protect red from the previous use
    if (lane == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < nw; ++i) r = IS_MIN ? fmin(r, red[i]) : r + red[i];
    return r;
}

// Folded constants narrowed once on the host and passed by value: they live in the kernel-parameter constant bank,
// so FFMA/FFMA2/DFMA read them as constant (uniform-register) operands instead of holding registers.
// double <- parity layout c, float / f2 <- throughput layout f (fold_env_params, capi.cu).
template <typename V>
static EnvConst<V> make_env(const EnvParamsHost &ep) {
    EnvConst<V> ec;
    for (int i = 0; i < 24; ++i) {
        if constexpr (std::is_same<V, double>::value)
            ec.c[i] = ep.c[i];
        else
            ec.c[i] = V((float)ep.f[i]);
    }
    for (int i = 0; i < 15; ++i) {
        if constexpr (std::is_same<V, double>::value)
            ec.w[i] = ep.w[i];
        else
            ec.w[i] = V((float)ep.w[i]);
    }
    return ec;
}

template <class Env>
__device__ __forceinline__ void add_model_noise(typename Env::S (&x)[Env::NS], double sigma, uint64_t seed,
                                                uint32_t stream, uint32_t draw, uint32_t k, uint32_t t) {
    typedef typename Env::S R;
    static_assert(Env::NS <= 4, "one Philox block of four normals per state");
    float z[4];
    philox_normal4(seed, stream, draw, k, t, z);
#pragma unroll
    for (int i = 0; i < Env::NS; ++i) x[i] += (R)sigma * NoiseUnit<Env>::unit(i) * (R)z[i];
}

// Philox streams of the model noise: sample rollouts of MPPI / CEM, nominal rollouts
enum { PHILOX_MODEL_MPPI = 4, PHILOX_MODEL_NOMINAL = 5, PHILOX_MODEL_CEM = 6 };

// Nominal rollout of one plan by a single thread: Agent.simulate_trajectory (agent.py:97-108).
// us: plan (already clipped by the caller if required); goal row of step t at g_z + t * g_stride (g_stride = 0: one
// row for every step); xs [H+1][ns] (nullable); returns the cost.  Callers keep us / g_z in shared memory: the loop is
// a serial dependency chain and must not wait on global loads.
template <class Env>
__device__ double nominal_rollout(const EnvConst<typename Env::Vec> &ec, int H, const double *x0, const double *us,
                                  const double *g_z, int g_stride, double *xs, double model_noise = 0.0,
                                  uint64_t seed = 0, uint32_t draw = 0, uint32_t which = 0) {
    constexpr int NS = Env::NS;
    typedef typename Env::S R;
    R x[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
        x[i] = (R)x0[i];
        if (xs) xs[i] = x0[i];
    }
    Env::enter(ec, x);
    double reward = 0.0;
    R gs[Env::NG];
    if (g_stride == 0) Env::prep_goal(ec, g_z, gs);
    for (int t = 0; t < H; ++t) {
        if (g_stride != 0) Env::prep_goal(ec, g_z + t * g_stride, gs);
        const R c = Env::step(ec, x, (R)us[t], gs);
        if (model_noise > 0.0) add_model_noise<Env>(x, model_noise, seed, PHILOX_MODEL_NOMINAL, draw, which, (uint32_t)t);
        reward += -(double)c;
        if (xs) Env::leave(ec, x, xs + (t + 1) * NS);
    }
    return -reward;
}

// mu + sd * z as numpy forms it (cem.py:58: np.random.normal(loc, scale) = loc + scale * z, two roundings): the float64
// unit is compiled with -fmad=false, so the product and the sum below stay separate there; float32 contracts to one FMA
__device__ __forceinline__ double sample_value(double sd, double z, double mu) { return mu + sd * z; }
__device__ __forceinline__ float sample_value(float sd, float z, float mu) { return fmaf(sd, z, mu); }

__device__ __forceinline__ float load_lanes(const float *p, float *) { return *p; }
__device__ __forceinline__ double load_lanes(const double *p, double *) { return *p; }
__device__ __forceinline__ f2 load_lanes(const float *p, f2 *) { return f2(*reinterpret_cast<const float2 *>(p)); }

}  // namespace mpc
#include "mppi.cuh"
namespace mpc {

// ================================================================================================ noise utilities
// [K_local][H] float64 (reference layout, du = 1) -> [H][Kpad] in the rollout dtype, padding zero-filled.
template <typename R>
__global__ void transpose_noise_kernel(const double *__restrict__ src, R *__restrict__ dst, int K_local, int Kpad,
                                       int H) {
    __shared__ double t[32][33];
    const int k0 = blockIdx.x * 32, h0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int k = k0 + r, h = h0 + threadIdx.x;
        t[r][threadIdx.x] = (k < K_local && h < H) ? src[(size_t)k * H + h] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int h = h0 + r, k = k0 + threadIdx.x;
        if (h < H && k < Kpad) dst[(size_t)h * Kpad + k] = (R)t[threadIdx.x][r];
    }
}

template <typename R>
__global__ void philox_fill_kernel(R *__restrict__ dst, int K_local, int k_begin, int Kpad, int H, double scale,
                                   uint64_t seed, uint32_t draw, uint32_t stream) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= Kpad) return;
    const bool valid = k < K_local;
    for (int t4 = 0; t4 < H; t4 += 4) {
        float z[4] = {0.f, 0.f, 0.f, 0.f};
        if (valid) philox_normal4(seed, stream, draw, (uint32_t)(k_begin + k), (uint32_t)(t4 >> 2), z);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (t4 + i < H) dst[(size_t)(t4 + i) * Kpad + k] = (R)scale * (R)z[i];
    }
}

// ================================================================================================ CEM
// sel[] layout
enum { SEL_DONE = 0, SEL_ITERS = 1, SEL_NOTSMALL = 2, SEL_BLOCKS = 3, SEL_WORDS = 8 };

__device__ __forceinline__ unsigned long long reward_key(double r) {
    // order-preserving map double -> u64 (larger reward => larger key); NaN sorts below everything
    if (r != r) return 0ull;
    unsigned long long u = (unsigned long long)__double_as_longlong(r);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}

// CEM.f (cem.py:42-48) for samples clip(mu + sd*z) (cem.py:58-60), one thread per sample.
template <class Env, int NOISE>
__global__ void __launch_bounds__(TILE) cem_rollout_kernel(const CemArgs a, const EnvConst<typename Env::Vec> ec,
                                                          const typename Env::S *__restrict__ zt,
                                                          const double *__restrict__ x0g,
                                                          const double *__restrict__ g_z,
                                                          const double *__restrict__ mu,
                                                          const double *__restrict__ sd, double *__restrict__ Rhist,
                                                          unsigned long long *__restrict__ keys,
                                                          const int *__restrict__ sel) {
    typedef typename Env::S R;
    constexpr int NS = Env::NS, NG = Env::NG;
    if (sel[SEL_DONE]) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int H = a.g.H, tid = threadIdx.x;
    R *smu = reinterpret_cast<R *>(smem_raw);
    R *ssd = smu + H;
    for (int t = tid; t < H; t += TILE) {
        smu[t] = (R)mu[t];
        ssd[t] = (R)sd[t];
    }
    R gs[NG], x[NS];
    Env::prep_goal(ec, g_z, gs);
#pragma unroll
    for (int i = 0; i < NS; ++i) x[i] = (R)x0g[i];
    Env::enter(ec, x);
    __syncthreads();
    const int k = blockIdx.x * TILE + tid;
    if (k >= a.g.K_local) return;
    const R lo = (R)a.lo, hi = (R)a.hi;
    double reward = 0.0;
    if (NOISE == NOISE_TENSOR) {
        for (int t = 0; t < H; ++t) {
            const R z = zt[(size_t)t * a.z_stride + a.z_off + k];
            const R u = vclip(sample_value(ssd[t], z, smu[t]), lo, hi);
            if constexpr (Env::GENERIC != 0)
                if (a.per_step_goal) Env::prep_goal(ec, g_z + (size_t)t * (NS + 1), gs);
            reward += -(double)Env::step(ec, x, u, gs);
            if (a.model_noise > 0.0)
                add_model_noise<Env>(x, a.model_noise, a.seed, PHILOX_MODEL_CEM, a.noise_draw, (uint32_t)(a.g.k_begin + k), (uint32_t)t);
        }
    } else {
        const uint32_t kglob = (uint32_t)(a.g.k_begin + k);
        for (int t4 = 0; t4 < H; t4 += 4) {
            float z[4];
            philox_normal4(a.seed, 1u, (uint32_t)a.iter, kglob, (uint32_t)(t4 >> 2), z);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int t = t4 + i;
                if (t < H) {
                    const R u = vclip(sample_value(ssd[t], (R)z[i], smu[t]), lo, hi);
                    if constexpr (Env::GENERIC != 0)
                        if (a.per_step_goal) Env::prep_goal(ec, g_z + (size_t)t * (NS + 1), gs);
                    reward += -(double)Env::step(ec, x, u, gs);
                    if (a.model_noise > 0.0)
                        add_model_noise<Env>(x, a.model_noise, a.seed, PHILOX_MODEL_CEM, a.noise_draw, kglob, (uint32_t)t);
                }
            }
        }
    }
    Rhist[(size_t)a.iter * a.g.Kpad + k] = reward;
    keys[k] = reward_key(reward);
}

// Elite refit (cem.py:69-76): one block per time step t.
template <typename R, int NOISE>
__global__ void __launch_bounds__(128) cem_refit_kernel(const CemArgs a, const R *__restrict__ zt,
                                                       const int *__restrict__ elite_hist, double *__restrict__ mu,
                                                       double *__restrict__ sd, int *__restrict__ sel) {
    if (sel[SEL_DONE]) return;
    __shared__ double red[8];
    const int t = blockIdx.x, tid = threadIdx.x, n = a.n_elite;
    const int *elite = elite_hist + (size_t)a.iter * n;
    const R m0 = (R)mu[t], s0 = (R)sd[t], lo = (R)a.lo, hi = (R)a.hi;
    auto sample = [&](int e) -> double {
        const int k = elite[e];  // GLOBAL sample index (== local index on one GPU)
        R z;
        if (NOISE == NOISE_TENSOR) {
            z = zt[(size_t)t * a.z_stride + (a.world > 1 ? k : a.z_off + k)];
        } else {
            float z4[4];
            philox_normal4(a.seed, 1u, (uint32_t)a.iter, (uint32_t)(a.world > 1 ? k : a.g.k_begin + k), (uint32_t)(t >> 2), z4);
            z = (R)z4[t & 3];
        }
        return (double)vclip(sample_value(s0, z, m0), lo, hi);  // identical expression to the rollout's
    };
    double s = 0.0;
    for (int e = tid; e < n; e += blockDim.x) s += sample(e);
    const double mean = block_reduce<false>(s, red) / n;       // np.mean(elites, axis=0)
    double v = 0.0;
    for (int e = tid; e < n; e += blockDim.x) {
        const double d = sample(e) - mean;
        v += d * d;
    }
    const double var = block_reduce<false>(v, red) / n;        // np.std: ddof = 0, two-pass
    if (tid == 0) {
        const double new_mu = a.alpha * mu[t] + (1 - a.alpha) * mean;     // cem.py:71-72
        const double new_sd = a.alpha * sd[t] + (1 - a.alpha) * sqrt(var);  // cem.py:73
        mu[t] = new_mu;
        sd[t] = new_sd;
        if (!(new_sd < a.epsilon)) atomicAdd(&sel[SEL_NOTSMALL], 1);
        __threadfence();
        const int arrived = atomicAdd(&sel[SEL_BLOCKS], 1);
        if (arrived == (int)gridDim.x - 1) {  // last block: evaluate (std < epsilon).all()   cem.py:75-76
            __threadfence();
            const int notsmall = atomicAdd(&sel[SEL_NOTSMALL], 0);
            sel[SEL_ITERS] = a.iter + 1;
            sel[SEL_NOTSMALL] = 0;
            sel[SEL_BLOCKS] = 0;
            if (notsmall == 0) sel[SEL_DONE] = 1;
        }
    }
}

// u0 = planned_us[0] (cem.py:77), plus the clipped plan's nominal rollout (agent.py:64-69).
template <class Env>
__global__ void cem_finish_kernel(const CemArgs a, const EnvConst<typename Env::Vec> ec, const double *__restrict__ x0g,
                                  const double *__restrict__ g_z, const double *__restrict__ mu,
                                  const int *__restrict__ sel, double *__restrict__ out) {
    constexpr int NS = Env::NS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *sUc = reinterpret_cast<double *>(smem_raw);
    const int H = a.g.H;
    double *sG = sUc + H;  // [H][NS + 1]
    for (int t = threadIdx.x; t < H; t += blockDim.x) {
        out[1 + t] = mu[t];
        sUc[t] = fmin(fmax(mu[t], a.lo), a.hi);
    }
    if (a.want_traj)
        for (int i = threadIdx.x; i < H * (NS + 1); i += blockDim.x) sG[i] = g_z[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        out[0] = mu[0];
        out[2 + H + (H + 1) * NS] = (double)sel[SEL_ITERS];
        if (a.want_traj) {
            out[1 + H] = nominal_rollout<Env>(ec, H, x0g, sUc, sG, NS + 1, out + 2 + H, a.model_noise, a.seed, a.noise_draw, 0u);
        }
    }
}

}  // namespace mpc
#include "cem_fused.cuh"
namespace mpc {

// ================================================================================================ small kernels
template <class Env>
__global__ void simulate_kernel(const EnvConst<typename Env::Vec> ec, int H, double lo, double hi, int do_clip,
                                const double *__restrict__ x0, const double *__restrict__ us,
                                const double *__restrict__ g_z, double *__restrict__ xs, double *__restrict__ cost) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *sU = reinterpret_cast<double *>(smem_raw);
    double *sG = sU + H;  // [H][NS + 1]
    for (int t = threadIdx.x; t < H; t += blockDim.x) sU[t] = do_clip ? fmin(fmax(us[t], lo), hi) : us[t];
    for (int i = threadIdx.x; i < H * (Env::NS + 1); i += blockDim.x) sG[i] = g_z[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        *cost = nominal_rollout<Env>(ec, H, x0, sU, sG, Env::NS + 1, xs);
    }
}

template <class Env>
__global__ void model_step_kernel(const EnvConst<typename Env::Vec> ec, int n, const double *__restrict__ xg,
                                  const double *__restrict__ ug, const double *__restrict__ gg,
                                  double *__restrict__ cost, double *__restrict__ xn) {
    typedef typename Env::S R;
    constexpr int NS = Env::NS, NZ = Env::NS + 1;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    R x[NS], gs[Env::NG];
#pragma unroll
    for (int j = 0; j < NS; ++j) x[j] = (R)xg[(size_t)i * NS + j];
    Env::prep_goal(ec, gg + (size_t)i * NZ, gs);
    Env::enter(ec, x);
    cost[i] = (double)Env::step(ec, x, (R)ug[i], gs);
    Env::leave(ec, x, xn + (size_t)i * NS);
}

// ------------------------------------------------------------------------------------------------ closed loop
// One control step of Experiment.run (Experiment.py:78-90) around the solve that has just finished on this stream,
// without the host: what Agent.predict_action does after _calc_action (agent.py:63-76) and Environment.step
// (env.py:35-50), one block.
//   out (the solve's outputs): u0, U[H] (updated, unclipped plan), cost and xs of the clipped plan's nominal rollout
//   plan: planned_us = clip(planned_us); log (planned_xs, planned_us, planned_costs); planned_us = roll(-1), last = 0
//   plant: u = clip(u0) (+ actuation noise); cost of (x_true, u) with a zero goal; x_true = model.predict(x_true, u);
//          observation = x_true (+ sensor noise) -> the next solve's x0
// The plant is always stepped in float64 with the reference's expression tree (this kernel is launched from the
// float64 translation unit whatever the rollout dtype): the environment is the ground truth of an episode.
// Noise: Philox streams 2 (actuation) and 3 (sensor), counter = control step.
template <class Env>
__global__ void episode_step_kernel(const EnvConst<double> ec, const EpisodeArgs a, const EpisodeBuffers b) {
    constexpr int NS = Env::NS;
    const int H = a.H, i = a.step, tid = threadIdx.x, nt = blockDim.x;
    const double *o = b.out;
    __shared__ double s_next[8];
    if (a.mode == 1) {  // Agent.warm_start: planned_us = clip(planned_us)   agent.py:45-47
        for (int t = tid; t < H; t += nt) b.U[t] = fmin(fmax(o[1 + t], a.lo), a.hi);
        return;
    }
    if (tid == 0) {
        double u = fmin(fmax(o[0], a.lo), a.hi);  // return np.clip(u0, ...)   agent.py:76
        float z[4];
        if (a.actuation_std > 0.0) {  // env.py:36-38
            philox_normal4(a.seed, 2u, (uint32_t)i, 0u, 0u, z);
            u += a.actuation_std * (double)z[0];
        }
        double x[NS], g[Env::NG], zero[NS + 1];
#pragma unroll
        for (int k = 0; k < NS; ++k) x[k] = b.true_state[k];
#pragma unroll
        for (int k = 0; k <= NS; ++k) zero[k] = 0.0;
        Env::prep_goal(ec, zero, g);  // Model.predict(x, u) with goal = None -> zeros   model.py:32-33
        const double c = Env::step(ec, x, u, g);
        if (a.sensor_std > 0.0) philox_normal4(a.seed, 3u, (uint32_t)i, 0u, 0u, z);  // env.py:42-44
#pragma unroll
        for (int k = 0; k < NS; ++k) {
            const double obs = x[k] + (a.sensor_std > 0.0 ? a.sensor_std * (double)z[k] : 0.0);
            b.true_state[k] = x[k];
            b.true_states[(size_t)(i + 1) * NS + k] = x[k];
            b.states[(size_t)(i + 1) * NS + k] = obs;
            s_next[k] = obs;
        }
        b.actions[i] = u;
        b.costs[i] = c;
        if (b.plan_costs) b.plan_costs[i] = o[1 + H];
    }
    // the log rows of this step (agent.py:67-69) and the receding-horizon shift (agent.py:70-71)
    for (int t = tid; t < H; t += nt) {
        const double uc = fmin(fmax(o[1 + t], a.lo), a.hi);
        if (b.plan_us) b.plan_us[(size_t)i * H + t] = uc;
        if (a.shift) {
            if (t > 0) b.U[t - 1] = uc;
            if (t == H - 1) b.U[t] = 0.0;
        } else {
            b.U[t] = uc;
        }
    }
    if (b.plan_xs)
        for (int k = tid; k < (H + 1) * NS; k += nt) b.plan_xs[(size_t)i * (H + 1) * NS + k] = o[2 + H + k];
    __syncthreads();
    if (tid < NS) b.x0[tid] = s_next[tid];  // the next predict_action sees env.state (Experiment.py:81-82)
}

// FMA-pipe peak: 16 independent chains per thread, no memory traffic.
template <typename R>
__global__ void __launch_bounds__(256) fma_peak_kernel(float *sink, int iters) {
    R a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = (R)(threadIdx.x + i) * (R)1e-3;
    const R m = (R)0.999, c = (R)1e-4;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = vfma(a[i], m, c);
    }
    R s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == (R)123456789) sink[0] = (float)s;
}

// ================================================================================================ per-(dtype, env) launchers
// This file is compiled once per (dtype, env) pair (build.py) so that the big solve-kernel instantiations build in
// parallel: -DMPC_REAL=float|double, -DMPC_ENV_ONLY=0|1|2 (Pendulum, CartPole, Acrobot), -DMPC_SUFFIX=_f32_e0 ...;
// the float64 units are built with -fmad=false so that the parity path keeps the reference's rounding sequence.
// The Acrobot float32 unit is split once more by trig flavour: -DMPC_TRIG_GROUP=0 (polynomial / MUFU) and 1 (the
// per-stage splits, suffix _f32_e2s).  Env-independent launchers live in the env-0 unit of each dtype.
#if !defined(MPC_REAL) || !defined(MPC_ENV_ONLY)
#error "compile with -DMPC_REAL=float|double -DMPC_ENV_ONLY=0|1|2 -DMPC_SUFFIX=_f32_e0|..."
#endif
#ifndef MPC_TRIG_GROUP
#define MPC_TRIG_GROUP 0
#endif
#define MPC_CAT_(a, b) a##b
#define MPC_CAT(a, b) MPC_CAT_(a, b)
#define FN(name) MPC_CAT(name, MPC_SUFFIX)
typedef MPC_REAL R_;

#if MPC_ENV_ONLY == 0
template <typename V, int TRIG, int GEN = 0>
using TheEnv = Pendulum<V, TRIG, GEN>;
#elif MPC_ENV_ONLY == 1
template <typename V, int TRIG, int GEN = 0>
using TheEnv = CartPole<V, TRIG, GEN>;
#else
template <typename V, int TRIG, int GEN = 0>
using TheEnv = Acrobot<V, TRIG, GEN>;
#endif
constexpr bool IS_ACROBOT = MPC_ENV_ONLY == 2;

// variant: bits 0-2 = trig flavour (0 polynomial, 1 MUFU; Acrobot 2..5: per-stage splits, envs.cuh), bit 3 = two samples
// per thread (packed f32x2), bit 4 = WIDE (32-warp whole-tile blocks, one sample per thread).  float64 has one flavour.
// bit 5 = generic cost (full W / per-step goals): one sample per thread, polynomial trig, whole tiles or Philox.
enum { VAR_TRIG_MASK = 7, VAR_PACK2 = 8, VAR_WIDE = 16, VAR_GEN = 32 };

// trig flavours this unit instantiates; everything else collapses to 0 / 1
static int unit_trig(int variant) {
    const int t = variant & VAR_TRIG_MASK;
    if (!IS_ACROBOT) return t ? 1 : 0;
    return t > 5 ? 1 : t;
}

// Calls f(Env{}) with this unit's env in the rollout flavour selected by `variant` (V = float / f2 x trig), or double.
// `splits` = the caller has kernels for the Acrobot split flavours (STREAM / TENSOR); otherwise 2..5 fall back to MUFU.
template <class F>
static void with_rollout_env(int variant, bool splits, F &&f) {
    if constexpr (std::is_same<R_, double>::value) {
        if (variant & VAR_GEN)
            f(TheEnv<double, 0, 1>{});
        else
            f(TheEnv<double, 0>{});
    } else {
#if MPC_TRIG_GROUP == 0
        if (variant & VAR_GEN) {
            f(TheEnv<float, 0, 1>{});
            return;
        }
#endif
        int trig = unit_trig(variant);
        if (trig > 1 && !splits) trig = 1;
        const bool pack = (variant & VAR_PACK2) != 0;
#if MPC_TRIG_GROUP == 0
        if (trig == 0) {
            if (pack) f(TheEnv<f2, 0>{}); else f(TheEnv<float, 0>{});
        } else {
            if (pack) f(TheEnv<f2, 1>{}); else f(TheEnv<float, 1>{});
        }
#else
        switch (trig) {
        case 2: if (pack) f(TheEnv<f2, 2>{}); else f(TheEnv<float, 2>{}); break;
        case 4: if (pack) f(TheEnv<f2, 4>{}); else f(TheEnv<float, 4>{}); break;
        default: if (pack) f(TheEnv<f2, 5>{}); else f(TheEnv<float, 5>{}); break;
        }
#endif
    }
}

// One sample per thread: CEM rollouts (accurate or fast trig), nominal rollouts and single steps (always accurate).
template <class F>
static void with_scalar_env(bool fast, F &&f, bool gen = false) {
    if constexpr (std::is_same<R_, double>::value) {
        if (gen)
            f(TheEnv<double, 0, 1>{});
        else
            f(TheEnv<double, 0>{});
    } else {
        if (gen)
            f(TheEnv<float, 0, 1>{});
        else if (fast)
            f(TheEnv<float, 1>{});
        else
            f(TheEnv<float, 0>{});
    }
}

template <class F>
static void with_scalar_env_gen(const EnvParamsHost &ep, bool fast, F &&f) {
    with_scalar_env(fast, f, ep.generic != 0);
}

template <class Env, int NOISE, bool NOISY = false, bool WIDE = false>
static cudaError_t mppi_solve_go(const MppiArgs &a, const EnvParamsHost &ep, const MppiBuffers &b, cudaStream_t s) {
    typedef typename Env::Vec V;
    auto kern = mppi_solve_kernel<Env, NOISE, NOISY, WIDE>;
    const size_t smem = a.g.smem_rollout;
    if (a.g.warps > SolveShape<V, NOISE, WIDE>::MAXW) return cudaErrorInvalidConfiguration;
    // per instantiation AND per device (the attribute is a per-device function property): raise the dynamic
    // shared memory limit only when it grows
    static size_t smem_set[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > smem_set[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        smem_set[dev & 63] = smem;
    }
    CUtensorMap map;
    if (b.noise_map)
        memcpy(&map, b.noise_map, sizeof map);
    else
        memset(&map, 0, sizeof map);  // PHILOX: never dereferenced
    PeerTable pt;
    memset(&pt, 0, sizeof pt);
    if (a.p2p)
        for (int r = 0; r < a.world && r < MAX_WORLD; ++r) pt.recs[r] = b.peer_recs[r];
    InlineInputs inl;
    if (b.inl)
        inl = *b.inl;
    else
        inl.on = 0;
    kern<<<a.g.grid, a.g.warps * 32, smem, s>>>(a, make_env<V>(ep), map, (const typename Env::S *)b.noise, b.x0, b.g_z, b.U,
                                                 b.S, b.part, b.rank_part, b.out, b.done_count, pt, inl);
    return cudaGetLastError();
}

#if MPC_TRIG_GROUP == 0
// shared-memory bytes of the solve kernel: records + ring (S stages of `rows` time steps, W warps) or records + the tail
// arena, whichever is larger
template <typename V>
static size_t mppi_smem_bytes(int H, int rows, int S, int W, int cap, int G, int nz, int uc) {
    const size_t ring = RingSmem<V>::ring_total(H, rows, S, W, uc ? uc : W), tail = RingSmem<V>::tail_total(H, W, cap, G, nz);
    return ring > tail ? ring : tail;
}
#endif

#if MPC_ENV_ONLY == 0
size_t FN(mppi_solve_smem)(int variant, int H, int rows, int S, int W, int cap, int G, int nz, int uc) {
    if constexpr (std::is_same<R_, double>::value)
        return mppi_smem_bytes<double>(H, rows, S, W, cap, G, nz, uc);
    else
        return (variant & VAR_PACK2) ? mppi_smem_bytes<f2>(H, rows, S, W, cap, G, nz, uc)
                                     : mppi_smem_bytes<float>(H, rows, S, W, cap, G, nz, uc);
}

size_t FN(mppi_stage_bytes)(int variant, int rows) {
    if constexpr (std::is_same<R_, double>::value)
        return RingSmem<double>::stage_bytes(rows);
    else
        return (variant & VAR_PACK2) ? RingSmem<f2>::stage_bytes(rows) : RingSmem<float>::stage_bytes(rows);
}

// warps per block the (noise mode, lanes, wide) flavour is compiled for (SolveShape)
int FN(mppi_max_warps)(int noise_kind, int variant) {
    if constexpr (std::is_same<R_, double>::value) return 16;
    const bool pack = (variant & VAR_PACK2) != 0, wide = (variant & VAR_WIDE) != 0;
    if (noise_kind == NOISE_STREAM) return pack ? SolveShape<f2, NOISE_STREAM, false>::MAXW : SolveShape<float, NOISE_STREAM, false>::MAXW;
    if (noise_kind == NOISE_TENSOR && wide && !pack) return SolveShape<float, NOISE_TENSOR, true>::MAXW;
    return 16;
}
#endif

cudaError_t FN(launch_mppi_solve)(int noise, int variant, const EnvParamsHost &ep, const MppiArgs &a,
                                  const MppiBuffers &b, cudaStream_t s) {
    cudaError_t e = cudaErrorInvalidValue;
#if MPC_TRIG_GROUP == 0
    if (a.model_noise > 0.0) {  // noisy model: one sample per thread (capi.cu forces lanes = 1, whole tiles)
        with_scalar_env(unit_trig(variant) != 0, [&](auto env) {
            typedef decltype(env) E;
            e = (noise == NOISE_PHILOX) ? mppi_solve_go<E, NOISE_PHILOX, true>(a, ep, b, s)
                                        : mppi_solve_go<E, NOISE_TENSOR, true>(a, ep, b, s);
        });
        return e;
    }
#endif
    const bool stream = noise != NOISE_PHILOX && a.g.stream_rows > 0;
    with_rollout_env(variant, stream, [&](auto env) {
        typedef decltype(env) E;
        typedef typename E::Vec V;
        if constexpr (E::GENERIC != 0) {  // generic cost: whole tiles or Philox only (capi.cu refuses STREAM shapes)
#if MPC_TRIG_GROUP == 0
            if (stream)
                e = cudaErrorInvalidConfiguration;
            else if (noise == NOISE_PHILOX)
                e = mppi_solve_go<E, NOISE_PHILOX>(a, ep, b, s);
            else
                e = mppi_solve_go<E, NOISE_TENSOR>(a, ep, b, s);
#endif
            return;
        }
        if (noise == NOISE_PHILOX) {
#if MPC_TRIG_GROUP == 0
            e = mppi_solve_go<E, NOISE_PHILOX>(a, ep, b, s);
#endif
        } else if (stream) {
            e = mppi_solve_go<E, NOISE_STREAM>(a, ep, b, s);
        } else {
#if MPC_TRIG_GROUP == 0
            if constexpr (std::is_same<V, float>::value) {
                if (variant & VAR_WIDE) {
                    e = mppi_solve_go<E, NOISE_TENSOR, false, true>(a, ep, b, s);
                    return;
                }
            }
            e = mppi_solve_go<E, NOISE_TENSOR>(a, ep, b, s);
#endif
        }
    });
    return e;
}

#if MPC_TRIG_GROUP == 0
cudaError_t FN(launch_mppi_rank_merge)(const EnvParamsHost &ep, const MppiArgs &a, const MppiBuffers &b,
                                       cudaStream_t s) {
    cudaError_t e = cudaSuccess;
    with_scalar_env_gen(ep, false, [&](auto env) {
        typedef decltype(env) E;
        auto kern = mppi_rank_merge_kernel<E>;
        const int H = a.g.H, nw = 256 / 32;
        const size_t smem = ((size_t)a.world + H + (size_t)H * (a.g.ns + 1) + (size_t)nw * H + H) * sizeof(double);
        static size_t smem_set[64] = {0};
        int dev = 0;
        cudaGetDevice(&dev);
        // (asked for well below the 48 KB default: static shared memory and the per-block reserve count against it too)
        if (smem > 32 * 1024 && smem > smem_set[dev & 63]) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return;
            smem_set[dev & 63] = smem;
        }
        kern<<<a.g.B, 256, smem, s>>>(a, make_env<typename E::Vec>(ep), b.rank_part, b.x0, b.g_z, b.U, b.out);
        e = cudaGetLastError();
    });
    return e;
}

#if MPC_ENV_ONLY == 0
cudaError_t FN(launch_transpose_noise)(const double *src, void *dst, int K_local, int Kpad, int H, cudaStream_t s) {
    dim3 grid((Kpad + 31) / 32, (H + 31) / 32), block(32, 8);
    transpose_noise_kernel<R_><<<grid, block, 0, s>>>(src, (R_ *)dst, K_local, Kpad, H);
    return cudaGetLastError();
}

cudaError_t FN(launch_philox_fill)(void *dst, int K_local, int k_begin, int Kpad, int H, double scale, uint64_t seed,
                                   uint32_t draw, cudaStream_t s) {
    philox_fill_kernel<R_><<<(Kpad + 127) / 128, 128, 0, s>>>((R_ *)dst, K_local, k_begin, Kpad, H, scale, seed, draw, 0u);
    return cudaGetLastError();
}

cudaError_t FN(launch_cem_refit)(int noise, const CemArgs &a, const CemBuffers &b, cudaStream_t s) {
    if (noise == NOISE_TENSOR)
        cem_refit_kernel<R_, NOISE_TENSOR><<<a.g.H, 128, 0, s>>>(a, (const R_ *)b.z, b.elite_idx, b.mu, b.sd, b.sel);
    else
        cem_refit_kernel<R_, NOISE_PHILOX><<<a.g.H, 128, 0, s>>>(a, (const R_ *)b.z, b.elite_idx, b.mu, b.sd, b.sel);
    return cudaGetLastError();
}

cudaError_t FN(launch_fma_peak)(float *sink, int blocks, int iters, cudaStream_t s) {
    fma_peak_kernel<R_><<<blocks, 256, 0, s>>>(sink, iters);
    return cudaGetLastError();
}
#endif

cudaError_t FN(launch_cem_rollout)(int noise, int variant, const EnvParamsHost &ep, const CemArgs &a,
                                   const CemBuffers &b, cudaStream_t s) {
    const size_t smem = 2 * (size_t)a.g.H * sizeof(double);
    const int grid = a.g.Kpad / TILE;
    with_scalar_env_gen(ep, unit_trig(variant) != 0, [&](auto env) {
        typedef decltype(env) E;
        typedef typename E::Vec V;
        if (noise == NOISE_TENSOR)
            cem_rollout_kernel<E, NOISE_TENSOR><<<grid, TILE, smem, s>>>(a, make_env<V>(ep), (const R_ *)b.z, b.x0, b.g_z,
                                                                         b.mu, b.sd, b.R, b.keys, b.sel);
        else
            cem_rollout_kernel<E, NOISE_PHILOX><<<grid, TILE, smem, s>>>(a, make_env<V>(ep), (const R_ *)b.z, b.x0, b.g_z,
                                                                         b.mu, b.sd, b.R, b.keys, b.sel);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_cem_fused)(int variant, const EnvParamsHost &ep, const CemFusedArgs &fa, const CemBuffers &b,
                                 const CemFusedScratch &sc, int grid, int coop, cudaStream_t s) {
    cudaError_t e = cudaErrorInvalidValue;
    with_scalar_env_gen(ep, unit_trig(variant) != 0, [&](auto env) {
        typedef decltype(env) E;
        typedef typename E::Vec V;
        auto kern = cem_solve_kernel<E>;
        const size_t smem = cem_fused_smem<R_>(fa.a.g.H, fa.a.g.ns, fa.cached != 0);
        static size_t smem_set[64] = {0};
        int dev = 0;
        cudaGetDevice(&dev);
        // (asked for well below the 48 KB default: static shared memory and the per-block reserve count against it too)
        if (smem > 32 * 1024 && smem > smem_set[dev & 63]) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return;
            smem_set[dev & 63] = smem;
        }
        // Every block must be resident (they meet at grid barriers).  The grid is at most one block per SM and a block
        // fits an empty SM by construction; that is checked against the occupancy calculator once per device and
        // shape.  `coop` (the default) launches with cudaLaunchCooperativeKernel, which fails instead of deadlocking
        // if the device cannot hold the grid; MPC_B200_CEM_COOP=0 selects a plain launch (measured: no difference in
        // host time, 118 vs 120 us p50 for BASELINE configs[1]).
        static size_t checked_smem[64] = {0};
        static int per_sm[64] = {0}, sms[64] = {0};
        if (per_sm[dev & 63] == 0 || checked_smem[dev & 63] != smem) {
            int n = 0, nsm = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, CEMF_THREADS, smem);
            if (e != cudaSuccess) return;
            cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
            per_sm[dev & 63] = n;
            sms[dev & 63] = nsm;
            checked_smem[dev & 63] = smem;
        }
        if (per_sm[dev & 63] < 1 || grid > per_sm[dev & 63] * sms[dev & 63]) {
            e = cudaErrorCooperativeLaunchTooLarge;
            return;
        }
        EnvConst<V> ecv = make_env<V>(ep);
        InlineInputs inl;
        if (b.inl)
            inl = *b.inl;
        else
            inl.on = 0;
        if (!coop) {
            kern<<<grid, CEMF_THREADS, smem, s>>>(fa, ecv, b.x0, b.g_z, b.mu, b.sd, b.R, b.keys, b.elite_idx, b.sel, b.out, sc,
                                                  inl);
            e = cudaGetLastError();
            return;
        }
        const double *x0 = b.x0, *g_z = b.g_z;
        double *mu = b.mu, *sd = b.sd, *R = b.R, *out = b.out;
        unsigned long long *keys = b.keys;
        int *elite = b.elite_idx, *sel = b.sel;
        void *args[] = {(void *)&fa, (void *)&ecv, (void *)&x0, (void *)&g_z, (void *)&mu, (void *)&sd, (void *)&R,
                        (void *)&keys, (void *)&elite, (void *)&sel, (void *)&out, (void *)&sc, (void *)&inl};
        e = cudaLaunchCooperativeKernel((void *)kern, dim3(grid), dim3(CEMF_THREADS), args, smem, s);
    });
    return e;
}

cudaError_t FN(launch_cem_finish)(const EnvParamsHost &ep, const CemArgs &a, const CemBuffers &b, cudaStream_t s) {
    const size_t smem = (size_t)a.g.H * (2 + a.g.ns) * sizeof(double);
    with_scalar_env_gen(ep, false, [&](auto env) {
        typedef decltype(env) E;
        cem_finish_kernel<E><<<1, 128, smem, s>>>(a, make_env<typename E::Vec>(ep), b.x0, b.g_z, b.mu, b.sel, b.out);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_simulate)(const EnvParamsHost &ep, int H, double lo, double hi, int do_clip, const double *x0,
                                const double *us, const double *g_z, double *xs, double *cost, cudaStream_t s) {
    const size_t smem = (size_t)H * 6 * sizeof(double);
    with_scalar_env_gen(ep, false, [&](auto env) {
        typedef decltype(env) E;
        simulate_kernel<E><<<1, 64, smem, s>>>(make_env<typename E::Vec>(ep), H, lo, hi, do_clip, x0, us, g_z, xs, cost);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_model_step)(const EnvParamsHost &ep, int n, const double *x, const double *u, const double *g,
                                  double *cost, double *xn, cudaStream_t s) {
    with_scalar_env_gen(ep, false, [&](auto env) {
        typedef decltype(env) E;
        model_step_kernel<E><<<(n + 127) / 128, 128, 0, s>>>(make_env<typename E::Vec>(ep), n, x, u, g, cost, xn);
    });
    return cudaGetLastError();
}

#ifdef MPC_IS_F64
// always the float64 plant, whatever the rollout dtype
cudaError_t FN(launch_episode_step)(const EnvParamsHost &ep, const EpisodeArgs &a, const EpisodeBuffers &b, cudaStream_t s) {
    if (ep.generic)
        episode_step_kernel<TheEnv<double, 0, 1>><<<1, 128, 0, s>>>(make_env<double>(ep), a, b);
    else
        episode_step_kernel<TheEnv<double, 0>><<<1, 128, 0, s>>>(make_env<double>(ep), a, b);
    return cudaGetLastError();
}
#endif
#endif  // MPC_TRIG_GROUP == 0

}  // namespace mpc

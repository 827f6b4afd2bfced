// kernels.cu — sm_100a device code of the sampling-MPC inner loop.
//
//   mppi_rollout_kernel   MPPI.f for every sample (mppi.py:35-48) fused with the softmin-weighted noise sum of
//                         MPPI._calc_action (mppi.py:59-67): the noise tile is staged ONCE in shared memory by
//                         bulk async copies (cp.async.bulk + mbarrier), rolled out with the state in registers,
//                         then reused for sum_k w_k eps[k][t] with an online (running-min) rescaling, so the noise
//                         tensor is read from HBM once per solve.
//   mppi_finalize_kernel  merges the per-block partials (log-sum-exp style), planned_us += erg (mppi.py:69), and the
//                         clipped plan's nominal rollout (agent.py:64-69,97-108).
//   cem_*                 CEM._calc_action (cem.py:54-78): rollout of clip(mu + sd*z), radix top-n_elite select,
//                         elite mean / std refit with alpha smoothing and the epsilon test.
#include "kernels.cuh"
#include "envs.cuh"
#include "philox.cuh"

#include <float.h>

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
    return ec;
}

// Nominal rollout of one plan by a single thread: Agent.simulate_trajectory (agent.py:97-108).
// us: plan (already clipped by the caller if required), g_z rows per step, xs [H+1][ns] (nullable), returns cost.
template <class Env>
__device__ double nominal_rollout(const EnvConst<typename Env::Vec> &ec, int H, const double *x0, const double *us,
                                  const double *g_z, double *xs) {
    constexpr int NS = Env::NS, NZ = Env::NS + 1;
    typedef typename Env::S R;
    R x[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
        x[i] = (R)x0[i];
        if (xs) xs[i] = x0[i];
    }
    double reward = 0.0;
    for (int t = 0; t < H; ++t) {
        R gs[Env::NG];
        Env::prep_goal(ec, g_z + t * NZ, gs);
        const R c = Env::step(ec, x, (R)us[t], gs);
        reward += -(double)c;
        if (xs) {
#pragma unroll
            for (int i = 0; i < NS; ++i) xs[(t + 1) * NS + i] = (double)x[i];
        }
    }
    return -reward;
}

// ================================================================================================ MPPI rollout
// One block = NT = TILE / PACK threads, each rolling out PACK samples of a 128-sample tile (PACK = 2: the two samples
// travel in one f2 and every FMA-pipe instruction is an FFMA2/FADD2/FMUL2).
// Shared memory: noise tile [H][TS] (row stride TS = TILE + pad keeps rows 16-byte aligned for the bulk copies and
// makes the LDS.128 row reads of the weighted-sum pass conflict-free) | sU[H] | sw[TILE] | sE[H*parts] (float64
// running sums) | red[8] | mbarrier
template <typename R>
struct RolloutSmem {
    static constexpr int PAD = 16 / sizeof(R);
    static constexpr int TS = TILE + PAD;
    static __host__ __device__ size_t align16(size_t b) { return (b + 15) / 16 * 16; }
    static __host__ __device__ size_t off_sU(int H) { return (size_t)H * TS * sizeof(R); }
    static __host__ __device__ size_t off_sw(int H) { return off_sU(H) + align16((size_t)H * sizeof(R)); }
    static __host__ __device__ size_t off_sE(int H) { return off_sw(H) + TILE * sizeof(R); }
    static __host__ __device__ size_t off_red(int H, int parts) { return off_sE(H) + (size_t)H * parts * sizeof(double); }
    static __host__ __device__ size_t off_bar(int H, int parts) { return off_red(H, parts) + 8 * sizeof(double); }
    static __host__ __device__ size_t total(int H, int parts) { return off_bar(H, parts) + 16; }
};

// sum_k w[k] * row[k] over n (multiple of 4) consecutive k: 128-bit shared loads, packed FMAs in float32
__device__ __forceinline__ double dot_row(const float *row, const float *w, int n) {
    f2 a0(0.f), a1(0.f);
#pragma unroll 4
    for (int i = 0; i < n; i += 4) {
        const float4 e = *reinterpret_cast<const float4 *>(row + i);
        const float4 q = *reinterpret_cast<const float4 *>(w + i);
        a0 = vfma(f2(q.x, q.y), f2(e.x, e.y), a0);
        a1 = vfma(f2(q.z, q.w), f2(e.z, e.w), a1);
    }
    const f2 a = a0 + a1;
    return (double)(a.v.x + a.v.y);
}
__device__ __forceinline__ double dot_row(const double *row, const double *w, int n) {
    double a0 = 0.0, a1 = 0.0;
#pragma unroll 4
    for (int i = 0; i < n; i += 2) {
        const double2 e = *reinterpret_cast<const double2 *>(row + i);
        const double2 q = *reinterpret_cast<const double2 *>(w + i);
        a0 = fma(q.x, e.x, a0);
        a1 = fma(q.y, e.y, a1);
    }
    return a0 + a1;
}

__device__ __forceinline__ float load_lanes(const float *p, float *) { return *p; }
__device__ __forceinline__ double load_lanes(const double *p, double *) { return *p; }
__device__ __forceinline__ f2 load_lanes(const float *p, f2 *) { return f2(*reinterpret_cast<const float2 *>(p)); }

template <class Env, int NOISE>
__global__ void __launch_bounds__(TILE / VT<typename Env::Vec>::N)
    mppi_rollout_kernel(const MppiArgs a, const EnvConst<typename Env::Vec> ec, const typename Env::S *__restrict__ noise,
                        const double *__restrict__ x0g, const double *__restrict__ g_z,
                        const double *__restrict__ Ug, double *__restrict__ S_out, double *__restrict__ part) {
    typedef typename Env::Vec V;
    typedef typename Env::S R;
    constexpr int NS = Env::NS, NZ = Env::NS + 1, NG = Env::NG, PACK = VT<V>::N, NT = TILE / PACK, CH = Env::CHUNK;
    constexpr int TS = RolloutSmem<R>::TS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int H = a.g.H, parts = a.g.parts;
    R *tile = reinterpret_cast<R *>(smem_raw);
    R *sU = reinterpret_cast<R *>(smem_raw + RolloutSmem<R>::off_sU(H));
    R *sw = reinterpret_cast<R *>(smem_raw + RolloutSmem<R>::off_sw(H));
    double *sE = reinterpret_cast<double *>(smem_raw + RolloutSmem<R>::off_sE(H));
    double *red = reinterpret_cast<double *>(smem_raw + RolloutSmem<R>::off_red(H, parts));
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + RolloutSmem<R>::off_bar(H, parts));

    const int tid = threadIdx.x;
    const int tile0 = blockIdx.x * a.g.chunk;
    const int tile1 = min(tile0 + a.g.chunk, a.g.total_tiles);
    if (tile0 >= tile1) return;

    if (NOISE == NOISE_TENSOR) {
        if (tid == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
        }
        __syncthreads();
    }

    // running softmin state of the current solve (identical in every thread); E lives in shared memory
    int cur = -1;
    double beta_run = INFINITY, eta_run = 0.0;
    R gs[NG];
    R xinit[NS];
    uint32_t phase = 0;
    const V pert = V((R)a.pert);
    const int items = H * parts, KP = TILE / parts;

    auto flush = [&]() {
        __syncthreads();  // every sE item of the last tile is written
        const int first_block = (int)(((long long)cur * a.g.tiles_per_solve) / a.g.chunk);
        const int slot = blockIdx.x - first_block;
        double *rec = part + ((size_t)cur * a.g.max_parts + slot) * (PART_HDR + H);
        if (tid == 0) {
            rec[0] = beta_run;
            rec[1] = eta_run;
        }
        for (int t = tid; t < H; t += NT) {
            double e = 0.0;
            for (int p = 0; p < parts; ++p) e += sE[p * H + t];
            rec[PART_HDR + t] = e;
        }
    };

    for (int tg = tile0; tg < tile1; ++tg) {
        const int b = tg / a.g.tiles_per_solve;
        const int kt = tg - b * a.g.tiles_per_solve;
        if (b != cur) {
            if (cur >= 0) flush();
            cur = b;
            beta_run = INFINITY;
            eta_run = 0.0;
            __syncthreads();  // nobody still reads sU / sE of the previous solve
            for (int t = tid; t < H; t += NT) sU[t] = (R)Ug[(size_t)b * H + t];
            for (int w = tid; w < items; w += NT) sE[w] = 0.0;
            Env::prep_goal(ec, g_z + (size_t)b * H * NZ, gs);  // goal row 0 (model.py:34,38)
#pragma unroll
            for (int i = 0; i < NS; ++i) xinit[i] = (R)x0g[(size_t)b * NS + i];
        }
        // ---- stage the [H][TILE] noise tile ------------------------------------------------------------
        if (NOISE == NOISE_TENSOR) {
            __syncthreads();  // the previous tile's readers are done (generic proxy) ...
            if (tid == 0) {
                fence_proxy_async();  // ... before the async proxy overwrites it
                const R *src = noise + (a.noise_batched ? (size_t)b * H * a.g.Kpad : 0) + (size_t)kt * TILE;
                mbar_expect_tx(bar, (uint32_t)(H * TILE * sizeof(R)));
                for (int t = 0; t < H; ++t)
                    bulk_g2s(tile + (size_t)t * TS, src + (size_t)t * a.g.Kpad, TILE * sizeof(R), bar);
            }
        }
        __syncthreads();  // sU visible; (PHILOX) the previous weighted-sum pass has finished reading the tile
        const int kl = kt * TILE + tid * PACK;  // first local sample index of this thread
        V x[NS];
#pragma unroll
        for (int i = 0; i < NS; ++i) x[i] = V(xinit[i]);
        double S[PACK];
#pragma unroll
        for (int l = 0; l < PACK; ++l) S[l] = 0.0;
        const R *col = tile + tid * PACK;
        // one rollout step: noise lanes e -> u = U_t + e (mppi.py:39), cost + step (mppi.py:43-44), perturbation
        // term lam * u * eps / sigma^2 (mppi.py:45-46)
        auto one = [&](int t, const V &e, V &acc) {
            const V u = V(sU[t]) + e;
            const V c = Env::step(ec, x, u, gs);
            acc = acc + vfma(u * e, pert, c);
        };
        auto fold = [&](const V &acc) {
#pragma unroll
            for (int l = 0; l < PACK; ++l) S[l] += (double)lane(acc, l);
        };
        if (NOISE == NOISE_TENSOR) {
            mbar_wait(bar, phase);
            phase ^= 1u;
            int t = 0;
            for (; t + CH <= H; t += CH) {
                V acc = V((R)0);
#pragma unroll
                for (int i = 0; i < CH; ++i) one(t + i, load_lanes(col + (size_t)(t + i) * TS, (V *)nullptr), acc);
                fold(acc);
            }
            for (; t < H; ++t) {
                V acc = V((R)0);
                one(t, load_lanes(col + (size_t)t * TS, (V *)nullptr), acc);
                fold(acc);
            }
        } else {
            // same chunking of the float32 cost sums as the tensor path (CH divides 4), so both noise modes give
            // bit-identical costs for identical noise values
            static_assert(4 % CH == 0, "CHUNK must divide the Philox group of 4 steps");
            const R sig = (R)a.sigma;
            for (int t4 = 0; t4 < H; t4 += 4) {
                float z[PACK][4];
#pragma unroll
                for (int l = 0; l < PACK; ++l)
                    philox_normal4(a.seed, 0u, a.draw, (uint32_t)(a.g.k_begin + kl + l), (uint32_t)(t4 >> 2), z[l]);
                auto noise_at = [&](int i, int t) {
                    R el[PACK];
#pragma unroll
                    for (int l = 0; l < PACK; ++l) {
                        el[l] = sig * (R)z[l][i];
                        tile[(size_t)t * TS + tid * PACK + l] = el[l];
                    }
                    if constexpr (PACK == 2)
                        return V(el[0], el[PACK - 1]);
                    else
                        return V(el[0]);
                };
#pragma unroll
                for (int c0 = 0; c0 < 4; c0 += CH) {
                    const int t = t4 + c0;
                    if (t + CH <= H) {
                        V acc = V((R)0);
#pragma unroll
                        for (int i = 0; i < CH; ++i) one(t + i, noise_at(c0 + i, t + i), acc);
                        fold(acc);
                    } else {
#pragma unroll
                        for (int i = 0; i < CH; ++i) {
                            if (t + i < H) {
                                V acc = V((R)0);
                                one(t + i, noise_at(c0 + i, t + i), acc);
                                fold(acc);
                            }
                        }
                    }
                }
            }
        }
        double Smin = INFINITY;
#pragma unroll
        for (int l = 0; l < PACK; ++l) {
            const bool valid = kl + l < a.g.K_local;
            if (!valid) S[l] = INFINITY;
            if (valid) S_out[(size_t)b * a.g.Kpad + kl + l] = S[l];
            Smin = fmin(Smin, S[l]);
        }
        // ---- online softmin: rescale the running sums to the new minimum (mppi.py:59-62) ------------------
        const double beta_tile = block_reduce<true>(Smin, red);
        const double beta_new = fmin(beta_run, beta_tile);
        const double scale_old = (eta_run == 0.0) ? 0.0 : exp(a.neg_inv_lam * (beta_run - beta_new));
        double wsum = 0.0;
#pragma unroll
        for (int l = 0; l < PACK; ++l) {
            const double w = (S[l] == INFINITY) ? 0.0 : exp(a.neg_inv_lam * (S[l] - beta_new));
            sw[tid * PACK + l] = (R)w;
            wsum += w;
        }
        const double eta_tile = block_reduce<false>(wsum, red);  // its barriers also publish sw and the Philox tile
        eta_run = eta_run * scale_old + eta_tile;
        beta_run = beta_new;
        // ---- E[t] += sum_k w_k eps[k][t]: item (part, t) = row t, k in [part*KP, (part+1)*KP) ------------------
        for (int w = tid; w < items; w += NT) {
            const int p = w / H, t = w - p * H;
            const double acc = dot_row(tile + (size_t)t * TS + p * KP, sw + p * KP, KP);
            sE[w] = sE[w] * scale_old + acc;
        }
    }
    flush();
}

// ================================================================================================ MPPI finalize
// One block of FIN_THREADS per solve.  src records [count][2+H] -> (beta, eta, E).  STAGE 0/2 apply, STAGE 1 writes a
// rank record.  The merge is latency-bound (count x H doubles scattered over L2), so it is spread over all warps:
// warp w owns records j = w, w + nw, ... (eight loads in flight per thread), lanes own time steps; the per-warp
// sums meet in shared memory in a fixed order, so repeated solves are bit-identical.
constexpr int FIN_THREADS = 1024;
template <class Env>
__global__ void __launch_bounds__(FIN_THREADS) mppi_finalize_kernel(const MppiArgs a, const EnvConst<typename Env::Vec> ec,
                                                                    const int stage, const double *__restrict__ part,
                                                                    double *__restrict__ rank_part,
                                                                    const double *__restrict__ x0g,
                                                                    const double *__restrict__ g_z,
                                                                    double *__restrict__ Ug, double *__restrict__ out) {
    constexpr int NS = Env::NS, NZ = Env::NS + 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int H = a.g.H, b = blockIdx.x, tid = threadIdx.x;
    const int lane = tid & 31, w = tid >> 5, nw = FIN_THREADS / 32;
    const int rec_len = PART_HDR + H;
    __shared__ double red[FIN_THREADS / 32];
    const int cap = (stage == 2) ? a.world : a.g.max_parts;
    double *sscale = reinterpret_cast<double *>(smem_raw);  // [cap]
    double *sUc = sscale + cap;                             // [H] clipped plan
    double *sE = sUc + H;                                   // [nw][H] per-warp sums

    const double *src;
    int count;
    size_t stride;  // between records of the same solve
    if (stage == 2) {
        src = rank_part + (size_t)b * rec_len;
        count = a.world;
        stride = (size_t)a.g.B * rec_len;
    } else {
        const int first = (int)(((long long)b * a.g.tiles_per_solve) / a.g.chunk);
        const int last = (int)((((long long)(b + 1)) * a.g.tiles_per_solve - 1) / a.g.chunk);
        count = last - first + 1;
        src = part + (size_t)b * a.g.max_parts * rec_len;
        stride = rec_len;
    }
    double m = INFINITY;
    for (int j = tid; j < count; j += FIN_THREADS) m = fmin(m, src[j * stride]);
    const double beta = block_reduce<true>(m, red);
    double es = 0.0;
    for (int j = tid; j < count; j += FIN_THREADS) {
        const double bj = src[j * stride];
        const double sc = (bj == INFINITY) ? 0.0 : exp(a.neg_inv_lam * (bj - beta));
        sscale[j] = sc;
        es += sc * src[j * stride + 1];
    }
    const double eta = block_reduce<false>(es, red);  // barriers publish sscale
    for (int t0 = 0; t0 < H; t0 += 32) {
        const int t = t0 + lane;
        double e[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) e[i] = 0.0;
        if (t < H) {
            const double *col = src + PART_HDR + t;
            int j = w;
            for (; j + 7 * nw < count; j += 8 * nw) {
                double v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = col[(size_t)(j + i * nw) * stride];
#pragma unroll
                for (int i = 0; i < 8; ++i) e[i] += sscale[j + i * nw] * v[i];
            }
            for (; j < count; j += nw) e[0] += sscale[j] * col[(size_t)j * stride];
            sE[w * H + t] = ((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[7]));
        }
    }
    __syncthreads();
    double *U = Ug + (size_t)b * H;
    double *o = out + (size_t)b * (size_t)(2 + H + (H + 1) * NS);
    for (int t = tid; t < H; t += FIN_THREADS) {
        double E = 0.0;
        for (int i = 0; i < nw; ++i) E += sE[i * H + t];
        if (stage == 1) {
            rank_part[((size_t)a.rank * a.g.B + b) * rec_len + PART_HDR + t] = E;
        } else {
            const double un = U[t] + (1.0 / eta) * E;  // w = (1/eta)*tmp ; planned_us += erg   mppi.py:62,69
            U[t] = un;
            o[1 + t] = un;
            sUc[t] = fmin(fmax(un, a.lo), a.hi);  // agent.py:64-65
        }
    }
    if (stage == 1) {
        if (tid == 0) {
            double *rec = rank_part + ((size_t)a.rank * a.g.B + b) * rec_len;
            rec[0] = beta;
            rec[1] = eta;
        }
        return;
    }
    __syncthreads();
    if (tid == 0) {
        o[0] = U[0];  // u0 = planned_us[0]  (mppi.py:71)
        if (a.want_traj) {
            o[1 + H] = nominal_rollout<Env>(ec, H, x0g + (size_t)b * NS, sUc, g_z + (size_t)b * H * NZ, o + 2 + H);
        }
    }
}

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
    __syncthreads();
    const int k = blockIdx.x * TILE + tid;
    if (k >= a.g.K_local) return;
    const R lo = (R)a.lo, hi = (R)a.hi;
    double reward = 0.0;
    if (NOISE == NOISE_TENSOR) {
        for (int t = 0; t < H; ++t) {
            const R z = zt[(size_t)t * a.g.Kpad + k];
            const R u = vclip(vfma(ssd[t], z, smu[t]), lo, hi);
            reward += -(double)Env::step(ec, x, u, gs);
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
                    const R u = vclip(vfma(ssd[t], (R)z[i], smu[t]), lo, hi);
                    reward += -(double)Env::step(ec, x, u, gs);
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
        const int k = elite[e];
        R z;
        if (NOISE == NOISE_TENSOR) {
            z = zt[(size_t)t * a.g.Kpad + k];
        } else {
            float z4[4];
            philox_normal4(a.seed, 1u, (uint32_t)a.iter, (uint32_t)(a.g.k_begin + k), (uint32_t)(t >> 2), z4);
            z = (R)z4[t & 3];
        }
        return (double)vclip(vfma(s0, z, m0), lo, hi);  // identical expression to the rollout's
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
    for (int t = threadIdx.x; t < H; t += blockDim.x) {
        out[1 + t] = mu[t];
        sUc[t] = fmin(fmax(mu[t], a.lo), a.hi);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        out[0] = mu[0];
        out[2 + H + (H + 1) * NS] = (double)sel[SEL_ITERS];
        if (a.want_traj) {
            out[1 + H] = nominal_rollout<Env>(ec, H, x0g, sUc, g_z, out + 2 + H);
        }
    }
}

// ================================================================================================ small kernels
template <class Env>
__global__ void simulate_kernel(const EnvConst<typename Env::Vec> ec, int H, double lo, double hi, int do_clip,
                                const double *__restrict__ x0, const double *__restrict__ us,
                                const double *__restrict__ g_z, double *__restrict__ xs, double *__restrict__ cost) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *sU = reinterpret_cast<double *>(smem_raw);
    for (int t = threadIdx.x; t < H; t += blockDim.x) sU[t] = do_clip ? fmin(fmax(us[t], lo), hi) : us[t];
    __syncthreads();
    if (threadIdx.x == 0) {
        *cost = nominal_rollout<Env>(ec, H, x0, sU, g_z, xs);
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
    cost[i] = (double)Env::step(ec, x, (R)ug[i], gs);
#pragma unroll
    for (int j = 0; j < NS; ++j) xn[(size_t)i * NS + j] = (double)x[j];
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

// ================================================================================================ per-dtype launchers
// This file is compiled twice (build.py): -DMPC_REAL=float -DMPC_SUFFIX=_f32 with FMA contraction on, and
// -DMPC_REAL=double -DMPC_SUFFIX=_f64 with -fmad=false so the float64 path keeps the reference's rounding sequence.
#ifndef MPC_REAL
#error "compile with -DMPC_REAL=float|double -DMPC_SUFFIX=_f32|_f64"
#endif
#define MPC_CAT_(a, b) a##b
#define MPC_CAT(a, b) MPC_CAT_(a, b)
#define FN(name) MPC_CAT(name, MPC_SUFFIX)
typedef MPC_REAL R_;

// variant: bit 0 = fast_math (MUFU trig), bit 1 = two samples per thread (packed f32x2).  float64 has one flavour.
enum { VAR_FAST = 1, VAR_PACK2 = 2 };

// Calls f(Env{}) with the rollout flavour selected by (env, variant): V = float / f2 x accurate / fast, or double.
template <class F>
static void with_rollout_env(int env, int variant, F &&f) {
    if constexpr (std::is_same<R_, double>::value) {
        switch (env) {
        case ENV_PENDULUM: f(Pendulum<double, false>{}); break;
        case ENV_CARTPOLE: f(CartPole<double, false>{}); break;
        default: f(Acrobot<double, false>{}); break;
        }
    } else {
#define MPC_ENV_CASES(E)                                      \
    switch (variant & 3) {                                    \
    case 0: f(E<float, false>{}); break;                      \
    case VAR_FAST: f(E<float, true>{}); break;                \
    case VAR_PACK2: f(E<f2, false>{}); break;                 \
    default: f(E<f2, true>{}); break;                         \
    }
        switch (env) {
        case ENV_PENDULUM: MPC_ENV_CASES(Pendulum) break;
        case ENV_CARTPOLE: MPC_ENV_CASES(CartPole) break;
        default: MPC_ENV_CASES(Acrobot) break;
        }
#undef MPC_ENV_CASES
    }
}

// One sample per thread: CEM rollouts (accurate or fast trig), nominal rollouts and single steps (always accurate).
template <class F>
static void with_scalar_env(int env, bool fast, F &&f) {
    if constexpr (std::is_same<R_, double>::value) {
        switch (env) {
        case ENV_PENDULUM: f(Pendulum<double, false>{}); break;
        case ENV_CARTPOLE: f(CartPole<double, false>{}); break;
        default: f(Acrobot<double, false>{}); break;
        }
    } else {
#define MPC_ENV_CASES(E)            \
    if (fast)                       \
        f(E<float, true>{});        \
    else                            \
        f(E<float, false>{});
        switch (env) {
        case ENV_PENDULUM: MPC_ENV_CASES(Pendulum) break;
        case ENV_CARTPOLE: MPC_ENV_CASES(CartPole) break;
        default: MPC_ENV_CASES(Acrobot) break;
        }
#undef MPC_ENV_CASES
    }
}

template <class Env, int NOISE>
static cudaError_t mppi_rollout_go(const MppiArgs &a, const EnvParamsHost &ep, const MppiBuffers &b, cudaStream_t s) {
    typedef typename Env::Vec V;
    auto kern = mppi_rollout_kernel<Env, NOISE>;
    const size_t smem = RolloutSmem<R_>::total(a.g.H, a.g.parts);
    static size_t smem_set = 0;  // per instantiation: raise the dynamic shared memory limit only when it grows
    if (smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        smem_set = smem;
    }
    kern<<<a.g.grid, TILE / VT<V>::N, smem, s>>>(a, make_env<V>(ep), (const R_ *)b.noise, b.x0, b.g_z, b.U, b.S, b.part);
    return cudaGetLastError();
}

template <class Env, int NOISE>
static int mppi_rollout_occ(int H, int parts) {
    auto kern = mppi_rollout_kernel<Env, NOISE>;
    const size_t smem = RolloutSmem<R_>::total(H, parts);
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, TILE / VT<typename Env::Vec>::N, smem) != cudaSuccess)
        return 0;
    return occ;
}

size_t FN(mppi_rollout_smem)(int H, int parts) { return RolloutSmem<R_>::total(H, parts); }

cudaError_t FN(launch_mppi_rollout)(int noise, int variant, const EnvParamsHost &ep, const MppiArgs &a,
                                    const MppiBuffers &b, cudaStream_t s) {
    cudaError_t e = cudaSuccess;
    with_rollout_env(ep.env, variant, [&](auto env) {
        typedef decltype(env) E;
        e = (noise == NOISE_TENSOR) ? mppi_rollout_go<E, NOISE_TENSOR>(a, ep, b, s)
                                    : mppi_rollout_go<E, NOISE_PHILOX>(a, ep, b, s);
    });
    return e;
}

int FN(mppi_rollout_occupancy)(int noise, int variant, int env, int H, int parts) {
    int occ = 0;
    with_rollout_env(env, variant, [&](auto e) {
        typedef decltype(e) E;
        occ = (noise == NOISE_TENSOR) ? mppi_rollout_occ<E, NOISE_TENSOR>(H, parts)
                                      : mppi_rollout_occ<E, NOISE_PHILOX>(H, parts);
    });
    return occ;
}

template <class Env>
static cudaError_t mppi_finalize_go(int stage, const MppiArgs &a, const EnvParamsHost &ep, const MppiBuffers &b,
                                    cudaStream_t s) {
    auto kern = mppi_finalize_kernel<Env>;
    const int cap = (stage == 2) ? a.world : a.g.max_parts;
    const size_t smem = ((size_t)cap + a.g.H + (size_t)(FIN_THREADS / 32) * a.g.H) * sizeof(double);
    static size_t smem_set = 48 * 1024;
    if (smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        smem_set = smem;
    }
    kern<<<a.g.B, FIN_THREADS, smem, s>>>(a, make_env<typename Env::Vec>(ep), stage, b.part, b.rank_part, b.x0, b.g_z,
                                          b.U, b.out);
    return cudaGetLastError();
}

cudaError_t FN(launch_mppi_finalize)(int stage, const EnvParamsHost &ep, const MppiArgs &a, const MppiBuffers &b,
                                     cudaStream_t s) {
    cudaError_t e = cudaSuccess;
    with_scalar_env(ep.env, false, [&](auto env) { e = mppi_finalize_go<decltype(env)>(stage, a, ep, b, s); });
    return e;
}

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

cudaError_t FN(launch_cem_rollout)(int noise, int variant, const EnvParamsHost &ep, const CemArgs &a,
                                   const CemBuffers &b, cudaStream_t s) {
    const size_t smem = 2 * (size_t)a.g.H * sizeof(double);
    const int grid = a.g.tiles_per_solve;
    with_scalar_env(ep.env, (variant & VAR_FAST) != 0, [&](auto env) {
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

cudaError_t FN(launch_cem_refit)(int noise, const CemArgs &a, const CemBuffers &b, cudaStream_t s) {
    if (noise == NOISE_TENSOR)
        cem_refit_kernel<R_, NOISE_TENSOR><<<a.g.H, 128, 0, s>>>(a, (const R_ *)b.z, b.elite_idx, b.mu, b.sd, b.sel);
    else
        cem_refit_kernel<R_, NOISE_PHILOX><<<a.g.H, 128, 0, s>>>(a, (const R_ *)b.z, b.elite_idx, b.mu, b.sd, b.sel);
    return cudaGetLastError();
}

cudaError_t FN(launch_cem_finish)(const EnvParamsHost &ep, const CemArgs &a, const CemBuffers &b, cudaStream_t s) {
    const size_t smem = (size_t)a.g.H * sizeof(double);
    with_scalar_env(ep.env, false, [&](auto env) {
        typedef decltype(env) E;
        cem_finish_kernel<E><<<1, 128, smem, s>>>(a, make_env<typename E::Vec>(ep), b.x0, b.g_z, b.mu, b.sel, b.out);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_simulate)(const EnvParamsHost &ep, int H, double lo, double hi, int do_clip, const double *x0,
                                const double *us, const double *g_z, double *xs, double *cost, cudaStream_t s) {
    const size_t smem = (size_t)H * sizeof(double);
    with_scalar_env(ep.env, false, [&](auto env) {
        typedef decltype(env) E;
        simulate_kernel<E><<<1, 64, smem, s>>>(make_env<typename E::Vec>(ep), H, lo, hi, do_clip, x0, us, g_z, xs, cost);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_model_step)(const EnvParamsHost &ep, int n, const double *x, const double *u, const double *g,
                                  double *cost, double *xn, cudaStream_t s) {
    with_scalar_env(ep.env, false, [&](auto env) {
        typedef decltype(env) E;
        model_step_kernel<E><<<(n + 127) / 128, 128, 0, s>>>(make_env<typename E::Vec>(ep), n, x, u, g, cost, xn);
    });
    return cudaGetLastError();
}

cudaError_t FN(launch_fma_peak)(float *sink, int blocks, int iters, cudaStream_t s) {
    fma_peak_kernel<R_><<<blocks, 256, 0, s>>>(sink, iters);
    return cudaGetLastError();
}

}  // namespace mpc

// cem_fused.cuh — CEM._calc_action (cem.py:54-78) as ONE persistent cooperative launch (included by kernels.cu).
//
// The per-iteration launches of the first implementation (rollout 12 us on 32 SMs, one-block radix select 12 us, refit
// 8 us: 17 launches and 171 us of latency-bound device time for BASELINE configs[1]) become phases of one kernel whose
// blocks stay resident and meet at a grid barrier twice per iteration:
//
//   every block owns a contiguous range of `spb` samples (K = 4096: 128 blocks x 32 samples)
//   [rollout]   clip(mu + sd z) with in-kernel Philox normals, reward, order-preserving key        cem.py:58-64
//   ---- grid barrier: all keys published
//   [select]    K <= 4096: every block ranks ITS samples against all K keys (all-pairs count from a shared-memory copy
//               of the keys: rank < n_elite <=> elite, ties by lowest index — np.argsort(-rewards)[:n_elite],
//               cem.py:67, with no serial select step at all); larger K: MSD radix passes with one global 256-bin
//               histogram per pass (blocks add their shared-memory histograms, a barrier, every block scans the same
//               256 bins), stopping as soon as the bin of the n_elite-th key is wanted entirely
//   [refit]     the block regenerates the samples of its elites (same Philox counters, same expression as the rollout)
//               and accumulates sum (x - c) and sum (x - c)^2 per time step, c = the current mean    cem.py:69-70
//   ---- grid barrier: all partial sums published
//   [update]    EVERY block reduces the partial sums in block order (identical result everywhere, no broadcast step),
//               applies the alpha smoothing and the epsilon test (cem.py:71-76); the elite indices go to the history
//   after the last iteration block 0 writes u0 / the plan and rolls out the clipped mean (agent.py:64-69,97-108).
//
// Deterministic: integer histogram atomics, fixed summation orders.  Single GPU, Philox noise; host-drawn normals and
// K sharded over several GPUs keep the per-iteration kernels (kernels.cu / dispatch.cu).
#pragma once

namespace mpc {

constexpr int CEMF_THREADS = CEM_FUSED_THREADS;
constexpr int CEMF_EC = 16;          // elites whose samples are regenerated per refit round

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu_u32(unsigned *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_add_acq_rel_gpu_u32(unsigned *p, unsigned v) {
    unsigned old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
// Generation barrier over the (co-resident: cooperative launch) blocks of the grid.  The block barriers make thread 0's
// acq_rel arrival cumulative over the whole block's earlier stores, and the other blocks' stores visible to every
// thread afterwards.  The last arriver resets the count before it releases the new generation, so the two words are
// ready for the next barrier and the next launch.  `gen` is the caller's copy of the generation: read once at kernel
// start (no block can have completed a barrier before every block has read it and arrived), incremented locally.
__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned nblocks, unsigned &gen) {
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atom_add_acq_rel_gpu_u32(&bar[0], 1u) == nblocks - 1) {
            bar[0] = 0u;
            st_release_gpu_u32(&bar[1], gen + 1u);
        } else {
            while (ld_acquire_gpu_u32(&bar[1]) == gen) {
            }
        }
    }
    ++gen;
    __syncthreads();
}

// exclusive scan of one int per thread over the block (blockDim <= 1024); *total = block sum.  swarp: >= 33 ints.
__device__ __forceinline__ int block_excl_scan(int v, int *swarp, int *total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    __syncthreads();  // swarp free
    if (lane == 31) swarp[w] = inc;
    __syncthreads();
    int base = 0, tot = 0;
    for (int i = 0; i < nw; ++i) {
        const int s = swarp[i];
        if (i < w) base += s;
        tot += s;
    }
    *total = tot;
    return base + inc - v;
}

template <class Env>
__global__ void __launch_bounds__(CEMF_THREADS, 1)
    cem_solve_kernel(const CemFusedArgs fa, const EnvConst<typename Env::Vec> ec, const double *__restrict__ x0g,
                     const double *__restrict__ g_z, double *__restrict__ mu_g, double *__restrict__ sd_g,
                     double *__restrict__ Rhist, unsigned long long *__restrict__ keys, int *__restrict__ elite_hist,
                     int *__restrict__ sel, double *__restrict__ out, const CemFusedScratch sc,
                     const __grid_constant__ InlineInputs inl) {  // batch-1 host call: state / plan as parameters
    typedef typename Env::S R;
    constexpr int NS = Env::NS, NG = Env::NG, NZ = Env::NS + 1;
    const CemArgs &a = fa.a;
    const int H = a.g.H, K = a.g.K_local, tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, w = tid >> 5;
    const int G = gridDim.x, blk = blockIdx.x;
    const int k0 = blk * fa.spb, k1 = min(k0 + fa.spb, K);
    const int n_elite = a.n_elite;
    // ---- shared memory
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *smu = reinterpret_cast<double *>(smem_raw);        // [H] mean (float64, the state of the iteration)
    double *ssd = smu + H;                                      // [H] std
    double *xs = ssd + H;                                       // [CEMF_EC][H] regenerated elite samples
    double *sUc = xs + (size_t)CEMF_EC * H;                     // [H] finish: clipped plan
    double *sG = sUc + H;                                       // [H][NZ] finish: goal trajectory
    R *smuR = reinterpret_cast<R *>(sG + (size_t)H * NZ);       // [H] mean / std in the rollout dtype
    R *ssdR = smuR + H;
    int *clist = reinterpret_cast<int *>(ssdR + H);             // [nt] elites of the current chunk, ascending
    int *swarp = clist + nt;                                    // [40] scans, flags
    int *shist = swarp + 40;                                    // [256] radix histogram / [32] ranks
    unsigned long long *smine = reinterpret_cast<unsigned long long *>(shist + 256);  // [32] all-pairs: this block's keys
    double *sred = reinterpret_cast<double *>(smine + 32);      // [2][nt] partial column sums of the update
    R *su = reinterpret_cast<R *>(sred + 2 * nt);               // [H][32] cached mode: this block's samples
    R *sz = su + (size_t)H * 32;                                // [H][32] cached mode: standard normals of the NEXT iteration
    __shared__ unsigned long long s_prefix;
    __shared__ int s_want, s_stop;

    unsigned bar_gen = ld_acquire_gpu_u32(&sc.bar[1]);
    // inline inputs: keep the device copy of the state current for device-resident solves enqueued later
    if ((inl.on & 1) && blk == 0 && tid < NS) const_cast<double *>(x0g)[tid] = inl.x0[tid];
    // ---- start of the solve: std = self.std.copy() (cem.py:56), mean = planned_us
    for (int t = tid; t < H; t += nt) {
        smu[t] = (inl.on & 2) ? inl.U[t] : mu_g[t];
        ssd[t] = a.sigma;
    }
    if (!fa.all_pairs)
        for (int i = blk * nt + tid; i < fa.max_iter * 8 * 256; i += G * nt) sc.hist[i] = 0;
    R gs[NG], xinit[NS];
    Env::prep_goal(ec, g_z, gs);
#pragma unroll
    for (int i = 0; i < NS; ++i) xinit[i] = (R)((inl.on & 1) ? inl.x0[i] : x0g[i]);
    Env::enter(ec, xinit);
    const R lo = (R)a.lo, hi = (R)a.hi;
    __syncthreads();

    // profiling: block 0 stamps the phases of iteration 1 (iteration 0 runs on cold caches) and the kernel's ends
    auto stamp = [&](int it, int i) {
        if (fa.trace && blk == 0 && tid == 0 && (i == 0 || i == 8 || i == 9 || it == (fa.max_iter > 1 ? 1 : 0))) fa.trace[i] = global_ns();
    };
    stamp(0, 0);
    int iters = 0, done = 0;
    for (int t = tid; t < H; t += nt) {  // mean / std in the rollout dtype (kept current by the update below)
        smuR[t] = (R)smu[t];
        ssdR[t] = (R)ssd[t];
    }
    // cached mode: the standard normals of iteration `draw` for the block's 32 samples, one Philox block per task,
    // drawn by threads first .. nt-1 (all of them before the first iteration; afterwards the 15 warps that have no
    // rollout to do draw the NEXT iteration's normals while warp 0 walks down its dependency chain)
    auto draw_normals = [&](const int draw, const int first) {
        const int nq = (H + 3) >> 2;
        for (int task = tid - first; task < 32 * nq; task += nt - first) {
            const int sidx = task & 31, tq = task >> 5;
            float z[4] = {0.f, 0.f, 0.f, 0.f};
            if (k0 + sidx < k1) philox_normal4(a.seed, 1u, (uint32_t)draw, (uint32_t)(a.g.k_begin + k0 + sidx), (uint32_t)tq, z);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int t = 4 * tq + i;
                if (t < H) sz[t * 32 + sidx] = (R)z[i];
            }
        }
    };
    if (fa.cached) draw_normals(0, 0);
    __syncthreads();
    for (int it = 0; it < fa.max_iter; ++it) {
        stamp(it, 1);
        // ---------------------------------------------------------------- rollout of this block's samples
        if (fa.cached) {
            // one warp's worth of samples per block: the samples clip(mu + sd z) are formed in shared memory from the
            // normals drawn ahead; warp 0 rolls them out (a pure dependency chain from here) and the refit reads the
            // elites' samples back from the same array
            for (int i = tid; i < 32 * H; i += nt) {
                const int t = i >> 5;
                su[i] = vclip(sample_value(ssdR[t], sz[i], smuR[t]), lo, hi);
            }
            __syncthreads();
            if (tid >= 32 && it + 1 < fa.max_iter) draw_normals(it + 1, 32);
            const int k = k0 + tid;
            if (tid < 32 && k < k1) {
                const uint32_t kglob = (uint32_t)(a.g.k_begin + k);
                R x[NS];
#pragma unroll
                for (int i = 0; i < NS; ++i) x[i] = xinit[i];
                double reward = 0.0;
#pragma unroll 2
                for (int t = 0; t < H; ++t) {
                    if constexpr (Env::GENERIC != 0)
                        if (a.per_step_goal) Env::prep_goal(ec, g_z + (size_t)t * NZ, gs);
                    reward += -(double)Env::step(ec, x, su[t * 32 + tid], gs);
                    if (a.model_noise > 0.0)
                        add_model_noise<Env>(x, a.model_noise, a.seed, PHILOX_MODEL_CEM, fa.noise_draw0 + (uint32_t)it, kglob,
                                             (uint32_t)t);
                }
                Rhist[(size_t)it * a.g.Kpad + k] = reward;
                keys[k] = reward_key(reward);
            }
        } else {
            for (int k = k0 + tid; k < k1; k += nt) {
                const uint32_t kglob = (uint32_t)(a.g.k_begin + k);
                R x[NS];
#pragma unroll
                for (int i = 0; i < NS; ++i) x[i] = xinit[i];
                double reward = 0.0;
                for (int t4 = 0; t4 < H; t4 += 4) {
                    float z[4];
                    philox_normal4(a.seed, 1u, (uint32_t)it, kglob, (uint32_t)(t4 >> 2), z);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int t = t4 + i;
                        if (t < H) {
                            const R u = vclip(sample_value(ssdR[t], (R)z[i], smuR[t]), lo, hi);
                            if constexpr (Env::GENERIC != 0)
                                if (a.per_step_goal) Env::prep_goal(ec, g_z + (size_t)t * NZ, gs);
                            reward += -(double)Env::step(ec, x, u, gs);
                            if (a.model_noise > 0.0)
                                add_model_noise<Env>(x, a.model_noise, a.seed, PHILOX_MODEL_CEM,
                                                     fa.noise_draw0 + (uint32_t)it, kglob, (uint32_t)t);
                        }
                    }
                }
                Rhist[(size_t)it * a.g.Kpad + k] = reward;
                keys[k] = reward_key(reward);
            }
        }
        stamp(it, 2);
        grid_barrier(sc.bar, (unsigned)G, bar_gen);
        stamp(it, 3);
        // ---------------------------------------------------------------- elite threshold
        unsigned long long prefix = 0ull, mask = 0ull;
        int want_ties = 0, tie_base = 0;
        bool split_ties = false;
        if (fa.all_pairs) {
            // rank of each of the block's (<= 32) samples among all K keys = number of keys that beat it.  "Beats" is a
            // strict total order on (key, index): larger key, or equal key at a LOWER index (np.argsort's order among
            // equal rewards here).  Warp 0 sorts the block's 32 samples by that order (rank counting among themselves);
            // every thread then locates each of its 8 keys (K <= 4096 = 8 x 512) among the sorted samples by binary
            // search — 5-6 compares instead of 32 — and adds one to the bin of that position; a suffix sum over the 33
            // bins gives every sample the number of keys above it.
            constexpr int KPT = CEM_FUSED_ALLPAIRS_K / CEMF_THREADS;
            unsigned long long kf[KPT];
#pragma unroll
            for (int i = 0; i < KPT; ++i) {
                const int j = tid + i * nt;
                kf[i] = j < K ? __ldcg(keys + j) : 0ull;
            }
            if (tid < 64) shist[tid] = 0;  // [0..32] position bins, [33..] spare
            if (tid < 32) {
                const int ks = k0 + tid;
                const unsigned long long mine = ks < k1 ? __ldcg(keys + ks) : ~0ull;  // padding lanes sort to the top
                int below = 0;  // samples of this block that `mine` beats
#pragma unroll
                for (int o = 0; o < 32; ++o) {
                    const unsigned long long other = __shfl_sync(0xffffffffu, mine, o);
                    below += (mine > other || (mine == other && tid < o)) ? 1 : 0;
                }
                smine[below] = mine;      // ascending in the total order: position 0 is beaten by every other sample
                clist[below] = ks;        // the sample at that position
                clist[32 + tid] = below;  // and the position of sample tid
            }
            __syncthreads();
            stamp(it, 10);
            // p = number of sorted samples that key j beats (they form a prefix of the ascending order): a branch-free
            // binary search, the 8 searches of a thread interleaved.  The keys alone decide it unless a probed sample's
            // key is EQUAL (exact ties: rare) — then the search is repeated with the index as the tie-break.
            int pos[KPT], selfpos[KPT];
            unsigned eqmask = 0u;
#pragma unroll
            for (int i = 0; i < KPT; ++i) {
                pos[i] = 0;
                const int own = tid + i * nt - k0;  // a key of this block's own samples equals itself: not a tie
                selfpos[i] = (own >= 0 && own < 32) ? clist[32 + own] : -1;
            }
#pragma unroll
            for (int step = 32; step >= 1; step >>= 1) {
#pragma unroll
                for (int i = 0; i < KPT; ++i) {
                    const int cand = pos[i] + step, at = min(cand, 32) - 1;
                    const unsigned long long m = smine[at];
                    eqmask |= (kf[i] == m && at != selfpos[i]) ? (1u << i) : 0u;
                    pos[i] = (cand <= 32 && kf[i] > m) ? cand : pos[i];
                }
            }
            if (eqmask) {
#pragma unroll
                for (int i = 0; i < KPT; ++i) {
                    if (eqmask & (1u << i)) {
                        const int j = tid + i * nt;
                        int lo_ = 0, hi_ = 32;
                        while (lo_ < hi_) {
                            const int mid = (lo_ + hi_) >> 1;
                            const unsigned long long m = smine[mid];
                            if (kf[i] > m || (kf[i] == m && j < clist[mid])) lo_ = mid + 1; else hi_ = mid;
                        }
                        pos[i] = lo_;
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < KPT; ++i)
                if (tid + i * nt < K) atomicAdd(&shist[pos[i]], 1);
            __syncthreads();
            stamp(it, 11);
            if (tid < 32) {
                // keys beating the sample at sorted position q: those with p > q
                int above = 0;
                for (int p_ = tid + 1; p_ <= 32; ++p_) above += shist[p_];
                shist[40 + tid] = above;
            }
            __syncthreads();
            if (tid < 32) shist[tid] = shist[40 + clist[32 + tid]];  // back to sample order: rank of sample tid
            __syncthreads();
        } else {
            int want = n_elite;
            bool stop = false;
            for (int pass = 0; pass < 8 && !stop; ++pass) {
                const int shift = 56 - 8 * pass;
                if (tid < 256) shist[tid] = 0;
                __syncthreads();
                for (int k = k0 + tid; k < k1; k += nt) {
                    const unsigned long long key = keys[k];
                    if ((key & mask) == prefix) atomicAdd(&shist[(int)((key >> shift) & 0xffull)], 1);
                }
                __syncthreads();
                int *gh = sc.hist + ((size_t)it * 8 + pass) * 256;
                if (tid < 256 && shist[tid]) atomicAdd(&gh[tid], shist[tid]);
                grid_barrier(sc.bar, (unsigned)G, bar_gen);
                if (w == 0) {
                    // descending scan over the 256 global bins: lane l owns bins 255 - 8 l ... 248 - 8 l
                    int c[8], sum = 0;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        c[i] = __ldcg(gh + 255 - 8 * lane - i);
                        sum += c[i];
                    }
                    int inc = sum;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int v = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += v;
                    }
                    int excl = inc - sum;
                    if (lane == 0) s_stop = 0;
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (excl < want && excl + c[i] >= want) {
                            s_prefix = prefix | ((unsigned long long)(255 - 8 * lane - i) << shift);
                            s_want = want - excl;
                            if (c[i] == want - excl) s_stop = 1;  // the whole bin is wanted: nothing left to decide
                        }
                        excl += c[i];
                    }
                }
                __syncthreads();
                prefix = s_prefix;
                want = s_want;
                stop = s_stop != 0;
                mask |= 0xffull << shift;
                __syncthreads();
            }
            want_ties = want;
            split_ties = !stop;  // eight digits and still more equal keys than wanted: exact ties, lowest indices win
            if (split_ties) {
                int c = 0;
                for (int k = k0 + tid; k < k1; k += nt) c += (keys[k] == prefix) ? 1 : 0;
                int tot;
                block_excl_scan(c, swarp, &tot);
                if (tid == 0) sc.tie_cnt[blk] = tot;
                grid_barrier(sc.bar, (unsigned)G, bar_gen);
                for (int b = 0; b < blk; ++b) tie_base += __ldcg(sc.tie_cnt + b);
            }
        }
        stamp(it, 4);
        // ---------------------------------------------------------------- this block's elites, refit partial sums
        double s1 = 0.0, s2 = 0.0;
        const double cshift = tid < H ? smu[tid] : 0.0;
        int n_el = 0;
        // E(t) partial sums over the elites listed in clist[0 .. tot)
        auto sum_elites = [&](const int tot) {
            if (fa.cached) {
                if (tid < H)
                    for (int e = 0; e < tot; ++e) {
                        const double d = (double)su[tid * 32 + (clist[e] - k0)] - cshift;
                        s1 += d;
                        s2 += d * d;
                    }
            } else {
                for (int e0 = 0; e0 < tot; e0 += CEMF_EC) {
                    const int ne = min(CEMF_EC, tot - e0), nq = (H + 3) >> 2;
                    for (int task = tid; task < ne * nq; task += nt) {
                        const int e = task / nq, tq = task - e * nq;
                        float z[4];
                        philox_normal4(a.seed, 1u, (uint32_t)it, (uint32_t)(a.g.k_begin + clist[e0 + e]), (uint32_t)tq, z);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int t = 4 * tq + i;
                            if (t < H)  // identical expression to the rollout's
                                xs[(size_t)e * H + t] = (double)vclip(sample_value(ssdR[t], (R)z[i], smuR[t]), lo, hi);
                        }
                    }
                    __syncthreads();
                    if (tid < H)
                        for (int e = 0; e < ne; ++e) {
                            const double d = xs[(size_t)e * H + tid] - cshift;
                            s1 += d;
                            s2 += d * d;
                        }
                    __syncthreads();
                }
            }
        };
        if (fa.all_pairs) {
            // one warp's worth of samples: warp 0 compacts the elites with a ballot (ascending index)
            if (tid < 32) {
                const int k = k0 + tid;
                const bool elite = k < k1 && shist[tid] < n_elite;
                const unsigned m = __ballot_sync(0xffffffffu, elite);
                if (elite) {
                    const int pos = __popc(m & ((1u << tid) - 1u));
                    clist[pos] = k;
                    sc.stage[k0 + pos] = k;
                }
                if (tid == 0) swarp[0] = __popc(m);
            }
            __syncthreads();
            n_el = swarp[0];
            sum_elites(n_el);
        } else {
            for (int c0 = k0; c0 < k1; c0 += nt) {
                const int k = c0 + tid;
                bool is_tie = false, elite = false;
                if (k < k1) {
                    const unsigned long long km = keys[k] & mask;
                    elite = km > prefix || (km == prefix && !split_ties);
                    is_tie = split_ties && km == prefix;
                }
                int tot;
                if (split_ties) {  // uniform over the grid
                    const int r = block_excl_scan(is_tie ? 1 : 0, swarp, &tot);
                    if (is_tie && tie_base + r < want_ties) elite = true;
                    tie_base += tot;
                }
                const int pos = block_excl_scan(elite ? 1 : 0, swarp, &tot);
                if (elite) {
                    clist[pos] = k;
                    sc.stage[k0 + n_el + pos] = k;
                }
                __syncthreads();
                sum_elites(tot);
                __syncthreads();  // clist is rewritten by the next chunk
                n_el += tot;
            }
        }
        double *slot = sc.slots + (size_t)blk * (2 + 2 * H);
        if (tid == 0) slot[0] = (double)n_el;
        if (tid < H) {
            slot[2 + tid] = s1;
            slot[2 + H + tid] = s2;
        }
        stamp(it, 5);
        grid_barrier(sc.bar, (unsigned)G, bar_gen);
        stamp(it, 6);
        // ---------------------------------------------------------------- every block: reduce, update, epsilon test
        // column t of the G slots is summed by P threads (slots p, p + P, ...: every load in flight at once), whose
        // partial sums meet in shared memory in a fixed order: identical in every block and from run to run
        int notsmall = 0;
        int before = 0;  // elites of the blocks before this one (loads issued ahead of the column loads)
        for (int b = lane; b < blk; b += 32) before += (int)__ldcg(sc.slots + (size_t)b * (2 + 2 * H));
        {
            int TT = 32;
            while (TT < H) TT <<= 1;
            const int P = nt / TT, t = tid & (TT - 1), pp = tid / TT;  // H <= nt
            double p1 = 0.0, p2 = 0.0;
            if (t < H && pp < P) {
                const double *col = sc.slots + 2 + t;
                constexpr int MAXR = 16;  // G <= 148 slots over P >= 2 partitions: one batch, every load in flight at once
                double v1[MAXR], v2[MAXR];
                for (int b0 = pp; b0 < G; b0 += P * MAXR) {
#pragma unroll
                    for (int i = 0; i < MAXR; ++i) {
                        const int b = b0 + i * P;
                        v1[i] = b < G ? __ldcg(col + (size_t)b * (2 + 2 * H)) : 0.0;
                        v2[i] = b < G ? __ldcg(col + (size_t)b * (2 + 2 * H) + H) : 0.0;
                    }
#pragma unroll
                    for (int i = 0; i < MAXR; ++i) {
                        p1 += v1[i];
                        p2 += v2[i];
                    }
                }
            }
            sred[tid] = p1;
            sred[nt + tid] = p2;
            __syncthreads();
        }
        if (tid < H) {
            int TT = 32;
            while (TT < H) TT <<= 1;
            const int P = nt / TT;
            double S1 = 0.0, S2 = 0.0;
            for (int pp = 0; pp < P; ++pp) {
                S1 += sred[pp * TT + tid];
                S2 += sred[nt + pp * TT + tid];
            }
            const double m1 = S1 / n_elite;
            const double mean = cshift + m1;                              // np.mean(elites, axis=0)
            const double var = fmax(S2 / n_elite - m1 * m1, 0.0);         // np.std: ddof = 0
            const double new_mu = a.alpha * smu[tid] + (1 - a.alpha) * mean;      // cem.py:71-72
            const double new_sd = a.alpha * ssd[tid] + (1 - a.alpha) * sqrt(var); // cem.py:73
            smu[tid] = new_mu;
            ssd[tid] = new_sd;
            smuR[tid] = (R)new_mu;
            ssdR[tid] = (R)new_sd;
            notsmall = !(new_sd < a.epsilon);
            if (blk == 0) {
                mu_g[tid] = new_mu;
                sd_g[tid] = new_sd;
            }
        }
        // the elite indices of this iteration, ascending: this block's go behind those of the blocks before it
        {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
            for (int i = tid; i < n_el; i += nt)
                if (before + i < n_elite) elite_hist[(size_t)it * n_elite + before + i] = sc.stage[k0 + i];
        }
        iters = it + 1;
        stamp(it, 7);
        if (!__syncthreads_or(notsmall)) {  // (std < epsilon).all()   cem.py:75-76
            done = 1;
            break;
        }
    }
    if (blk != 0) return;
    stamp(0, 8);
    // ---------------------------------------------------------------- u0 = planned_us[0] (cem.py:77), nominal rollout
    for (int t = tid; t < H; t += nt) {
        out[1 + t] = smu[t];
        sUc[t] = fmin(fmax(smu[t], a.lo), a.hi);
    }
    if (a.want_traj)
        for (int i = tid; i < H * NZ; i += nt) sG[i] = g_z[i];
    __syncthreads();
    const size_t stride = (size_t)(2 + H + (H + 1) * NS);
    if (tid == 0) {
        out[0] = smu[0];
        out[stride] = (double)iters;
        sel[SEL_ITERS] = iters;
        sel[SEL_DONE] = done;
        if (a.want_traj)
            out[1 + H] = nominal_rollout<Env>(ec, H, (inl.on & 1) ? inl.x0 : x0g, sUc, sG, NZ, out + 2 + H, a.model_noise,
                                              a.seed, fa.noise_draw0, 0u);
    }
    stamp(0, 9);
    if (fa.host_out) {
        __syncthreads();
        const int n_copy = a.want_traj ? (int)stride : 1 + H;
        for (int i = tid; i < n_copy; i += nt) fa.host_out[i] = out[i];
        if (tid == 0) {
            fa.host_out[stride] = out[stride];
            fa.host_out[stride + 1] = 0.0;
        }
        __threadfence_system();
        __syncthreads();
        if (tid == 0) st_release_sys(fa.host_flag, fa.host_epoch);
    }
}

// dynamic shared memory of cem_solve_kernel
template <typename R>
static size_t cem_fused_smem(int H, int ns, bool cached) {
    size_t b = (size_t)(2 * H + (size_t)CEMF_EC * H + H + (size_t)H * (ns + 1)) * sizeof(double);
    b += 2 * (size_t)H * sizeof(R);
    b = (b + 7) / 8 * 8;
    b += (size_t)(CEMF_THREADS + 40 + 256) * sizeof(int);
    b = (b + 7) / 8 * 8;
    b += 32 * sizeof(unsigned long long) + 2 * (size_t)CEMF_THREADS * sizeof(double);
    if (cached) b += 2 * (size_t)H * 32 * sizeof(R);
    return b + 16;
}

}  // namespace mpc

// mppi.cuh — the MPPI solve as ONE persistent sm_100a kernel (included by kernels.cu, compiled per dtype).
//
//   MPPI.f for every sample (mppi.py:35-48)  +  softmin weights and weighted noise sum (mppi.py:59-67)
//   +  planned_us += erg (mppi.py:69)  +  the clipped plan's nominal rollout (agent.py:64-69,97-108)
//
// Shape: one block per SM, W warps, a ring of S shared-memory stages.  A stage holds the [H][WT] noise tile of one
// WARP (WT = 32 lanes x PACK samples, dense rows of 128/256 B), filled by ONE TMA tensor copy
// (cp.async.bulk.tensor.3d, box [H][WT] of the [H][Kpad] noise tensor) that completes on the stage's mbarrier.  Tile j of the block goes to warp j % W and stage j % S; the warp that finishes a tile
// re-arms its stage with tile j + S, so S - W tiles are always in flight ahead of the consumers and no warp ever
// waits on another warp: there is NO block barrier inside the tile loop.  Everything per tile is warp-local:
//   rollout (state in registers, noise read from the stage) -> warp-shuffle min -> softmin weights -> the weighted
//   sum sum_k w_k eps[k][t] from the SAME stage (the noise tensor is read from HBM once per solve), accumulated in
//   registers with the running-minimum rescaling (lane owns rows t = lane, lane + 32, ...).
// Tail: the warps' (beta, eta, E[H]) records meet in shared memory, one record per block goes to global memory and
// the last block to arrive (atomic ticket per solve) merges them, updates the plan and rolls out the nominal
// trajectory — no second launch.  With many small solves per launch (batch) a block owns whole solves, keeps
// per-warp records and finalises its solves itself, one thread per nominal rollout.
#pragma once

namespace mpc {

constexpr int MAX_WARPS = 32;   // warps per rollout block (the register budget of a flavour may allow fewer: SolveShape)
constexpr int MAX_ROWS = 8;     // time-step rows per lane in the weighted sum: H <= 32 * MAX_ROWS

// Warps per block (= per SM: the grid is persistent, one block per SM) a flavour is compiled for.  The bound is the
// register file: 64 K registers / (32 W) per thread.  Whole-tile flavours are limited by shared memory to
// 227 KB / (H * WT * sizeof(R)) resident tiles anyway (17 at H = 50, packed lanes), so they keep 16 warps x 128
// registers; the STREAM flavour's per-warp chunk ring is a few KB, so it runs 24 (packed lanes, 80 registers) or 32
// (one sample per thread, 64 registers) warps: the rollout is a dependent chain per sample and the XU / FMA pipes only
// fill up with 6-8 chains per scheduler.  WIDE asks for the 32-warp whole-tile flavour (one sample per thread, short
// horizons).
template <typename V, int NOISE, bool WIDE>
struct SolveShape {
    static constexpr bool is_f64 = sizeof(typename VT<V>::S) == 8;
    static constexpr int PACK = VT<V>::N;
    static constexpr int MAXW = is_f64 ? 16 : (NOISE == NOISE_STREAM ? (PACK == 2 ? 24 : 32) : (WIDE ? 32 : 16));
};

template <typename V>
struct WarpTile {
    typedef typename VT<V>::S R;
    static constexpr int PACK = VT<V>::N;
    static constexpr int WT = 32 * PACK;                 // samples per warp tile
    static constexpr int CE = 16 / sizeof(R);            // elements per 16-byte chunk
    static constexpr int NCHUNK = WT / CE;               // chunks per row (power of two)
};

// Shared-memory layout.  Offset 0: the warps' running records [W][2 + H] doubles (beta, eta, E[H]) — they live through
// the whole kernel (the weighted sum of every tile folds into them) — and the plan in float64 [H].  Then the tile loop's ring | U | weights |
// mbarriers | gen; the tail reuses everything behind the records after a block barrier.
template <typename V>
struct RingSmem {
    typedef typename VT<V>::S R;
    static constexpr int WT = WarpTile<V>::WT;
    static __host__ __device__ size_t a16(size_t b) { return (b + 15) / 16 * 16; }
    static __host__ __device__ size_t a128(size_t b) { return (b + 127) / 128 * 128; }
    // [H] doubles behind the records: the block's float64 copy of the plan (the finalising block applies the update
    // to it without another trip to global memory)
    static __host__ __device__ size_t off_Ud(int H, int W) { return (size_t)W * (PART_HDR + H) * sizeof(double); }
    static __host__ __device__ size_t off_ring(int H, int W) { return a128(off_Ud(H, W) + (size_t)H * sizeof(double)); }
    // a stage holds `rows` time steps of one warp tile: rows = H (whole tiles), or the STREAM chunk
    static __host__ __device__ size_t stage_bytes(int rows) { return (size_t)rows * WT * sizeof(R); }
    static __host__ __device__ size_t off_sU(int H, int rows, int S, int W) { return off_ring(H, W) + S * stage_bytes(rows); }
    // uc = copies of the plan in the rollout dtype: 1 (a block works on one solve: fewer solves than SMs) or W (one
    // per warp: a block may own several solves)
    static __host__ __device__ size_t off_sw(int H, int rows, int S, int W, int uc) { return off_sU(H, rows, S, W) + a16((size_t)uc * H * sizeof(R)); }
    static __host__ __device__ size_t off_bar(int H, int rows, int S, int W, int uc) { return off_sw(H, rows, S, W, uc) + (size_t)W * WT * sizeof(R); }
    static __host__ __device__ size_t off_gen(int H, int rows, int S, int W, int uc) { return off_bar(H, rows, S, W, uc) + (size_t)S * 8; }
    static __host__ __device__ size_t ring_total(int H, int rows, int S, int W, int uc) { return off_gen(H, rows, S, W, uc) + (size_t)S * 4 + 16; }
    // tail arena behind the records (doubles): scale[cap] | clipped plans [G][H] | goals [G][H*NZ] | group sums [W][H] | E
    static __host__ __device__ size_t tail_total(int H, int W, int cap, int G, int nz) {
        return off_ring(H, W) + ((size_t)cap + (size_t)G * H + (size_t)G * H * nz + (size_t)W * H + H + 64) * sizeof(double);
    }
};

// sum_k w[k] * row[k] over one dense row of WT elements, 16-byte chunks visited in an order rotated by the lane so
// that the eight lanes of a quarter warp (eight different rows, same bank offset) hit eight different bank groups
__device__ __forceinline__ double dot_row_rot(const float *row, const float *w, int lane, f2 *) {
    f2 a0(0.f), a1(0.f);
#pragma unroll
    for (int it = 0; it < 16; ++it) {
        const int c = ((it + lane) & 15) * 4;
        const float4 e = *reinterpret_cast<const float4 *>(row + c);
        const float4 q = *reinterpret_cast<const float4 *>(w + c);
        a0 = vfma(f2(q.x, q.y), f2(e.x, e.y), a0);
        a1 = vfma(f2(q.z, q.w), f2(e.z, e.w), a1);
    }
    const f2 a = a0 + a1;
    return (double)(a.v.x + a.v.y);
}
__device__ __forceinline__ double dot_row_rot(const float *row, const float *w, int lane, float *) {
    f2 a0(0.f), a1(0.f);
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int c = ((it + lane) & 7) * 4;
        const float4 e = *reinterpret_cast<const float4 *>(row + c);
        const float4 q = *reinterpret_cast<const float4 *>(w + c);
        a0 = vfma(f2(q.x, q.y), f2(e.x, e.y), a0);
        a1 = vfma(f2(q.z, q.w), f2(e.z, e.w), a1);
    }
    const f2 a = a0 + a1;
    return (double)(a.v.x + a.v.y);
}
__device__ __forceinline__ double dot_row_rot(const double *row, const double *w, int lane, double *) {
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int it = 0; it < 16; ++it) {
        const int c = ((it + lane) & 15) * 2;
        const double2 e = *reinterpret_cast<const double2 *>(row + c);
        const double2 q = *reinterpret_cast<const double2 *>(w + c);
        a0 = fma(q.x, e.x, a0);
        a1 = fma(q.y, e.y, a1);
    }
    return a0 + a1;
}

// ---- STREAM flavour of the weighted sum: rows come straight from the (L2-resident) noise tensor, one coalesced row
// per load, weights stay in registers.  v[r] = this lane's share of row r; the butterfly below is a reduce-scatter
// over the warp: 31 shuffles for 32 rows, and lane l ends up with the complete sum of row l in v[0].
__device__ __forceinline__ float lane_dot(const f2 &w, const f2 &e) { return fmaf(w.v.y, e.v.y, w.v.x * e.v.x); }
__device__ __forceinline__ float lane_dot(float w, float e) { return w * e; }
__device__ __forceinline__ double lane_dot(double w, double e) { return w * e; }
template <typename R>
__device__ __forceinline__ R warp_rows_reduce_scatter(R (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const R keep = up ? v[i + off] : v[i];
            const R send = up ? v[i] : v[i + off];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

// softmin scale of a record with minimum bj relative to the merged minimum beta (empty records carry +inf)
__device__ __forceinline__ double rec_scale(double neg_inv_lam, double bj, double beta) {
    return (bj == INFINITY) ? 0.0 : exp(neg_inv_lam * (bj - beta));
}

// ------------------------------------------------------------------------------------------------ peer exchange
// Multi-GPU (peer.cuh): every rank's solve kernel writes its (beta, eta, E[H]) record straight into the exchange
// buffer of EVERY rank over NVLink, raises that rank's flag for (writer, solve) to the solve's epoch and waits for the
// flags of all writers in its own buffer: an all-gather inside the kernel, no NCCL launch and no second kernel.
// The buffer is DOUBLE-BUFFERED by epoch parity: records [2][world][B][2+H] doubles, then u64 flags [2][world][B].
// A rank that has seen every flag of epoch e may already be publishing epoch e + 1 while a slower peer still reads its
// epoch-e slots; they are different slots.  It cannot publish e + 2 before every peer has published e + 1, and a peer
// publishes e + 1 only after its epoch-e kernel has finished: two parities are race-free with no acknowledgement.
__device__ __forceinline__ double *peer_recs(double *buf, int world, int B, int rec_len, unsigned long long epoch) {
    return buf + (size_t)(epoch & 1ull) * world * B * rec_len;
}
__device__ __forceinline__ unsigned long long *peer_flags(double *buf, int world, int B, int rec_len,
                                                          unsigned long long epoch) {
    return reinterpret_cast<unsigned long long *>(buf + (size_t)2 * world * B * rec_len) + (size_t)(epoch & 1ull) * world * B;
}
// phase stamps of the finalising block (nanoseconds of %globaltimer), read back with mpc_get_trace
enum { TR_START = 0, TR_TILES_DONE = 1, TR_TICKET = 2, TR_MERGED = 3, TR_PUBLISHED = 4, TR_PEERS_IN = 5,
       TR_MERGED2 = 6, TR_APPLIED = 7, TR_END = 8, TR_WORDS = 16 };
__device__ __forceinline__ void stamp(unsigned long long *trace, int i) {
    if (trace && threadIdx.x == 0) trace[i] = global_ns();
}

// thread q < world: raise flag (rank, b) in rank q's buffer; then wait until writer q has raised (q, b) in ours
__device__ __forceinline__ void peer_signal_and_wait(const PeerTable &pt, const MppiArgs &a, int b, int q, int rec_len,
                                                     double *err_word) {
    const size_t slot = (size_t)a.rank * a.g.B + b;
    peer_flag_exchange(peer_flags(pt.recs[q], a.world, a.g.B, rec_len, a.epoch) + slot,
                       peer_flags(pt.recs[a.rank], a.world, a.g.B, rec_len, a.epoch) + (size_t)q * a.g.B + b, a.epoch,
                       err_word);
}

// ---- float32 flavour of the per-tile softmin bookkeeping.  The reference minimum beta only has to be NEAR the minimum
// (softmin weights are invariant under a shift of beta: it cancels between w_k and eta), so the warp minimum is taken on
// the float32-rounded costs with five 32-bit shuffles instead of ten 64-bit ones, and the weights exp(-(S - beta)/lam)
// are one MUFU.EX2 of the float32-rounded exponent (the difference S - beta itself is formed in float64): relative
// weight error <= 2^-22 + 6e-8 |exponent|, far inside the float32 flavour's cost error divided by lam.
__device__ __forceinline__ float warp_min_f32(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum_f32(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// arrival ticket of a block: acq_rel at gpu scope.  With the block barrier before it, this publishes every thread's
// record stores (release, cumulative over the barrier) and, for the last arriver, makes the other blocks' records
// visible to the loads that follow the next barrier (acquire) — no separate __threadfence by all threads.
__device__ __forceinline__ int ticket_acq_rel(int *p) {
    int old;
    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(old) : "l"(p) : "memory");
    return old;
}

// ------------------------------------------------------------------------------------------------ record merging
// Block-wide merge of `count` records (beta_j, eta_j, E_j[H]) at src + j * stride: beta = min beta_j,
// eta = sum s_j eta_j, Eout[t] = sum s_j E_j[t] with s_j = exp(-(beta_j - beta)/lam).  The records were written by
// other blocks (or ranks): they are read through L2 (__ldcg).  The merge is latency bound (a few KB per thread
// block, on the critical path of every solve), so ALL loads of the common case are issued before the first
// reduction: the headers (beta_j, eta_j) by thread j, and the first MERGE_BATCH records of every column by the
// thread that owns (column t, record group g) — the scales are applied to registers once they are known.  Threads
// are laid out as G groups x TT columns (TT = power of two >= H); group g owns records g, g + G, ...; the group sums
// meet in shared memory in a fixed order, so repeated solves are bit-identical.
// Scratch: sscale[count], sEw[>= G][H] (callers size it [warps][H]; G <= warps), red[32].  All threads return beta / eta.
constexpr int MERGE_BATCH = 20;
constexpr int MERGE_SMALL = MERGE_GROUP;  // up to this many records are merged warp-locally (no block reductions)
__device__ __forceinline__ void block_merge(const double neg_inv_lam, const int H, const double *src, const int count,
                                            const size_t stride, double *sscale, double *sEw, double *red,
                                            double *Eout, double *beta_out, double *eta_out) {
    const int tid = threadIdx.x, nt = blockDim.x;
    if (count <= MERGE_SMALL) {
        // Few records (the rank records of a multi-GPU solve, the block records of a small solve): no block
        // reductions at all.  Every warp reads the headers (lane j = record j) and derives beta, the scales and eta
        // with shuffles — redundantly, identically; thread t owns column t and has its `count` loads in flight
        // before the scales exist.  One L2 round trip, one barrier.
        const int lane = tid & 31;
        const double bj = (lane < count) ? __ldcg(src + (size_t)lane * stride) : INFINITY;
        const double ej = (lane < count) ? __ldcg(src + (size_t)lane * stride + 1) : 0.0;
        double e[MERGE_SMALL];
        const bool own = tid < H;
        if (own) {
#pragma unroll
            for (int j = 0; j < MERGE_SMALL; ++j)
                e[j] = (j < count) ? __ldcg(src + (size_t)j * stride + PART_HDR + tid) : 0.0;
        }
        const double beta = warp_min(bj);
        const double sc = rec_scale(neg_inv_lam, bj, beta);  // 0 for lanes >= count (bj = inf)
        const double eta = warp_sum(sc * ej);
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < MERGE_SMALL; ++j) {
            const double sj = __shfl_sync(0xffffffffu, sc, j);
            if (own) acc += sj * e[j];
        }
        if (own) Eout[tid] = acc;
        for (int t0 = nt; t0 < H; t0 += nt) {  // blocks narrower than H: the remaining columns, loads not hoisted
            const int t = t0 + tid;            // (the shuffles stay warp-uniform: the bound is t0, not t)
            double a2 = 0.0;
            for (int j = 0; j < count; ++j) {
                const double sj = __shfl_sync(0xffffffffu, sc, j);
                if (t < H) a2 += sj * __ldcg(src + (size_t)j * stride + PART_HDR + t);
            }
            if (t < H) Eout[t] = a2;
        }
        __syncthreads();
        *beta_out = beta;
        *eta_out = eta;
        return;
    }
    int TT = 32;
    while (TT < H && TT < nt) TT <<= 1;
    const bool by_column = TT >= H;  // every column has its own thread in each group (else: blocks narrower than H)
    const int G = nt / TT, t = tid & (TT - 1), g = tid / TT;
    const bool single = count <= nt;  // one header per thread
    // ---- every load of the common case, issued back to back
    double b0 = INFINITY, e0 = 0.0, m = INFINITY;
    if (single) {
        if (tid < count) {
            b0 = __ldcg(src + (size_t)tid * stride);
            e0 = __ldcg(src + (size_t)tid * stride + 1);
        }
        m = b0;
    } else {
        for (int j = tid; j < count; j += nt) m = fmin(m, __ldcg(src + (size_t)j * stride));
    }
    double ereg[MERGE_BATCH];
    const double *col = src + PART_HDR + t;
    if (by_column && t < H) {
#pragma unroll
        for (int i = 0; i < MERGE_BATCH; ++i) {
            const int j = g + i * G;
            ereg[i] = (j < count) ? __ldcg(col + (size_t)j * stride) : 0.0;
        }
    }
    // ---- beta, scales, eta
    const double beta = block_reduce<true>(m, red);
    double es = 0.0;
    if (single) {
        if (tid < count) {
            const double sc = rec_scale(neg_inv_lam, b0, beta);
            sscale[tid] = sc;
            es = sc * e0;
        }
    } else {
        for (int j = tid; j < count; j += nt) {
            const double sc = rec_scale(neg_inv_lam, __ldcg(src + (size_t)j * stride), beta);
            sscale[j] = sc;
            es += sc * __ldcg(src + (size_t)j * stride + 1);
        }
    }
    const double eta = block_reduce<false>(es, red);  // its barriers publish sscale
    // ---- E
    if (by_column) {
        if (t < H) {
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < MERGE_BATCH; ++i) {
                const int j = g + i * G;
                if (j < count) acc += sscale[j] * ereg[i];
            }
            int j = g + MERGE_BATCH * G;
            for (; j + 7 * G < count; j += 8 * G) {  // records beyond the preloaded batch: eight loads in flight
                double v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = __ldcg(col + (size_t)(j + i * G) * stride);
#pragma unroll
                for (int i = 0; i < 8; ++i) acc += sscale[j + i * G] * v[i];
            }
            for (; j < count; j += G) acc += sscale[j] * __ldcg(col + (size_t)j * stride);
            sEw[g * H + t] = acc;
        }
        __syncthreads();
        for (int tt = tid; tt < H; tt += nt) {
            double E = 0.0;
            for (int i = 0; i < G; ++i) E += sEw[i * H + tt];
            Eout[tt] = E;
        }
    } else {
        // narrow blocks (fewer threads than H rounded up): every thread walks its columns, all records
        for (int tt = tid; tt < H; tt += nt) {
            double E = 0.0;
            for (int j = 0; j < count; ++j) E += sscale[j] * __ldcg(src + (size_t)j * stride + PART_HDR + tt);
            Eout[tt] = E;
        }
    }
    __syncthreads();
    *beta_out = beta;
    *eta_out = eta;
}

// What follows the merge of one solve, by the whole block.  record_only (multi-GPU) and !final: publish this rank's record
// (the exchange and the final merge follow); otherwise planned_us += erg (mppi.py:62,69), clip (agent.py:64-65),
// outputs, and the nominal rollout of the clipped plan by thread 0 from shared memory (agent.py:97-108).
// Scratch: sUc[H], sG[H * NZ] (only touched for a non-zero goal), zero[NZ].
template <class Env>
__device__ __forceinline__ void block_apply(const MppiArgs &a, const EnvConst<typename Env::Vec> &ec, const int b,
                                            const bool final_stage, const double beta, const double eta,
                                            const double *Eout, double *rank_part, const double *x0g,
                                            const double *g_z, double *Ug, double *out, double *sUc, double *sG,
                                            const double *Usrc = nullptr) {  // the plan before the update, if the
                                                                              // caller holds a copy of Ug[b] (shared memory)
    constexpr int NS = Env::NS, NZ = Env::NS + 1;
    const int H = a.g.H, tid = threadIdx.x, nt = blockDim.x, rec_len = PART_HDR + H;
    if (a.record_only && !final_stage) {
        double *rec = rank_part + ((size_t)a.rank * a.g.B + b) * rec_len;
        for (int t = tid; t < H; t += nt) rec[PART_HDR + t] = Eout[t];
        if (tid == 0) {
            rec[0] = beta;
            rec[1] = eta;
        }
        return;
    }
    double *U = Ug + (size_t)b * H;
    const double *Uold = Usrc ? Usrc : U;
    double *o = out + (size_t)b * (size_t)(2 + H + (H + 1) * NS);
    const double inv_eta = 1.0 / eta;
    for (int t = tid; t < H; t += nt) {
        const double un = Uold[t] + inv_eta * Eout[t];  // w = (1/eta)*tmp ; planned_us += erg   mppi.py:62,69
        U[t] = un;
        o[1 + t] = un;
        if (t == 0) o[0] = un;                       // u0 = planned_us[0]  (mppi.py:71)
        sUc[t] = fmin(fmax(un, a.lo), a.hi);         // agent.py:64-65
    }
    if (a.want_traj && !a.goal_zero)
        for (int i = tid; i < H * NZ; i += nt) sG[i] = g_z[(size_t)b * H * NZ + i];
    if (a.want_traj && a.goal_zero && tid < NZ) sG[tid] = 0.0;
    __syncthreads();
    stamp(a.trace, TR_APPLIED);
    if (tid == 0 && a.want_traj)
        o[1 + H] = nominal_rollout<typename Env::Scalar>(scalar_consts(ec), H, x0g + (size_t)b * NS, sUc, sG,
                                                         a.goal_zero ? 0 : NZ, o + 2 + H, a.model_noise, a.seed,
                                                         a.noise_draw, (uint32_t)b);
}

// ================================================================================================ the solve kernel
// Geometry (capi.cu): tiles are WARP tiles.  mode A (solves_per_block == 0): blocks_per_solve blocks share one solve,
// block blk works on solve blk / blocks_per_solve.  mode B: a block owns solves_per_block whole solves.
template <class Env, int NOISE, bool NOISY = false, bool WIDE = false>  // NOISY: the model has sensor noise (one sample per thread only)
__global__ void __launch_bounds__(SolveShape<typename Env::Vec, NOISE, WIDE>::MAXW * 32, 1)
    mppi_solve_kernel(const MppiArgs a, const EnvConst<typename Env::Vec> ec,
                      const __grid_constant__ CUtensorMap noise_map, const typename Env::S *__restrict__ noise_g,
                      const double *__restrict__ x0g, const double *__restrict__ g_z, double *__restrict__ Ug,
                      double *__restrict__ S_out, double *__restrict__ part, double *__restrict__ rank_part,
                      double *__restrict__ out, int *__restrict__ done_count,
                      const __grid_constant__ PeerTable pt,     // indexed by rank at run time: stays in parameter space
                      const __grid_constant__ InlineInputs inl) {  // batch-1 host call: state / plan as parameters
    typedef typename Env::Vec V;
    typedef typename Env::S R;
    constexpr int NS = Env::NS, NZ = Env::NS + 1, NG = Env::NG, PACK = VT<V>::N, CH = Env::CHUNK;
    constexpr int WT = WarpTile<V>::WT;
    constexpr bool F32 = sizeof(R) == 4;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr bool STREAM = NOISE == NOISE_STREAM;
    // the descriptor's address in parameter space (the lambdas below capture the pointer, not the 128-byte object)
    const CUtensorMap *const map_ptr = &noise_map;
    const int H = a.g.H, W = blockDim.x >> 5, S = a.g.stages, tps = a.g.tiles_per_solve;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const bool mode_b = a.g.solves_per_block != 0;
    // STREAM: the ring is W private rings of `depth` chunk stages of `rows` time steps (S = W * depth)
    const int rows = STREAM ? a.g.stream_rows : H, depth = STREAM ? a.g.stream_depth : 0;
    const int nchunk = STREAM ? (H + rows - 1) / rows : 1;
    const size_t stage_elems = (size_t)rows * WT;
    const int rec_len = PART_HDR + H;
    double *wrec_all = reinterpret_cast<double *>(smem_raw);  // [W][rec_len]: the warps' running (beta, eta, E[H])
    double *wrec = wrec_all + (size_t)w * rec_len;
    double *sUd = reinterpret_cast<double *>(smem_raw + RingSmem<V>::off_Ud(H, W));  // mode A: the plan in float64
    R *stages = reinterpret_cast<R *>(smem_raw + RingSmem<V>::off_ring(H, W));
    // the plan in the rollout dtype: one copy per block (mode A: one solve per block) or per warp (mode B)
    R *sU = reinterpret_cast<R *>(smem_raw + RingSmem<V>::off_sU(H, rows, S, W)) + (mode_b ? (size_t)w * H : 0);
    const int uc = a.g.u_copies;
    R *sw = reinterpret_cast<R *>(smem_raw + RingSmem<V>::off_sw(H, rows, S, W, uc)) + (size_t)w * WT;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + RingSmem<V>::off_bar(H, rows, S, W, uc));
    // gen[st] = tile whose load the stage's mbarrier currently tracks.  An mbarrier carries one bit of phase: a
    // warp that ran a whole ring turn ahead of a slow warp would see the parity of an OLDER completed phase and
    // read a stale tile.  Waiting for gen[st] == j first makes the wait exact for any interleaving of the warps.
    volatile int *gen = reinterpret_cast<volatile int *>(smem_raw + RingSmem<V>::off_gen(H, rows, S, W, uc));

    if (blockIdx.x == 0) stamp(a.trace, TR_START);
    // inline inputs: keep the device copy of the state current (device-resident solves enqueued later read it; every
    // block of THIS launch reads the parameters, so there is no race)
    if ((inl.on & 1) && blockIdx.x == 0 && threadIdx.x < NS) const_cast<double *>(x0g)[threadIdx.x] = inl.x0[threadIdx.x];
    // ---- this block's tile range ----------------------------------------------------------------------------
    int tile0, tile1, b_lo, b_hi;
    if (!mode_b) {
        b_lo = b_hi = blockIdx.x / a.g.blocks_per_solve;
        const int part_idx = blockIdx.x - b_lo * a.g.blocks_per_solve;
        tile0 = b_lo * tps + part_idx * a.g.chunk;
        tile1 = min(tile0 + a.g.chunk, (b_lo + 1) * tps);
    } else {
        b_lo = blockIdx.x * a.g.solves_per_block;
        b_hi = min(b_lo + a.g.solves_per_block, a.g.B) - 1;
        tile0 = b_lo * tps;
        tile1 = (b_hi + 1) * tps;
    }

    // one TMA tensor copy (box [H][WT] of the [Bn][H][Kpad] noise tensor) of tile j = (solve b, warp tile kt) into
    // stage st, by lane 0; callers have passed a __syncwarp after the warp's last read of the stage
    auto issue = [&](int j, int st, int b, int kt) {
        if (lane == 0) {
            mbar_expect_tx(&bars[st], (uint32_t)(stage_elems * sizeof(R)));
            __threadfence_block();
            gen[st] = j;
            fence_proxy_async();  // generic-proxy reads of the stage precede the async-proxy writes
            tma_load_3d(stages + st * stage_elems, map_ptr, kt * WT, 0, a.noise_batched ? b : 0, &bars[st]);
        }
    };

    // STREAM: chunk q of this warp's sequence (tile tile0 + w + (q / nchunk) W, rows (q % nchunk) * rows ...) goes to
    // stage q % depth of the warp's private ring; the same warp issues and consumes in order, so the phase parity of
    // a stage is (q / depth) & 1 and no cross-warp bookkeeping is needed
    R *wring = stages + (size_t)w * depth * stage_elems;
    uint64_t *wbars = bars + (size_t)w * depth;
    int iss_st = 0, use_st = 0;           // ring stage of the next issue / the next use
    uint32_t use_phase = 0;               // phase parity of the next use: flips each time use_st wraps
    int iss_tile = tile0 + w, iss_c = 0;  // tile and chunk the next issue loads (kept incrementally: no divisions)
    int iss_kt = iss_tile % tps;
    auto issue_chunk = [&]() {
        if (iss_tile < tile1) {
            if (lane == 0) {
                mbar_expect_tx(&wbars[iss_st], (uint32_t)(stage_elems * sizeof(R)));  // rows past H: zero-filled, counted
                fence_proxy_async();
                tma_load_3d(wring + (size_t)iss_st * stage_elems, map_ptr, iss_kt * WT, iss_c * rows, 0, &wbars[iss_st]);
            }
            if (++iss_c == nchunk) {
                iss_c = 0;
                iss_tile += W;
                iss_kt += W;
                while (iss_kt >= tps) iss_kt -= tps;
            }
        }
        if (++iss_st == depth) iss_st = 0;
    };

    if (NOISE == NOISE_TENSOR || STREAM) {
        if (tid == 0) {
            for (int i = 0; i < S; ++i) {
                mbar_init(&bars[i], 1);
                gen[i] = -1;
            }
            fence_mbar_init();
        }
    }
    // mode A: the block's single solve's plan, once per block (every warp reads the same copy)
    if (!mode_b)
        for (int t = tid; t < H; t += blockDim.x) {
            const double u = (inl.on & 2) ? inl.U[t] : Ug[(size_t)b_lo * H + t];
            sUd[t] = u;
            sU[t] = (R)u;
        }
    for (int t = lane; t < rec_len; t += 32) wrec[t] = 0.0;  // E starts at zero (the first tile's scale_old is 0, 0 * NaN is not)
    __syncthreads();
    if (NOISE == NOISE_TENSOR || STREAM) {
        if (STREAM) {
            for (int i = 0; i < depth; ++i) issue_chunk();
        } else {
            for (int i = w; i < S && tile0 + i < tile1; i += W) {
                const int j = tile0 + i, b = j / tps;
                issue(j, i, b, j - b * tps);
            }
        }
    }

    // ---- per-warp running softmin state of the current solve; E rows lane, lane + 32, ... live in wrec (shared) ----
    int cur = b_lo;
    double beta_run = INFINITY, eta_run = 0.0;
    R gs[NG];
    R xinit[NS];
    const V pert = V((R)a.pert);

    auto load_solve = [&](int b) {
        if (mode_b) {
            __syncwarp();
            for (int t = lane; t < H; t += 32) sU[t] = (R)Ug[(size_t)b * H + t];
        }
        Env::prep_goal(ec, g_z + (size_t)b * H * NZ, gs);  // goal row 0 (model.py:34,38)
#pragma unroll
        for (int i = 0; i < NS; ++i) xinit[i] = (R)((inl.on & 1) ? inl.x0[i] : x0g[(size_t)b * NS + i]);
        Env::enter(ec, xinit);
        __syncwarp();
    };
    // mode B only: the warp's record of solve `cur` goes to global memory (slot = warp)
    auto flush_global = [&]() {
        double *rec = part + ((size_t)cur * a.g.part_stride + w) * rec_len;
        __syncwarp();
        if (lane == 0) {
            rec[0] = beta_run;
            rec[1] = eta_run;
        }
        for (int t = lane; t < H; t += 32) {
            rec[PART_HDR + t] = wrec[PART_HDR + t];
            wrec[PART_HDR + t] = 0.0;
        }
        beta_run = INFINITY;
        eta_run = 0.0;
        __syncwarp();
    };
    load_solve(cur);

    // tile j = tile0 + w + i W of the block goes to ring stage st = (j - tile0) % S on turn q = (j - tile0) / S of the
    // ring; (solve b, warp tile kt) = (j / tps, j % tps).  All four are carried incrementally: no division per tile.
    int st = w, q = 0;  // w < W <= S
    int b = (tile0 + w) / tps, kt = (tile0 + w) - b * tps;
    for (int j = tile0 + w; j < tile1; j += W) {
        while (cur < b) {  // mode B: every warp leaves a record (possibly empty) for every solve of the block
            flush_global();
            ++cur;
            if (cur == b) load_solve(cur);
        }
        R *tile = (NOISE == NOISE_TENSOR) ? stages + st * stage_elems : stages + w * stage_elems;  // STREAM: unused
        const int kl = kt * WT + lane * PACK;  // first local sample index of this lane
        V x[NS];
#pragma unroll
        for (int i = 0; i < NS; ++i) x[i] = V(xinit[i]);
        double Sk[PACK];
#pragma unroll
        for (int l = 0; l < PACK; ++l) Sk[l] = 0.0;
        const R *col = tile + lane * PACK;
        // one rollout step: u = U_t + eps (mppi.py:39), cost + step (mppi.py:43-44), lam * u * eps / sigma^2 (:45-46)
        auto one = [&](int t, const V &e, V &acc) {
            const V u = V(sU[t]) + e;
            if constexpr (Env::GENERIC != 0)  // opt-in extension: goal row t at step t (the reference's rollouts: row 0)
                if (a.per_step_goal) Env::prep_goal(ec, g_z + ((size_t)b * H + t) * NZ, gs);
            const V c = Env::step(ec, x, u, gs);
            if constexpr (NOISY) {  // newstate += N(0, sensor_std)   model.py:42-43
                static_assert(PACK == 1, "the noisy-model flavour rolls out one sample per thread");
                add_model_noise<Env>(x, a.model_noise, a.seed, PHILOX_MODEL_MPPI, a.noise_draw,
                                     (uint32_t)(a.g.k_begin + kl), (uint32_t)t);
            }
            acc = acc + vfma(u * e, pert, c);
        };
        auto fold = [&](const V &acc) {
#pragma unroll
            for (int l = 0; l < PACK; ++l) Sk[l] += (double)lane_of(acc, l);
        };
        if (NOISE == NOISE_TENSOR) {
            while (gen[st] != j) {
            }
            __threadfence_block();
            mbar_wait(&bars[st], (uint32_t)(q & 1));
            int t = 0;
            for (; t + CH <= H; t += CH) {
                V acc = V((R)0);
#pragma unroll
                for (int i = 0; i < CH; ++i) one(t + i, load_lanes(col + (size_t)(t + i) * WT, (V *)nullptr), acc);
                fold(acc);
            }
            for (; t < H; ++t) {
                V acc = V((R)0);
                one(t, load_lanes(col + (size_t)t * WT, (V *)nullptr), acc);
                fold(acc);
            }
        } else if (STREAM) {
            // chunk by chunk through the warp's private ring; the grouping of the float32 cost sums is the tensor
            // path's (rows is a multiple of CH), so both flavours give bit-identical costs
            for (int c = 0; c < nchunk; ++c) {
                mbar_wait(&wbars[use_st], use_phase);
                const R *ccol = wring + (size_t)use_st * stage_elems + lane * PACK;
                const int t0 = c * rows, n = min(rows, H - t0);
                int r = 0;
                for (; r + CH <= n; r += CH) {
                    V acc = V((R)0);
#pragma unroll
                    for (int i = 0; i < CH; ++i) one(t0 + r + i, load_lanes(ccol + (size_t)(r + i) * WT, (V *)nullptr), acc);
                    fold(acc);
                }
                for (; r < n; ++r) {
                    V acc = V((R)0);
                    one(t0 + r, load_lanes(ccol + (size_t)r * WT, (V *)nullptr), acc);
                    fold(acc);
                }
                if (++use_st == depth) {
                    use_st = 0;
                    use_phase ^= 1u;
                }
                __syncwarp();    // every lane has read the stage
                issue_chunk();   // re-arm it with the chunk `depth` places ahead (possibly of the warp's next tile)
            }
        } else {
            // same chunking of the float32 cost sums as the tensor path (CH divides 4), so both noise modes give
            // bit-identical costs for identical noise values
            static_assert(4 % CH == 0, "CHUNK must divide the Philox group of 4 steps");
            const R sig = (R)a.sigma;
            __syncwarp();  // the previous tile's weighted sum has finished reading this warp's buffer
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
                        tile[(size_t)t * WT + lane * PACK + l] = el[l];
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
#pragma unroll
        for (int l = 0; l < PACK; ++l)
            if (!(kl + l < a.g.K_local)) Sk[l] = INFINITY;
        if (!a.store_costs) {
        } else if (PACK == 2) {
            if (kl + 1 < a.g.K_local)
                *reinterpret_cast<double2 *>(S_out + (size_t)b * a.g.Kpad + kl) = make_double2(Sk[0], Sk[PACK - 1]);
            else if (kl < a.g.K_local)
                S_out[(size_t)b * a.g.Kpad + kl] = Sk[0];
        } else if (kl < a.g.K_local) {
            S_out[(size_t)b * a.g.Kpad + kl] = Sk[0];
        }
        // ---- online softmin: rescale the running sums to the new minimum (mppi.py:59-62) ------------------------
        double beta_new, scale_old;
        R wl[PACK];
        if constexpr (F32) {
            float m = INFINITY;
#pragma unroll
            for (int l = 0; l < PACK; ++l) m = fminf(m, (float)Sk[l]);
            beta_new = fmin(beta_run, (double)warp_min_f32(m));
            // warp-uniform: the running minimum only moves O(log tiles) times per warp
            scale_old = (eta_run == 0.0) ? 0.0 : (beta_new == beta_run ? 1.0 : exp(a.neg_inv_lam * (beta_run - beta_new)));
            float wsum = 0.f;
#pragma unroll
            for (int l = 0; l < PACK; ++l) {
                const float ex = (float)(a.neg_inv_lam * (Sk[l] - beta_new)) * 1.44269504f;  // -inf for padding samples
                wl[l] = (Sk[l] == INFINITY) ? 0.f : ex2_approx(ex);
                wsum += wl[l];
            }
            eta_run = eta_run * scale_old + (double)warp_sum_f32(wsum);
        } else {
            double Smin = INFINITY;
#pragma unroll
            for (int l = 0; l < PACK; ++l) Smin = fmin(Smin, Sk[l]);
            beta_new = fmin(beta_run, warp_min(Smin));
            scale_old = (eta_run == 0.0) ? 0.0 : exp(a.neg_inv_lam * (beta_run - beta_new));
            double wsum = 0.0;
#pragma unroll
            for (int l = 0; l < PACK; ++l) {
                const double wk = (Sk[l] == INFINITY) ? 0.0 : exp(a.neg_inv_lam * (Sk[l] - beta_new));
                wl[l] = (R)wk;
                wsum += wk;
            }
            eta_run = eta_run * scale_old + warp_sum(wsum);
        }
        beta_run = beta_new;
        if constexpr (STREAM) {
            // ---- E[t] += sum_k w_k eps[k][t] with the rows re-read from the noise tensor (L2; the tile went through
            // this SM a tile time ago), weights in registers, 32 rows per reduce-scatter --------------------------
            V wv;
            if constexpr (PACK == 2)
                wv = V(wl[0], wl[PACK - 1]);
            else
                wv = V(wl[0]);
            const R *grow = noise_g + (size_t)kt * WT + lane * PACK;
#pragma unroll
            for (int i = 0; i < MAX_ROWS; ++i) {
                if (32 * i < H) {
                    R v[32];
#pragma unroll
                    for (int r = 0; r < 32; ++r) {
                        const int t = 32 * i + r;
                        v[r] = (t < H) ? lane_dot(wv, load_lanes(grow + (size_t)t * a.g.Kpad, (V *)nullptr)) : (R)0;
                    }
                    const double acc = (double)warp_rows_reduce_scatter(v, lane);
                    const int r = lane + 32 * i;
                    if (r < H) wrec[PART_HDR + r] = wrec[PART_HDR + r] * scale_old + acc;
                }
            }
        } else {
            if constexpr (PACK == 2)
                *reinterpret_cast<float2 *>(sw + lane * 2) = make_float2(wl[0], wl[PACK - 1]);
            else
                sw[lane] = wl[0];
            __syncwarp();  // weights (and the Philox tile) visible to the whole warp
            // ---- E[t] += sum_k w_k eps[k][t], lane owns rows t = lane + 32 i -------------------------------------
#pragma unroll
            for (int i = 0; i < MAX_ROWS; ++i) {
                if (32 * i < H) {
                    const int r = lane + 32 * i;
                    if (r < H) {
                        const double acc = dot_row_rot(tile + (size_t)r * WT, sw, lane, (V *)nullptr);
                        wrec[PART_HDR + r] = wrec[PART_HDR + r] * scale_old + acc;
                    }
                }
            }
        }
        // ---- re-arm the stage with the tile S places ahead, then step (st, q, b, kt) to this warp's next tile -------
        if (NOISE == NOISE_TENSOR) {
            __syncwarp();
            if (j + S < tile1) {
                int b2 = b, kt2 = kt + S;
                while (kt2 >= tps) {
                    kt2 -= tps;
                    ++b2;
                }
                issue(j + S, st, b2, kt2);
            }
        }
        st += W;
        if (st >= S) {
            st -= S;
            ++q;
        }
        kt += W;
        while (kt >= tps) {
            kt -= tps;
            ++b;
        }
    }
    if (mode_b) {
        while (cur <= b_hi) {
            flush_global();
            ++cur;
        }
    }

    // ---- tail: every stage is free after this barrier; the arena behind the records is reused ----------------------
    __shared__ unsigned long long s_wexit[MAX_WARPS];  // profiling: when each warp left the tile loop
    if (a.trace && lane == 0) s_wexit[w] = global_ns();
    if (!mode_b && lane == 0) {
        wrec[0] = beta_run;
        wrec[1] = eta_run;
    }
    __syncthreads();
    unsigned long long t_wmin = ~0ull, t_wmax = 0ull;
    if (a.trace && tid == 0)
        for (int i = 0; i < W; ++i) {
            t_wmin = s_wexit[i] < t_wmin ? s_wexit[i] : t_wmin;
            t_wmax = s_wexit[i] > t_wmax ? s_wexit[i] : t_wmax;
        }
    __shared__ double red[32];
    __shared__ int s_last;
    double *arena = reinterpret_cast<double *>(smem_raw + RingSmem<V>::off_ring(H, W));
    const int nt = blockDim.x;
    if (!mode_b) {
        // mode A: warp records (shared) -> one block record in global memory -> ticket -> last block merges
        const int b = b_lo, nb = a.g.blocks_per_solve, part_idx = blockIdx.x - b * nb;
        double *sscale = arena;                    // [max(W, nb)]
        double *sUc = sscale + max(W, nb);         // [H]
        double *sG = sUc + H;                      // [H * NZ]
        double *sEw = sG + (size_t)H * NZ;         // [W][H]
        double *Eout = sEw + (size_t)W * H;        // [H]
        const unsigned long long t_a = a.trace ? global_ns() : 0ull;
        double bb = INFINITY;
        for (int i = 0; i < W; ++i) bb = fmin(bb, wrec_all[(size_t)i * rec_len]);
        if (tid < W) sscale[tid] = rec_scale(a.neg_inv_lam, wrec_all[(size_t)tid * rec_len], bb);
        __syncthreads();
        // records of solve b: [max_parts] block records, then [n_groups] group records
        double *const solve_part = part + (size_t)b * a.g.part_stride * rec_len;
        int *const tickets = done_count + (size_t)b * (1 + a.g.n_groups);  // [0] the solve's, [1 + g] group g's
        double *grec = solve_part + (size_t)part_idx * rec_len;
        for (int t = tid; t < H; t += nt) {
            double e = 0.0;
            for (int i = 0; i < W; ++i) e += sscale[i] * wrec_all[(size_t)i * rec_len + PART_HDR + t];
            grec[PART_HDR + t] = e;
        }
        if (tid == 0) {
            double eta = 0.0;
            for (int i = 0; i < W; ++i) eta += sscale[i] * wrec_all[(size_t)i * rec_len + 1];
            grec[0] = bb;
            grec[1] = eta;
        }
        const unsigned long long t_b = a.trace ? global_ns() : 0ull;
        // arrive: the block barrier orders every thread's record stores before thread 0's acq_rel ticket; the block
        // whose ticket completes the count sees the other blocks' records after the second barrier.  Self-cleaning.
        auto arrive_last = [&](int *counter, int expected) {
            __syncthreads();
            if (tid == 0) {
                const int ticket = ticket_acq_rel(counter);
                s_last = (ticket == expected - 1);
                if (s_last) *counter = 0;
            }
            __syncthreads();
            return s_last != 0;
        };
        double beta, eta;
        const double *final_src = solve_part;
        int final_count = nb;
        if (a.g.n_groups > 0) {
            // Two-level merge: a merge of many records is a chain of dependent L2 round trips in ONE block on the
            // critical path of every solve (5-8 us for 148 records).  Groups of MERGE_SMALL consecutive blocks: the last
            // block of a group to arrive merges the group's records warp-locally (every load in flight at once, one
            // round trip) into a group record; the last GROUP to arrive merges the <= MERGE_SMALL group records the same
            // way.  Membership and summation order are fixed, so repeated solves stay bit-identical.
            const int g = part_idx / MERGE_SMALL, gcount = min(MERGE_SMALL, nb - g * MERGE_SMALL);
            if (!arrive_last(&tickets[1 + g], gcount)) return;
            block_merge(a.neg_inv_lam, H, solve_part + (size_t)g * MERGE_SMALL * rec_len, gcount, (size_t)rec_len, sscale,
                        sEw, red, Eout, &beta, &eta);
            double *rec2 = solve_part + (size_t)(a.g.max_parts + g) * rec_len;
            for (int t = tid; t < H; t += nt) rec2[PART_HDR + t] = Eout[t];
            if (tid == 0) {
                rec2[0] = beta;
                rec2[1] = eta;
            }
            final_src = solve_part + (size_t)a.g.max_parts * rec_len;
            final_count = a.g.n_groups;
        }
        if (!arrive_last(&tickets[0], a.g.n_groups > 0 ? a.g.n_groups : nb)) return;
        if (a.trace && tid == 0) {
            a.trace[TR_TILES_DONE] = t_wmax;  // the last warp's exit from the tile loop (of the finalising block)
            a.trace[9] = t_a;
            a.trace[10] = t_b;
            a.trace[12] = t_wmin;
            a.trace[13] = t_wmax;
        }
        stamp(a.trace, TR_TICKET);
        block_merge(a.neg_inv_lam, H, final_src, final_count, (size_t)rec_len, sscale, sEw, red, Eout, &beta, &eta);
        stamp(a.trace, TR_MERGED);
        double *const err_word = out + (size_t)a.g.B * (2 + H + (H + 1) * NS);
        const double *const x0p = (inl.on & 1) ? inl.x0 : x0g;  // inline inputs come with batch 1: b == 0
        if (a.p2p) {
            // in-kernel all-gather of the rank records over NVLink, then the same merge over `world` records
            for (int qr = 0; qr < a.world; ++qr) {
                double *dst = peer_recs(pt.recs[qr], a.world, a.g.B, rec_len, a.epoch) + ((size_t)a.rank * a.g.B + b) * rec_len;
                for (int t = tid; t < H; t += nt) dst[PART_HDR + t] = Eout[t];
                if (tid == 0) {
                    dst[0] = beta;
                    dst[1] = eta;
                }
            }
            // no system fence by every thread here: the block barrier orders all the record stores before the flag
            // stores below, and those are st.release.sys (cumulative) — one fence per flag writer, not two in a row
            __syncthreads();
            stamp(a.trace, TR_PUBLISHED);
            if (tid < a.world) peer_signal_and_wait(pt, a, b, tid, rec_len, err_word);
            __syncthreads();
            stamp(a.trace, TR_PEERS_IN);
            block_merge(a.neg_inv_lam, H, peer_recs(pt.recs[a.rank], a.world, a.g.B, rec_len, a.epoch) + (size_t)b * rec_len,
                        a.world, (size_t)a.g.B * rec_len, sscale, sEw, red, Eout, &beta, &eta);
            stamp(a.trace, TR_MERGED2);
            block_apply<Env>(a, ec, b, true, beta, eta, Eout, rank_part, x0p, g_z, Ug, out, sUc, sG, sUd);
        } else {
            block_apply<Env>(a, ec, b, false, beta, eta, Eout, rank_part, x0p, g_z, Ug, out, sUc, sG, sUd);
        }
        if (a.host_out && !(a.record_only)) {
            // outputs straight into the caller-visible pinned buffer, then the completion flag (system-scope release)
            __syncthreads();
            const size_t stride = (size_t)(2 + H + (H + 1) * NS);
            const double *o = out + (size_t)b * stride;
            const int n_copy = a.want_traj ? (int)stride : 1 + H;
            for (int i = tid; i < n_copy; i += nt) a.host_out[i] = o[i];
            if (tid < 2) a.host_out[stride + tid] = err_word[tid];  // peer-wait error words
            __threadfence_system();
            __syncthreads();
            if (tid == 0) st_release_sys(a.host_flag, a.host_epoch);
        }
        stamp(a.trace, TR_END);
    } else {
        // mode B: this block owns whole solves; its warps' records are in global memory (slot = warp).  One warp
        // merges one solve; then one THREAD per solve rolls out the nominal trajectory, G solves at a time.
        const int nsolve = b_hi - b_lo + 1, G = a.g.fin_group;
        double *sUc = arena;                       // [G][H]
        double *sG = sUc + (size_t)G * H;          // [G][H * NZ], or one zero row
        double *const err_word = out + (size_t)a.g.B * (2 + H + (H + 1) * NS);
        for (int g0 = 0; g0 < nsolve; g0 += G) {
            const int ng = min(G, nsolve - g0);
            for (int sl = w; sl < ng; sl += W) {
                const int b = b_lo + g0 + sl;
                // one warp, one solve: merge `cnt` (<= 32) records, lane j holds record j's scale
                const double *recs = part + (size_t)b * a.g.part_stride * rec_len;
                size_t rstride = rec_len;
                int cnt = W;
                double bj = (lane < cnt) ? recs[(size_t)lane * rstride] : INFINITY;
                double beta = warp_min(bj);
                double sc = rec_scale(a.neg_inv_lam, bj, beta);
                double eta = warp_sum((lane < cnt) ? sc * recs[(size_t)lane * rstride + 1] : 0.0);
                auto merged_row = [&](int t) {
                    double e = 0.0;
                    for (int i = 0; i < cnt; ++i) {
                        const double si = __shfl_sync(0xffffffffu, sc, i);
                        if (t < H) e += si * __ldcg(recs + (size_t)i * rstride + PART_HDR + t);
                    }
                    return e;
                };
                if (a.p2p || a.record_only) {
                    // publish this rank's record: into every rank's exchange buffer (p2p) or into rank_part
                    const int nq = a.p2p ? a.world : 1;
                    for (int t0 = 0; t0 < H; t0 += 32) {
                        const int t = t0 + lane;
                        const double e = merged_row(t);
                        for (int qr = 0; qr < nq; ++qr) {
                            double *dst = (a.p2p ? peer_recs(pt.recs[qr], a.world, a.g.B, rec_len, a.epoch) : rank_part) +
                                          ((size_t)a.rank * a.g.B + b) * rec_len;
                            if (t < H) dst[PART_HDR + t] = e;
                            if (t0 == 0 && lane == 0) {
                                dst[0] = beta;
                                dst[1] = eta;
                            }
                        }
                    }
                }
                if (a.p2p) {
                    __threadfence_system();
                    __syncwarp();
                    if (lane < a.world) peer_signal_and_wait(pt, a, b, lane, rec_len, err_word);
                    __syncwarp();
                    recs = peer_recs(pt.recs[a.rank], a.world, a.g.B, rec_len, a.epoch) + (size_t)b * rec_len;
                    rstride = (size_t)a.g.B * rec_len;
                    cnt = a.world;
                    bj = (lane < cnt) ? __ldcg(recs + (size_t)lane * rstride) : INFINITY;
                    beta = warp_min(bj);
                    sc = rec_scale(a.neg_inv_lam, bj, beta);
                    eta = warp_sum((lane < cnt) ? sc * __ldcg(recs + (size_t)lane * rstride + 1) : 0.0);
                }
                if (!a.record_only) {
                    const double inv_eta = 1.0 / eta;
                    double *U = Ug + (size_t)b * H;
                    double *o = out + (size_t)b * (size_t)(2 + H + (H + 1) * NS);
                    for (int t0 = 0; t0 < H; t0 += 32) {
                        const int t = t0 + lane;
                        const double e = merged_row(t);
                        if (t < H) {
                            const double un = U[t] + inv_eta * e;  // mppi.py:62,69
                            U[t] = un;
                            o[1 + t] = un;
                            if (t == 0) o[0] = un;
                            sUc[(size_t)sl * H + t] = fmin(fmax(un, a.lo), a.hi);
                        }
                    }
                }
                if (!a.record_only && a.want_traj && !a.goal_zero)
                    for (int i = lane; i < H * NZ; i += 32) sG[(size_t)sl * H * NZ + i] = g_z[(size_t)b * H * NZ + i];
            }
            if (a.goal_zero && tid < NZ) sG[tid] = 0.0;
            __syncthreads();
            if (!a.record_only && a.want_traj && tid < ng) {
                const int b = b_lo + g0 + tid;
                double *o = out + (size_t)b * (size_t)(2 + H + (H + 1) * NS);
                o[1 + H] = nominal_rollout<typename Env::Scalar>(scalar_consts(ec), H, x0g + (size_t)b * NS, sUc + (size_t)tid * H,
                                                a.goal_zero ? sG : sG + (size_t)tid * H * NZ, a.goal_zero ? 0 : NZ,
                                                o + 2 + H, a.model_noise, a.seed, a.noise_draw, (uint32_t)b);
            }
            __syncthreads();
        }
    }
}

// ================================================================================================ rank merge
// Multi-GPU second half: one block per solve merges the `world` rank records gathered in rank_part
// ([world][B][2+H]) and applies them exactly as the single-GPU tail does.
template <class Env>
__global__ void __launch_bounds__(256) mppi_rank_merge_kernel(const MppiArgs a, const EnvConst<typename Env::Vec> ec,
                                                              double *__restrict__ rank_part,
                                                              const double *__restrict__ x0g,
                                                              const double *__restrict__ g_z, double *Ug,
                                                              double *__restrict__ out) {
    constexpr int NZ = Env::NS + 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ double red[32];
    const int H = a.g.H, b = blockIdx.x, rec_len = PART_HDR + H, nw = blockDim.x >> 5;
    double *sscale = reinterpret_cast<double *>(smem_raw);  // [world]
    double *sUc = sscale + a.world;                         // [H]
    double *sG = sUc + H;                                   // [H * NZ]
    double *sEw = sG + (size_t)H * NZ;                      // [nw][H]
    double *Eout = sEw + (size_t)nw * H;                    // [H]
    double beta, eta;
    block_merge(a.neg_inv_lam, H, rank_part + (size_t)b * rec_len, a.world, (size_t)a.g.B * rec_len, sscale, sEw, red,
                Eout, &beta, &eta);
    block_apply<Env>(a, ec, b, true, beta, eta, Eout, rank_part, x0g, g_z, Ug, out, sUc, sG);
}

}  // namespace mpc

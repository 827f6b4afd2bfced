// kernels.cuh — launch-side declarations shared by kernels.cu (device code) and capi.cu (the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mpc {

constexpr int TILE = 128;       // padding unit of K (row stride of the noise tensor) and CEM block size
constexpr int PART_HDR = 2;     // partial record = [beta, eta, E[0..H-1]]
constexpr int MERGE_GROUP = 16; // records merged warp-locally in one L2 round trip; block records are merged in groups of it

// NOISE_STREAM: the resident tensor again, but staged in small per-warp chunk rings and re-read (L2) by the weighted sum
enum { NOISE_TENSOR = 0, NOISE_PHILOX = 1, NOISE_STREAM = 2 };
enum { DT_F64 = 0, DT_F32 = 1 };

struct EnvParamsHost {
    int env;
    double c[24];  // folded constants, float64 parity layout (see fold_env_params)
    double f[24];  // folded constants, float32 throughput layout (narrowed when the kernel is launched)
    const char *f32_unsupported;  // null, or why these parameters have no float32 flavour
    int generic;   // generic-cost flavour: full (non-diagonal) W and / or per-step goal rows in the sample rollouts
    double w[15];  // its cost matrix: upper triangle (row major) of the symmetrised effective matrix (envs.cuh)
};

// Geometry of one handle's shard; identical for every solve.  MPPI tiles are WARP tiles of wt samples (mppi.cuh).
struct Geometry {
    int K_local;          // samples of each solve owned by this handle
    int k_begin;          // global index of the first one
    int Kpad;             // K_local rounded up to a multiple of TILE (row stride of the noise tensor)
    int H;
    int B;                // batch of independent solves
    int wt;               // samples per warp tile: 32 lanes x samples per thread
    int tiles_per_solve;  // Kpad / wt
    int total_tiles;      // B * tiles_per_solve
    int warps;            // W: warps per solve-kernel block
    int stages;           // S: shared-memory ring stages (S >= W; S - W tiles prefetched ahead)
    int solves_per_block; // 0 = mode A (a solve is shared by blocks_per_solve blocks); > 0 = mode B (whole solves)
    int blocks_per_solve; // mode A
    int chunk;            // mode A: consecutive tiles per block
    int grid;             // solve-kernel blocks
    int max_parts;        // partial records per solve: blocks_per_solve (A) or warps (B)
    int n_groups;         // mode A, more than MERGE_GROUP blocks per solve: groups of the two-level merge; else 0
    int part_stride;      // records per solve in `part`: max_parts + n_groups (the group records follow the block records)
    int fin_group;        // mode B: solves finalised together (one thread per nominal rollout)
    int stream_rows;      // 0 = whole [H][wt] tiles in the ring; > 0 = STREAM: time steps per chunk of a warp's private ring
    int stream_depth;     // STREAM: chunk stages per warp
    int u_copies;         // copies of the plan in shared memory (rollout dtype): 1 when a block works on one solve, else W
    int variant;          // rollout flavour: bits 0-2 trig (0 polynomial, 1 MUFU, Acrobot 2/4/5 per-stage splits), bit 3 two
                          // samples per thread (f32x2), bit 4 WIDE (32-warp whole-tile blocks, one sample per thread)
    int ns;
    size_t smem_rollout;  // dynamic shared memory of the solve kernel
};

// State and plan of a batch-1 host call passed as kernel parameters (constant bank) instead of through a host-to-device
// copy that the launch would have to wait for: ~5 us off every predict_action-sized call.
constexpr int INLINE_MAX_H = 128;
struct InlineInputs {
    int on;                    // bit 0: x0 is valid, bit 1: U is valid
    double x0[4];
    double U[INLINE_MAX_H];
};

struct MppiArgs {
    Geometry g;
    double neg_inv_lam;   // -(1/lam)                       mppi.py:60
    double pert;          // lam * pinv(std^2)              mppi.py:45-46
    double sigma;
    double lo, hi;        // action bounds (agent.py:64-65)
    uint64_t seed;
    uint32_t draw;        // Philox draw index
    int noise_batched;    // TENSOR: 1 = one tensor per solve of the batch, 0 = shared
    int want_traj;        // compute the clipped plan's nominal rollout in the tail of the solve kernel
    int store_costs;      // write the per-sample costs S[k] (agent.sample_costs before mppi.py:73 zeroes them)
    int goal_zero;        // the goal trajectory is all zeros (what every reference script passes)
    int per_step_goal;    // generic-cost flavour only: sample rollouts use goal row t at step t (an extension: the
                          // reference's rollouts only ever see row 0, model.py:34,38)
    int record_only;      // the tail publishes this rank's (beta, eta, E) record instead of applying it
    int p2p;              // multi-GPU: all-gather the rank records inside the kernel over peer memory, then apply
    unsigned long long epoch;  // p2p: flag value of this solve (monotonic per handle, equal on all ranks)
    unsigned long long *trace; // null, or 16 words for the finalising block's phase stamps (mpc_get_trace)
    int world;            // ranks whose records the final merge combines
    int rank;
    // host-flag completion (batch = 1, host-buffer solve call): the finalising block also writes the solve's outputs
    // into mapped pinned host memory and then raises host_flag to host_epoch — the host polls the flag instead of
    // submitting a D2H copy and a stream synchronise
    double model_noise;   // std of the noise Model.predict adds to every new state (model.py:42-43); 0 = off
    uint32_t noise_draw;  // Philox draw of the model noise: the solve counter (fresh per solve)
    double *host_out;                 // null = off; same layout as `out`
    unsigned long long *host_flag;
    unsigned long long host_epoch;
};

struct CemArgs {
    Geometry g;
    double sigma, lo, hi, alpha, epsilon;
    int n_elite;
    int iter;             // iteration index (Philox draw, history slot)
    uint64_t seed;
    int want_traj;
    // K sharded over several GPUs: elite indices are GLOBAL sample indices
    int world, rank;
    int n_cand;           // candidates a rank contributes per iteration: min(n_elite, largest K_local)
    int z_stride;         // row stride of the standard-normal tensor (TENSOR noise): Kpad, or the padded FULL K
    int z_off;            // column of this rank's first sample in it: 0, or k_begin
    unsigned long long epoch;  // flag value of this iteration's candidate exchange
    double model_noise;   // as MppiArgs
    uint32_t noise_draw;  // solve counter * max_iter + iteration
    int per_step_goal;    // as MppiArgs
};

// CEM as one persistent cooperative launch (cem_fused.cuh): single GPU, Philox noise
struct CemFusedArgs {
    CemArgs a;            // iter unused
    int max_iter;
    int spb;              // samples per block (multiple of 32)
    int all_pairs;        // K <= CEM_FUSED_ALLPAIRS_K: rank by counting, no radix passes
    int cached;           // one warp's worth of samples per block (spb = 32): the samples live in shared memory
    uint32_t noise_draw0; // model-noise draw of iteration 0: solve counter * max_iter
    unsigned long long *trace;  // null, or 16 words: %globaltimer stamps of block 0 (mpc_get_trace)
    double *host_out;     // null, or mapped pinned memory receiving the outputs + completion flag (batch-1 host call)
    unsigned long long *host_flag;
    unsigned long long host_epoch;
};
struct CemFusedScratch {
    double *slots;        // [grid][2 + 2 H]: elite count, -, sum (x - c) [H], sum (x - c)^2 [H]
    int *stage;           // [Kpad] a block's elite indices before their global offset is known
    int *hist;            // [max_iter][8][256] radix histograms (zeroed by the kernel)
    int *tie_cnt;         // [grid] keys equal to the threshold key per block (exact-tie split, rare)
    unsigned *bar;        // [2] grid barrier: arrival count, generation (persist across launches)
};

struct MppiBuffers {
    const void *noise;     // R [Bn][H][Kpad]
    const void *noise_map; // host copy of the CUtensorMap over `noise` (rank 3: Kpad, H, Bn; box wt x H x 1)
    const double *x0;      // [B][ns]
    const double *g_z;     // [B][H][nz]
    double *U;             // [B][H]   device-resident plan (in: before, out: after)
    double *S;             // [B][Kpad]
    double *part;          // [B][part_stride][2+H]
    double *rank_part;     // [world][B][2+H]  (all-gather buffer; this rank writes slot `rank`)
    double *out;           // [B][out_stride]: u0, U[H], cost, xs[(H+1)*ns]
    int *done_count;       // [B][1 + n_groups] arrival tickets of the solve kernel's blocks (self-cleaning)
    double *const *peer_recs;  // host array [world] of the ranks' exchange buffers (p2p), else null
    const InlineInputs *inl;   // host pointer or null: state / plan of this solve as kernel parameters
};

struct CemBuffers {
    const void *z;         // R [H][Kpad] standard normals of this iteration (TENSOR)
    const double *x0;      // [ns]
    const double *g_z;     // [H][nz]
    double *mu;            // [H]
    double *sd;            // [H]
    double *R;             // [max_iter][Kpad] rewards history
    unsigned long long *keys;  // [Kpad] order-preserving keys of this iteration
    int *elite_idx;        // [max_iter][n_elite]
    int *sel;              // scratch: [0]=done flag, [1]=iters executed, [2]=elite counter, ...
    double *out;           // u0, U[H], cost, xs
    double *const *peer_recs;  // host array [world] of the ranks' exchange buffers (world > 1)
    double *err_word;      // set when a peer did not deliver in time
    const InlineInputs *inl;   // single-launch solve only: host pointer or null
};

// closed-loop control step on the device (episode_step_kernel, kernels.cu)
struct EpisodeArgs {
    int H, step, n_steps, shift, mode;  // mode 0 = control step, 1 = warm-start step (clip the plan only)
    double lo, hi, actuation_std, sensor_std;
    uint64_t seed;
};
struct EpisodeBuffers {
    const double *out;    // solve outputs (see above)
    double *U;            // device-resident plan of the agent (MPPI planned_us / CEM mean)
    double *x0;           // the solver's state input: receives the next observation
    double *true_state;   // [ns] plant state
    double *states;       // [n+1][ns] observations   (row 0 written by the host)
    double *true_states;  // [n+1][ns]
    double *actions;      // [n] applied controls (after actuation noise: env.py history "action")
    double *costs;        // [n]
    double *plan_us;      // [n][H]        nullable
    double *plan_xs;      // [n][H+1][ns]  nullable
    double *plan_costs;   // [n]           nullable
};

// launches (kernels.cu); all return cudaGetLastError()
// the whole solve (rollout + softmin update + plan update + nominal rollout); world > 1: up to this rank's record
cudaError_t launch_mppi_solve(int dtype, int noise, int variant, const EnvParamsHost &ep, const MppiArgs &a,
                              const MppiBuffers &b, cudaStream_t s);
// multi GPU second half: merge the gathered rank records, apply, nominal rollout
cudaError_t launch_mppi_rank_merge(int dtype, const EnvParamsHost &ep, const MppiArgs &a, const MppiBuffers &b,
                                   cudaStream_t s);
cudaError_t launch_transpose_noise(int dtype, const double *src_KH, void *dst_HK, int K_local, int Kpad, int H,
                                   cudaStream_t s);
cudaError_t launch_philox_fill(int dtype, void *dst_HK, int K_local, int k_begin, int Kpad, int H, double scale,
                               uint64_t seed, uint32_t draw, cudaStream_t s);
cudaError_t launch_philox_materialize(double *dst_KH, int K_local, int k_begin, int H, uint64_t seed, uint32_t draw,
                                      cudaStream_t s);
cudaError_t launch_philox_blocks(double *dst_K_n_4, int K_local, int k_begin, int n_blocks, uint64_t seed, uint32_t draw,
                                 uint32_t stream, cudaStream_t s);
cudaError_t launch_cem_rollout(int dtype, int noise, int variant, const EnvParamsHost &ep, const CemArgs &a, const CemBuffers &b,
                               cudaStream_t s);
cudaError_t launch_cem_select(const CemArgs &a, const CemBuffers &b, cudaStream_t s);
cudaError_t launch_cem_refit(int dtype, int noise, const CemArgs &a, const CemBuffers &b, cudaStream_t s);
cudaError_t launch_cem_finish(int dtype, const EnvParamsHost &ep, const CemArgs &a, const CemBuffers &b, cudaStream_t s);
cudaError_t launch_cem_reset(const CemArgs &a, const CemBuffers &b, cudaStream_t s);
// the whole CEM solve in one cooperative launch of `grid` blocks (cem_fused.cuh); smem = launch_cem_fused_smem(...)
cudaError_t launch_cem_fused(int dtype, int variant, const EnvParamsHost &ep, const CemFusedArgs &fa, const CemBuffers &b,
                             const CemFusedScratch &sc, int grid, int coop, cudaStream_t s);
constexpr int CEM_FUSED_THREADS = 512;
constexpr int CEM_FUSED_ALLPAIRS_K = 4096;  // up to this K the elites are found by all-pairs rank counting
constexpr int CEM_FUSED_MAX_H = 512;
cudaError_t launch_simulate(int dtype, const EnvParamsHost &ep, int H, double lo, double hi, int do_clip,
                            const double *x0, const double *us, const double *g_z, double *xs, double *cost,
                            cudaStream_t s);
cudaError_t launch_model_step(int dtype, const EnvParamsHost &ep, int n, const double *x, const double *u,
                              const double *g, double *cost, double *xn, cudaStream_t s);
cudaError_t launch_fma_peak(int dtype, float *sink, int blocks, int iters, cudaStream_t s);
cudaError_t launch_fill(float *p, size_t n, float v, cudaStream_t s);
// always the float64 plant (compiled in the float64 translation unit only)
cudaError_t launch_episode_step(const EnvParamsHost &ep, const EpisodeArgs &a, const EpisodeBuffers &b, cudaStream_t s);

// rows = time steps per ring stage (H, or the STREAM chunk); S = ring stages of the block (STREAM: W * depth)
// uc = copies of the plan in the rollout dtype: 1, or 0 for "one per warp"
size_t mppi_solve_smem(int dtype, int variant, int H, int rows, int S, int W, int cap, int G, int nz, int uc);
size_t mppi_stage_bytes(int dtype, int variant, int rows);
// warps per block the solve kernel of (dtype, noise kind, variant) is compiled for (SolveShape in mppi.cuh)
int mppi_max_warps(int dtype, int noise_kind, int variant);
constexpr int MPPI_MAX_H = 256;      // 32 * MAX_ROWS

}  // namespace mpc

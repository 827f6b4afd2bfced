/*
 * mpc_b200.h — C ABI of libmpcb200.so: the sampling-MPC inner loop (MPPI / CEM rollout, cost accumulation and
 * control update) of MPCBenchmark as hand-written sm_100a CUDA.
 *
 * This is the drop-in boundary.  The reference is pure Python and has no FFI; the entry points below are what a
 * ctypes binding placed inside the reference's agents would call.  Each one names the reference code it replaces
 * (paths relative to /root/reference/MPCBenchmark).  The binding a maintainer would add is in INTEGRATION.md; the
 * repo's own binding is benchmarking-mpc-solvers_b200/capi.py.
 *
 * Conventions
 *   - plain pointers and sizes only; every host pointer is borrowed for the duration of the call;
 *   - all host-side arrays are float64, C-contiguous, in the reference's own layouts
 *     (delta_u [K][H][du], planned_us [H][du], g_z [H][nz], xs [H+1][ns]); du = 1 for the three models;
 *   - every function returns MPC_OK (0) or a negative mpc_status; mpc_last_error() gives the message;
 *   - a handle is not re-entrant: one call in flight per handle (same as the reference agents, which mutate
 *     planned_us and model.last_reward);  calls block until results are in the caller's buffers unless the name
 *     ends in _async;
 *   - there is NO CPU fallback: without a CUDA device mpc_create fails with MPC_ECUDA.
 */
#ifndef MPC_B200_H
#define MPC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPC_B200_ABI_VERSION 5

typedef struct mpc_handle mpc_handle;

typedef enum {
    MPC_OK = 0,
    MPC_EINVAL = -1, /* bad argument / unsupported configuration */
    MPC_ECUDA = -2,  /* CUDA runtime error, or no device */
    MPC_ENCCL = -3,  /* NCCL error */
    MPC_ENOMEM = -4, /* host or device allocation failed */
    MPC_ESTATE = -5  /* call sequence error (e.g. solve before noise was set) */
} mpc_status;

typedef enum { MPC_AGENT_MPPI = 0, MPC_AGENT_CEM = 1 } mpc_agent;
/* models/pendulum_model.py, models/cartpole_swingup_model.py, models/acrobot_model.py */
typedef enum { MPC_ENV_PENDULUM = 0, MPC_ENV_CARTPOLE = 1, MPC_ENV_ACROBOT = 2 } mpc_env;
/* Arithmetic of the rollout.  MPC_F64: everything float64 (reference precision, parity 1e-12).
 * MPC_F32: dynamics and per-step cost in float32; per-sample cost accumulation, softmin weights, elite keys and
 * all reductions stay float64 (parity 1e-5, see DESIGN.md "precision"). */
typedef enum { MPC_F64 = 0, MPC_F32 = 1 } mpc_dtype;
/* MPC_NOISE_TENSOR: a resident noise tensor (MPPI: agent.delta_u, mppi.py:20-21; CEM: the standard normals behind
 * np.random.normal, cem.py:58-59) uploaded by the caller.  MPC_NOISE_PHILOX: generated in-kernel, Philox4x32-10,
 * counter = (global sample index, t/4, draw, stream) so the stream does not depend on the sharding. */
typedef enum { MPC_NOISE_TENSOR = 0, MPC_NOISE_PHILOX = 1 } mpc_noise;

#define MPC_MAX_ENV_PARAMS 24

/* Model parameters, read from the reference model instance by the host binding:
 *  pendulum: g m l dt max_speed max_torque W0 W1 W2                        (pendulum_model.py:8-22)
 *  cartpole: g m_c m_p l force_mag dt b x_threshold W0..W4 (diag of self.W)  (cartpole_swingup_model.py:23-46)
 *  acrobot : dt m1 m2 l1 lc1 lc2 I1 I2 g max_vel1 max_vel2 W0..W4 nips     (acrobot_model.py:60-97,107; nips != 0
 *            selects the book_or_nips == "nips" equations of :124-128, MPC_F64 only; 0 / absent = "book") */
typedef struct {
    uint32_t struct_size; /* sizeof(mpc_config), for ABI checking */
    int32_t agent;        /* mpc_agent */
    int32_t env;          /* mpc_env */
    int32_t dtype;        /* mpc_dtype */
    int32_t noise;        /* mpc_noise */
    int32_t device;       /* CUDA device ordinal */
    int32_t K;            /* samples per solve, GLOBAL count (params["K"], mppi.py:12 / cem.py:20) */
    int32_t H;            /* horizon (params["T"], mppi.py:13 / cem.py:21) */
    int32_t batch;        /* independent solves per call (>= 1); each has its own state, nominal and goal */
    int32_t rank;         /* this handle rolls out samples [K*rank/world, K*(rank+1)/world) of every solve */
    int32_t world;
    int32_t max_iter;     /* CEM (cem.py:22) */
    int32_t n_elite;      /* CEM (cem.py:24) */
    int32_t resample;     /* PHILOX only: 0 = one fixed noise draw reused by every solve (the reference's MPPI
                             semantics, mppi.py:20-21), 1 = fresh draw per solve */
    int32_t fast_math;    /* MPC_F32 only: 0 = polynomial sincos (1.5e-7 abs, default); 1 = MUFU.SIN/COS in the dynamics
                             (~4e-7 abs per call plus the rounding of the radian argument).  Acrobot (four RK4 stages per
                             step, models/acrobot_model.py:270-282) also takes per-stage splits between the FMA pipe
                             (polynomial) and the XU pipe (MUFU): 2 = stage 1 polynomial, stages 2-4 MUFU; 4 = MUFU with
                             the unwrapped stage angles first reduced modulo one turn; 5 = 2 with the reduction of 4 */
    int32_t lanes;        /* MPC_F32 MPPI only: samples rolled out per thread, 1 or 2 (2 = packed f32x2 arithmetic,
                             FFMA2); 0 = library default.  Results do not depend on it. */
    double lam;           /* MPPI temperature (mppi.py:17) */
    double sigma;         /* params["std"] (mppi.py:16 / cem.py:28) */
    double epsilon;       /* CEM (cem.py:25) */
    double alpha;         /* CEM (cem.py:26) */
    double bounds_low;    /* model.bounds_low[0] / bounds_high[0] (agent.py:9-10) */
    double bounds_high;
    uint64_t seed;        /* PHILOX key */
    double env_params[MPC_MAX_ENV_PARAMS];
    double model_noise_std; /* Model.set_sensor_noise (model.py:20-24): every Model.predict INSIDE the rollouts (and the
                             nominal rollout) adds N(0, std) to each component of the new state (model.py:42-43).  0 = off.
                             The draws are Philox (fresh per solve): statistically the reference's, not its numpy stream
                             (which is not reproducible there either: forked workers share one RNG state). */
    /* Generic cost.  full_W != 0: W holds the model's whole cost matrix self.W, row major nz x nz
     * (nz = state_size + 1), used as einsum("bi,ij,bj->b") does (pendulum_model.py:46-50,
     * cartpole_swingup_model.py:95-104, acrobot_model.py:172-176) — any matrix, negative or non-symmetric; the W_i
     * entries of env_params are then ignored.  per_step_goal != 0: the sample rollouts use goal row t at step t (an
     * extension: the reference's rollouts only ever see row 0, model.py:34,38; its nominal rollout already uses row t,
     * agent.py:103-104, and so does this library's).  Either switch selects the generic flavour of the kernels: one
     * sample per thread, polynomial trig, whole-tile or Philox noise. */
    int32_t full_W;
    int32_t per_step_goal;
    double W[25];
} mpc_config;

/* ---- lifetime ------------------------------------------------------------------------------------------ */
/* Replaces MPPI.__init__ (mppi.py:10-24) / CEM.__init__ (cem.py:18-31): allocates every device buffer, the stream
 * and pinned staging.  The nominal plan starts at zeros like planned_us. */
int mpc_create(const mpc_config *cfg, mpc_handle **out);
/* Replaces MPPI.close / CEM.close (mppi.py:31-33, cem.py:38-40): frees device memory, stream, communicator. */
int mpc_destroy(mpc_handle *h);
/* Message of the last failure on this handle (h == NULL: last failure of mpc_create in this thread). */
const char *mpc_last_error(const mpc_handle *h);
int mpc_abi_version(void);

/* ---- noise --------------------------------------------------------------------------------------------- */
/* MPPI, TENSOR mode: upload agent.delta_u (mppi.py:20-21), host layout [K][H] float64 (GLOBAL K; the handle takes
 * its own shard) and re-lay it out as [H][K_local] tiles in the rollout dtype. */
int mpc_set_noise(mpc_handle *h, const double *delta_u_KH);
/* TENSOR mode without a host tensor: fill the resident tensor on the device with sigma * N(0,1) from Philox
 * (same values mpc_philox_noise(h, 0, ...) returns, scaled by sigma). */
int mpc_fill_noise(mpc_handle *h);
/* Materialise what the in-kernel generator produces for this handle's shard: out [K_local][H] float64 standard
 * normals for draw index `draw` (MPPI: solve counter if resample else 0; CEM: iteration counter).  Lets a checker
 * feed the identical noise to a CPU implementation. */
int mpc_philox_noise(mpc_handle *h, uint64_t draw, double *out_KlocalH);
/* The standard normals the noisy-model flavour (mpc_config.model_noise_std > 0) adds, times the std, to the new state
 * of sample k at step t during MPPI solve number `solve_index` (0 = the handle's first solve): out [K_local][H][4],
 * components 0..ns-1 used (model.py:42-43).  For checkers. */
int mpc_philox_model_noise(mpc_handle *h, uint64_t solve_index, double *out_KlocalH4);

/* ---- MPPI ---------------------------------------------------------------------------------------------- */
/* One MPPI._calc_action (mppi.py:54-75) for `batch` independent solves, fused with what Agent.predict_action does
 * right after it (agent.py:63-69): per-sample rollout costs (MPPI.f, mppi.py:35-48, Model.predict model.py:26-44),
 * softmin weights, weighted noise sum, planned_us += erg, then the clipped plan's nominal rollout
 * (Agent.simulate_trajectory, agent.py:97-108).
 *   x        [batch][ns]      in   state
 *   g_z      [batch][H][nz]   in   goal trajectory; MPPI.f uses row 0 (model.py:34,38), simulate uses row t
 *   U_in     [batch][H]       in   planned_us before the solve; NULL = keep the device-resident plan
 *   u0       [batch]          out  planned_us[0] after the update (not clipped, as _calc_action returns it)
 *   U_out    [batch][H]       out  planned_us after the update (not clipped); may be NULL
 *   xs_out   [batch][H+1][ns] out  nominal rollout of clip(U_out); may be NULL (then it is not computed)
 *   cost_out [batch]          out  its cost; may be NULL */
int mpc_mppi_solve(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *u0, double *U_out,
                   double *xs_out, double *cost_out);
/* Per-sample costs of the last solve, this handle's shard: out [batch][K_local] (the reference's
 * agent.sample_costs before it is zeroed at mppi.py:73). */
int mpc_mppi_get_costs(mpc_handle *h, double *S_out);

/* ---- CEM ----------------------------------------------------------------------------------------------- */
/* CEM._calc_action (cem.py:54-78) split so that the host can keep the reference's RNG contract (one
 * np.random.normal draw per executed iteration).
 * begin: std = self.std.copy() (cem.py:56); mean = planned_us (U_in [batch][H], NULL = device-resident). */
int mpc_cem_begin(mpc_handle *h, const double *x, const double *g_z, const double *U_in);
/* One refit iteration (cem.py:58-76).  z_KH: TENSOR mode, standard normals [K][H] float64 such that
 * np.random.normal(mean, std) == mean + std*z; NULL in PHILOX mode.  *done = 1 when every std < epsilon;
 * done == NULL only enqueues the iteration (no synchronisation; later iterations are no-ops once the device-side
 * flag is set).  With world > 1 (K sharded, mpc_p2p_* connected) pass the FULL [K][H] array on every rank: elites
 * are global and the refit reads any sample's normals; elite indices reported by mpc_cem_get_elites are global. */
int mpc_cem_iterate(mpc_handle *h, const double *z_KH, int32_t *done);
/* Result: u0 = planned_us[0] (cem.py:77), U_out = planned_us, xs/cost as for MPPI. */
int mpc_cem_end(mpc_handle *h, double *u0, double *U_out, double *xs_out, double *cost_out);
/* PHILOX mode convenience: begin + up to max_iter iterations with the early break evaluated on the device + end,
 * one host synchronisation.  iters_out receives the number of executed iterations. */
int mpc_cem_solve(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *u0, double *U_out,
                  double *xs_out, double *cost_out, int32_t *iters_out);
/* Rewards of iteration `iter` of the last solve (solve 0 of a batch): out [K]; elite indices: out [n_elite],
 * ascending.  (cem.py:64,67) */
int mpc_cem_get_rewards(mpc_handle *h, int32_t iter, double *R_out);
int mpc_cem_get_elites(mpc_handle *h, int32_t iter, int32_t *idx_out);
/* Iterations the last solve executed before all(std < epsilon) stopped it (cem.py:76-77), at most max_iter. */
int mpc_cem_get_iterations(mpc_handle *h, int32_t *iters_out);
/* Current std of solve 0 (cem.py:73): out [H]. */
int mpc_cem_get_std(mpc_handle *h, double *std_out);

/* ---- closed-loop episode, resident on the device ----------------------------------------------------------------
 * Replaces the control loop of Experiment.run (ExperimentCore/Experiment.py:67-90): env.reset(start), agent.reset(),
 * agent.warm_start(env.true_state, warm_start_iters, goal = 0) (agents/agent.py:30-48), then n_steps times
 * action = agent.predict_action(env.state, goal = 0) (agent.py:50-76, logging branch included when `shift`) followed
 * by env.step(action) (envs/env.py:35-50) with optional actuation / sensor noise.  Every solve, the plan clip / log /
 * shift and the plant step are kernels enqueued back to back on the handle's stream; the host synchronises ONCE, at
 * the end, and reads the histories.  The plant is stepped in float64 whatever the rollout dtype.  Goal = zeros (what
 * every reference script passes); batch = 1, world = 1.  MPPI: resident noise tensor or Philox; CEM: Philox noise.
 * Environment noise is Philox (streams 2, 3; key `seed`), so noisy episodes are reproducible — the reference's are
 * not (global numpy RNG shared with forked workers). */
typedef struct mpc_episode_opts {
    uint32_t struct_size;      /* sizeof(mpc_episode_opts) */
    int32_t n_steps;           /* Experiment.experiment_length */
    int32_t warm_start_iters;  /* 0 = no warm start */
    int32_t shift;             /* 1 = agent.logging (default of the reference): log rows + receding-horizon shift */
    double actuation_std;      /* Environment.set_actuation_noise, 0 = off */
    double sensor_std;         /* Environment.set_sensor_noise, 0 = off */
    uint64_t seed;
} mpc_episode_opts;
/* x0 [ns] start state; U_in [H] initial plan (null = keep the resident one; agent.reset does not clear it).
 * Outputs (caller-owned): states [n+1][ns] observations (env.history "state", row 0 = x0), true_states [n+1][ns],
 * actions [n] (env.history "action": after actuation noise), costs [n] (env.history "cost"); nullable: plan_us
 * [n][H] (planning_action_history), plan_xs [n][H+1][ns] (planning_state_history), plan_costs [n]
 * (planning_costs_history), U_final [H] (agent.planned_us after the last step). */
int mpc_episode_run(mpc_handle *h, const double *x0, const double *U_in, const mpc_episode_opts *opts, double *states,
                    double *true_states, double *actions, double *costs, double *plan_us, double *plan_xs,
                    double *plan_costs, double *U_final);

/* ---- nominal rollout ------------------------------------------------------------------------------------- */
/* Agent.simulate_trajectory (agent.py:97-108) for one plan: us [H], g_z [H][nz] -> xs [H+1][ns], cost. */
int mpc_simulate(mpc_handle *h, const double *x, const double *us, const double *g_z, double *xs_out,
                 double *cost_out);
/* Model.predict + get_reward (model.py:26-44,62-63) for n independent (x,u) pairs with goal rows g [n][nz]:
 * cost_out [n], xn_out [n][ns]. */
int mpc_model_step(mpc_handle *h, int32_t n, const double *x, const double *u, const double *g, double *cost_out,
                   double *xn_out);

/* ---- multi-GPU (one process per GPU; only the sample dimension shards) ------------------------------------ */
/* Replaces the scatter/gather of Pool.map (mppi.py:55-57, cem.py:61-64) across devices.  Rank 0 creates an id,
 * the host side broadcasts its bytes (MPC_COMM_ID_BYTES), every rank calls mpc_comm_init.  After that every solve
 * performs ONE small collective: all-gather of (min cost, sum w, sum w*eps[H]) per rank. */
#define MPC_COMM_ID_BYTES 128
int mpc_comm_create_id(void *id_out);
int mpc_comm_init(mpc_handle *h, const void *id);
/* The same exchange INSIDE the solve kernel, over NVLink peer memory: each rank's kernel stores its record into every
 * rank's exchange buffer, raises a flag with a system-scope release and waits (bounded, 4 s) for the others' flags —
 * one launch per solve, no NCCL call, no second kernel.  One process per GPU: every rank exports the 64-byte IPC
 * handle of its buffer, the host side all-gathers them (rank-major) and every rank imports.  One process driving
 * several GPUs: mpc_p2p_link over the `world` handles (rank-ordered).  Takes precedence over mpc_comm_init.
 * Requires the whole grid resident (always true for batch = 1). */
#define MPC_P2P_HANDLE_BYTES 64
int mpc_p2p_export(mpc_handle *h, void *handle_out);
int mpc_p2p_import(mpc_handle *h, const void *handles);
int mpc_p2p_link(mpc_handle *const *handles, int32_t world);
/* The same exchange with a caller-supplied transport (MPI, gloo, sockets ...), in two halves:
 * rollout_partial: rollout of this handle's K shard, block partials merged into ONE record per solve,
 *                  rec_out [batch][2+H] = (min cost, sum w, sum w*eps[0..H-1]) relative to the shard's own minimum;
 * apply_partials:  recs [world][batch][2+H] gathered from every rank (rank-major) -> rescale to the global minimum,
 *                  planned_us += erg, nominal rollout; outputs as mpc_mppi_solve.  Every rank gets the same plan. */
int mpc_mppi_rollout_partial(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *rec_out);
int mpc_mppi_apply_partials(mpc_handle *h, const double *recs, double *u0, double *U_out, double *xs_out,
                            double *cost_out);

/* ---- measurement helpers --------------------------------------------------------------------------------- */
/* Device-resident solve: inputs uploaded once, then `mpc_solve_enqueue` launches the solve kernels on the handle's
 * stream without any host<->device copy or synchronisation. */
int mpc_upload_inputs(mpc_handle *h, const double *x, const double *g_z, const double *U_in);
int mpc_solve_enqueue(mpc_handle *h);
/* The same with the nominal rollout of the clipped plan (Agent.simulate_trajectory, agent.py:97-108: the logging
 * branch of predict_action, not part of _calc_action) switched on or off. */
int mpc_solve_enqueue_ex(mpc_handle *h, int32_t want_traj);
/* n device-resident solves, each between its own pair of CUDA events on the handle's stream, optionally with an
 * (untimed) L2 flush after each, all enqueued before the single synchronisation at the end: ms_each [n] = the device
 * time of every solve's own launches, free of host launch latency.  world > 1: one untimed solve is enqueued first so
 * that its in-kernel exchange aligns the ranks' device timelines. */
int mpc_time_solves(mpc_handle *h, int32_t n, int32_t want_traj, int32_t flush_l2, float *ms_each);
int mpc_download_outputs(mpc_handle *h, double *u0, double *U_out, double *xs_out, double *cost_out);
/* CUDA events on the handle's stream (torch.cuda.Event only sees torch's current stream). */
int mpc_timer_start(mpc_handle *h);
int mpc_timer_stop(mpc_handle *h, float *ms_out); /* records, synchronises, returns elapsed ms */
int mpc_sync(mpc_handle *h);
/* Profiling switches.  bit 0: the finalising block stamps its phases (mpc_get_trace); bit 1: do NOT store the
 * per-sample costs (then mpc_mppi_get_costs fails with MPC_ESTATE instead of returning stale data); bit 2: every MPPI
 * solve brackets its kernel with its own event pair — two more stream operations per solve — and mpc_get_rollout_ms
 * returns that kernel's device time for the last solve. */
int mpc_set_profiling(mpc_handle *h, int32_t on);
int mpc_get_rollout_ms(mpc_handle *h, float *ms_out);
/* With profiling on, the block that finalises an MPPI solve stamps its phases with %globaltimer (ns): out[0] kernel
 * start, [1] all tiles done, [2] last block's ticket, [3] block records merged, [4] rank record published to the
 * peers, [5] peers' records in, [6] rank records merged, [7] plan applied, [8] nominal rollout done (batch = 1). */
int mpc_get_trace(mpc_handle *h, uint64_t out[16]);
/* Number of kernel launches issued by this handle so far. */
int64_t mpc_launch_count(const mpc_handle *h);
/* Self-check (tests): number of non-zero arrival tickets of the MPPI solve kernel; they clean themselves, so it is 0
 * between solves. */
int mpc_debug_tickets(mpc_handle *h, int32_t *nonzero_out);
/* Geometry chosen for the rollout kernel: out[0]=grid, out[1]=block, out[2]=dynamic smem bytes, out[3]=tiles,
 * out[4]=K_local, out[5]=k_begin, out[6]=ring stages, out[7]=time steps per stage of the STREAM flavour (0: whole
 * [H][tile] stages). */
int mpc_get_geometry(const mpc_handle *h, int32_t out[8]);
/* FP32 / FP64 FMA-pipe peak of the device measured with a dependent-free FMA chain (TFLOP/s, 2 flops per FMA). */
int mpc_measure_fma_peak(int32_t device, int32_t dtype, double *tflops_out);
/* Page-lock / release a caller-owned host buffer (cudaHostRegister).  Batched solves (batch > 1) copy every output
 * field straight from the device into the caller's arrays, one strided copy per field; arrays registered once receive
 * them by direct DMA instead of through the driver's staging. */
int mpc_pin_host_buffer(void *p, size_t bytes);
int mpc_unpin_host_buffer(void *p);
/* Flush L2 by writing a buffer larger than it. */
int mpc_flush_l2(mpc_handle *h);

/* ---- allocation-free control update ------------------------------------------------------------------------ */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The arithmetic is the
 * control update of MPPI._calc_action (mppi.py:59-71): beta = min S, w = exp(-(S - beta)/lam) / eta,
 * U_out = U_in + sum_k w_k eps[k] — the same update mpc_mppi_solve fuses behind its rollout — as a variant of that
 * kernel's block-record merge whose scratch (one record per block and the ticket word) lives in a file-scope
 * `static __device__` array instead of memory allocated by mpc_create.  The call needs no handle and calls no
 * cudaMalloc / cudaFree / allocating library routine; it enqueues one 256-byte memset and one kernel on `stream`
 * and returns without synchronising.
 *   device    CUDA device ordinal the pointers belong to
 *   K, H      samples and horizon: 1 <= H <= MPC_STATIC_MAX_H, 1 <= K <= MPC_STATIC_MAX_K; anything larger is
 *             rejected with MPC_EINVAL (the array cannot grow) — use mpc_mppi_solve for those
 *   lam       MPPI temperature (mppi.py:17), > 0
 *   S_dev     [K]     in   per-sample costs, float64, DEVICE memory
 *   eps_dev   [K][H]  in   noise (delta_u, du = 1), float64, DEVICE memory
 *   U_in_dev  [H]     in   planned_us before the update, float64, DEVICE memory
 *   U_out_dev [H]     out  planned_us after the update (not clipped), DEVICE memory; may alias U_in_dev
 *   stream    cudaStream_t (NULL = the legacy default stream)
 *   err_out   receives the message of a failure (err_cap bytes, NUL-terminated); may be NULL
 * Limits: the array is about 0.6 MB (296 block records of 2 + 256 doubles) and every GPU that loads the library holds
 * it until the process exits.  Every call on a device shares it and it keeps the last call's contents (the ticket
 * word is cleared on `stream` before each launch): calls that OVERLAP on one GPU — different streams or threads
 * without an ordering between them — race on it and produce garbage; calls on one stream are safe. */
#define MPC_STATIC_MAX_H 256
#define MPC_STATIC_MAX_K (1 << 24)
int mpc_mppi_update_static(int32_t device, int32_t K, int32_t H, double lam, const double *S_dev, const double *eps_dev,
                           const double *U_in_dev, double *U_out_dev, void *stream, char *err_out, size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* MPC_B200_H */

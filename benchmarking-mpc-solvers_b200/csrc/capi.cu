// capi.cu — the C ABI declared in include/mpc_b200.h: handle, buffers, staging, launch sequencing, NCCL exchange.
// No torch types, no CPU fallback: every entry point either runs the CUDA path or returns an error.
#include "../../include/mpc_b200.h"
#include "kernels.cuh"

#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

using namespace mpc;

// ---------------------------------------------------------------------------------------------- NCCL (lazy dlopen)
// Only the five entry points the exchange needs; resolved on first use so that single-GPU users never load NCCL.
namespace {
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclFloat64 = 8 };
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
} g_nccl;

bool nccl_load(std::string *err) {
    if (g_nccl.ok) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) {
        *err = std::string("cannot load libnccl: ") + dlerror();
        return false;
    }
#define LOADSYM(field, name)                                              \
    g_nccl.field = (decltype(g_nccl.field))dlsym(g_nccl.lib, name);       \
    if (!g_nccl.field) {                                                  \
        *err = std::string("libnccl lacks ") + name;                      \
        return false;                                                     \
    }
    LOADSYM(GetUniqueId, "ncclGetUniqueId");
    LOADSYM(CommInitRank, "ncclCommInitRank");
    LOADSYM(AllGather, "ncclAllGather");
    LOADSYM(CommDestroy, "ncclCommDestroy");
    LOADSYM(GetErrorString, "ncclGetErrorString");
#undef LOADSYM
    g_nccl.ok = true;
    return true;
}
thread_local std::string g_create_error;
}  // namespace

// ---------------------------------------------------------------------------------------------- handle
struct mpc_handle {
    mpc_config cfg;
    EnvParamsHost ep;
    Geometry g;
    int ns, nz;
    int num_sms;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evr0 = nullptr, evr1 = nullptr;
    bool profiling = false;
    bool rollout_events = false;  // mpc_set_profiling bit 2: an event pair around the solve kernel (mpc_get_rollout_ms)
    bool store_costs = true;      // mpc_set_profiling(h, 2) turns the per-sample cost store off (tools/trace_probe.py)
    bool prefer_stream = false;   // STREAM flavour whenever it gives more resident warps than whole tiles (see mpc_create)
    // device
    void *d_noise = nullptr;      // R [Bn][H][Kpad]          (MPPI tensor / CEM z of the current iteration)
    CUtensorMap noise_map;        // TMA descriptor over d_noise: box [H][wt] per warp tile (MPPI, TENSOR mode)
    bool has_noise_map = false;
    double *d_in = nullptr;       // x0 [B][ns] | g_z [B][H][nz] | U [B][H]
    double *d_S = nullptr;        // [B][Kpad]
    double *d_part = nullptr;     // [B][part_stride][2+H]: block (or warp) records, then the group records
    double *d_rank_part = nullptr;  // [world][B][2+H]
    int *d_done = nullptr;        // [B][1 + n_groups] block arrival tickets of the solve kernel
    // peer exchange (multi-GPU, in-kernel all-gather): this rank's buffer and every rank's mapped buffer
    double *d_xchg = nullptr;     // [2][world][B][2+H] doubles | u64 flags [2][world][B] (double-buffered by epoch parity)
    double *peer_recs[16] = {nullptr};
    bool peer_opened[16] = {false};  // mapped through cudaIpcOpenMemHandle (to be closed)
    bool p2p_ready = false;
    unsigned long long *d_trace = nullptr;  // phase stamps of the last solve's finalising block (when profiling)
    unsigned long long epoch = 0;
    double *d_out = nullptr;      // [B][out_stride] (+1)
    double *d_scratch = nullptr;  // staging for uploads: max(K_local*H, misc)
    size_t scratch_elems = 0;
    // CEM
    double *d_sd = nullptr, *d_R = nullptr;
    unsigned long long *d_keys = nullptr;
    int *d_elite = nullptr, *d_sel = nullptr;
    int cem_iter = 0;
    bool cem_open = false;
    int cem_n_cand = 0;           // candidates per rank and iteration (multi-GPU)
    // the whole solve as one cooperative launch (cem_fused.cuh): single GPU, Philox noise
    bool cem_fused = false;
    bool cem_coop = true;         // MPC_B200_CEM_COOP=0: a plain launch instead of cudaLaunchCooperativeKernel
    int cem_grid = 0, cem_spb = 0;
    double *d_cem_slots = nullptr;
    int *d_cem_stage = nullptr, *d_cem_hist = nullptr, *d_cem_tie = nullptr;
    unsigned *d_cem_bar = nullptr;
    int cem_z_stride = 0;         // row stride of the standard-normal tensor
    // host (pinned)
    double *h_in = nullptr, *h_out = nullptr;
    unsigned long long *h_flag = nullptr;  // pinned completion word the solve kernel raises (host-flag path)
    unsigned long long host_epoch = 0;
    bool host_flag_on = true;              // MPC_B200_HOSTFLAG=0 selects the D2H copy + stream synchronise instead
    bool inline_inputs = true;             // MPC_B200_INLINE=0: always stage state / plan through the pinned buffer
    size_t in_elems = 0, out_elems = 0, out_stride = 0;
    float *d_flush = nullptr;
    double *d_episode = nullptr;  // histories of mpc_episode_run
    size_t ep_elems = 0;
    size_t flush_elems = 0;
    // state
    bool noise_ready = false;
    bool g_zero_on_device = false;
    uint64_t solves = 0;
    int64_t launches = 0;
    int last_iters = 0;
    ncclComm_t comm = nullptr;
    std::vector<cudaEvent_t> ev_pool;  // mpc_time_solves: one pair per timed solve
    std::string err;
};

static int fail(mpc_handle *h, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h)
        h->err = buf;
    else
        g_create_error = buf;
    return code;
}

#define CU(h, call)                                                                                       \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return fail(h, MPC_ECUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                           __FILE__, __LINE__);                                           \
    } while (0)

// Fold the reference models' parameters into the constants envs.cuh consumes — in float64 and in the reference's
// own association order, so the float64 path differs from numpy only by FMA contraction and libm ulps.
static int fold_env_params(int env, const double *p, EnvParamsHost *out) {
    memset(out, 0, sizeof *out);
    out->env = env;
    double *c = out->c, *f = out->f;
    bool neg = false;
    if (env == MPC_ENV_PENDULUM) {  // pendulum_model.py:14-21,28-38
        const double g = p[0], m = p[1], l = p[2];
        c[0] = -3 * g / (2 * l);
        c[1] = 3. / (m * (l * l));
        c[2] = p[3];
        c[3] = p[4];
        c[4] = p[5];
        c[5] = p[6]; c[6] = p[7]; c[7] = p[8];
        f[0] = -c[0] * p[3];
        f[1] = c[1] * p[3];
        f[2] = p[3];
        f[3] = p[4];
        f[4] = p[5];
        for (int i = 0; i < 3; ++i) {
            if (p[6 + i] < 0) neg = true;
            f[5 + i] = sqrt(fabs(p[6 + i]));
        }
    } else if (env == MPC_ENV_CARTPOLE) {  // cartpole_swingup_model.py:25-33,63-85
        const double g = p[0], m_c = p[1], m_p = p[2], l = p[3], b = p[6];
        const double total_m = m_p + m_c, m_p_l = m_p * l;
        c[0] = 1.0 / p[7];
        c[1] = p[4];
        c[2] = p[5];
        c[3] = b;
        c[4] = -2 * m_p_l;
        c[5] = 3 * m_p * g;
        c[6] = 4 * b;
        c[7] = 4 * total_m;
        c[8] = 3 * m_p;
        c[9] = -3 * m_p_l;
        c[10] = 6 * total_m * g;
        c[11] = 4 * l * total_m;
        c[12] = 3 * m_p_l;
        for (int i = 0; i < 5; ++i) c[13 + i] = -p[8 + i];
        c[18] = p[7];
        f[0] = 1.0 / p[7];
        f[1] = 4 * p[4];
        f[2] = p[5];
        f[3] = 6 * p[4] / l;
        f[4] = c[4]; f[5] = c[5]; f[6] = c[6]; f[7] = c[7]; f[8] = c[8];
        f[9] = c[9] / l;
        f[10] = c[10] / l;
        f[11] = 6 * b / l;
        f[12] = 1.0;
        for (int i = 0; i < 5; ++i) {
            if (-p[8 + i] < 0) neg = true;
            f[13 + i] = sqrt(fabs(p[8 + i]));
        }
    } else if (env == MPC_ENV_ACROBOT) {  // acrobot_model.py:100-132
        const double m1 = p[1], m2 = p[2], l1 = p[3], lc1 = p[4], lc2 = p[5], I1 = p[6], I2 = p[7], g = p[8];
        c[0] = p[0];
        c[1] = m1 * (lc1 * lc1) + m2 * ((l1 * l1) + (lc2 * lc2)) + I1 + I2;
        c[2] = 2 * m2 * l1 * lc2;
        c[3] = m2 * (lc2 * lc2) + I2;
        c[4] = m2 * l1 * lc2;
        c[5] = (m1 * lc1 + m2 * l1) * g;
        c[6] = m2 * lc2 * g;
        c[7] = p[9];
        c[8] = p[10];
        for (int i = 0; i < 5; ++i) c[9 + i] = p[11 + i];
        c[14] = m1 * (lc1 * lc1);
        c[15] = (l1 * l1) + (lc2 * lc2);
        c[16] = 2 * l1 * lc2;
        c[17] = m2;
        c[18] = I1;
        c[19] = I2;
        c[20] = lc2 * lc2;
        c[21] = l1 * lc2;
        c[22] = p[16] != 0.0 ? 0.0 : c[4];  // book_or_nips == "nips" drops the w1^2 sin(th2) term (acrobot_model.py:124-128)
        if (p[16] != 0.0) out->f32_unsupported = "the 'nips' Acrobot equations are implemented in float64 mode only";
        for (int i = 0; i < 9; ++i) f[i] = c[i];
        f[9] = p[0] / 2;
        f[10] = p[0] / 6;
        if (p[11] < 0 || p[15] < 0) neg = true;
        f[11] = sqrt(fabs(p[11]));
        f[12] = sqrt(fabs(p[15]));
        f[13] = p[12]; f[14] = p[13]; f[15] = p[14];
        f[16] = p[0] / (2 * M_PI);  // the float flavour integrates the angles in half-turns (envs.cuh)
        f[17] = p[0] / M_PI;
        f[18] = p[0] / (6 * M_PI);
    } else {
        return -1;
    }
    return neg ? -2 : 0;  // -2: a negative cost weight, which the folded float32 cost (sqrt(W_i) residuals) cannot take
}


static size_t off_x(const mpc_handle *) { return 0; }
static size_t off_gz(const mpc_handle *h) { return (size_t)h->g.B * h->ns; }
static size_t off_U(const mpc_handle *h) { return off_gz(h) + (size_t)h->g.B * h->g.H * h->nz; }

static MppiArgs mppi_args(const mpc_handle *h, int want_traj) {
    MppiArgs a;
    memset(&a, 0, sizeof a);
    a.g = h->g;
    a.neg_inv_lam = -(1 / h->cfg.lam);
    a.pert = h->cfg.lam * (1.0 / (h->cfg.sigma * h->cfg.sigma));
    a.sigma = h->cfg.sigma;
    a.lo = h->cfg.bounds_low;
    a.hi = h->cfg.bounds_high;
    a.seed = h->cfg.seed;
    a.draw = h->cfg.resample ? (uint32_t)(h->solves & 0x7fffffffu) : 0u;
    a.noise_batched = 0;
    a.want_traj = want_traj;
    a.goal_zero = h->g_zero_on_device ? 1 : 0;
    a.per_step_goal = h->cfg.per_step_goal ? 1 : 0;
    a.store_costs = h->store_costs ? 1 : 0;
    a.p2p = (h->cfg.world > 1 && h->p2p_ready) ? 1 : 0;
    a.record_only = (h->cfg.world > 1 && !a.p2p) ? 1 : 0;
    a.epoch = h->epoch;
    a.trace = h->profiling ? h->d_trace : nullptr;
    a.world = h->cfg.world;
    a.rank = h->cfg.rank;
    a.model_noise = h->cfg.model_noise_std > 0 ? h->cfg.model_noise_std : 0.0;
    a.noise_draw = (uint32_t)(h->solves & 0xffffffffu);
    return a;
}

static MppiBuffers mppi_bufs(const mpc_handle *h) {
    MppiBuffers b;
    b.noise = h->d_noise;
    b.noise_map = h->has_noise_map ? &h->noise_map : nullptr;
    b.x0 = h->d_in + off_x(h);
    b.g_z = h->d_in + off_gz(h);
    b.U = h->d_in + off_U(h);
    b.S = h->d_S;
    b.part = h->d_part;
    b.rank_part = h->d_rank_part;
    b.out = h->d_out;
    b.done_count = h->d_done;
    b.peer_recs = h->peer_recs;
    b.inl = nullptr;
    return b;
}

static CemArgs cem_args(const mpc_handle *h, int iter, int want_traj) {
    CemArgs a;
    memset(&a, 0, sizeof a);
    a.g = h->g;
    a.sigma = h->cfg.sigma;
    a.lo = h->cfg.bounds_low;
    a.hi = h->cfg.bounds_high;
    a.alpha = h->cfg.alpha;
    a.epsilon = h->cfg.epsilon;
    a.n_elite = h->cfg.n_elite;
    a.iter = iter;
    // fresh standard normals every solve, as np.random.normal draws them (cem.py:58-59): the solve counter is folded
    // into the Philox key unconditionally (`resample` only concerns MPPI, whose reference noise IS fixed, mppi.py:20-21)
    a.seed = h->cfg.seed + 0x9E3779B97F4A7C15ull * h->solves;
    a.want_traj = want_traj;
    a.world = h->cfg.world;
    a.rank = h->cfg.rank;
    a.n_cand = h->cem_n_cand;
    a.z_stride = h->cem_z_stride;
    a.z_off = h->cfg.world > 1 ? h->g.k_begin : 0;
    a.epoch = h->epoch;
    a.model_noise = h->cfg.model_noise_std > 0 ? h->cfg.model_noise_std : 0.0;
    a.noise_draw = (uint32_t)((h->solves * (uint64_t)h->cfg.max_iter + (uint64_t)iter) & 0xffffffffu);
    a.per_step_goal = h->cfg.per_step_goal ? 1 : 0;
    return a;
}

static CemBuffers cem_bufs(const mpc_handle *h) {
    CemBuffers b;
    b.z = h->d_noise;
    b.x0 = h->d_in + off_x(h);
    b.g_z = h->d_in + off_gz(h);
    b.mu = h->d_in + off_U(h);
    b.sd = h->d_sd;
    b.R = h->d_R;
    b.keys = h->d_keys;
    b.elite_idx = h->d_elite;
    b.sel = h->d_sel;
    b.out = h->d_out;
    b.peer_recs = h->peer_recs;
    b.err_word = h->d_out + (size_t)h->g.B * h->out_stride + 1;
    b.inl = nullptr;
    return b;
}

// TMA descriptor of the resident noise tensor.  cuTensorMapEncodeTiled lives in the driver library; it is resolved
// through the runtime (no link against libcuda).
static int encode_noise_map(mpc_handle *h) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            return fail(nullptr, MPC_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
        encode = (encode_fn)fn;
    }
    const Geometry &g = h->g;
    const bool f32 = h->cfg.dtype == MPC_F32;
    const cuuint64_t esz = f32 ? 4 : 8;
    const cuuint64_t gdim[3] = {(cuuint64_t)g.Kpad, (cuuint64_t)g.H, 1};
    const cuuint64_t gstride[2] = {(cuuint64_t)g.Kpad * esz, (cuuint64_t)g.Kpad * esz * g.H};
    const cuuint32_t box[3] = {(cuuint32_t)g.wt, (cuuint32_t)(g.stream_rows ? g.stream_rows : g.H), 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(&h->noise_map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3,
                        h->d_noise, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(nullptr, MPC_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    h->has_noise_map = true;
    return MPC_OK;
}

extern "C" {

int mpc_abi_version(void) { return MPC_B200_ABI_VERSION; }

const char *mpc_last_error(const mpc_handle *h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int mpc_destroy(mpc_handle *h) {
    if (!h) return MPC_OK;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->comm && g_nccl.ok) g_nccl.CommDestroy(h->comm);
    cudaFree(h->d_noise);
    cudaFree(h->d_in);
    cudaFree(h->d_S);
    cudaFree(h->d_part);
    cudaFree(h->d_rank_part);
    cudaFree(h->d_done);
    cudaFree(h->d_trace);
    for (int r = 0; r < 16; ++r)
        if (h->peer_opened[r] && h->peer_recs[r]) cudaIpcCloseMemHandle(h->peer_recs[r]);
    cudaFree(h->d_xchg);
    cudaFree(h->d_out);
    cudaFree(h->d_scratch);
    cudaFree(h->d_sd);
    cudaFree(h->d_R);
    cudaFree(h->d_keys);
    cudaFree(h->d_elite);
    cudaFree(h->d_sel);
    cudaFree(h->d_cem_slots);
    cudaFree(h->d_cem_stage);
    cudaFree(h->d_cem_hist);
    cudaFree(h->d_cem_tie);
    cudaFree(h->d_cem_bar);
    cudaFree(h->d_flush);
    cudaFree(h->d_episode);
    if (h->h_in) cudaFreeHost(h->h_in);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->h_flag) cudaFreeHost(h->h_flag);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->evr0) cudaEventDestroy(h->evr0);
    if (h->evr1) cudaEventDestroy(h->evr1);
    for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return MPC_OK;
}

int mpc_create(const mpc_config *cfg, mpc_handle **out) {
    if (!cfg || !out) return fail(nullptr, MPC_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->struct_size != sizeof(mpc_config))
        return fail(nullptr, MPC_EINVAL, "mpc_config size mismatch: caller %u, library %zu", cfg->struct_size,
                    sizeof(mpc_config));
    if (cfg->env < 0 || cfg->env > 2) return fail(nullptr, MPC_EINVAL, "unknown env %d", cfg->env);
    if (cfg->agent != MPC_AGENT_MPPI && cfg->agent != MPC_AGENT_CEM)
        return fail(nullptr, MPC_EINVAL, "unknown agent %d", cfg->agent);
    if (cfg->dtype != MPC_F64 && cfg->dtype != MPC_F32) return fail(nullptr, MPC_EINVAL, "unknown dtype %d", cfg->dtype);
    if (cfg->noise != MPC_NOISE_TENSOR && cfg->noise != MPC_NOISE_PHILOX)
        return fail(nullptr, MPC_EINVAL, "unknown noise mode %d", cfg->noise);
    if (cfg->K < 1 || cfg->H < 1 || cfg->batch < 1) return fail(nullptr, MPC_EINVAL, "K, H, batch must be >= 1");
    if (cfg->lanes < 0 || cfg->lanes > 2) return fail(nullptr, MPC_EINVAL, "lanes must be 0 (default), 1 or 2");
    if (cfg->world < 1 || cfg->rank < 0 || cfg->rank >= cfg->world)
        return fail(nullptr, MPC_EINVAL, "bad rank/world %d/%d", cfg->rank, cfg->world);
    if (cfg->K < cfg->world) return fail(nullptr, MPC_EINVAL, "K = %d smaller than world = %d", cfg->K, cfg->world);
    if (!(cfg->sigma > 0)) return fail(nullptr, MPC_EINVAL, "std must be > 0");
    if (!(cfg->model_noise_std >= 0)) return fail(nullptr, MPC_EINVAL, "model_noise_std must be >= 0");
    if (cfg->agent == MPC_AGENT_MPPI && !(cfg->lam > 0)) return fail(nullptr, MPC_EINVAL, "lam must be > 0");
    if (cfg->agent == MPC_AGENT_CEM) {
        if (cfg->batch != 1) return fail(nullptr, MPC_EINVAL, "CEM supports batch = 1");
        if (cfg->world > 16) return fail(nullptr, MPC_EINVAL, "CEM shards over at most 16 GPUs");
        if (cfg->n_elite < 1 || cfg->n_elite > cfg->K)
            return fail(nullptr, MPC_EINVAL, "n_elite = %d must be in [1, K]", cfg->n_elite);
        if (cfg->max_iter < 1) return fail(nullptr, MPC_EINVAL, "max_iter must be >= 1");
    }
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(nullptr, MPC_ECUDA, "no CUDA device (%s); this library has no CPU path",
                    ce == cudaSuccess ? "device count 0" : cudaGetErrorString(ce));
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, MPC_EINVAL, "device %d of %d", cfg->device, ndev);

    mpc_handle *h = new (std::nothrow) mpc_handle();
    if (!h) return fail(nullptr, MPC_ENOMEM, "out of host memory");
    h->cfg = *cfg;
    const int fold_rc = fold_env_params(cfg->env, cfg->env_params, &h->ep);
    if (fold_rc == -1) {
        delete h;
        return fail(nullptr, MPC_EINVAL, "bad env");
    }
    // generic cost: the symmetrised upper triangle of the effective cost matrix (sign: CartPole's _state_cost returns
    // -costs, cartpole_swingup_model.py:104)
    h->ep.generic = (cfg->full_W || cfg->per_step_goal) ? 1 : 0;
    if (h->ep.generic) {
        const int nz = (cfg->env == MPC_ENV_PENDULUM) ? 3 : 5;
        const double sgn = cfg->env == MPC_ENV_CARTPOLE ? -1.0 : 1.0;
        const int d0 = cfg->env == MPC_ENV_PENDULUM ? 6 : (cfg->env == MPC_ENV_CARTPOLE ? 8 : 11);  // W_i in env_params
        int idx = 0;
        for (int i = 0; i < nz; ++i)
            for (int j = i; j < nz; ++j) {
                double v;
                if (cfg->full_W)
                    v = (i == j) ? cfg->W[i * nz + i] : cfg->W[i * nz + j] + cfg->W[j * nz + i];
                else
                    v = (i == j) ? cfg->env_params[d0 + i] : 0.0;
                h->ep.w[idx++] = sgn * v;
            }
        if (cfg->model_noise_std > 0) {
            delete h;
            return fail(nullptr, MPC_EINVAL, "the generic-cost flavour (full_W / per_step_goal) does not combine with model noise");
        }
        h->ep.f32_unsupported = nullptr;  // nothing is folded through sqrt(W) here
    }
    if (cfg->dtype == MPC_F32 && !h->ep.generic && fold_rc == -2) {
        delete h;
        return fail(nullptr, MPC_EINVAL, "the float32 flavour folds sqrt(W_i) into the cost residuals and needs non-negative "
                    "cost weights; use MPC_F64 or the generic cost (full_W)");
    }
    if (cfg->dtype == MPC_F32 && h->ep.f32_unsupported) {
        const char *why = h->ep.f32_unsupported;
        delete h;
        return fail(nullptr, MPC_EINVAL, "%s", why);
    }
    h->ns = (cfg->env == MPC_ENV_PENDULUM) ? 2 : 4;
    h->nz = h->ns + 1;

#define CUC(call)                                                                                    \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            int rc_ = fail(nullptr, e_ == cudaErrorMemoryAllocation ? MPC_ENOMEM : MPC_ECUDA,        \
                           "%s: %s", #call, cudaGetErrorString(e_));                                 \
            mpc_destroy(h);                                                                          \
            return rc_;                                                                              \
        }                                                                                            \
    } while (0)

    CUC(cudaSetDevice(cfg->device));
    CUC(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, cfg->device));
    CUC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CUC(cudaEventCreate(&h->ev0));
    CUC(cudaEventCreate(&h->ev1));
    CUC(cudaEventCreate(&h->evr0));
    CUC(cudaEventCreate(&h->evr1));

    // ---- geometry: contiguous K shard, padded to whole tiles, tiles dealt to a one-wave grid ----------------
    Geometry &g = h->g;
    const long long K = cfg->K;
    g.k_begin = (int)(K * cfg->rank / cfg->world);
    g.K_local = (int)(K * (cfg->rank + 1) / cfg->world) - g.k_begin;
    g.Kpad = (g.K_local + TILE - 1) / TILE * TILE;
    g.H = cfg->H;
    g.B = cfg->batch;
    // rollout flavour: float32 rolls out two samples per thread (packed f32x2) unless asked otherwise
    const int lanes = (cfg->dtype == MPC_F32 && cfg->agent == MPC_AGENT_MPPI && !(cfg->model_noise_std > 0) && !h->ep.generic)
                          ? (cfg->lanes ? cfg->lanes : 2) : 1;  // noisy-model and generic-cost flavours: one sample per thread
    // fast_math: 0 polynomial trig, 1 MUFU; Acrobot also 2, 4, 5 = per-stage splits (envs.cuh, STREAM flavour only)
    int trig = cfg->dtype == MPC_F32 ? cfg->fast_math : 0;
    if (trig < 0 || trig > 5 || trig == 3) trig = trig ? 1 : 0;
    if (cfg->env != MPC_ENV_ACROBOT && trig > 1) trig = 1;
    if (h->ep.generic) trig = 0;
    g.variant = trig | (lanes == 2 ? 8 : 0) | (h->ep.generic ? 32 : 0);
    g.wt = 32 * lanes;
    g.tiles_per_solve = g.Kpad / g.wt;
    g.total_tiles = g.B * g.tiles_per_solve;
    g.ns = h->ns;
    if (cfg->agent == MPC_AGENT_MPPI) {
        // ---- solve-kernel shape (mppi.cuh): W warps, ring of S stages, one block per SM -------------------------
        if (cfg->H > MPPI_MAX_H) {
            int rc = fail(nullptr, MPC_EINVAL, "H = %d exceeds the %d steps the weighted-sum registers hold", cfg->H,
                          MPPI_MAX_H);
            mpc_destroy(h);
            return rc;
        }
        int max_smem = 0;
        CUC(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device));
        max_smem -= 640;  // the solve kernel's static shared memory (reduction scratch, profiling stamps: 532 B) shares the limit
        const int uc = g.B < h->num_sms ? 1 : 0;  // fewer solves than SMs: a block works on ONE solve, one copy of its plan
        const int nblk_cap = h->num_sms > 32 ? h->num_sms : 32;  // records the last block may have to merge
        // does a block of W warps with a ring of S stages of `rows` time steps fit in shared memory?
        auto fits = [&](int W, int S, int rows) {
            return mppi_solve_smem(cfg->dtype, g.variant, cfg->H, rows, S, W, W > nblk_cap ? W : nblk_cap, 1, h->nz, uc) <=
                   (size_t)max_smem;
        };
        // whole [H][wt] tiles: the most stages that fit next to 16 warps' worth of records (more stages than warps
        // = tiles prefetched ahead)
        int s_cap = 0;
        while (s_cap < 40 && fits(s_cap + 1 < 16 ? s_cap + 1 : 16, s_cap + 1, cfg->H)) ++s_cap;
        if (s_cap < 1) {
            int rc = fail(nullptr, MPC_EINVAL, "H = %d needs %zu B of shared memory per warp tile (limit %d)", cfg->H,
                          mppi_stage_bytes(cfg->dtype, g.variant, cfg->H), max_smem);
            mpc_destroy(h);
            return rc;
        }
        int W, S, maxW = 16;
        if (cfg->noise == MPC_NOISE_PHILOX) {
            W = S = s_cap < maxW ? s_cap : maxW;  // one private buffer per warp, nothing to prefetch
        } else {
            // WIDE (MPC_B200_WIDE=1, one sample per thread only): the 32-warp whole-tile flavour at 64 registers
            if (const char *e = getenv("MPC_B200_WIDE"))
                if (atoi(e) != 0 && lanes == 1 && cfg->dtype == MPC_F32 && !(cfg->model_noise_std > 0) && !h->ep.generic)
                    g.variant |= 16;
            maxW = mppi_max_warps(cfg->dtype, NOISE_TENSOR, g.variant);
            if (maxW > 16) {  // recount the stages next to 32 warps' records
                s_cap = 0;
                while (s_cap < 40 && fits(s_cap + 1 < maxW ? s_cap + 1 : maxW, s_cap + 1, cfg->H)) ++s_cap;
            }
            // measured (tools/tune_probe.py, profiles/): more resident warps beat deeper prefetch, and warp counts that
            // fill the four SM sub-partitions evenly (multiples of 4) beat the ones in between
            S = s_cap;
            W = S < maxW ? S : maxW;
            if (W >= 8) W -= W % 4;
        }
        // STREAM (mppi.cuh): the rollout reads the noise through small per-warp chunk rings and the weighted sum
        // re-reads the rows from L2, so the resident warps are bounded by registers (24 packed / 32 scalar), not by
        // H * wt * 4 bytes per warp.  Default: whenever whole tiles leave room for fewer than 8 warps (long horizons);
        // for Acrobot (a long dependent chain per step: the XU and FMA pipes only fill up with 6-8 warps per scheduler,
        // profiles/r2*) whenever it gives more warps than whole tiles do.  MPC_B200_STREAM=0|1 overrides (profiling).
        g.stream_rows = g.stream_depth = 0;
        if (cfg->noise == MPC_NOISE_TENSOR) {
            const int stream_maxW = mppi_max_warps(cfg->dtype, NOISE_STREAM, g.variant);
            bool stream = s_cap < 8 || (h->prefer_stream && stream_maxW > W);
            if (const char *e = getenv("MPC_B200_STREAM")) stream = atoi(e) != 0;
            if (cfg->model_noise_std > 0) stream = false;  // the noisy-model flavour exists for whole tiles only
            if (h->ep.generic) stream = false;  // the generic-cost flavour exists for whole tiles only (fewer warps at long H)
            if (stream) {
                g.variant &= ~16;
                g.stream_rows = 8;
                g.stream_depth = 4;
                if (const char *e = getenv("MPC_B200_STREAM_ROWS")) { int v = atoi(e); if (v >= 4 && v <= 64 && v % 4 == 0) g.stream_rows = v; }
                if (const char *e = getenv("MPC_B200_STREAM_DEPTH")) { int v = atoi(e); if (v >= 2 && v <= 16) g.stream_depth = v; }
                maxW = stream_maxW;
                W = maxW;
                while (W > 1 && !fits(W, W * g.stream_depth, g.stream_rows)) --W;
                if (W >= 8) W -= W % 4;
                s_cap = maxW;
            }
        }
        // latency regime: when the solve is at most one tile per warp in one wave of blocks, fewer warps per block
        // spread it over more SMs, so each warp's dependent chain shares an issue port with fewer others — the
        // smallest multiple of 4 warps (one per SM sub-partition) that still fits one wave.  Measured
        // (profiles/r1j_warps_latency.txt): K=4096 p50 45 -> 39 µs (Pendulum), 53 -> 42 (CartPole); K=65536 48 -> 44.
        if (g.B < h->num_sms) {
            const int wave = h->num_sms / g.B;
            for (int w = 4; w < W; w += 4)
                if ((g.tiles_per_solve + w - 1) / w <= wave) { W = w; break; }
        }
        // tuning knobs (profiling only): MPC_B200_WARPS, MPC_B200_STAGES
        if (const char *e = getenv("MPC_B200_WARPS")) { int v = atoi(e); if (v >= 1 && v <= maxW && v <= s_cap) W = v; }
        if (const char *e = getenv("MPC_B200_STAGES")) { int v = atoi(e); if (v >= W && v <= s_cap) S = v; }
        if (cfg->noise == MPC_NOISE_PHILOX || S < W) S = W;
        const int tps = g.tiles_per_solve;
        if (g.B >= h->num_sms && tps <= 4 * W) {  // mode B: many small solves, a block owns whole solves
            g.solves_per_block = (g.B + h->num_sms - 1) / h->num_sms;
            g.grid = (g.B + g.solves_per_block - 1) / g.solves_per_block;
            g.blocks_per_solve = 1;
            g.chunk = tps;
            g.max_parts = W;
            g.n_groups = 0;
            const long long per = (long long)cfg->H * (1 + h->nz) * 8;
            long long G = (96 * 1024) / per;
            if (G > g.solves_per_block) G = g.solves_per_block;
            if (G > W * 32) G = W * 32;
            g.fin_group = G < 1 ? 1 : (int)G;
        } else {  // mode A: blocks_per_solve blocks share a solve
            int nb = g.B >= h->num_sms ? 1 : h->num_sms / g.B;
            const int want = (tps + W - 1) / W;
            if (nb > want) nb = want;
            if (nb < 1) nb = 1;
            g.chunk = (tps + nb - 1) / nb;
            nb = (tps + g.chunk - 1) / g.chunk;
            if (g.chunk < W) { W = g.chunk; if (cfg->noise == MPC_NOISE_PHILOX) S = W; }
            g.solves_per_block = 0;
            g.blocks_per_solve = nb;
            g.grid = g.B * nb;
            g.max_parts = nb;
            g.n_groups = nb > MERGE_GROUP ? (nb + MERGE_GROUP - 1) / MERGE_GROUP : 0;  // two-level merge (mppi.cuh tail)
            if (const char *e = getenv("MPC_B200_MERGE_FLAT"))  // profiling: one ticket, the last block merges all records
                if (atoi(e) != 0) g.n_groups = 0;
            g.fin_group = 1;
        }
        if (g.stream_rows) S = W * g.stream_depth;
        g.part_stride = g.max_parts + g.n_groups;
        g.warps = W;
        g.stages = S;
        const int cap = g.max_parts > W ? g.max_parts : W;
        g.u_copies = uc ? 1 : W;
        g.smem_rollout = mppi_solve_smem(cfg->dtype, g.variant, cfg->H, g.stream_rows ? g.stream_rows : cfg->H, S, W, cap,
                                         g.fin_group, h->nz, uc);
        if (g.smem_rollout > (size_t)max_smem) {
            int rc = fail(nullptr, MPC_EINVAL, "solve kernel needs %zu B of shared memory (limit %d)", g.smem_rollout,
                          max_smem);
            mpc_destroy(h);
            return rc;
        }
    }

    // ---- buffers ------------------------------------------------------------------------------------------
    const size_t esz = cfg->dtype == MPC_F32 ? 4 : 8;
    const size_t rec = PART_HDR + g.H;
    h->out_stride = 2 + g.H + (size_t)(g.H + 1) * h->ns;
    h->in_elems = off_U(h) + (size_t)g.B * g.H;
    h->out_elems = (size_t)g.B * h->out_stride + 2;  // + CEM iterations / MPPI peer error, + CEM peer error
    CUC(cudaMalloc(&h->d_in, h->in_elems * sizeof(double)));
    CUC(cudaMemsetAsync(h->d_in, 0, h->in_elems * sizeof(double), h->stream));  // planned_us = zeros (mppi.py:14)
    CUC(cudaMalloc(&h->d_out, h->out_elems * sizeof(double)));
    CUC(cudaMemsetAsync(h->d_out, 0, h->out_elems * sizeof(double), h->stream));
    CUC(cudaMallocHost(&h->h_in, h->in_elems * sizeof(double)));
    CUC(cudaMallocHost(&h->h_out, h->out_elems * sizeof(double)));
    CUC(cudaMallocHost(&h->h_flag, 64));
    *h->h_flag = 0ull;
    if (const char *e = getenv("MPC_B200_HOSTFLAG")) h->host_flag_on = atoi(e) != 0;
    if (const char *e = getenv("MPC_B200_INLINE")) h->inline_inputs = atoi(e) != 0;
    memset(h->h_in, 0, h->in_elems * sizeof(double));
    h->g_zero_on_device = true;
    // CEM over several GPUs: elites are global, so the standard normals of an iteration (TENSOR noise) are resident in
    // full on every rank; candidates per rank and iteration = min(n_elite, largest shard)
    h->cem_z_stride = g.Kpad;
    if (cfg->agent == MPC_AGENT_CEM && cfg->world > 1) {
        h->cem_z_stride = (cfg->K + TILE - 1) / TILE * TILE;
        const int shard_max = (cfg->K + cfg->world - 1) / cfg->world;
        h->cem_n_cand = cfg->n_elite < shard_max ? cfg->n_elite : shard_max;
    }
    if (cfg->noise == MPC_NOISE_TENSOR) {
        const size_t cols = cfg->agent == MPC_AGENT_CEM ? (size_t)h->cem_z_stride : (size_t)g.Kpad;
        CUC(cudaMalloc(&h->d_noise, (size_t)g.H * cols * esz));
        CUC(cudaMemsetAsync(h->d_noise, 0, (size_t)g.H * cols * esz, h->stream));
        if (cfg->agent == MPC_AGENT_MPPI) {
            const int rc = encode_noise_map(h);
            if (rc) {
                mpc_destroy(h);
                return rc;
            }
        }
    }
    if (cfg->agent == MPC_AGENT_MPPI) {
        CUC(cudaMalloc(&h->d_S, (size_t)g.B * g.Kpad * sizeof(double)));
        CUC(cudaMalloc(&h->d_part, (size_t)g.B * g.part_stride * rec * sizeof(double)));
        CUC(cudaMalloc(&h->d_rank_part, (size_t)cfg->world * g.B * rec * sizeof(double)));
        CUC(cudaMalloc(&h->d_trace, 16 * sizeof(unsigned long long)));
        CUC(cudaMemsetAsync(h->d_trace, 0, 16 * sizeof(unsigned long long), h->stream));
        CUC(cudaMalloc(&h->d_done, (size_t)g.B * (1 + g.n_groups) * sizeof(int)));
        CUC(cudaMemsetAsync(h->d_done, 0, (size_t)g.B * (1 + g.n_groups) * sizeof(int), h->stream));
    } else {
        CUC(cudaMalloc(&h->d_sd, (size_t)g.H * sizeof(double)));
        CUC(cudaMalloc(&h->d_R, (size_t)cfg->max_iter * g.Kpad * sizeof(double)));
        CUC(cudaMalloc(&h->d_keys, (size_t)g.Kpad * sizeof(unsigned long long)));
        CUC(cudaMalloc(&h->d_elite, (size_t)cfg->max_iter * cfg->n_elite * sizeof(int)));
        CUC(cudaMalloc(&h->d_sel, 8 * sizeof(int)));
        CUC(cudaMemsetAsync(h->d_sel, 0, 8 * sizeof(int), h->stream));
        CUC(cudaMemsetAsync(h->d_R, 0, (size_t)cfg->max_iter * g.Kpad * sizeof(double), h->stream));
        CUC(cudaMemsetAsync(h->d_elite, 0, (size_t)cfg->max_iter * cfg->n_elite * sizeof(int), h->stream));
        // One persistent cooperative launch per solve (cem_fused.cuh) when the normals are generated in the kernel and
        // K is not sharded: blocks own contiguous sample ranges, at most one block per SM.  MPC_B200_CEM_FUSED=0 keeps
        // the per-iteration kernels (profiling / comparison).
        h->cem_fused = cfg->noise == MPC_NOISE_PHILOX && cfg->world == 1 && cfg->H <= CEM_FUSED_MAX_H;
        if (const char *e = getenv("MPC_B200_CEM_FUSED")) h->cem_fused = h->cem_fused && atoi(e) != 0;
        if (const char *e = getenv("MPC_B200_CEM_COOP")) h->cem_coop = atoi(e) != 0;
        if (h->cem_fused) {
            int coop = 0;
            CUC(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, cfg->device));
            if (!coop) h->cem_fused = false;
        }
        if (h->cem_fused) {
            int grid = (g.K_local + 31) / 32;
            if (grid > h->num_sms) grid = h->num_sms;
            int spb = ((g.K_local + grid - 1) / grid + 31) / 32 * 32;
            grid = (g.K_local + spb - 1) / spb;
            h->cem_grid = grid;
            h->cem_spb = spb;
            CUC(cudaMalloc(&h->d_cem_slots, (size_t)grid * (2 + 2 * g.H) * sizeof(double)));
            CUC(cudaMalloc(&h->d_cem_stage, (size_t)g.Kpad * sizeof(int)));
            CUC(cudaMalloc(&h->d_cem_hist, (size_t)cfg->max_iter * 8 * 256 * sizeof(int)));
            CUC(cudaMalloc(&h->d_cem_tie, (size_t)grid * sizeof(int)));
            CUC(cudaMalloc(&h->d_trace, 16 * sizeof(unsigned long long)));
            CUC(cudaMemsetAsync(h->d_trace, 0, 16 * sizeof(unsigned long long), h->stream));
            CUC(cudaMalloc(&h->d_cem_bar, 2 * sizeof(unsigned)));
            CUC(cudaMemsetAsync(h->d_cem_bar, 0, 2 * sizeof(unsigned), h->stream));
        }
    }
    CUC(cudaStreamSynchronize(h->stream));
#undef CUC
    h->noise_ready = (cfg->noise == MPC_NOISE_PHILOX);
    *out = h;
    return MPC_OK;
}

static int ensure_scratch(mpc_handle *h, size_t elems) {
    if (h->scratch_elems >= elems) return MPC_OK;
    if (h->d_scratch) cudaFree(h->d_scratch);
    h->d_scratch = nullptr;
    h->scratch_elems = 0;
    cudaError_t e = cudaMalloc(&h->d_scratch, elems * sizeof(double));
    if (e != cudaSuccess) return fail(h, MPC_ENOMEM, "cudaMalloc scratch of %zu doubles: %s", elems, cudaGetErrorString(e));
    h->scratch_elems = elems;
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- noise
int mpc_set_noise(mpc_handle *h, const double *delta_u_KH) {
    if (!h || !delta_u_KH) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.noise != MPC_NOISE_TENSOR) return fail(h, MPC_ESTATE, "handle is in PHILOX noise mode");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    const size_t n = (size_t)g.K_local * g.H;
    int rc = ensure_scratch(h, n);
    if (rc) return rc;
    CU(h, cudaMemcpyAsync(h->d_scratch, delta_u_KH + (size_t)g.k_begin * g.H, n * sizeof(double),
                          cudaMemcpyHostToDevice, h->stream));
    CU(h, launch_transpose_noise(h->cfg.dtype, h->d_scratch, h->d_noise, g.K_local, g.Kpad, g.H, h->stream));
    h->launches++;
    CU(h, cudaStreamSynchronize(h->stream));
    h->noise_ready = true;
    return MPC_OK;
}

int mpc_fill_noise(mpc_handle *h) {
    if (!h) return MPC_EINVAL;
    if (h->cfg.noise != MPC_NOISE_TENSOR) return fail(h, MPC_ESTATE, "handle is in PHILOX noise mode");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    CU(h, launch_philox_fill(h->cfg.dtype, h->d_noise, g.K_local, g.k_begin, g.Kpad, g.H, h->cfg.sigma, h->cfg.seed, 0u,
                             h->stream));
    h->launches++;
    CU(h, cudaStreamSynchronize(h->stream));
    h->noise_ready = true;
    return MPC_OK;
}

int mpc_philox_noise(mpc_handle *h, uint64_t draw, double *out) {
    if (!h || !out) return fail(h, MPC_EINVAL, "null argument");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    const size_t n = (size_t)g.K_local * g.H;
    int rc = ensure_scratch(h, n);
    if (rc) return rc;
    const uint32_t stream_bit = (h->cfg.agent == MPC_AGENT_CEM) ? 0x80000000u : 0u;
    uint64_t seed = h->cfg.seed;
    if (h->cfg.agent == MPC_AGENT_CEM) seed += 0x9E3779B97F4A7C15ull * (h->solves ? h->solves - 1 : 0);  // the last solve's
    CU(h, launch_philox_materialize(h->d_scratch, g.K_local, g.k_begin, g.H, seed,
                                    ((uint32_t)draw & 0x7fffffffu) | stream_bit, h->stream));
    h->launches++;
    CU(h, cudaMemcpyAsync(out, h->d_scratch, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_philox_model_noise(mpc_handle *h, uint64_t solve_index, double *out) {
    if (!h || !out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_MPPI) return fail(h, MPC_ESTATE, "not an MPPI handle");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    const size_t n = (size_t)g.K_local * g.H * 4;
    int rc = ensure_scratch(h, n);
    if (rc) return rc;
    CU(h, launch_philox_blocks(h->d_scratch, g.K_local, g.k_begin, g.H, h->cfg.seed, (uint32_t)(solve_index & 0xffffffffu),
                               4u /* PHILOX_MODEL_MPPI */, h->stream));
    h->launches++;
    CU(h, cudaMemcpyAsync(out, h->d_scratch, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- staging
// h_in layout = device layout.  g_z == NULL means a zero goal (what every reference script passes,
// Experiment.py:81-82): the device copy is zeroed once and not re-sent.
static int stage_inputs(mpc_handle *h, const double *x, const double *g_z, const double *U_in) {
    const Geometry &g = h->g;
    if (!x) return fail(h, MPC_EINVAL, "state pointer is null");
    memcpy(h->h_in + off_x(h), x, (size_t)g.B * h->ns * sizeof(double));
    size_t upto = off_gz(h);
    const size_t gz_elems = (size_t)g.B * g.H * h->nz;
    if (g_z) {
        memcpy(h->h_in + off_gz(h), g_z, gz_elems * sizeof(double));
        h->g_zero_on_device = false;
        upto = off_U(h);
    } else if (!h->g_zero_on_device) {
        memset(h->h_in + off_gz(h), 0, gz_elems * sizeof(double));
        h->g_zero_on_device = true;
        upto = off_U(h);
    }
    if (U_in) {
        memcpy(h->h_in + off_U(h), U_in, (size_t)g.B * g.H * sizeof(double));
        if (upto == off_U(h)) {
            upto = h->in_elems;
        } else {
            CU(h, cudaMemcpyAsync(h->d_in + off_U(h), h->h_in + off_U(h), (size_t)g.B * g.H * sizeof(double),
                                  cudaMemcpyHostToDevice, h->stream));
        }
    }
    CU(h, cudaMemcpyAsync(h->d_in, h->h_in, upto * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    return MPC_OK;
}

// Batch-1 host calls with the usual zero goal: state and plan travel as kernel parameters (InlineInputs) and no copy is
// enqueued at all.  Returns false when the call does not qualify (the caller stages through the pinned buffer then).
static bool stage_inline(mpc_handle *h, const double *x, const double *g_z, const double *U_in, InlineInputs *inl) {
    const Geometry &g = h->g;
    if (!h->inline_inputs || g.B != 1 || g_z || !h->g_zero_on_device || !x || g.H > INLINE_MAX_H || h->ns > 4) return false;
    inl->on = 1;
    for (int i = 0; i < h->ns; ++i) inl->x0[i] = x[i];
    if (U_in) {
        inl->on |= 2;
        memcpy(inl->U, U_in, (size_t)g.H * sizeof(double));
    }
    return true;
}

// A peer never delivered within the bounded wait: report MPC_ENCCL once and clear the device-side error words, so that
// the handle is usable again after the caller has dealt with it.  The resident plan U was updated from a partial
// merge by the timed-out kernel: the caller must pass U_in on its next solve (every reference-style caller does).
static int peer_timeout(mpc_handle *h, const char *what) {
    cudaMemsetAsync(h->d_out + (size_t)h->g.B * h->out_stride, 0, 2 * sizeof(double), h->stream);
    cudaStreamSynchronize(h->stream);
    h->cem_open = false;
    return fail(h, MPC_ENCCL, "peer exchange timed out: a rank did not deliver its %s within 4 s; the resident plan is "
                "undefined until the next call passes U_in", what);
}

static int fetch_outputs(mpc_handle *h, int want_traj, double *u0, double *U_out, double *xs_out, double *cost_out) {
    const Geometry &g = h->g;
    // the peer waits' error words live behind the last solve's outputs: slot 0 MPPI (CEM: iterations), slot 1 CEM
    const size_t err_at = (size_t)g.B * h->out_stride + (h->cfg.agent == MPC_AGENT_CEM ? 1 : 0);
    if (h->p2p_ready)
        CU(h, cudaMemcpyAsync(h->h_out + err_at, h->d_out + err_at, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (g.B == 1) {
        const size_t per = want_traj ? h->out_stride : (size_t)(1 + g.H);
        CU(h, cudaMemcpyAsync(h->h_out, h->d_out, (per + (want_traj ? 1 : 0)) * sizeof(double), cudaMemcpyDeviceToHost,
                              h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        if (h->p2p_ready && h->h_out[err_at] != 0.0) return peer_timeout(h, "record");
        const double *o = h->h_out;
        if (u0) u0[0] = o[0];
        if (U_out) memcpy(U_out, o + 1, g.H * sizeof(double));
        if (want_traj) {
            if (cost_out) cost_out[0] = o[1 + g.H];
            if (xs_out) memcpy(xs_out, o + 2 + g.H, (size_t)(g.H + 1) * h->ns * sizeof(double));
        }
        return MPC_OK;
    }
    // Batches: megabytes of outputs.  The per-solve records [u0 | U[H] | cost | xs] are de-interleaved BY THE COPY
    // ENGINE straight into the caller's arrays (one strided copy per field) — no staging buffer, no host-side scatter.
    // Pinned caller arrays (mpc_pin_host_buffer) make these true DMA transfers.
    const size_t sp = h->out_stride * sizeof(double);
    auto field = [&](double *dst, size_t off, size_t n) -> int {
        if (dst)
            CU(h, cudaMemcpy2DAsync(dst, n * sizeof(double), h->d_out + off, sp, n * sizeof(double), g.B,
                                    cudaMemcpyDeviceToHost, h->stream));
        return MPC_OK;
    };
    int rc;
    if ((rc = field(u0, 0, 1))) return rc;
    if ((rc = field(U_out, 1, g.H))) return rc;
    if (want_traj) {
        if ((rc = field(cost_out, 1 + g.H, 1))) return rc;
        if ((rc = field(xs_out, 2 + g.H, (size_t)(g.H + 1) * h->ns))) return rc;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->p2p_ready && h->h_out[err_at] != 0.0) return peer_timeout(h, "record");
    return MPC_OK;
}

// Host-flag completion of a batch-1 solve: poll the pinned word the finalising block raises.  The stream is queried
// now and then so that a faulted launch turns into an error instead of an endless spin.
static int await_host_flag(mpc_handle *h, int want_traj, double *u0, double *U_out, double *xs_out, double *cost_out) {
    volatile unsigned long long *flag = h->h_flag;
    const unsigned long long want = h->host_epoch;
    unsigned spins = 0;
    while (*flag != want) {
        if ((++spins & 0x3fffu) == 0) {
            cudaError_t q = cudaStreamQuery(h->stream);
            if (q == cudaSuccess) {
                if (*flag == want) break;
                return fail(h, MPC_ECUDA, "solve kernel finished without raising its completion flag");
            }
            if (q != cudaErrorNotReady) return fail(h, MPC_ECUDA, "solve kernel: %s", cudaGetErrorString(q));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    __atomic_thread_fence(__ATOMIC_ACQUIRE);
    const Geometry &g = h->g;
    const double *o = h->h_out;
    if (h->p2p_ready && o[h->out_stride] != 0.0) return peer_timeout(h, "record");
    if (u0) u0[0] = o[0];
    if (U_out) memcpy(U_out, o + 1, g.H * sizeof(double));
    if (want_traj) {
        if (cost_out) cost_out[0] = o[1 + g.H];
        if (xs_out) memcpy(xs_out, o + 2 + g.H, (size_t)(g.H + 1) * h->ns * sizeof(double));
    }
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- MPPI
static int mppi_enqueue(mpc_handle *h, int want_traj, bool host_flag = false, const InlineInputs *inl = nullptr) {
    if (h->cfg.agent != MPC_AGENT_MPPI) return fail(h, MPC_ESTATE, "not an MPPI handle");
    if (!h->noise_ready) return fail(h, MPC_ESTATE, "noise tensor not set: call mpc_set_noise or mpc_fill_noise first");
    if (h->cfg.world > 1 && !h->comm && !h->p2p_ready)
        return fail(h, MPC_ESTATE, "world > 1 but neither mpc_p2p_import/mpc_p2p_link nor mpc_comm_init was called");
    if (h->p2p_ready) h->epoch++;
    MppiArgs a = mppi_args(h, want_traj);
    if (host_flag) {
        a.host_out = h->h_out;  // pinned host memory is device-addressable at the same address (UVA)
        a.host_flag = h->h_flag;
        a.host_epoch = ++h->host_epoch;
    }
    MppiBuffers b = mppi_bufs(h);
    b.inl = inl;
    if (h->rollout_events) CU(h, cudaEventRecord(h->evr0, h->stream));
    CU(h, launch_mppi_solve(h->cfg.dtype, h->cfg.noise, h->g.variant, h->ep, a, b, h->stream));
    if (h->rollout_events) CU(h, cudaEventRecord(h->evr1, h->stream));
    h->launches++;
    if (h->cfg.world > 1 && !a.p2p) {
        const size_t cnt = (size_t)h->g.B * (PART_HDR + h->g.H);
        ncclResult_t r = g_nccl.AllGather(h->d_rank_part + (size_t)h->cfg.rank * cnt, h->d_rank_part, cnt, ncclFloat64,
                                          h->comm, h->stream);
        if (r != 0) return fail(h, MPC_ENCCL, "ncclAllGather: %s", g_nccl.GetErrorString(r));
        CU(h, launch_mppi_rank_merge(h->cfg.dtype, h->ep, a, b, h->stream));
        h->launches += 2;
    }
    h->solves++;
    return MPC_OK;
}

int mpc_mppi_solve(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *u0, double *U_out,
                   double *xs_out, double *cost_out) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    const int want_traj = (xs_out || cost_out) ? 1 : 0;
    InlineInputs inl;
    // (not on the NCCL transport: its second kernel applies the update to the plan in global memory)
    const bool inlined = h->g.solves_per_block == 0 && (h->cfg.world == 1 || h->p2p_ready) &&
                         stage_inline(h, x, g_z, U_in, &inl);
    int rc = inlined ? MPC_OK : stage_inputs(h, x, g_z, U_in);
    if (rc) return rc;
    // host-flag completion: batch = 1 and the tail that applies the update runs inside the solve kernel (one GPU,
    // or the in-kernel peer exchange) — the kernel writes the outputs into pinned host memory and raises a flag
    const bool use_flag = h->host_flag_on && h->g.B == 1 && (h->cfg.world == 1 || h->p2p_ready);
    rc = mppi_enqueue(h, want_traj, use_flag, inlined ? &inl : nullptr);
    if (rc) return rc;
    if (!use_flag) return fetch_outputs(h, want_traj, u0, U_out, xs_out, cost_out);
    return await_host_flag(h, want_traj, u0, U_out, xs_out, cost_out);
}

int mpc_mppi_rollout_partial(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *rec_out) {
    if (!h || !rec_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_MPPI) return fail(h, MPC_ESTATE, "not an MPPI handle");
    if (!h->noise_ready) return fail(h, MPC_ESTATE, "noise tensor not set: call mpc_set_noise or mpc_fill_noise first");
    CU(h, cudaSetDevice(h->cfg.device));
    int rc = stage_inputs(h, x, g_z, U_in);
    if (rc) return rc;
    MppiArgs a = mppi_args(h, 0);
    a.record_only = 1;
    const MppiBuffers b = mppi_bufs(h);
    CU(h, launch_mppi_solve(h->cfg.dtype, h->cfg.noise, h->g.variant, h->ep, a, b, h->stream));
    h->launches++;
    const size_t cnt = (size_t)h->g.B * (PART_HDR + h->g.H);
    CU(h, cudaMemcpyAsync(rec_out, h->d_rank_part + (size_t)h->cfg.rank * cnt, cnt * sizeof(double),
                          cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_mppi_apply_partials(mpc_handle *h, const double *recs, double *u0, double *U_out, double *xs_out,
                            double *cost_out) {
    if (!h || !recs) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_MPPI) return fail(h, MPC_ESTATE, "not an MPPI handle");
    CU(h, cudaSetDevice(h->cfg.device));
    const int want_traj = (xs_out || cost_out) ? 1 : 0;
    const size_t cnt = (size_t)h->cfg.world * h->g.B * (PART_HDR + h->g.H);
    CU(h, cudaMemcpyAsync(h->d_rank_part, recs, cnt * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CU(h, launch_mppi_rank_merge(h->cfg.dtype, h->ep, mppi_args(h, want_traj), mppi_bufs(h), h->stream));
    h->launches++;
    h->solves++;
    return fetch_outputs(h, want_traj, u0, U_out, xs_out, cost_out);
}

int mpc_mppi_get_costs(mpc_handle *h, double *S_out) {
    if (!h || !S_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_MPPI) return fail(h, MPC_ESTATE, "not an MPPI handle");
    if (!h->store_costs) return fail(h, MPC_ESTATE, "the per-sample cost store is switched off (mpc_set_profiling bit 1)");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    CU(h, cudaMemcpy2DAsync(S_out, (size_t)g.K_local * sizeof(double), h->d_S, (size_t)g.Kpad * sizeof(double),
                            (size_t)g.K_local * sizeof(double), g.B, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- CEM
static int cem_enqueue_iter(mpc_handle *h, int iter) {
    if (h->cfg.world > 1) {
        if (!h->p2p_ready)
            return fail(h, MPC_ESTATE, "world > 1: connect the ranks' exchange buffers first (mpc_p2p_import / mpc_p2p_link)");
        h->epoch++;
    }
    const CemArgs a = cem_args(h, iter, 0);
    const CemBuffers b = cem_bufs(h);
    CU(h, launch_cem_rollout(h->cfg.dtype, h->cfg.noise, h->g.variant, h->ep, a, b, h->stream));
    CU(h, launch_cem_select(a, b, h->stream));
    CU(h, launch_cem_refit(h->cfg.dtype, h->cfg.noise, a, b, h->stream));
    h->launches += 3;
    return MPC_OK;
}

// The whole solve as one cooperative launch (cem_fused.cuh).  host_flag: the kernel also writes the outputs into the
// pinned host buffer and raises the completion word (batch-1 host call: no D2H copy, no stream synchronise).
static int cem_fused_enqueue(mpc_handle *h, int want_traj, bool host_flag, const InlineInputs *inl = nullptr) {
    CemFusedArgs fa;
    memset(&fa, 0, sizeof fa);
    fa.a = cem_args(h, 0, want_traj);
    fa.max_iter = h->cfg.max_iter;
    fa.spb = h->cem_spb;
    fa.all_pairs = (h->g.K_local <= CEM_FUSED_ALLPAIRS_K && h->cem_spb == 32) ? 1 : 0;
    fa.cached = (h->cem_spb == 32 && (size_t)h->g.H * 32 * (h->cfg.dtype == MPC_F32 ? 4 : 8) <= 64 * 1024) ? 1 : 0;
    fa.noise_draw0 = (uint32_t)((h->solves * (uint64_t)h->cfg.max_iter) & 0xffffffffu);
    fa.trace = h->profiling ? h->d_trace : nullptr;
    if (host_flag) {
        fa.host_out = h->h_out;
        fa.host_flag = h->h_flag;
        fa.host_epoch = ++h->host_epoch;
    }
    CemFusedScratch sc;
    sc.slots = h->d_cem_slots;
    sc.stage = h->d_cem_stage;
    sc.hist = h->d_cem_hist;
    sc.tie_cnt = h->d_cem_tie;
    sc.bar = h->d_cem_bar;
    CemBuffers cb = cem_bufs(h);
    cb.inl = inl;
    CU(h, launch_cem_fused(h->cfg.dtype, h->g.variant, h->ep, fa, cb, sc, h->cem_grid, h->cem_coop ? 1 : 0, h->stream));
    h->launches++;
    h->solves++;
    h->cem_open = false;
    return MPC_OK;
}

static int cem_await_host_flag(mpc_handle *h, int want_traj, double *u0, double *U_out, double *xs_out, double *cost_out,
                               int32_t *iters_out) {
    volatile unsigned long long *flag = h->h_flag;
    const unsigned long long want = h->host_epoch;
    unsigned spins = 0;
    while (*flag != want) {
        if ((++spins & 0x3fffu) == 0) {
            cudaError_t q = cudaStreamQuery(h->stream);
            if (q == cudaSuccess) {
                if (*flag == want) break;
                return fail(h, MPC_ECUDA, "CEM solve kernel finished without raising its completion flag");
            }
            if (q != cudaErrorNotReady) return fail(h, MPC_ECUDA, "CEM solve kernel: %s", cudaGetErrorString(q));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    __atomic_thread_fence(__ATOMIC_ACQUIRE);
    const Geometry &g = h->g;
    const double *o = h->h_out;
    if (u0) u0[0] = o[0];
    if (U_out) memcpy(U_out, o + 1, g.H * sizeof(double));
    if (want_traj) {
        if (cost_out) cost_out[0] = o[1 + g.H];
        if (xs_out) memcpy(xs_out, o + 2 + g.H, (size_t)(g.H + 1) * h->ns * sizeof(double));
    }
    h->last_iters = (int)o[h->out_stride];
    if (iters_out) *iters_out = h->last_iters;
    return MPC_OK;
}

int mpc_cem_begin(mpc_handle *h, const double *x, const double *g_z, const double *U_in) {
    if (!h) return MPC_EINVAL;
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    CU(h, cudaSetDevice(h->cfg.device));
    int rc = stage_inputs(h, x, g_z, U_in);
    if (rc) return rc;
    CU(h, launch_cem_reset(cem_args(h, 0, 0), cem_bufs(h), h->stream));
    h->launches++;
    h->cem_iter = 0;
    h->cem_open = true;
    return MPC_OK;
}

int mpc_cem_iterate(mpc_handle *h, const double *z_KH, int32_t *done) {
    if (!h) return MPC_EINVAL;
    if (!h->cem_open) return fail(h, MPC_ESTATE, "mpc_cem_iterate without mpc_cem_begin");
    if (h->cem_iter >= h->cfg.max_iter) return fail(h, MPC_ESTATE, "more than max_iter = %d iterations", h->cfg.max_iter);
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    if (h->cfg.noise == MPC_NOISE_TENSOR) {
        if (!z_KH) return fail(h, MPC_EINVAL, "TENSOR noise mode needs the standard normals of this iteration");
        // one GPU: this handle's shard; several GPUs: all K samples (the elite refit reads any sample's normals)
        const int cols = h->cfg.world > 1 ? h->cfg.K : g.K_local;
        const size_t n = (size_t)cols * g.H;
        int rc = ensure_scratch(h, n);
        if (rc) return rc;
        CU(h, cudaMemcpyAsync(h->d_scratch, z_KH + (h->cfg.world > 1 ? 0 : (size_t)g.k_begin * g.H), n * sizeof(double),
                              cudaMemcpyHostToDevice, h->stream));
        CU(h, launch_transpose_noise(h->cfg.dtype, h->d_scratch, h->d_noise, cols, h->cem_z_stride, g.H, h->stream));
        h->launches++;
    }
    int rc = cem_enqueue_iter(h, h->cem_iter);
    if (rc) return rc;
    h->cem_iter++;
    if (!done) return MPC_OK;  // enqueue only: the caller synchronises later (ranks driven by one thread)
    int sel[2] = {0, 0};
    CU(h, cudaMemcpyAsync(sel, h->d_sel, sizeof sel, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *done = sel[0];
    return MPC_OK;
}

static int cem_finish(mpc_handle *h, int want_traj, double *u0, double *U_out, double *xs_out, double *cost_out,
                      int32_t *iters_out) {
    CU(h, launch_cem_finish(h->cfg.dtype, h->ep, cem_args(h, 0, want_traj), cem_bufs(h), h->stream));
    h->launches++;
    CU(h, cudaMemcpyAsync(h->h_out, h->d_out, h->out_elems * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    const Geometry &g = h->g;
    const double *o = h->h_out;
    if (u0) u0[0] = o[0];
    if (U_out) memcpy(U_out, o + 1, g.H * sizeof(double));
    if (want_traj) {
        if (cost_out) cost_out[0] = o[1 + g.H];
        if (xs_out) memcpy(xs_out, o + 2 + g.H, (size_t)(g.H + 1) * h->ns * sizeof(double));
    }
    h->last_iters = (int)o[h->out_stride];
    if (iters_out) *iters_out = h->last_iters;
    if (o[h->out_stride + 1] != 0.0) return peer_timeout(h, "elite candidates");
    h->cem_open = false;
    h->solves++;
    return MPC_OK;
}

int mpc_cem_end(mpc_handle *h, double *u0, double *U_out, double *xs_out, double *cost_out) {
    if (!h) return MPC_EINVAL;
    if (!h->cem_open) return fail(h, MPC_ESTATE, "mpc_cem_end without mpc_cem_begin");
    CU(h, cudaSetDevice(h->cfg.device));
    return cem_finish(h, (xs_out || cost_out) ? 1 : 0, u0, U_out, xs_out, cost_out, nullptr);
}

int mpc_cem_solve(mpc_handle *h, const double *x, const double *g_z, const double *U_in, double *u0, double *U_out,
                  double *xs_out, double *cost_out, int32_t *iters_out) {
    if (!h) return MPC_EINVAL;
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    if (h->cfg.noise != MPC_NOISE_PHILOX)
        return fail(h, MPC_ESTATE, "mpc_cem_solve needs PHILOX noise; use begin/iterate/end with host noise");
    if (h->cem_fused) {
        CU(h, cudaSetDevice(h->cfg.device));
        const int want_traj = (xs_out || cost_out) ? 1 : 0;
        InlineInputs inl;
        const bool inlined = stage_inline(h, x, g_z, U_in, &inl);
        int rc = inlined ? MPC_OK : stage_inputs(h, x, g_z, U_in);
        if (rc) return rc;
        rc = cem_fused_enqueue(h, want_traj, h->host_flag_on, inlined ? &inl : nullptr);
        if (rc) return rc;
        if (h->host_flag_on) return cem_await_host_flag(h, want_traj, u0, U_out, xs_out, cost_out, iters_out);
        CU(h, cudaMemcpyAsync(h->h_out, h->d_out, h->out_elems * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        const double *o = h->h_out;
        if (u0) u0[0] = o[0];
        if (U_out) memcpy(U_out, o + 1, h->g.H * sizeof(double));
        if (want_traj) {
            if (cost_out) cost_out[0] = o[1 + h->g.H];
            if (xs_out) memcpy(xs_out, o + 2 + h->g.H, (size_t)(h->g.H + 1) * h->ns * sizeof(double));
        }
        h->last_iters = (int)o[h->out_stride];
        if (iters_out) *iters_out = h->last_iters;
        return MPC_OK;
    }
    int rc = mpc_cem_begin(h, x, g_z, U_in);
    if (rc) return rc;
    for (int it = 0; it < h->cfg.max_iter; ++it) {  // later iterations return at once when the done flag is set
        rc = cem_enqueue_iter(h, it);
        if (rc) return rc;
    }
    h->cem_iter = h->cfg.max_iter;
    return cem_finish(h, (xs_out || cost_out) ? 1 : 0, u0, U_out, xs_out, cost_out, iters_out);
}

int mpc_cem_get_rewards(mpc_handle *h, int32_t iter, double *R_out) {
    if (!h || !R_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    if (iter < 0 || iter >= h->cfg.max_iter) return fail(h, MPC_EINVAL, "iteration %d out of range", iter);
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaMemcpyAsync(R_out, h->d_R + (size_t)iter * h->g.Kpad, (size_t)h->g.K_local * sizeof(double),
                          cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_cem_get_elites(mpc_handle *h, int32_t iter, int32_t *idx_out) {
    if (!h || !idx_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    if (iter < 0 || iter >= h->cfg.max_iter) return fail(h, MPC_EINVAL, "iteration %d out of range", iter);
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaMemcpyAsync(idx_out, h->d_elite + (size_t)iter * h->cfg.n_elite, (size_t)h->cfg.n_elite * sizeof(int),
                          cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_cem_get_iterations(mpc_handle *h, int32_t *iters_out) {
    if (!h || !iters_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    CU(h, cudaSetDevice(h->cfg.device));
    double it = 0.0;  // cem_finish_kernel leaves the executed iteration count behind the solve's outputs
    CU(h, cudaMemcpyAsync(&it, h->d_out + h->out_stride, sizeof it, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *iters_out = h->last_iters = (int)it;
    return MPC_OK;
}

int mpc_cem_get_std(mpc_handle *h, double *std_out) {
    if (!h || !std_out) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.agent != MPC_AGENT_CEM) return fail(h, MPC_ESTATE, "not a CEM handle");
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaMemcpyAsync(std_out, h->d_sd, (size_t)h->g.H * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- nominal rollout
// ---------------------------------------------------------------------------------------------- closed-loop episode
int mpc_episode_run(mpc_handle *h, const double *x0, const double *U_in, const mpc_episode_opts *opts, double *states,
                    double *true_states, double *actions, double *costs, double *plan_us, double *plan_xs,
                    double *plan_costs, double *U_final) {
    if (!h) return MPC_EINVAL;
    if (!opts || opts->struct_size != sizeof(mpc_episode_opts))
        return fail(h, MPC_EINVAL, "mpc_episode_opts.struct_size mismatch (header/library versions differ)");
    if (!x0 || !states || !true_states || !actions || !costs) return fail(h, MPC_EINVAL, "null output or state pointer");
    if (opts->n_steps < 1 || opts->warm_start_iters < 0) return fail(h, MPC_EINVAL, "n_steps >= 1, warm_start_iters >= 0");
    if (h->g.B != 1 || h->cfg.world != 1) return fail(h, MPC_ESTATE, "episodes run on batch = 1, world = 1 handles");
    const bool mppi = h->cfg.agent == MPC_AGENT_MPPI;
    if (mppi && !h->noise_ready) return fail(h, MPC_ESTATE, "noise tensor not set: call mpc_set_noise or mpc_fill_noise first");
    if (!mppi && h->cfg.noise != MPC_NOISE_PHILOX)
        return fail(h, MPC_ESTATE, "a device-resident CEM episode needs PHILOX noise (host draws cannot be resident)");
    CU(h, cudaSetDevice(h->cfg.device));
    const Geometry &g = h->g;
    const int n = opts->n_steps, ns = h->ns, H = g.H;
    // device histories: true_state[ns] | states | true_states | actions | costs | plan_us | plan_xs | plan_costs
    const size_t o_true = 0, o_states = o_true + ns, o_tstates = o_states + (size_t)(n + 1) * ns,
                 o_act = o_tstates + (size_t)(n + 1) * ns, o_cost = o_act + n, o_pus = o_cost + n,
                 o_pxs = o_pus + (plan_us ? (size_t)n * H : 0), o_pc = o_pxs + (plan_xs ? (size_t)n * (H + 1) * ns : 0),
                 total = o_pc + (plan_costs ? n : 0);
    if (total > h->ep_elems) {
        if (h->d_episode) cudaFree(h->d_episode);
        h->d_episode = nullptr;
        h->ep_elems = 0;
        CU(h, cudaMalloc(&h->d_episode, total * sizeof(double)));
        h->ep_elems = total;
    }
    double *d = h->d_episode;
    int rc = stage_inputs(h, x0, nullptr, U_in);  // state + zero goal (+ plan) in one copy
    if (rc) return rc;
    CU(h, cudaMemcpyAsync(d + o_true, h->d_in + off_x(h), ns * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    CU(h, cudaMemcpyAsync(d + o_states, h->d_in + off_x(h), ns * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    CU(h, cudaMemcpyAsync(d + o_tstates, h->d_in + off_x(h), ns * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    EpisodeArgs ea;
    memset(&ea, 0, sizeof ea);
    ea.H = H;
    ea.n_steps = n;
    ea.shift = opts->shift ? 1 : 0;
    ea.lo = h->cfg.bounds_low;
    ea.hi = h->cfg.bounds_high;
    ea.actuation_std = opts->actuation_std;
    ea.sensor_std = opts->sensor_std;
    ea.seed = opts->seed;
    EpisodeBuffers eb;
    eb.out = h->d_out;
    eb.U = h->d_in + off_U(h);
    eb.x0 = h->d_in + off_x(h);
    eb.true_state = d + o_true;
    eb.states = d + o_states;
    eb.true_states = d + o_tstates;
    eb.actions = d + o_act;
    eb.costs = d + o_cost;
    eb.plan_us = plan_us ? d + o_pus : nullptr;
    eb.plan_xs = plan_xs ? d + o_pxs : nullptr;
    eb.plan_costs = plan_costs ? d + o_pc : nullptr;
    // one solve of either agent, enqueued only
    auto enqueue_solve = [&](int want_traj) -> int {
        if (mppi) return mppi_enqueue(h, want_traj);
        if (h->cem_fused) return cem_fused_enqueue(h, want_traj, false);
        CU(h, launch_cem_reset(cem_args(h, 0, 0), cem_bufs(h), h->stream));
        h->launches++;
        for (int it = 0; it < h->cfg.max_iter; ++it) {
            const int r = cem_enqueue_iter(h, it);
            if (r) return r;
        }
        CU(h, launch_cem_finish(h->cfg.dtype, h->ep, cem_args(h, 0, want_traj), cem_bufs(h), h->stream));
        h->launches++;
        h->solves++;
        return MPC_OK;
    };
    for (int w = 0; w < opts->warm_start_iters; ++w) {  // Agent.warm_start (agent.py:42-47)
        rc = enqueue_solve(0);
        if (rc) return rc;
        ea.mode = 1;
        CU(h, launch_episode_step(h->ep, ea, eb, h->stream));
        h->launches++;
    }
    ea.mode = 0;
    const int want_traj = (ea.shift || plan_xs || plan_costs) ? 1 : 0;
    for (int i = 0; i < n; ++i) {
        rc = enqueue_solve(want_traj);
        if (rc) return rc;
        ea.step = i;
        CU(h, launch_episode_step(h->ep, ea, eb, h->stream));
        h->launches++;
    }
    // one synchronisation: histories back through the pinned staging buffer when they fit, else straight to the caller
    auto fetch = [&](double *dst, size_t off, size_t cnt) -> int {
        if (dst) CU(h, cudaMemcpyAsync(dst, d + off, cnt * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        return MPC_OK;
    };
    if ((rc = fetch(states, o_states, (size_t)(n + 1) * ns))) return rc;
    if ((rc = fetch(true_states, o_tstates, (size_t)(n + 1) * ns))) return rc;
    if ((rc = fetch(actions, o_act, n))) return rc;
    if ((rc = fetch(costs, o_cost, n))) return rc;
    if ((rc = fetch(plan_us, o_pus, (size_t)n * H))) return rc;
    if ((rc = fetch(plan_xs, o_pxs, (size_t)n * (H + 1) * ns))) return rc;
    if ((rc = fetch(plan_costs, o_pc, n))) return rc;
    if (U_final)
        CU(h, cudaMemcpyAsync(U_final, h->d_in + off_U(h), H * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_simulate(mpc_handle *h, const double *x, const double *us, const double *g_z, double *xs_out, double *cost_out) {
    if (!h || !x || !us) return fail(h, MPC_EINVAL, "null argument");
    CU(h, cudaSetDevice(h->cfg.device));
    const int H = h->g.H, ns = h->ns, nz = h->nz;
    const size_t n_in = ns + (size_t)H + (size_t)H * nz, n_out = 1 + (size_t)(H + 1) * ns;
    int rc = ensure_scratch(h, n_in + n_out);
    if (rc) return rc;
    std::string tmp((n_in + n_out) * sizeof(double), '\0');
    double *hb = reinterpret_cast<double *>(&tmp[0]);
    memcpy(hb, x, ns * sizeof(double));
    memcpy(hb + ns, us, H * sizeof(double));
    if (g_z) memcpy(hb + ns + H, g_z, (size_t)H * nz * sizeof(double));
    double *d = h->d_scratch;
    CU(h, cudaMemcpyAsync(d, hb, n_in * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CU(h, launch_simulate(h->cfg.dtype, h->ep, H, h->cfg.bounds_low, h->cfg.bounds_high, 0, d, d + ns, d + ns + H,
                          d + n_in + 1, d + n_in, h->stream));
    h->launches++;
    CU(h, cudaMemcpyAsync(hb + n_in, d + n_in, n_out * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    if (cost_out) *cost_out = hb[n_in];
    if (xs_out) memcpy(xs_out, hb + n_in + 1, (size_t)(H + 1) * ns * sizeof(double));
    return MPC_OK;
}

int mpc_model_step(mpc_handle *h, int32_t n, const double *x, const double *u, const double *g, double *cost_out,
                   double *xn_out) {
    if (!h || !x || !u || n < 1) return fail(h, MPC_EINVAL, "bad argument");
    CU(h, cudaSetDevice(h->cfg.device));
    const int ns = h->ns, nz = h->nz;
    const size_t n_in = (size_t)n * (ns + 1 + nz), n_out = (size_t)n * (1 + ns);
    int rc = ensure_scratch(h, n_in + n_out);
    if (rc) return rc;
    double *d = h->d_scratch;
    double *dx = d, *du = d + (size_t)n * ns, *dg = du + n, *dc = d + n_in, *dxn = dc + n;
    CU(h, cudaMemcpyAsync(dx, x, (size_t)n * ns * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(du, u, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (g)
        CU(h, cudaMemcpyAsync(dg, g, (size_t)n * nz * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    else
        CU(h, cudaMemsetAsync(dg, 0, (size_t)n * nz * sizeof(double), h->stream));
    CU(h, launch_model_step(h->cfg.dtype, h->ep, n, dx, du, dg, dc, dxn, h->stream));
    h->launches++;
    if (cost_out) CU(h, cudaMemcpyAsync(cost_out, dc, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (xn_out) CU(h, cudaMemcpyAsync(xn_out, dxn, (size_t)n * ns * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- multi-GPU
int mpc_comm_create_id(void *id_out) {
    if (!id_out) return MPC_EINVAL;
    std::string err;
    if (!nccl_load(&err)) return fail(nullptr, MPC_ENCCL, "%s", err.c_str());
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != 0) return fail(nullptr, MPC_ENCCL, "ncclGetUniqueId: %s", g_nccl.GetErrorString(r));
    static_assert(sizeof(ncclUniqueId) == MPC_COMM_ID_BYTES, "id size");
    memcpy(id_out, &id, sizeof id);
    return MPC_OK;
}

int mpc_comm_init(mpc_handle *h, const void *id_bytes) {
    if (!h || !id_bytes) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.world == 1) return MPC_OK;
    std::string err;
    if (!nccl_load(&err)) return fail(h, MPC_ENCCL, "%s", err.c_str());
    CU(h, cudaSetDevice(h->cfg.device));
    ncclUniqueId id;
    memcpy(&id, id_bytes, sizeof id);
    ncclResult_t r = g_nccl.CommInitRank(&h->comm, h->cfg.world, id, h->cfg.rank);
    if (r != 0) return fail(h, MPC_ENCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
    return MPC_OK;
}

// Peer exchange: the in-kernel all-gather of the solve kernel (mppi.cuh) needs every rank's exchange buffer mapped
// into every rank.  One process per GPU: export -> host all-gather of the 64-byte handles -> import.  One process
// driving several GPUs: mpc_p2p_link.
static int xchg_alloc(mpc_handle *h) {
    if (h->d_xchg) return MPC_OK;
    if (h->cfg.world > 16) return fail(h, MPC_EINVAL, "peer exchange supports up to 16 ranks");
    if (h->cfg.agent == MPC_AGENT_CEM) {
        // double-buffered by epoch parity: candidate pairs [2][world][n_cand] of (u64 key, u64 global index), then u64
        // flags [2][world]
        CU(h, cudaSetDevice(h->cfg.device));
        const size_t n = (size_t)h->cfg.world * (h->cem_n_cand > 0 ? h->cem_n_cand : 1);
        const size_t bytes = 2 * (n * 16 + (size_t)h->cfg.world * 8);
        CU(h, cudaMalloc(&h->d_xchg, bytes));
        CU(h, cudaMemset(h->d_xchg, 0, bytes));
        return MPC_OK;
    }
    // mode A blocks wait for their peers while holding an SM: every block of the grid must be resident
    if (h->g.solves_per_block == 0 && h->g.grid > h->num_sms)
        return fail(h, MPC_ESTATE, "peer exchange needs the whole grid resident (batch %d too large for K %d); use NCCL",
                    h->g.B, h->cfg.K);
    CU(h, cudaSetDevice(h->cfg.device));
    const size_t slots = 2 * (size_t)h->cfg.world * h->g.B;  // double-buffered by epoch parity (mppi.cuh)
    const size_t bytes = slots * (PART_HDR + h->g.H) * sizeof(double) + slots * sizeof(unsigned long long);
    CU(h, cudaMalloc(&h->d_xchg, bytes));
    CU(h, cudaMemset(h->d_xchg, 0, bytes));
    return MPC_OK;
}

int mpc_p2p_export(mpc_handle *h, void *handle_out) {
    if (!h || !handle_out) return fail(h, MPC_EINVAL, "null argument");
    int rc = xchg_alloc(h);
    if (rc) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == MPC_P2P_HANDLE_BYTES, "ipc handle size");
    cudaIpcMemHandle_t ih;
    CU(h, cudaIpcGetMemHandle(&ih, h->d_xchg));
    memcpy(handle_out, &ih, sizeof ih);
    return MPC_OK;
}

int mpc_p2p_import(mpc_handle *h, const void *handles) {
    if (!h || !handles) return fail(h, MPC_EINVAL, "null argument");
    if (h->cfg.world == 1) return MPC_OK;
    int rc = xchg_alloc(h);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->cfg.device));
    for (int r = 0; r < h->cfg.world; ++r) {
        if (r == h->cfg.rank) {
            h->peer_recs[r] = h->d_xchg;
            continue;
        }
        cudaIpcMemHandle_t ih;
        memcpy(&ih, (const char *)handles + (size_t)r * MPC_P2P_HANDLE_BYTES, sizeof ih);
        void *p = nullptr;
        CU(h, cudaIpcOpenMemHandle(&p, ih, cudaIpcMemLazyEnablePeerAccess));
        h->peer_recs[r] = (double *)p;
        h->peer_opened[r] = true;
    }
    h->p2p_ready = true;
    return MPC_OK;
}

int mpc_p2p_link(mpc_handle *const *hs, int32_t world) {
    if (!hs || world < 1) return MPC_EINVAL;
    for (int r = 0; r < world; ++r) {
        if (!hs[r]) return MPC_EINVAL;
        if (hs[r]->cfg.world != world || hs[r]->cfg.rank != r)
            return fail(hs[r], MPC_EINVAL, "mpc_p2p_link: handle %d has rank %d of %d", r, hs[r]->cfg.rank, hs[r]->cfg.world);
        int rc = xchg_alloc(hs[r]);
        if (rc) return rc;
    }
    for (int r = 0; r < world; ++r) {
        mpc_handle *h = hs[r];
        CU(h, cudaSetDevice(h->cfg.device));
        for (int q = 0; q < world; ++q) {
            if (q != r && hs[q]->cfg.device != h->cfg.device) {
                int can = 0;
                CU(h, cudaDeviceCanAccessPeer(&can, h->cfg.device, hs[q]->cfg.device));
                if (!can) return fail(h, MPC_ECUDA, "device %d cannot access device %d", h->cfg.device, hs[q]->cfg.device);
                cudaError_t e = cudaDeviceEnablePeerAccess(hs[q]->cfg.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled)
                    cudaGetLastError();
                else if (e != cudaSuccess)
                    return fail(h, MPC_ECUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            }
            h->peer_recs[q] = hs[q]->d_xchg;
        }
        h->p2p_ready = true;
    }
    return MPC_OK;
}

// ---------------------------------------------------------------------------------------------- measurement
int mpc_upload_inputs(mpc_handle *h, const double *x, const double *g_z, const double *U_in) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    int rc = stage_inputs(h, x, g_z, U_in);
    if (rc) return rc;
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

extern "C" int mpc_flush_l2(mpc_handle *h);
static int solve_enqueue(mpc_handle *h, int want_traj) {
    if (h->cfg.agent == MPC_AGENT_MPPI) return mppi_enqueue(h, want_traj);
    if (h->cfg.noise != MPC_NOISE_PHILOX) return fail(h, MPC_ESTATE, "device-resident CEM solve needs PHILOX noise");
    if (h->cem_fused) return cem_fused_enqueue(h, want_traj, false);
    CU(h, launch_cem_reset(cem_args(h, 0, 0), cem_bufs(h), h->stream));
    h->launches++;
    for (int it = 0; it < h->cfg.max_iter; ++it) {
        int rc = cem_enqueue_iter(h, it);
        if (rc) return rc;
    }
    CU(h, launch_cem_finish(h->cfg.dtype, h->ep, cem_args(h, 0, want_traj), cem_bufs(h), h->stream));
    h->launches++;
    h->solves++;
    return MPC_OK;
}

int mpc_solve_enqueue(mpc_handle *h) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    return solve_enqueue(h, 1);
}

int mpc_solve_enqueue_ex(mpc_handle *h, int32_t want_traj) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    return solve_enqueue(h, want_traj ? 1 : 0);
}

int mpc_time_solves(mpc_handle *h, int32_t n, int32_t want_traj, int32_t flush_l2, float *ms_each) {
    if (!h || !ms_each || n < 1) return fail(h, MPC_EINVAL, "bad argument");
    CU(h, cudaSetDevice(h->cfg.device));
    while ((int)h->ev_pool.size() < 2 * n) {
        cudaEvent_t e;
        CU(h, cudaEventCreate(&e));
        h->ev_pool.push_back(e);
    }
    // every solve between its own pair of events, nothing but the (untimed) L2 flush between two pairs, and no host
    // synchronisation until all n are enqueued: the device queue never runs dry, so an event pair brackets the solve's
    // own launches and not the host's launch latency
    if (h->cfg.world > 1) {
        // K sharded over several GPUs: one untimed solve first.  Its in-kernel exchange lines the ranks' device
        // timelines up (a host barrier leaves them tens of microseconds apart, which the first timed solve of the
        // early ranks would otherwise spend waiting for the late ones).
        int rc = solve_enqueue(h, want_traj ? 1 : 0);
        if (rc) return rc;
        if (flush_l2 && (rc = mpc_flush_l2(h))) return rc;
    }
    for (int i = 0; i < n; ++i) {
        CU(h, cudaEventRecord(h->ev_pool[2 * i], h->stream));
        int rc = solve_enqueue(h, want_traj ? 1 : 0);
        if (rc) return rc;
        CU(h, cudaEventRecord(h->ev_pool[2 * i + 1], h->stream));
        if (flush_l2) {
            rc = mpc_flush_l2(h);
            if (rc) return rc;
        }
    }
    CU(h, cudaStreamSynchronize(h->stream));
    for (int i = 0; i < n; ++i) CU(h, cudaEventElapsedTime(&ms_each[i], h->ev_pool[2 * i], h->ev_pool[2 * i + 1]));
    return MPC_OK;
}

int mpc_download_outputs(mpc_handle *h, double *u0, double *U_out, double *xs_out, double *cost_out) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    return fetch_outputs(h, 1, u0, U_out, xs_out, cost_out);
}

int mpc_timer_start(mpc_handle *h) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaEventRecord(h->ev0, h->stream));
    return MPC_OK;
}

int mpc_timer_stop(mpc_handle *h, float *ms_out) {
    if (!h || !ms_out) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaEventRecord(h->ev1, h->stream));
    CU(h, cudaEventSynchronize(h->ev1));
    CU(h, cudaEventElapsedTime(ms_out, h->ev0, h->ev1));
    return MPC_OK;
}

int mpc_set_profiling(mpc_handle *h, int32_t on) {
    if (!h) return MPC_EINVAL;
    h->profiling = (on & 1) != 0;
    h->store_costs = (on & 2) == 0;
    h->rollout_events = (on & 4) != 0;
    return MPC_OK;
}

int mpc_get_rollout_ms(mpc_handle *h, float *ms_out) {
    if (!h || !ms_out) return MPC_EINVAL;
    if (!h->rollout_events) return fail(h, MPC_ESTATE, "rollout events are off: call mpc_set_profiling(h, 4) before the solve");
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaEventSynchronize(h->evr1));
    CU(h, cudaEventElapsedTime(ms_out, h->evr0, h->evr1));
    return MPC_OK;
}

int mpc_get_trace(mpc_handle *h, uint64_t out[16]) {
    if (!h || !out) return MPC_EINVAL;
    if (!h->d_trace) return fail(h, MPC_ESTATE, "no trace buffer (MPPI handles and single-launch CEM handles only)");
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaMemcpyAsync(out, h->d_trace, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int mpc_sync(mpc_handle *h) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    CU(h, cudaStreamSynchronize(h->stream));
    return MPC_OK;
}

int64_t mpc_launch_count(const mpc_handle *h) { return h ? h->launches : 0; }

// Self-check of the solve kernel's arrival tickets: they are self-cleaning, so between solves every counter is zero.
int mpc_debug_tickets(mpc_handle *h, int32_t *nonzero_out) {
    if (!h || !nonzero_out) return MPC_EINVAL;
    *nonzero_out = 0;
    if (!h->d_done) return MPC_OK;
    CU(h, cudaSetDevice(h->cfg.device));
    const size_t n = (size_t)h->g.B * (1 + h->g.n_groups);
    std::vector<int> v(n);
    CU(h, cudaMemcpyAsync(v.data(), h->d_done, n * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    for (int x : v) *nonzero_out += x != 0;
    return MPC_OK;
}

int mpc_get_geometry(const mpc_handle *h, int32_t out[8]) {
    if (!h || !out) return MPC_EINVAL;
    out[0] = h->g.grid;
    out[1] = h->g.warps * 32;
    out[2] = (int32_t)h->g.smem_rollout;
    out[3] = h->g.total_tiles;
    out[4] = h->g.K_local;
    out[5] = h->g.k_begin;
    out[6] = h->g.stages;
    out[7] = h->g.stream_rows;
    return MPC_OK;
}

// Page-lock / release a caller-owned host buffer (cudaHostRegister): output arrays of batched solves registered once
// receive their results by direct DMA.
int mpc_pin_host_buffer(void *p, size_t bytes) {
    if (!p || !bytes) return MPC_EINVAL;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return MPC_OK;
    }
    if (e != cudaSuccess) return fail(nullptr, MPC_ECUDA, "cudaHostRegister: %s", cudaGetErrorString(e));
    return MPC_OK;
}

int mpc_unpin_host_buffer(void *p) {
    if (!p) return MPC_EINVAL;
    cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, MPC_ECUDA, "cudaHostUnregister: %s", cudaGetErrorString(e));
    }
    return MPC_OK;
}

int mpc_flush_l2(mpc_handle *h) {
    if (!h) return MPC_EINVAL;
    CU(h, cudaSetDevice(h->cfg.device));
    if (!h->d_flush) {
        h->flush_elems = (size_t)256 << 20 >> 2;  // 256 MiB > 126 MB of L2
        CU(h, cudaMalloc(&h->d_flush, h->flush_elems * sizeof(float)));
    }
    CU(h, launch_fill(h->d_flush, h->flush_elems, (float)h->solves, h->stream));
    h->launches++;
    return MPC_OK;
}

int mpc_measure_fma_peak(int32_t device, int32_t dtype, double *tflops_out) {
    if (!tflops_out) return MPC_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(nullptr, MPC_ECUDA, "no such CUDA device %d", device);
    CU(nullptr, cudaSetDevice(device));
    int sms = 0;
    CU(nullptr, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    float *sink = nullptr;
    CU(nullptr, cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    CU(nullptr, cudaEventCreate(&e0));
    CU(nullptr, cudaEventCreate(&e1));
    const int blocks = sms * 8, iters = dtype == MPC_F32 ? 8192 : 4096;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(nullptr, cudaEventRecord(e0, 0));
        CU(nullptr, launch_fma_peak(dtype == MPC_F32 ? DT_F32 : DT_F64, sink, blocks, iters, 0));
        CU(nullptr, cudaEventRecord(e1, 0));
        CU(nullptr, cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(nullptr, cudaEventElapsedTime(&ms, e0, e1));
        const double flops = (double)blocks * 256.0 * 16.0 * (double)iters * 2.0;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tflops_out = best;
    return MPC_OK;
}

}  // extern "C"

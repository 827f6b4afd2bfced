// noalloc.cu — mpc_mppi_update_static: the MPPI control update (softmin weights + weighted noise sum + planned_us += erg,
// agents/mppi.py:59-71) as a stand-alone entry point that ALLOCATES NOTHING.  It has no counterpart in the reference;
// it was added for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).
//
// It is a variant of the tail of the solve kernel (mppi.cuh: per-block records (min cost, sum w, sum w*eps[t]) merged by
// the block that takes the last ticket).  Where the solve kernel keeps those records and the ticket in buffers that
// mpc_create allocates, this one keeps them in a file-scope `static __device__` array of fixed size (NA_SCRATCH_BYTES,
// ~0.6 MB: every GPU that loads the library holds it until the process exits).  Consequences, all documented in
// include/mpc_b200.h:
//   - sizes beyond the fixed maximum (H > 256, more than 2^24 samples) are rejected, not grown into;
//   - every call on a device shares the one array: two calls that overlap on that GPU (different streams, different
//     host threads) race on it.  Calls on ONE stream are ordered and safe;
//   - the array keeps the last call's contents; the ticket word is cleared on the caller's stream before every launch.
// In host code `g_na_scratch` is a host-side stand-in for the symbol, not a device address: the device address is taken
// once per device with cudaGetSymbolAddress (device current, return code checked) and that pointer is what the kernel
// receives.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/mpc_b200.h"

namespace {

constexpr int NA_MAX_H = MPC_STATIC_MAX_H;
constexpr int NA_MAX_K = MPC_STATIC_MAX_K;
constexpr int NA_MAX_BLOCKS = 296;  // two blocks per SM of a B200
constexpr int NA_THREADS = 256;     // 8 warps: one sample row per warp at a time, lanes across the horizon
constexpr int NA_WARPS = NA_THREADS / 32;
constexpr int NA_COLS = NA_MAX_H / 32;  // horizon columns per lane
constexpr int NA_REC = 2 + NA_MAX_H;    // doubles per block record: min cost, sum w, sum w*eps[0..H-1]
constexpr int NA_HEAD = 256;            // bytes in front of the records: the ticket word (rest unused)
constexpr size_t NA_SCRATCH_BYTES = NA_HEAD + sizeof(double) * (size_t)NA_MAX_BLOCKS * NA_REC;
constexpr int NA_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];

// S [K] per-sample costs, eps [K][H] (the reference's delta_u layout, du = 1), U_in / U_out [H]; all float64, device.
__global__ void __launch_bounds__(NA_THREADS)
    mppi_update_static_kernel(const int K, const int H, const double inv_lam, const double *__restrict__ S,
                              const double *__restrict__ eps, const double *U_in, double *U_out /* may alias */,
                              unsigned char *__restrict__ scratch) {
    unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch);
    double *recs = reinterpret_cast<double *>(scratch + NA_HEAD);
    __shared__ double s_red[NA_WARPS];
    __shared__ double s_acc[NA_WARPS][NA_MAX_H];
    __shared__ double s_scale[NA_MAX_BLOCKS];
    __shared__ double s_min, s_eta;
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, G = gridDim.x, b = blockIdx.x;
    // this block's contiguous share of the samples
    const int per = (K + G - 1) / G, k0 = min(K, b * per), k1 = min(K, k0 + per);

    // ---- block minimum (beta of mppi.py:60, per block; rescaled to the global minimum in the merge)
    double m = INFINITY;
    for (int k = k0 + tid; k < k1; k += NA_THREADS) m = fmin(m, S[k]);
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_red[w] = m;
    __syncthreads();
    if (tid == 0) {
        double mm = s_red[0];
        for (int i = 1; i < NA_WARPS; ++i) mm = fmin(mm, s_red[i]);
        s_min = mm;
    }
    __syncthreads();
    const double mb = s_min;

    // ---- weights and weighted noise sum: warp w takes rows k0+w, k0+w+8, ...; lane l owns columns l, l+32, ...
    double acc[NA_COLS], wsum = 0.0;
#pragma unroll
    for (int c = 0; c < NA_COLS; ++c) acc[c] = 0.0;
    for (int k = k0 + w; k < k1; k += NA_WARPS) {
        const double wk = exp(-inv_lam * (S[k] - mb));  // mppi.py:61
        wsum += wk;
        const double *row = eps + (size_t)k * H;
#pragma unroll
        for (int c = 0; c < NA_COLS; ++c) {
            const int t = lane + 32 * c;
            if (t < H) acc[c] += wk * row[t];
        }
    }
    __syncthreads();  // s_red free
    if (lane == 0) s_red[w] = wsum;
#pragma unroll
    for (int c = 0; c < NA_COLS; ++c) {
        const int t = lane + 32 * c;
        if (t < H) s_acc[w][t] = acc[c];
    }
    __syncthreads();
    double *rec = recs + (size_t)b * NA_REC;
    for (int t = tid; t < H; t += NA_THREADS) {
        double a = 0.0;
#pragma unroll
        for (int i = 0; i < NA_WARPS; ++i) a += s_acc[i][t];  // fixed order: the result does not depend on timing
        rec[2 + t] = a;
    }
    if (tid == 0) {
        double a = 0.0;
        for (int i = 0; i < NA_WARPS; ++i) a += s_red[i];
        rec[0] = mb;
        rec[1] = a;
    }
    // ---- ticket: the block that arrives last merges the records
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(ticket, 1u) == (unsigned)(G - 1));
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid == 0) {
        double mm = INFINITY;
        for (int i = 0; i < G; ++i) mm = fmin(mm, __ldcg(recs + (size_t)i * NA_REC));
        s_min = mm;
    }
    __syncthreads();
    for (int i = tid; i < G; i += NA_THREADS) {
        const double mi = __ldcg(recs + (size_t)i * NA_REC);
        s_scale[i] = (mi == INFINITY) ? 0.0 : exp(-inv_lam * (mi - s_min));  // an empty block's record carries no weight
    }
    __syncthreads();
    if (tid == 0) {
        double eta = 0.0;  // mppi.py:62
        for (int i = 0; i < G; ++i) eta += s_scale[i] * __ldcg(recs + (size_t)i * NA_REC + 1);
        s_eta = eta;
    }
    __syncthreads();
    const double inv_eta = 1.0 / s_eta;
    for (int t = tid; t < H; t += NA_THREADS) {
        double a = 0.0;
        for (int i = 0; i < G; ++i) a += s_scale[i] * __ldcg(recs + (size_t)i * NA_REC + 2 + t);
        U_out[t] = U_in[t] + inv_eta * a;  // mppi.py:63-71: planned_us += sum_k w_k delta_u[k]
    }
    if (tid == 0) *ticket = 0u;  // leave the word at zero (the host clears it before every launch as well)
}

std::mutex g_na_mu;
unsigned char *g_na_ptr[NA_MAX_DEVICES] = {};  // device address of g_na_scratch, per device

int na_fail(char *err, size_t cap, int code, const char *fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int mpc_mppi_update_static(int32_t device, int32_t K, int32_t H, double lam, const double *S_dev,
                                      const double *eps_dev, const double *U_in_dev, double *U_out_dev, void *stream,
                                      char *err_out, size_t err_cap) {
    if (err_out && err_cap) err_out[0] = 0;
    if (K < 1 || H < 1 || !(lam > 0.0) || !S_dev || !eps_dev || !U_in_dev || !U_out_dev)
        return na_fail(err_out, err_cap, MPC_EINVAL, "mpc_mppi_update_static: K >= 1, H >= 1, lam > 0 and four device pointers are required");
    if (H > NA_MAX_H)
        return na_fail(err_out, err_cap, MPC_EINVAL,
                       "mpc_mppi_update_static: H = %d exceeds MPC_STATIC_MAX_H = %d (the scratch array is static and "
                       "cannot grow; use mpc_mppi_solve)", H, NA_MAX_H);
    if (K > NA_MAX_K)
        return na_fail(err_out, err_cap, MPC_EINVAL,
                       "mpc_mppi_update_static: K = %d exceeds MPC_STATIC_MAX_K = %d (use mpc_mppi_solve)", K, NA_MAX_K);
    if (device < 0 || device >= NA_MAX_DEVICES)
        return na_fail(err_out, err_cap, MPC_EINVAL, "mpc_mppi_update_static: device %d out of range", device);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(err_out, err_cap, MPC_ECUDA, "mpc_mppi_update_static: %s", cudaGetErrorString(e));
    unsigned char *scratch;
    {
        // set-up, once per device, with that device current: the device address of the static array
        std::lock_guard<std::mutex> lk(g_na_mu);
        if (!g_na_ptr[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_na_scratch);
            if (e != cudaSuccess || !p) {
                if (prev != device && prev >= 0) cudaSetDevice(prev);
                return na_fail(err_out, err_cap, MPC_ECUDA, "mpc_mppi_update_static: cudaGetSymbolAddress: %s",
                               cudaGetErrorString(e));
            }
            g_na_ptr[device] = static_cast<unsigned char *>(p);
        }
        scratch = g_na_ptr[device];
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int grid = (K + NA_THREADS - 1) / NA_THREADS < NA_MAX_BLOCKS ? (K + NA_THREADS - 1) / NA_THREADS : NA_MAX_BLOCKS;
    e = cudaMemsetAsync(scratch, 0, NA_HEAD, s);  // the ticket must start at zero whatever the last call left
    if (e == cudaSuccess) {
        mppi_update_static_kernel<<<grid, NA_THREADS, 0, s>>>(K, H, 1.0 / lam, S_dev, eps_dev, U_in_dev, U_out_dev, scratch);
        e = cudaGetLastError();
    }
    if (prev != device && prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) return na_fail(err_out, err_cap, MPC_ECUDA, "mpc_mppi_update_static: %s", cudaGetErrorString(e));
    return MPC_OK;
}

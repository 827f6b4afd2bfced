// peer.cuh — system-scope primitives of the multi-GPU exchanges that run INSIDE kernels over NVLink peer memory.
//
// Every rank owns an exchange buffer that every other rank has mapped (cudaIpcOpenMemHandle, or plain peer access
// when one process drives all GPUs).  A rank stores its contribution into slot [rank] of every buffer, raises a flag
// there with a system-scope release store, and waits (bounded) for the flags of the other writers in its own buffer.
#pragma once
#include <cuda_runtime.h>

namespace mpc {

constexpr int MAX_WORLD = 16;
struct PeerTable {
    double *recs[MAX_WORLD];  // recs[r]: rank r's exchange buffer [world][B][2+H] doubles, then u64 flags [world][B]
};
constexpr unsigned long long PEER_TIMEOUT_NS = 4000000000ull;

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)::"memory");  // pinned: a phase stamp must not float past barriers
    return t;
}
// raise `theirs` (a flag in a peer's buffer) to `epoch`, then wait until the peer has raised `mine` (in our buffer)
__device__ __forceinline__ void peer_flag_exchange(unsigned long long *theirs, const unsigned long long *mine,
                                                   unsigned long long epoch, double *err_word) {
    st_release_sys(theirs, epoch);
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(mine) < epoch) {
        if (global_ns() - t0 > PEER_TIMEOUT_NS) {
            *err_word = 1.0;  // the peer never delivered: the host turns this into MPC_ENCCL
            break;
        }
    }
}

}  // namespace mpc

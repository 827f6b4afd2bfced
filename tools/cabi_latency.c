/* Development probe (GPU box): per-solve latency of BASELINE configs[1] (CEM CartPole, K = 4096, H = 50, elite ratio
 * 0.1, 5 iterations, float32 rollout, in-kernel Philox normals) through the C ABI with HOST buffers, from plain C — the
 * call a maintainer's binding makes, without the ~8 us of ctypes / numpy that the Python wrapper adds to bench.py's
 * p50.  Every call sends the state and the plan and receives action, plan, logging trajectory and cost.
 *   gcc -O2 -std=c99 -I include tools/cabi_latency.c -o /tmp/cabi_latency -L benchmarking-mpc-solvers_b200 -lmpcb200 \
 *       -Wl,-rpath,$PWD/benchmarking-mpc-solvers_b200 -lm
 *   /tmp/cabi_latency [n_solves] [want_traj 0|1]                                                                      */
#define _POSIX_C_SOURCE 199309L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "mpc_b200.h"

static double now_us(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return 1e6 * (double)ts.tv_sec + 1e-3 * (double)ts.tv_nsec;
}
static int cmp(const void *a, const void *b) {
    const double x = *(const double *)a, y = *(const double *)b;
    return x < y ? -1 : x > y;
}

int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 500, want_traj = argc > 2 ? atoi(argv[2]) : 1;
    enum { K = 4096, H = 50 };
    /* g m_c m_p l force_mag dt b x_threshold W0..W4   (cartpole_swingup_model.py:23-46, the reference's defaults) */
    const double p[13] = {9.82, 0.5, 0.5, 0.6, 10.0, 0.05, 0.1, 2.4, -1.0, -0.0, -5.0, -0.0, -0.0};
    mpc_config c;
    memset(&c, 0, sizeof c);
    c.struct_size = sizeof c;
    c.agent = MPC_AGENT_CEM;
    c.env = MPC_ENV_CARTPOLE;
    c.dtype = MPC_F32;
    c.noise = MPC_NOISE_PHILOX;
    c.K = K;
    c.H = H;
    c.batch = 1;
    c.world = 1;
    c.max_iter = 5;
    c.n_elite = K / 10;
    c.sigma = 1.0;
    c.epsilon = 1e-5;
    c.alpha = 0.2;
    c.bounds_low = -1.0;
    c.bounds_high = 1.0;
    c.seed = 7;
    memcpy(c.env_params, p, sizeof p);
    mpc_handle *h = NULL;
    if (mpc_create(&c, &h) != MPC_OK) {
        fprintf(stderr, "mpc_create: %s\n", mpc_last_error(NULL));
        return 1;
    }
    const double x[4] = {0.01, 0.01, 3.151592653589793, 0.01};
    double U[H], Uo[H], xs[(H + 1) * 4], u0 = 0.0, cost = 0.0;
    memset(U, 0, sizeof U);
    int32_t iters = 0;
    double *lat = (double *)malloc(sizeof(double) * (size_t)n);
    for (int i = 0; i < 20 + n; ++i) {
        const double t0 = now_us();
        const int rc = mpc_cem_solve(h, x, NULL, U, &u0, Uo, want_traj ? xs : NULL, want_traj ? &cost : NULL, &iters);
        const double t1 = now_us();
        if (rc != MPC_OK) {
            fprintf(stderr, "mpc_cem_solve: %s\n", mpc_last_error(h));
            return 1;
        }
        if (i >= 20) lat[i - 20] = t1 - t0;
    }
    qsort(lat, (size_t)n, sizeof(double), cmp);
    printf("{\"workload\": \"cem_cartpole_k4096_h50_5it\", \"caller\": \"plain C, host buffers\", \"want_traj\": %d, "
           "\"n\": %d, \"iters\": %d, \"p50_us\": %.1f, \"p10_us\": %.1f, \"p90_us\": %.1f, \"u0\": %.6f}\n",
           want_traj, n, (int)iters, lat[n / 2], lat[n / 10], lat[(9 * n) / 10], u0);
    free(lat);
    mpc_destroy(h);
    return 0;
}

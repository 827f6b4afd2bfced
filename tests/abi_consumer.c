/* A plain-C consumer of include/mpc_b200.h: what a maintainer's binding of MPPI._calc_action (agents/mppi.py:54-75)
 * does, without Python.  Built by tests/test_abi_consumer.py with gcc against libmpcb200.so.
 *   abi_consumer probe           -> prints the ABI version; exit 0 when mpc_create fails loudly without a device
 *   abi_consumer solve K H seed  -> Pendulum MPPI, float64, noise = LCG-driven Box-Muller (printed so that the checker
 *                                   can rebuild the identical tensor), prints u0, the plan and the nominal cost */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "mpc_b200.h"

static unsigned long long lcg_state;
static double lcg_uniform(void) {
    lcg_state = lcg_state * 6364136223846793005ULL + 1442695040888963407ULL;
    return ((double)(lcg_state >> 11) + 0.5) / 9007199254740992.0;
}
static double lcg_normal(void) {
    const double u1 = lcg_uniform(), u2 = lcg_uniform();
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

static void pendulum_config(mpc_config *c, int K, int H) {
    const double p[9] = {10.0, 1.0, 1.0, 0.05, 8.0, 2.0, 1.0, 0.1, 0.001}; /* g m l dt max_speed max_torque W0 W1 W2 */
    memset(c, 0, sizeof *c);
    c->struct_size = sizeof *c;
    c->agent = MPC_AGENT_MPPI;
    c->env = MPC_ENV_PENDULUM;
    c->dtype = MPC_F64;
    c->noise = MPC_NOISE_TENSOR;
    c->K = K;
    c->H = H;
    c->batch = 1;
    c->world = 1;
    c->lam = 0.5;
    c->sigma = 1.0;
    c->bounds_low = -2.0;
    c->bounds_high = 2.0;
    memcpy(c->env_params, p, sizeof p);
}

int main(int argc, char **argv) {
    if (argc >= 2 && strcmp(argv[1], "probe") == 0) {
        mpc_config c;
        mpc_handle *h = NULL;
        pendulum_config(&c, 64, 10);
        printf("abi %d\n", mpc_abi_version());
        const int rc = mpc_create(&c, &h);
        if (rc == MPC_OK) {
            printf("device present\n");
            mpc_destroy(h);
            return 0;
        }
        printf("create failed: %d: %s\n", rc, mpc_last_error(NULL));
        return rc == MPC_ECUDA ? 0 : 2; /* no device => a loud MPC_ECUDA, never a CPU fallback */
    }
    if (argc < 5 || strcmp(argv[1], "solve") != 0) return 64;
    const int K = atoi(argv[2]), H = atoi(argv[3]);
    lcg_state = strtoull(argv[4], NULL, 10);
    mpc_config c;
    mpc_handle *h = NULL;
    pendulum_config(&c, K, H);
    if (mpc_create(&c, &h) != MPC_OK) {
        fprintf(stderr, "mpc_create: %s\n", mpc_last_error(NULL));
        return 1;
    }
    double *eps = (double *)malloc(sizeof(double) * (size_t)K * H);
    double *U = (double *)calloc((size_t)H, sizeof(double)), *Uo = (double *)malloc(sizeof(double) * H);
    double *xs = (double *)malloc(sizeof(double) * (size_t)(H + 1) * 2);
    for (long i = 0; i < (long)K * H; ++i) eps[i] = lcg_normal();
    const double x[2] = {3.0, -0.5};
    double u0 = 0.0, cost = 0.0;
    int rc = mpc_set_noise(h, eps);
    if (rc == MPC_OK) rc = mpc_mppi_solve(h, x, NULL, U, &u0, Uo, xs, &cost);
    if (rc != MPC_OK) {
        fprintf(stderr, "solve: %s\n", mpc_last_error(h));
        return 1;
    }
    printf("u0 %.17g\ncost %.17g\nU", u0, cost);
    for (int t = 0; t < H; ++t) printf(" %.17g", Uo[t]);
    printf("\neps0 %.17g %.17g\n", eps[0], eps[(long)K * H - 1]);
    mpc_destroy(h);
    free(eps), free(U), free(Uo), free(xs);
    return 0;
}

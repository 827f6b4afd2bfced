/*
 * mpc_oracle.c — TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C, float64 restatement of the reference's sampling-MPC inner loop, written to follow the reference's
 * numpy arithmetic operation by operation (same association order, pow() where numpy calls pow(), the floored
 * remainder, the repeated +-2*pi wrap, numpy's pairwise summation).  Paths under /root/reference/MPCBenchmark:
 *
 *   orc_step            models/model.py:26-44,62-63  (cost of (x_t,u_t) with goal row 0, THEN the step)
 *     pendulum          models/pendulum_model.py:28-50,58-59
 *     cartpole          models/cartpole_swingup_model.py:63-104
 *     acrobot           models/acrobot_model.py:99-176,185-220,270-282
 *   orc_mppi_costs      agents/mppi.py:35-48     (MPPI.f for every k; the Pool.map fan-out of mppi.py:55-57)
 *   orc_mppi_update     agents/mppi.py:59-71
 *   orc_cem_solve       agents/cem.py:42-48,54-78
 *   orc_simulate        agents/agent.py:97-108
 *
 * Pinned against the tests/golden npz fixtures, which oracle/gen_golden.py produced by executing the unmodified
 * reference (the reference ships no golden vectors of its own): see tests/test_oracle_golden.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * Build: gcc -O2 -ffp-contract=off -pthread -shared -fPIC (oracle/Makefile).  -ffp-contract=off keeps gcc from
 * fusing a*b+c so the rounding sequence stays numpy's.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define ORC_PENDULUM 0
#define ORC_CARTPOLE 1
#define ORC_ACROBOT 2

/* Parameter vectors (filled by oracle/oracle.py from the reference constructors' defaults):
 *  pendulum: g m l dt max_speed max_torque W0 W1 W2                       pendulum_model.py:8-22
 *  cartpole: g m_c m_p l force_mag dt b x_threshold W0..W4 (self.W diag)  cartpole_swingup_model.py:23-46
 *  acrobot : dt m1 m2 l1 lc1 lc2 I1 I2 g max_vel1 max_vel2 W0..W4 nips   acrobot_model.py:60-97,107,124 */
typedef struct {
    int env;
    int ns;
    double p[24];
    int full_w;      /* != 0: W holds the model's whole cost matrix self.W (nz x nz, row major); the W_i of p are ignored */
    int goal_stride; /* 0: one goal row for every rollout step (the reference: model.py:34,38 uses row 0);
                        nz: row t at step t in the SAMPLE rollouts too (an extension of this repo, opt-in) */
    double W[25];
} orc_model;

static double clipd(double a, double lo, double hi) { return fmin(fmax(a, lo), hi); }

/* numpy floored remainder (npy_divmod): result takes the sign of the divisor */
static double np_mod(double a, double b) {
    double r = fmod(a, b);
    if (r != 0.0) {
        if ((b < 0) != (r < 0)) r += b;
    } else {
        r = copysign(0.0, b);
    }
    return r;
}

/* _state_cost: einsum("bi,ij,bj->b", zd, W, zd) for one row   pendulum_model.py:46-50 (same in the other two models);
 * diagonal W (the reference's three models): sum_i (zd_i W_ii) zd_i */
static double state_cost(const orc_model *m, const double *zd, int nz, const double *wdiag) {
    double acc = 0.0;
    if (m->full_w) {
        for (int i = 0; i < nz; ++i)
            for (int j = 0; j < nz; ++j) acc += (zd[i] * m->W[i * nz + j]) * zd[j];
    } else {
        for (int i = 0; i < nz; ++i) acc += (zd[i] * wdiag[i]) * zd[i];
    }
    return acc;
}

/* ---------------------------------------------------------------- pendulum */
static void pendulum_step(const orc_model *mdl, const double *x, double u, const double *g, double *cost, double *xn) {
    const double *p = mdl->p;
    const double grav = p[0], m = p[1], l = p[2], dt = p[3], max_speed = p[4], max_torque = p[5];
    const double th = x[0], thdot = x[1];
    /* _transform: z = -[angle_normalize(th), thdot, u]   (u NOT clipped)          pendulum_model.py:40-44 */
    double thn = np_mod(th + M_PI, 2 * M_PI) - M_PI;
    double z0 = -thn, z1 = -thdot, z2 = -u;
    double zd[3] = {z0 - g[0], z1 - g[1], z2 - g[2]};
    *cost = state_cost(mdl, zd, 3, p + 6);
    /* _dynamics                                                               pendulum_model.py:28-38 */
    double uc = clipd(u, -max_torque, max_torque);
    double newthdot = thdot + (-3 * grav / (2 * l) * sin(th + M_PI) + 3. / (m * (l * l)) * uc) * dt;
    double newth = th + newthdot * dt;
    newthdot = clipd(newthdot, -max_speed, max_speed);
    xn[0] = newth;
    xn[1] = newthdot;
}

/* ---------------------------------------------------------------- cartpole */
static void cartpole_step(const orc_model *mdl, const double *x, double u, const double *g, double *cost, double *xn) {
    const double *p = mdl->p;
    const double grav = p[0], m_c = p[1], m_p = p[2], l = p[3], force_mag = p[4], dt = p[5], b = p[6],
                 x_threshold = p[7];
    const double total_m = m_p + m_c, m_p_l = m_p * l;
    const double pos = x[0], x_dot = x[1], theta = x[2], theta_dot = x[3];
    /* _transform                                                     cartpole_swingup_model.py:87-93 */
    double q = pos / x_threshold;
    double z[5] = {q * q + pow(q, 10.0), x_dot, 1 - cos(theta), theta_dot, u};
    /* _state_cost: -(einsum(zd W zd)) with W = self.W (negative diag)   cartpole_swingup_model.py:95-104 */
    double zd[5];
    for (int i = 0; i < 5; ++i) zd[i] = z[i] - g[i];
    *cost = -state_cost(mdl, zd, 5, p + 8);
    /* _dynamics                                                      cartpole_swingup_model.py:63-85 */
    double action = clipd(u, -1.0, 1.0);
    action *= force_mag;
    double s = sin(theta), c = cos(theta);
    double xdot_update = (-2 * m_p_l * (theta_dot * theta_dot) * s + 3 * m_p * grav * s * c + 4 * action -
                          4 * b * x_dot) /
                         (4 * total_m - 3 * m_p * (c * c));
    double thetadot_update = (-3 * m_p_l * (theta_dot * theta_dot) * s * c + 6 * total_m * grav * s +
                              6 * (action - b * x_dot) * c) /
                             (4 * l * total_m - 3 * m_p_l * (c * c));
    xn[0] = pos + x_dot * dt;
    xn[2] = theta + theta_dot * dt;
    xn[1] = x_dot + xdot_update * dt;
    xn[3] = theta_dot + thetadot_update * dt;
}

/* ---------------------------------------------------------------- acrobot */
static void acrobot_dsdt(const double *p, const double *s, double *ds) {
    /* _dsdt, "book" branch (default) or "nips" branch (p[16] != 0)          acrobot_model.py:99-132 */
    const double m1 = p[1], m2 = p[2], l1 = p[3], lc1 = p[4], lc2 = p[5], I1 = p[6], I2 = p[7], g = p[8];
    const double a = s[4];
    const double theta1 = s[0], theta2 = s[1], dtheta1 = s[2], dtheta2 = s[3];
    double d1 = m1 * (lc1 * lc1) + m2 * ((l1 * l1) + (lc2 * lc2) + 2 * l1 * lc2 * cos(theta2)) + I1 + I2;
    double d2 = m2 * ((lc2 * lc2) + l1 * lc2 * cos(theta2)) + I2;
    double phi2 = m2 * lc2 * g * cos(theta1 + theta2 - M_PI / 2.);
    double phi1 = -m2 * l1 * lc2 * (dtheta2 * dtheta2) * sin(theta2) -
                  2 * m2 * l1 * lc2 * dtheta2 * dtheta1 * sin(theta2) +
                  (m1 * lc1 + m2 * l1) * g * cos(theta1 - M_PI / 2) + phi2;
    double ddtheta2;
    if (p[16] != 0.0) /* book_or_nips == "nips"                              acrobot_model.py:124-128 */
        ddtheta2 = (a + d2 / d1 * phi1 - phi2) / (m2 * (lc2 * lc2) + I2 - (d2 * d2) / d1);
    else
        ddtheta2 = (a + d2 / d1 * phi1 - m2 * l1 * lc2 * (dtheta1 * dtheta1) * sin(theta2) - phi2) /
                   (m2 * (lc2 * lc2) + I2 - (d2 * d2) / d1);
    double ddtheta1 = -(d2 * ddtheta2 + phi1) / d1;
    ds[0] = dtheta1;
    ds[1] = dtheta2;
    ds[2] = ddtheta1;
    ds[3] = ddtheta2;
    ds[4] = 0.;
}

static double wrap_angle(double x, double m, double M) { /* acrobot_model.py:185-203 */
    double diff = M - m;
    while (x > M) x = x - diff;
    while (x < m) x = x + diff;
    return x;
}

static void acrobot_step(const orc_model *mdl, const double *x, double u, const double *g, double *cost, double *xn) {
    const double *p = mdl->p;
    /* _transform + _state_cost                                              acrobot_model.py:162-176 */
    double z[5] = {-cos(x[0]) - cos(x[1] + x[0]) - 2, 0.0, 0.0, 0.0, u};
    double zd[5];
    for (int i = 0; i < 5; ++i) zd[i] = z[i] - g[i];
    double acc = state_cost(mdl, zd, 5, p + 11);
    *cost = acc;
    /* _dynamics: one rk4 step over t = [0, dt]                    acrobot_model.py:134-160,270-282 */
    const double dt = p[0] - 0.0, dt2 = dt / 2.0;
    double y0[5] = {x[0], x[1], x[2], x[3], u}, k1[5], k2[5], k3[5], k4[5], y[5];
    acrobot_dsdt(p, y0, k1);
    for (int i = 0; i < 5; ++i) y[i] = y0[i] + dt2 * k1[i];
    acrobot_dsdt(p, y, k2);
    for (int i = 0; i < 5; ++i) y[i] = y0[i] + dt2 * k2[i];
    acrobot_dsdt(p, y, k3);
    for (int i = 0; i < 5; ++i) y[i] = y0[i] + dt * k3[i];
    acrobot_dsdt(p, y, k4);
    for (int i = 0; i < 4; ++i) y[i] = y0[i] + dt / 6.0 * (k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i]);
    xn[0] = wrap_angle(y[0], -M_PI, M_PI);
    xn[1] = wrap_angle(y[1], -M_PI, M_PI);
    xn[2] = fmin(fmax(y[2], -p[9]), p[9]);
    xn[3] = fmin(fmax(y[3], -p[10]), p[10]);
}

/* Model.predict + get_reward: cost of (x,u) against goal row 0, then next state. */
void orc_step(const orc_model *m, const double *x, double u, const double *goal_nz, double *cost, double *xn) {
    switch (m->env) {
    case ORC_PENDULUM: pendulum_step(m, x, u, goal_nz, cost, xn); break;
    case ORC_CARTPOLE: cartpole_step(m, x, u, goal_nz, cost, xn); break;
    default: acrobot_step(m, x, u, goal_nz, cost, xn); break;
    }
}

/* numpy's pairwise summation (pairwise_sum_DOUBLE, blocks of 128, 8 accumulators), stride in elements */
static double pairwise_sum(const double *a, long n, long stride) {
    if (n < 8) {
        double res = 0.;
        for (long i = 0; i < n; i++) res += a[i * stride];
        return res;
    } else if (n <= 128) {
        double r[8], res;
        long i;
        for (i = 0; i < 8; ++i) r[i] = a[i * stride];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[(i + j) * stride];
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; i++) res += a[i * stride];
        return res;
    } else {
        long n2 = n / 2;
        n2 -= n2 % 8;
        return pairwise_sum(a, n2, stride) + pairwise_sum(a + n2 * stride, n - n2, stride);
    }
}

/* ---- a minimal pthread parallel-for over samples (stands in for the reference's Pool.map, mppi.py:57) ---- */
typedef void (*orc_range_fn)(void *ctx, int k0, int k1);
typedef struct {
    orc_range_fn fn;
    void *ctx;
    int k0, k1;
} orc_task;

static void *orc_task_main(void *arg) {
    orc_task *t = (orc_task *)arg;
    t->fn(t->ctx, t->k0, t->k1);
    return NULL;
}

int orc_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n < 1 ? 1 : (int)n;
}

static void orc_parallel_for(orc_range_fn fn, void *ctx, int K, int nthreads) {
    if (nthreads < 1) nthreads = orc_max_threads();
    if (nthreads > K) nthreads = K > 0 ? K : 1;
    if (nthreads > 256) nthreads = 256;
    if (nthreads == 1) {
        fn(ctx, 0, K);
        return;
    }
    pthread_t th[256];
    orc_task tasks[256];
    for (int i = 0; i < nthreads; ++i) {
        tasks[i].fn = fn;
        tasks[i].ctx = ctx;
        tasks[i].k0 = (int)((long)K * i / nthreads);
        tasks[i].k1 = (int)((long)K * (i + 1) / nthreads);
        pthread_create(&th[i], NULL, orc_task_main, &tasks[i]);
    }
    for (int i = 0; i < nthreads; ++i) pthread_join(th[i], NULL);
}

typedef struct {
    const orc_model *m;
    int H;
    const double *x0, *U, *eps, *goal;
    double lam, pinv;
    double *S;
    const double *xnoise; /* NULL, or [K][H][ns]: the draws of model.py:42-43 when the model has sensor noise */
} mppi_ctx;

static void mppi_range(void *vctx, int k0, int k1) {
    const mppi_ctx *c = (const mppi_ctx *)vctx;
    const int ns = c->m->ns, H = c->H;
    for (int k = k0; k < k1; ++k) {
        double x[4], xn[4], sample_cost = 0.0, cost;
        for (int i = 0; i < ns; ++i) x[i] = c->x0[i];
        for (int t = 0; t < H; ++t) {
            double e = c->eps[(long)k * H + t];
            double test_u = c->U[t] + e;                             /* mppi.py:39 */
            orc_step(c->m, x, test_u, c->goal + (long)t * c->m->goal_stride, &cost, xn); /* mppi.py:43-44 */
            sample_cost += cost + ((c->lam * test_u) * c->pinv) * e; /* mppi.py:45-46 */
            for (int i = 0; i < ns; ++i) x[i] = xn[i];
            if (c->xnoise) /* newstate += np.random.normal(0, sensor_std, state_size)   model.py:42-43 */
                for (int i = 0; i < ns; ++i) x[i] += c->xnoise[((long)k * H + t) * ns + i];
        }
        c->S[k] = sample_cost;
    }
}

/* MPPI.f for all K samples.  eps is the reference layout [K][H] (du = 1); goal = g_z[0]. */
void orc_mppi_costs(const orc_model *m, int K, int H, const double *x0, const double *U, const double *eps,
                    const double *goal, double lam, double sigma, double *S, int nthreads) {
    mppi_ctx c = {m, H, x0, U, eps, goal, lam, 1.0 / (sigma * sigma) /* np.linalg.pinv(std**2), 1x1 */, S, NULL};
    orc_parallel_for(mppi_range, &c, K, nthreads);
}

/* The same with a noisy model (Model.set_sensor_noise, model.py:20-24,42-43): xnoise [K][H][ns] holds the normal draws,
 * already scaled by sensor_std, that predict adds to each new state.  The reference draws them from numpy's global RNG
 * inside forked workers (not reproducible); handing them in makes the restatement deterministic. */
void orc_mppi_costs_noisy(const orc_model *m, int K, int H, const double *x0, const double *U, const double *eps,
                          const double *goal, double lam, double sigma, const double *xnoise, double *S, int nthreads) {
    mppi_ctx c = {m, H, x0, U, eps, goal, lam, 1.0 / (sigma * sigma), S, xnoise};
    orc_parallel_for(mppi_range, &c, K, nthreads);
}

/* Softmin update; U is updated in place (planned_us += erg).  w (nullable) receives the normalised weights. */
void orc_mppi_update(int K, int H, const double *S, const double *eps, double lam, double *U, double *w_out) {
    double beta = S[0];
    for (int k = 1; k < K; ++k) beta = fmin(beta, S[k]);
    double *tmp = (double *)malloc(sizeof(double) * K);
    double *p1 = (double *)malloc(sizeof(double) * K);
    for (int k = 0; k < K; ++k) tmp[k] = exp(-(1 / lam) * (S[k] - beta));
    double eta = pairwise_sum(tmp, K, 1);
    for (int k = 0; k < K; ++k) tmp[k] = (1 / eta) * tmp[k];
    for (int t = 0; t < H; ++t) {
        for (int k = 0; k < K; ++k) p1[k] = tmp[k] * eps[(long)k * H + t];
        U[t] += pairwise_sum(p1, K, 1);
    }
    if (w_out) memcpy(w_out, tmp, sizeof(double) * K);
    free(tmp);
    free(p1);
}

/* Agent.simulate_trajectory: goal row t per step (agent.py:103-104), returns xs[H+1][ns] and -sum(reward). */
void orc_simulate(const orc_model *m, int H, const double *x0, const double *us, const double *g_z, double *xs,
                  double *cost_out) {
    const int ns = m->ns, nz = m->ns + 1;
    double reward_sum = 0.0, c;
    for (int i = 0; i < ns; ++i) xs[i] = x0[i];
    for (int t = 0; t < H; ++t) {
        orc_step(m, xs + (long)t * ns, us[t], g_z + (long)t * nz, &c, xs + (long)(t + 1) * ns);
        reward_sum += -c;
    }
    *cost_out = -reward_sum;
}

typedef struct {
    const orc_model *m;
    int H;
    const double *x0, *samples, *goal;
    double *R;
} cem_ctx;

static void cem_range(void *vctx, int k0, int k1) {
    const cem_ctx *c = (const cem_ctx *)vctx;
    const int ns = c->m->ns, H = c->H;
    for (int k = k0; k < k1; ++k) {
        double x[4], xn[4], reward = 0.0, cost;
        for (int i = 0; i < ns; ++i) x[i] = c->x0[i];
        for (int t = 0; t < H; ++t) {
            orc_step(c->m, x, c->samples[(long)k * H + t], c->goal + (long)t * c->m->goal_stride, &cost, xn);
            reward += -cost; /* cem.py:46-47: reward += model.get_reward() */
            for (int i = 0; i < ns; ++i) x[i] = xn[i];
        }
        c->R[k] = reward;
    }
}

/* CEM.f rewards for K (already clipped) samples laid out [K][H]. */
void orc_cem_rewards(const orc_model *m, int K, int H, const double *x0, const double *samples, const double *goal,
                     double *R, int nthreads) {
    cem_ctx c = {m, H, x0, samples, goal, R};
    orc_parallel_for(cem_range, &c, K, nthreads);
}

typedef struct {
    double key;
    int idx;
} orc_kv;

static int cmp_desc(const void *a, const void *b) {
    const orc_kv *x = (const orc_kv *)a, *y = (const orc_kv *)b;
    if (x->key > y->key) return -1;
    if (x->key < y->key) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx); /* ties: lower index first (reference: unspecified) */
}

/* CEM._calc_action.  z: standard normals [max_iter][K][H] (np.random.normal(mu, sd) == mu + sd*z bitwise).
 * mu (planned_us) is updated in place.  Optional per-iteration outputs (nullable): rewards[it][K],
 * elites[it][n_elite] (sorted ascending), mu_hist[it][H], sd_hist[it][H].  Returns iterations executed. */
int orc_cem_solve(const orc_model *m, int K, int H, int max_iter, int n_elite, double epsilon, double alpha,
                  double sigma, double lo, double hi, const double *x0, const double *goal, const double *z,
                  double *mu, double *rewards, int *elites, double *mu_hist, double *sd_hist, int nthreads) {
    double *sd = (double *)malloc(sizeof(double) * H);
    double *samples = (double *)malloc(sizeof(double) * (size_t)K * H);
    double *R = (double *)malloc(sizeof(double) * K);
    orc_kv *kv = (orc_kv *)malloc(sizeof(orc_kv) * K);
    for (int t = 0; t < H; ++t) sd[t] = sigma; /* std = self.std.copy(), cem.py:56 */
    int it;
    for (it = 0; it < max_iter; ++it) {
        const double *zi = z + (size_t)it * K * H;
        for (int k = 0; k < K; ++k)
            for (int t = 0; t < H; ++t)
                samples[(size_t)k * H + t] = clipd(mu[t] + sd[t] * zi[(size_t)k * H + t], lo, hi);
        orc_cem_rewards(m, K, H, x0, samples, goal, R, nthreads);
        for (int k = 0; k < K; ++k) {
            kv[k].key = R[k];
            kv[k].idx = k;
        }
        qsort(kv, K, sizeof(orc_kv), cmp_desc); /* np.argsort(-rewards)[:n_elite], cem.py:67 */
        int done = 1;
        for (int t = 0; t < H; ++t) {
            /* np.mean / np.std over axis 0 add the elite rows sequentially, in sorted order */
            double s = 0.0;
            for (int e = 0; e < n_elite; ++e) s += samples[(size_t)kv[e].idx * H + t];
            double mean = s / n_elite;
            double v = 0.0;
            for (int e = 0; e < n_elite; ++e) {
                double d = samples[(size_t)kv[e].idx * H + t] - mean;
                v += d * d;
            }
            double new_std = sqrt(v / n_elite);
            mu[t] = alpha * mu[t] + (1 - alpha) * mean;
            sd[t] = alpha * sd[t] + (1 - alpha) * new_std;
            if (!(sd[t] < epsilon)) done = 0;
        }
        if (rewards) memcpy(rewards + (size_t)it * K, R, sizeof(double) * K);
        if (elites) {
            int *dst = elites + (size_t)it * n_elite;
            for (int e = 0; e < n_elite; ++e) dst[e] = kv[e].idx;
            for (int a = 1; a < n_elite; ++a) { /* insertion sort ascending */
                int v = dst[a], b = a - 1;
                while (b >= 0 && dst[b] > v) {
                    dst[b + 1] = dst[b];
                    --b;
                }
                dst[b + 1] = v;
            }
        }
        if (mu_hist) memcpy(mu_hist + (size_t)it * H, mu, sizeof(double) * H);
        if (sd_hist) memcpy(sd_hist + (size_t)it * H, sd, sizeof(double) * H);
        if (done) {
            ++it;
            break;
        }
    }
    free(sd);
    free(samples);
    free(R);
    free(kv);
    return it;
}

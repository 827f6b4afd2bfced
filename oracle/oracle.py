"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/liboracle.so (the plain-C restatement in mpc_oracle.c).

May be imported only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package never imports this module; it fails loudly when its CUDA library is missing instead.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")

ENV_IDS = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
STATE_SIZE = {"pendulum": 2, "cartpole": 4, "acrobot": 4}
BOUNDS = {"pendulum": (-2.0, 2.0), "cartpole": (-1.0, 1.0), "acrobot": (-1.0, 1.0)}


class OrcModel(C.Structure):
    _fields_ = [("env", C.c_int), ("ns", C.c_int), ("p", C.c_double * 24), ("full_w", C.c_int),
                ("goal_stride", C.c_int), ("W", C.c_double * 25)]


def default_params(env, g=10.0):
    """Constructor defaults of the three reference models (pendulum_model.py:8-22,
    cartpole_swingup_model.py:23-46, acrobot_model.py:60-97,107)."""
    if env == "pendulum":
        return [g, 1.0, 1.0, 0.05, 8.0, 2.0, 1.0, 0.1, 0.001]
    if env == "cartpole":
        return [9.82, 0.5, 0.5, 0.6, 10.0, 0.05, 0.1, 2.4, -1.0, -0.0, -5.0, -0.0, -0.0]
    if env == "acrobot":
        return [0.2, 1.0, 1.0, 1.0, 0.5, 0.5, 1.0, 1.0, 9.8, 4 * np.pi, 9 * np.pi, 1.0, 0.0, 0.0, 0.0, 0.0,
                0.0]  # last: book_or_nips == "nips" flag (acrobot_model.py:124), "book" by default
    raise ValueError(env)


def build(force=False):
    src = os.path.join(HERE, "mpc_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        mp = C.POINTER(OrcModel)
        L.orc_step.argtypes = [mp, dp, C.c_double, dp, dp, dp]
        L.orc_step.restype = None
        L.orc_mppi_costs.argtypes = [mp, C.c_int, C.c_int, dp, dp, dp, dp, C.c_double, C.c_double, dp, C.c_int]
        L.orc_mppi_costs.restype = None
        L.orc_mppi_costs_noisy.argtypes = [mp, C.c_int, C.c_int, dp, dp, dp, dp, C.c_double, C.c_double, dp, dp, C.c_int]
        L.orc_mppi_costs_noisy.restype = None
        L.orc_mppi_update.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, dp, dp]
        L.orc_mppi_update.restype = None
        L.orc_simulate.argtypes = [mp, C.c_int, dp, dp, dp, dp, dp]
        L.orc_simulate.restype = None
        L.orc_cem_rewards.argtypes = [mp, C.c_int, C.c_int, dp, dp, dp, dp, C.c_int]
        L.orc_cem_rewards.restype = None
        L.orc_cem_solve.argtypes = [mp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                    C.c_double, C.c_double, dp, dp, dp, dp, dp, ip, dp, dp, C.c_int]
        L.orc_cem_solve.restype = C.c_int
        L.orc_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


def model(env, params=None, W=None, per_step_goal=False, **kw):
    """W: the model's whole cost matrix self.W (any nz x nz matrix); per_step_goal: the sample rollouts use goal row t
    at step t (callers then pass goal as [H][nz]) — an extension, the reference's rollouts use row 0."""
    m = OrcModel()
    m.env = ENV_IDS[env]
    m.ns = STATE_SIZE[env]
    p = list(params) if params is not None else default_params(env, **kw)
    for i, v in enumerate(p):
        m.p[i] = float(v)
    if W is not None:
        m.full_w = 1
        for i, v in enumerate(np.asarray(W, dtype=np.float64).reshape(-1)):
            m.W[i] = float(v)
    m.goal_stride = (m.ns + 1) if per_step_goal else 0
    m._env_name = env
    return m


def max_threads():
    return int(lib().orc_max_threads())


def step(m, x, u, goal=None):
    """Model.predict + get_reward: returns (cost, next_state)."""
    ns = m.ns
    x_, xp = _d(x)
    g_, gp = _d(np.zeros(ns + 1) if goal is None else goal)
    cost = C.c_double()
    xn = np.zeros(ns)
    lib().orc_step(C.byref(m), xp, float(u), gp, C.byref(cost), xn.ctypes.data_as(C.POINTER(C.c_double)))
    return cost.value, xn


def mppi_costs(m, x0, U, eps, lam, sigma, goal=None, nthreads=1):
    """eps: [K,H] or [K,H,1] (reference layout).  Returns S[K]."""
    eps = np.asarray(eps, dtype=np.float64)
    K, H = eps.shape[0], eps.shape[1]
    x_, xp = _d(x0)
    U_, Up = _d(np.asarray(U).reshape(-1))
    e_, ep = _d(eps.reshape(K, H))
    g_, gp = _d(np.zeros(m.ns + 1) if goal is None else goal)
    S = np.zeros(K)
    lib().orc_mppi_costs(C.byref(m), K, H, xp, Up, ep, gp, float(lam), float(sigma),
                         S.ctypes.data_as(C.POINTER(C.c_double)), int(nthreads))
    return S


def mppi_costs_noisy(m, x0, U, eps, lam, sigma, xnoise, goal=None, nthreads=1):
    """MPPI.f with a model that has sensor noise (model.py:42-43): xnoise [K,H,ns] = the scaled normal draws."""
    eps = np.asarray(eps, dtype=np.float64)
    K, H = eps.shape[0], eps.shape[1]
    x_, xp = _d(x0)
    U_, Up = _d(np.asarray(U).reshape(-1))
    e_, ep = _d(eps.reshape(K, H))
    n_, np_ = _d(np.asarray(xnoise, dtype=np.float64).reshape(K, H, m.ns))
    g_, gp = _d(np.zeros(m.ns + 1) if goal is None else goal)
    S = np.zeros(K)
    lib().orc_mppi_costs_noisy(C.byref(m), K, H, xp, Up, ep, gp, float(lam), float(sigma), np_,
                               S.ctypes.data_as(C.POINTER(C.c_double)), int(nthreads))
    return S


def mppi_update(S, eps, lam, U):
    """Returns (U_new[H], w[K])."""
    eps = np.asarray(eps, dtype=np.float64)
    K, H = eps.shape[0], eps.shape[1]
    S_, Sp = _d(S)
    e_, ep = _d(eps.reshape(K, H))
    U2 = np.array(np.asarray(U).reshape(-1), dtype=np.float64)
    w = np.zeros(K)
    lib().orc_mppi_update(K, H, Sp, ep, float(lam), U2.ctypes.data_as(C.POINTER(C.c_double)),
                          w.ctypes.data_as(C.POINTER(C.c_double)))
    return U2, w


def mppi_solve(m, x0, U, eps, lam, sigma, goal=None, nthreads=1):
    """One MPPI._calc_action: returns (S[K], U_new[H], u0)."""
    S = mppi_costs(m, x0, U, eps, lam, sigma, goal, nthreads)
    U2, _ = mppi_update(S, eps, lam, U)
    return S, U2, U2[0]


def simulate(m, x0, us, g_z=None):
    us = np.asarray(us, dtype=np.float64).reshape(-1)
    H = us.shape[0]
    ns = m.ns
    x_, xp = _d(x0)
    u_, up = _d(us)
    g_, gp = _d(np.zeros((H, ns + 1)) if g_z is None else g_z)
    xs = np.zeros((H + 1, ns))
    cost = C.c_double()
    lib().orc_simulate(C.byref(m), H, xp, up, gp, xs.ctypes.data_as(C.POINTER(C.c_double)), C.byref(cost))
    return xs, cost.value


def cem_rewards(m, x0, samples, goal=None, nthreads=1):
    samples = np.asarray(samples, dtype=np.float64)
    K, H = samples.shape[0], samples.shape[1]
    x_, xp = _d(x0)
    s_, sp = _d(samples.reshape(K, H))
    g_, gp = _d(np.zeros(m.ns + 1) if goal is None else goal)
    R = np.zeros(K)
    lib().orc_cem_rewards(C.byref(m), K, H, xp, sp, gp, R.ctypes.data_as(C.POINTER(C.c_double)), int(nthreads))
    return R


def cem_solve(m, x0, mu, z, n_elite, epsilon, alpha, sigma, lo=None, hi=None, goal=None, nthreads=1):
    """z: [max_iter, K, H] standard normals.  Returns dict(iters, mu, rewards, elites, mu_hist, sd_hist, u0)."""
    z = np.asarray(z, dtype=np.float64)
    max_iter, K, H = z.shape[0], z.shape[1], z.shape[2]
    if lo is None:
        lo, hi = BOUNDS[m._env_name]
    x_, xp = _d(x0)
    z_, zp = _d(z.reshape(max_iter, K, H))
    g_, gp = _d(np.zeros(m.ns + 1) if goal is None else goal)
    mu2 = np.array(np.asarray(mu).reshape(-1), dtype=np.float64)
    rewards = np.zeros((max_iter, K))
    elites = np.zeros((max_iter, n_elite), dtype=np.int32)
    mu_hist = np.zeros((max_iter, H))
    sd_hist = np.zeros((max_iter, H))
    dp = C.POINTER(C.c_double)
    iters = lib().orc_cem_solve(C.byref(m), K, H, max_iter, int(n_elite), float(epsilon), float(alpha), float(sigma),
                                float(lo), float(hi), xp, gp, zp, mu2.ctypes.data_as(dp),
                                rewards.ctypes.data_as(dp), elites.ctypes.data_as(C.POINTER(C.c_int)),
                                mu_hist.ctypes.data_as(dp), sd_hist.ctypes.data_as(dp), int(nthreads))
    return dict(iters=iters, mu=mu2, rewards=rewards[:iters], elites=elites[:iters], mu_hist=mu_hist[:iters],
                sd_hist=sd_hist[:iters], u0=mu2[0])

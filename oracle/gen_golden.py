"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz by EXECUTING THE UNMODIFIED REFERENCE
(/root/reference/MPCBenchmark, imported through oracle/shims.py) on seeded inputs.

Run in the build container (the reference tree is not on the GPU box):
    python oracle/gen_golden.py

The reference has no tests / golden vectors of its own, so these files are what pins the
oracle (oracle/mpc_oracle.c) and, through it, the CUDA path.

How internals are observed without touching reference code:
  * the agents' `self.pool` (a multiprocessing.Pool, mppi.py:24 / cem.py:31) is replaced after construction by a
    serial, recording stand-in with the same `map` signature — exactly the commented-out serial path at
    mppi.py:58 / cem.py:65 — so per-sample costs (zeroed at mppi.py:73) and per-iteration CEM rewards and sample
    tensors are captured;
  * CEM's noise: np.random.normal(loc, scale, size) == loc + scale*standard_normal(size) bitwise for the legacy
    global RNG, so z = (samples_unclipped - mean)/std is recovered by replaying the seed with standard_normal and
    asserting the reconstruction reproduces the recorded (clipped) samples exactly.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import shims  # noqa: E402

shims.load_reference()
from MPCBenchmark.agents import CEM, MPPI  # noqa: E402
from MPCBenchmark.models import AcrobotModel, CartPoleSwingUpModel, PendulumModel  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
ENVS = {"pendulum": PendulumModel, "cartpole": CartPoleSwingUpModel, "acrobot": AcrobotModel}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}


class RecordingPool:
    """Serial stand-in for multiprocessing.Pool (same .map/.close/.join), recording every map call."""

    def __init__(self):
        self.calls = []

    def map(self, fn, inputs):
        out = [fn(i) for i in inputs]
        self.calls.append((inputs, out))
        return out

    def close(self):
        pass

    def join(self):
        pass


def make_agent(cls, model, params):
    agent = cls(model, params, cores=1)
    agent.pool.close()
    agent.pool.join()
    agent.pool = RecordingPool()
    return agent


def rand_state(env, rng, n):
    if env == "pendulum":
        return np.stack([rng.uniform(-3 * np.pi, 3 * np.pi, n), rng.uniform(-10, 10, n)], 1)
    if env == "cartpole":
        return np.stack([rng.uniform(-4, 4, n), rng.uniform(-5, 5, n),
                         rng.uniform(-2 * np.pi, 2 * np.pi, n), rng.uniform(-10, 10, n)], 1)
    return np.stack([rng.uniform(-3 * np.pi, 3 * np.pi, n), rng.uniform(-3 * np.pi, 3 * np.pi, n),
                     rng.uniform(-15, 15, n), rng.uniform(-30, 30, n)], 1)


def gen_step_kat():
    """Single Model.predict / get_reward known answers (model.py:26-44,62-63)."""
    rng = np.random.RandomState(1234)
    out = {}
    for env, cls in ENVS.items():
        model = cls()
        n = 256
        xs = rand_state(env, rng, n)
        # a few exact edge values: angle exactly at +-pi / 0, saturating controls, zero state
        xs[0] = 0.0
        xs[1, 0] = np.pi
        xs[2, 0] = -np.pi
        us = rng.uniform(-4, 4, (n, 1))
        us[3] = 0.0
        nz = xs.shape[1] + 1
        gs = rng.uniform(-1, 1, (n, nz))
        gs[: n // 2] = 0.0  # every reference script passes a zero goal
        cost = np.zeros(n)
        xn = np.zeros_like(xs)
        for i in range(n):
            # goal handed over like MPPI.f does: a [H, nz] array whose row 0 is used (model.py:34,38)
            g = np.repeat(gs[i][None, :], 3, axis=0)
            g[1:] += 7.0  # rows != 0 must be ignored
            xn[i] = model.predict(xs[i].copy(), us[i].copy(), goal=g)
            cost[i] = -model.get_reward()
        out.update({env + "_x": xs, env + "_u": us, env + "_g": gs, env + "_cost": cost, env + "_xn": xn})
    np.savez(os.path.join(OUT, "step_kat.npz"), **out)
    print("step_kat: ok")


def gen_step_kat_nips():
    """Acrobot with book_or_nips = "nips" (acrobot_model.py:124-128) on the states / controls / goals of step_kat.npz."""
    d = np.load(os.path.join(OUT, "step_kat.npz"))
    model = AcrobotModel()
    model.book_or_nips = "nips"
    xs, us, gs = d["acrobot_x"], d["acrobot_u"], d["acrobot_g"]
    cost, xn = np.zeros(len(xs)), np.zeros_like(xs)
    for i in range(len(xs)):
        xn[i] = model.predict(xs[i].copy(), us[i].copy(), goal=np.repeat(gs[i][None, :], 2, axis=0))
        cost[i] = -model.get_reward()
    np.savez(os.path.join(OUT, "step_kat_nips.npz"), acrobot_cost=cost, acrobot_xn=xn)
    print("step_kat_nips: ok, differs from book in %d of %d next states" % (
        int(np.any(xn != d["acrobot_xn"], axis=1).sum()), len(xs)))


def gen_fullw_cases():
    """A full (non-diagonal, non-symmetric, indefinite) cost matrix self.W — einsum("bi,ij,bj->b") in _state_cost
    (pendulum_model.py:46-50, cartpole_swingup_model.py:95-104, acrobot_model.py:172-176): single predicts, MPPI.f
    per-sample costs (mppi.py:35-48), CEM.f rewards (cem.py:42-48) and Agent.simulate_trajectory with a per-step goal
    trajectory (agent.py:97-108), all from the unmodified reference with only model.W replaced."""
    out = {}
    for ei, (env, cls) in enumerate(ENVS.items()):
        rng = np.random.RandomState(500 + ei)
        model = cls()
        ns = model.state_size
        nz = ns + 1
        W = np.array(model.W, dtype=np.float64) + rng.normal(0, 0.3, (nz, nz))
        model.W = W
        n = 64
        xs = rand_state(env, rng, n)
        us = rng.uniform(-2, 2, (n, 1))
        gs = rng.uniform(-1, 1, (n, nz))
        cost, xn = np.zeros(n), np.zeros_like(xs)
        for i in range(n):
            xn[i] = model.predict(xs[i].copy(), us[i].copy(), goal=gs[i][None, :])
            cost[i] = -model.get_reward()
        K, H, lam, std = 48, 15, 0.5, 1.0
        x = START[env] + rng.normal(0, 0.3, ns)
        U = rng.uniform(-0.5, 0.5, (H, 1))
        delta_u = rng.normal(0, std, (K, H, 1))
        g = rng.uniform(-0.5, 0.5, nz)
        g_z = np.repeat(g[None, :], H, axis=0)
        S = np.array([MPPI.f(model, x.copy(), U, delta_u[k], g_z, H, lam, np.ones((1, 1)) * std) for k in range(K)],
                     dtype=np.float64).reshape(K)
        lo, hi = model.bounds_low, model.bounds_high
        samples = np.clip(rng.normal(0, 1.0, (K, H, 1)), lo, hi)
        R = np.array([CEM.f(model, x.copy(), samples[k], g_z) for k in range(K)], dtype=np.float64).reshape(K)
        # nominal rollout against a goal TRAJECTORY (row t at step t): Agent.simulate_trajectory
        agent = make_agent(MPPI, model, {"K": 2, "T": H, "std": 1.0, "lam": 1.0})
        g_traj = rng.uniform(-0.5, 0.5, (H, nz))
        us_nom = np.clip(rng.uniform(-1.5, 1.5, (H, 1)), lo, hi)
        xs_nom, c_nom = agent.simulate_trajectory(x.copy(), us_nom, g_traj)
        agent.close()
        out.update({env + "_W": W, env + "_x": xs, env + "_u": us, env + "_g": gs, env + "_cost": cost, env + "_xn": xn,
                    env + "_x0": x, env + "_U": U, env + "_delta_u": delta_u, env + "_goal": g, env + "_S": S,
                    env + "_samples": samples, env + "_R": R, env + "_g_traj": g_traj, env + "_us_nom": us_nom,
                    env + "_xs_nom": xs_nom, env + "_c_nom": np.array(float(c_nom))})
    out["meta"] = np.array(json.dumps(dict(K=48, H=15, lam=0.5, std=1.0)))
    np.savez(os.path.join(OUT, "fullw_cases.npz"), **out)
    print("fullw_cases: ok")


def gen_mppi_cases():
    """MPPI._calc_action (mppi.py:54-75) with the agent's own delta_u; costs captured from the pool."""
    cases = [
        # env, K, H, lam, std, seed, nominal ("zero" | "rand"), goal ("zero" | "rand"), n_solves
        ("pendulum", 20, 50, 0.01, 1.0, 0, "zero", "zero", 3),
        ("pendulum", 257, 20, 0.5, 1.0, 1, "rand", "zero", 2),
        ("pendulum", 64, 5, 10.0, 0.5, 2, "rand", "rand", 1),
        ("pendulum", 1024, 50, 0.5, 1.0, 3, "zero", "zero", 1),
        ("pendulum", 128, 25, 0.001, 1.0, 4, "rand", "zero", 2),
        ("cartpole", 20, 50, 0.01, 1.0, 5, "zero", "zero", 3),
        ("cartpole", 300, 25, 0.5, 1.0, 6, "rand", "zero", 2),
        ("cartpole", 64, 10, 1.0, 0.5, 7, "rand", "rand", 1),
        ("cartpole", 1024, 50, 0.5, 1.0, 8, "zero", "zero", 1),
        ("acrobot", 20, 50, 0.01, 1.0, 9, "zero", "zero", 3),
        ("acrobot", 200, 100, 0.5, 1.0, 10, "rand", "zero", 2),
        ("acrobot", 64, 10, 0.1, 0.5, 11, "rand", "rand", 1),
        ("acrobot", 512, 50, 0.5, 1.0, 12, "zero", "zero", 1),
        ("pendulum", 1, 7, 0.5, 1.0, 13, "rand", "zero", 1),     # K = 1: weight is exactly 1
        ("cartpole", 33, 1, 0.5, 1.0, 14, "rand", "zero", 1),    # H = 1
    ]
    out, meta = {}, []
    for ci, (env, K, H, lam, std, seed, nominal, goal, n_solves) in enumerate(cases):
        np.random.seed(seed)
        model = ENVS[env]()
        agent = make_agent(MPPI, model, {"K": K, "T": H, "std": std, "lam": lam})
        rng = np.random.RandomState(1000 + seed)
        ns = model.state_size
        x = START[env] + rng.normal(0, 0.3, ns)
        if nominal == "rand":
            agent.planned_us[:] = rng.uniform(-1.5, 1.5, (H, 1))
        g = rng.uniform(-0.5, 0.5, ns + 1) if goal == "rand" else np.zeros(ns + 1)
        g_z = np.repeat(g[None, :], H, axis=0)
        p = "c%d_" % ci
        out[p + "x"] = x
        out[p + "goal"] = g
        out[p + "delta_u"] = agent.delta_u.copy()
        out[p + "U_in"] = agent.planned_us.copy()
        costs, U_out, u0s = [], [], []
        for _ in range(n_solves):
            u0 = agent._calc_action(x, g_z)
            u0s.append(np.array(u0, dtype=np.float64).copy())
            costs.append(np.array(agent.pool.calls[-1][1], dtype=np.float64))
            U_out.append(agent.planned_us.copy())   # NOT clipped between solves (bare _calc_action)
        out[p + "costs"] = np.stack(costs)
        out[p + "U_out"] = np.stack(U_out)
        out[p + "u0"] = np.stack(u0s)
        meta.append(dict(env=env, K=K, H=H, lam=lam, std=std, n_solves=n_solves))
    out["meta"] = np.array(json.dumps(meta))
    np.savez(os.path.join(OUT, "mppi_cases.npz"), **out)
    print("mppi_cases: ok (%d cases)" % len(cases))


def gen_cem_cases():
    """CEM._calc_action (cem.py:54-78); per-iteration rewards / samples captured from the pool."""
    cases = [
        # env, K, H, max_iter, n_elite, eps, alpha, std, seed, nominal
        ("pendulum", 64, 20, 5, 6, 1e-5, 0.2, 1.0, 0, "zero"),
        ("pendulum", 200, 10, 10, 50, 1e-5, 0.2, 1.0, 1, "rand"),
        ("cartpole", 128, 25, 5, 12, 1e-5, 0.2, 1.0, 2, "zero"),
        ("cartpole", 409, 50, 3, 40, 1e-5, 0.2, 1.0, 3, "rand"),
        ("acrobot", 64, 20, 5, 6, 1e-5, 0.2, 1.0, 4, "zero"),
        ("acrobot", 100, 50, 3, 100, 1e-5, 0.0, 0.5, 5, "rand"),       # n_elite == K, alpha = 0
        ("pendulum", 50, 8, 10, 5, 0.6, 0.5, 1.0, 6, "zero"),          # large epsilon: early break (cem.py:75)
        ("cartpole", 31, 5, 4, 1, 1e-5, 0.2, 1.0, 7, "rand"),          # n_elite = 1: std of one sample = 0
    ]
    out, meta = {}, []
    for ci, (env, K, H, max_iter, n_elite, eps, alpha, std, seed, nominal) in enumerate(cases):
        model = ENVS[env]()
        agent = make_agent(CEM, model, {"K": K, "T": H, "max_iter": max_iter, "n_elite": n_elite,
                                        "epsilon": eps, "alpha": alpha, "std": std})
        rng = np.random.RandomState(2000 + seed)
        ns = model.state_size
        x = START[env] + rng.normal(0, 0.3, ns)
        if nominal == "rand":
            agent.planned_us[:] = rng.uniform(-0.8, 0.8, (H, 1))
        g_z = np.zeros((H, ns + 1))
        U_in = agent.planned_us.copy()
        np.random.seed(seed)
        u0 = np.array(agent._calc_action(x, g_z)).copy()
        iters = len(agent.pool.calls)
        # replay the RNG to recover the standard normals and cross-check cem.py:58-60,67-73 step by step
        np.random.seed(seed)
        mu, sd = U_in.copy(), np.ones((H, 1)) * std
        lo, hi = agent.bounds_low, agent.bounds_high
        zs, rewards, elites, mus, sds = [], [], [], [], []
        for it in range(iters):
            z = np.random.standard_normal((K, H, 1))
            samples = np.clip(mu + sd * z, lo, hi)
            rec_samples = np.array([inp[2] for inp in agent.pool.calls[it][0]])
            assert np.array_equal(samples, rec_samples), "normal() != loc + scale*standard_normal()"
            r = np.array(agent.pool.calls[it][1], dtype=np.float64)
            order = np.argsort(-r)[:n_elite]
            el = samples[order]
            mu = alpha * mu + (1 - alpha) * np.mean(el, axis=0)
            sd = alpha * sd + (1 - alpha) * np.std(el, axis=0)
            zs.append(z[:, :, 0]); rewards.append(r); elites.append(np.sort(order)); mus.append(mu.copy()); sds.append(sd.copy())
        assert np.array_equal(mu, agent.planned_us), "replayed CEM refit differs from the reference"
        assert (iters == max_iter) or (sd < eps).all()
        p = "c%d_" % ci
        out[p + "x"] = x
        out[p + "U_in"] = U_in
        out[p + "z"] = np.stack(zs)                  # [iters, K, H]
        out[p + "rewards"] = np.stack(rewards)       # [iters, K]
        out[p + "elites"] = np.stack(elites)         # [iters, n_elite] sorted ascending
        out[p + "mu"] = np.stack(mus)                # [iters, H, 1]
        out[p + "sd"] = np.stack(sds)
        out[p + "u0"] = u0
        meta.append(dict(env=env, K=K, H=H, max_iter=max_iter, n_elite=n_elite, epsilon=eps, alpha=alpha,
                         std=std, iters=iters))
    out["meta"] = np.array(json.dumps(meta))
    np.savez(os.path.join(OUT, "cem_cases.npz"), **out)
    print("cem_cases: ok (%d cases)" % len(cases))


def gen_episode_cases():
    """Agent.warm_start + predict_action closed loop (agent.py:30-76, Experiment.py:70-90) with the plant stepped
    by a second Model instance (env.py:41 without the pandas history)."""
    cases = [
        # agent, env, params, warm, steps, seed
        ("MPPI", "pendulum", {"K": 20, "T": 50, "std": 1.0, "lam": 0.01}, 5, 12, 0),
        ("MPPI", "acrobot", {"K": 20, "T": 50, "std": 1.0, "lam": 0.01}, 5, 8, 1),   # experiment_test.py:23-28
        ("MPPI", "cartpole", {"K": 50, "T": 25, "std": 1.0, "lam": 0.5}, 3, 8, 2),
        ("CEM", "pendulum", {"K": 40, "T": 15, "max_iter": 4, "n_elite": 8, "epsilon": 1e-5, "alpha": 0.2,
                             "std": 1.0}, 2, 6, 3),
        ("CEM", "cartpole", {"K": 50, "T": 20, "max_iter": 3, "n_elite": 5, "epsilon": 1e-5, "alpha": 0.2,
                             "std": 1.0}, 2, 5, 4),
    ]
    out, meta = {}, []
    for ci, (agent_name, env, params, warm, steps, seed) in enumerate(cases):
        np.random.seed(seed)
        model, plant = ENVS[env](), ENVS[env]()
        agent = make_agent(MPPI if agent_name == "MPPI" else CEM, model, params)
        ns = model.state_size
        goal = np.zeros(ns + 1)
        state = START[env].copy()
        p = "c%d_" % ci
        if agent_name == "MPPI":
            out[p + "delta_u"] = agent.delta_u.copy()
        out[p + "rng_state_seed"] = np.array(seed)
        agent.warm_start(state, warm, goal_state=goal)
        out[p + "U_after_warm"] = agent.planned_us.copy()
        acts, Us, xs_hist, cost_hist, states = [], [], [], [], [state.copy()]
        for _ in range(steps):
            a = agent.predict_action(state, goal_state=goal)
            acts.append(np.array(a).copy())
            Us.append(agent.planned_us.copy())
            xs_hist.append(agent.planned_xs.copy())
            cost_hist.append(float(agent.planned_costs))
            state = plant.predict(state, np.array(a)).copy()
            states.append(state.copy())
        out[p + "actions"] = np.stack(acts)
        out[p + "U"] = np.stack(Us)
        out[p + "planned_xs"] = np.stack(xs_hist)
        out[p + "planned_costs"] = np.array(cost_hist)
        out[p + "states"] = np.stack(states)
        meta.append(dict(agent=agent_name, env=env, params=params, warm=warm, steps=steps, seed=seed))
    out["meta"] = np.array(json.dumps(meta))
    np.savez(os.path.join(OUT, "episode_cases.npz"), **out)
    print("episode_cases: ok (%d cases)" % len(cases))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] in ("nips", "fullw"):  # only an added fixture; the others stay byte-identical
        {"nips": gen_step_kat_nips, "fullw": gen_fullw_cases}[sys.argv[1]]()
        sys.exit(0)
    gen_step_kat()
    gen_step_kat_nips()
    gen_fullw_cases()
    gen_mppi_cases()
    gen_cem_cases()
    gen_episode_cases()

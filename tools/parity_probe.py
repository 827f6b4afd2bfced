#!/usr/bin/env python
"""Development probe (GPU box): the float32 flavours' per-sample cost error against the CPU oracle (float64, the pinned
restatement of the reference) on the identical noise tensor, at the shapes and seeds the parity tests use — printed as
p50 / p95 / p99 / MAX and the number of samples beyond 1e-5, so that the asserted bars are the measured ones.
Output: one JSON line per case (profiles/r2*_parity_probe.jsonl).
    python tools/parity_probe.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402,F401
from mpc_b200 import capi  # noqa: E402
from oracle import oracle as orc  # noqa: E402

ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}


def rand_case(env, K, H, seed, sigma=1.0):
    rng = np.random.RandomState(seed)
    x = START[env] + rng.normal(0, 0.2, orc.STATE_SIZE[env])
    U = rng.uniform(-0.5, 0.5, H)
    eps = rng.normal(0, sigma, (K, H))
    return x, U, eps


def stats(rel):
    return dict(p50=float(np.quantile(rel, .5)), p95=float(np.quantile(rel, .95)), p99=float(np.quantile(rel, .99)),
                max=float(rel.max()), n_over_1e5=int((rel > 1e-5).sum()), n=int(rel.size))


def mppi_case(env, K, H, lam, seed, flavours):
    x, U, eps = rand_case(env, K, H, seed)
    S_ref, U_ref, _ = orc.mppi_solve(orc.model(env), x, U, eps, lam, 1.0, nthreads=0)
    for trig, lanes in flavours:
        rec = dict(kind="mppi", env=env, K=K, H=H, lam=lam, seed=seed, trig=trig, lanes=lanes)
        try:
            if trig > 1:
                os.environ["MPC_B200_STREAM"] = "1"  # the per-stage splits exist for the STREAM flavour only
            else:
                os.environ.pop("MPC_B200_STREAM", None)
            s = capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env],
                            dtype="f32", lam=lam, sigma=1.0, fast_math=trig, lanes=lanes)
            s.set_noise(eps)
            u0, Un, xs, cost = s.mppi_solve(x, None, U)
            S = s.mppi_costs()[0]
            rec["cost_rel"] = stats(np.abs(S - S_ref) / np.abs(S_ref))
            rec["plan_abs_max"] = float(np.abs(Un[0] - U_ref).max())
            rec["plan_bound"] = 6e-8 * float(np.max(np.abs(S_ref))) / lam * 4 + 1e-5
            rec["stream_rows"] = s.geometry["stream_rows"]
            s.close()
        except Exception as e:
            rec["error"] = repr(e)
        print(json.dumps(rec), flush=True)


def cem_case(env, K, H, iters, n_elite, seed, trig):
    rng = np.random.RandomState(seed)
    x = START[env] + rng.normal(0, 0.2, orc.STATE_SIZE[env])
    z = rng.standard_normal((iters, K, H))
    ref = orc.cem_solve(orc.model(env), x, np.zeros(H), z, n_elite, 1e-5, 0.2, 1.0, nthreads=0)
    rec = dict(kind="cem", env=env, K=K, H=H, iters=iters, n_elite=n_elite, seed=seed, trig=trig)
    try:
        s = capi.Solver("CEM", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], dtype="f32",
                        sigma=1.0, max_iter=iters, n_elite=n_elite, epsilon=1e-5, alpha=0.2, fast_math=trig)
        s.cem_begin(x, None, np.zeros(H))
        for it in range(iters):
            s.cem_iterate(z[it])
        u0, U, xs, cost = s.cem_end()
        its = []
        for it in range(iters):
            R = s.cem_rewards(it)
            el = np.sort(s.cem_elites(it))
            Rr = ref["rewards"][it]
            common = np.intersect1d(el, ref["elites"][it]).size
            # how far from the elite threshold (in units of the reward) are the samples the two sets disagree on?
            diff = np.setxor1d(el, ref["elites"][it])
            thr = np.sort(Rr)[::-1][n_elite - 1]
            gap = float(np.max(np.abs(Rr[diff] - thr) / abs(thr))) if diff.size else 0.0
            its.append(dict(reward_rel=stats(np.abs(R - Rr) / np.abs(Rr)), overlap=common / n_elite,
                            max_rel_gap_of_disputed=gap))
        rec["iters_stats"] = its
        rec["mu_abs_max"] = float(np.abs(U - ref["mu"]).max())
        rec["sd_abs_max"] = float(np.abs(s.cem_std() - ref["sd_hist"][-1]).max())
        s.close()
    except Exception as e:
        rec["error"] = repr(e)
    print(json.dumps(rec), flush=True)


if __name__ == "__main__" and len(sys.argv) == 1:
    orc.build()
    ALL = [(0, 1), (0, 2), (1, 1), (1, 2)]
    for env in ("pendulum", "cartpole", "acrobot"):
        for K, H, lam in ((4096, 50, 0.5), (4100, 37, 1.0)):
            mppi_case(env, K, H, lam, K + H, ALL)
    for env, H in (("pendulum", 50), ("acrobot", 100), ("cartpole", 50)):
        for seed in (17, 18, 19):
            fl = ALL + ([(2, 1), (2, 2), (4, 1), (4, 2), (5, 2)] if env == "acrobot" else [])
            mppi_case(env, 8192, H, 0.5, seed, fl)
    mppi_case("acrobot", 65536, 100, 0.5, 3, ALL)
    mppi_case("pendulum", 65536, 50, 0.5, 3, ALL)
    mppi_case("pendulum", 65536, 50, 0.001, 3, [(0, 2), (1, 2)])
    for env in ("pendulum", "cartpole", "acrobot"):
        for trig in (0, 1):
            cem_case(env, 4096, 50, 5, 409, 11, trig)


def golden_cases():
    """The committed reference fixtures (tests/golden/mppi_cases.npz) through both float32 trig flavours."""
    d = np.load(os.path.join(ROOT, "tests", "golden", "mppi_cases.npz"))
    meta = json.loads(str(d["meta"]))
    for ci, mt in enumerate(meta):
        p = "c%d_" % ci
        K, H = mt["K"], mt["H"]
        for trig in (0, 1):
            s = capi.Solver("MPPI", ENV_ID[mt["env"]], orc.default_params(mt["env"]), K=K, H=H, bounds=orc.BOUNDS[mt["env"]],
                            dtype="f32", lam=mt["lam"], sigma=mt["std"], fast_math=trig)
            s.set_noise(d[p + "delta_u"].reshape(K, H))
            g_z = np.repeat(d[p + "goal"][None, :], H, axis=0)
            U = d[p + "U_in"].reshape(-1).copy()
            for sv in range(mt["n_solves"]):
                s.mppi_solve(d[p + "x"], g_z, U)
                S, Sref = s.mppi_costs()[0], d[p + "costs"][sv]
                print(json.dumps(dict(kind="golden", case=ci, solve=sv, trig=trig, **mt,
                                      cost_rel=stats(np.abs(S - Sref) / np.abs(Sref)))), flush=True)
                U = d[p + "U_out"][sv].reshape(-1).copy()
            s.close()


def energetic(env, K, H, seed, umax, flavours):
    """Plans of large amplitude (the reference does not clip the plan inside a rollout): the energetic regime."""
    rng = np.random.RandomState(seed)
    x = START[env] + rng.normal(0, 0.5, orc.STATE_SIZE[env])
    U = rng.uniform(-umax, umax, H)
    eps = rng.normal(0, 1.0, (K, H))
    S_ref, U_ref, _ = orc.mppi_solve(orc.model(env), x, U, eps, 0.5, 1.0, nthreads=0)
    for trig, lanes in flavours:
        s = capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], dtype="f32",
                        lam=0.5, sigma=1.0, fast_math=trig, lanes=lanes)
        s.set_noise(eps)
        s.mppi_solve(x, None, U)
        S = s.mppi_costs()[0]
        print(json.dumps(dict(kind="energetic", env=env, K=K, H=H, seed=seed, umax=umax, trig=trig, lanes=lanes,
                              cost_rel=stats(np.abs(S - S_ref) / np.abs(S_ref)))), flush=True)
        s.close()


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "more":
    golden_cases()
    for env, H in (("pendulum", 50), ("acrobot", 50), ("acrobot", 25), ("acrobot", 100)):
        for seed in (1, 2, 3):
            energetic(env, 16384, H, seed, 1.5, [(0, 2), (1, 2)])

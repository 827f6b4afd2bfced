#!/usr/bin/env python
"""Development probe (GPU box): solve-kernel time over the ring shape (warps W x stages S, set through the
MPC_B200_WARPS / MPC_B200_STAGES tuning knobs) for the three K = 2^20 workloads, plus the fixed per-solve latency
at K = 64 (one warp tile: launch + load + rollout + tail).
    python tools/tune_probe.py [env ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from mpc_b200.models import ENV_NAMES, env_spec  # noqa: E402

START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}


def solver(env, K, H, **kw):
    model = ENV_NAMES[env]()
    env_id, env_params = env_spec(model)
    return capi.Solver("MPPI", env_id, env_params, K=K, H=H, bounds=(float(model.bounds_low[0]), float(model.bounds_high[0])),
                       lam=0.5, sigma=1.0, **kw)


def time_solver(s, env, H, n=8, flush=True):
    s.set_profiling(True, rollout_events=True)
    s.upload_inputs(START[env] + 0.01, None, np.zeros(H))
    for _ in range(3):
        s.solve_enqueue()
        if flush:
            s.flush_l2()
    s.sync()
    roll = []
    for _ in range(n):
        s.solve_enqueue()
        roll.append(s.rollout_ms())
        if flush:
            s.flush_l2()
    s.sync()
    return float(np.median(roll))


def main():
    envs = sys.argv[1:] or ["pendulum", "cartpole", "acrobot"]
    mpc_b200.build_library()
    for env in envs:
        H = 100 if env == "acrobot" else 50
        for noise in ("tensor", "philox"):
            shapes = [(0, 0)]
            if noise == "tensor":
                shapes += [(w, s) for s in ((8, 7, 6) if H == 100 else (17, 16, 14, 12)) for w in range(4, 17) if w <= s]
            else:
                shapes += [(w, w) for w in ((4, 6, 8) if H == 100 else (8, 12, 16))]
            for (w, st) in shapes:
                os.environ.pop("MPC_B200_WARPS", None)
                os.environ.pop("MPC_B200_STAGES", None)
                if w:
                    os.environ["MPC_B200_WARPS"] = str(w)
                    os.environ["MPC_B200_STAGES"] = str(st)
                s = solver(env, 1 << 20, H, dtype="f32", noise=noise)
                if noise == "tensor":
                    s.fill_noise()
                ms = time_solver(s, env, H)
                print("%-9s %-6s W=%2d S=%2d  solve kernel %.4f ms  %.3e steps/s  geo=%s" % (
                    env, noise, w, st, ms, (1 << 20) * H / (ms * 1e-3), s.geometry), flush=True)
                s.close()
        os.environ.pop("MPC_B200_WARPS", None)
        os.environ.pop("MPC_B200_STAGES", None)
        for K in (64, 4096, 65536):
            s = solver(env, K, H, dtype="f32", noise="tensor")
            s.fill_noise()
            ms = time_solver(s, env, H, n=20, flush=False)
            print("%-9s K=%6d fixed-latency probe: solve kernel %.4f ms  geo=%s" % (env, K, ms, s.geometry), flush=True)
            s.close()


if __name__ == "__main__":
    main()

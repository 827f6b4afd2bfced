#!/usr/bin/env python
"""Development probe (GPU box): Acrobot H=100 K=2^20 solve time for the whole-tile vs STREAM flavour of the solve
kernel, accurate vs MUFU trig, a few ring shapes.   python tools/stream_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def run(env, K, H, envvars, **kw):
    for k in ("MPC_B200_STREAM", "MPC_B200_STREAM_ROWS", "MPC_B200_STREAM_DEPTH", "MPC_B200_WARPS"):
        os.environ.pop(k, None)
    os.environ.update(envvars)
    s = capi.Solver("MPPI", {"pendulum": 0, "cartpole": 1, "acrobot": 2}[env], orc.default_params(env), K=K, H=H,
                    bounds=orc.BOUNDS[env], dtype="f32", lam=0.5, sigma=1.0, **kw)
    s.fill_noise()
    s.set_profiling(True)
    ns = 2 if env == "pendulum" else 4
    s.upload_inputs(np.zeros(ns) + 0.01, None, np.zeros(H))
    for _ in range(3):
        s.solve_enqueue()
    s.sync()
    tot, n = 0.0, 10
    for _ in range(n):
        s.timer_start()
        s.solve_enqueue()
        tot += s.timer_stop()
    s.sync()
    g = s.geometry
    print("%-9s H=%3d %-40s %-22s  %.4f ms  %.3e steps/s  block=%d smem=%d stages=%d rows=%d" % (
        env, H, str(envvars), str(kw), tot / n, K * H / (tot / n * 1e-3), g["block"], g["smem"], g["stages"],
        g["stream_rows"]), flush=True)
    s.close()


if __name__ == "__main__":
    mpc_b200.build_library()
    K = 1 << 20
    for fast in (0, 1):
        for lanes in (2, 1):
            run("acrobot", K, 100, {"MPC_B200_STREAM": "0"}, fast_math=fast, lanes=lanes)
            run("acrobot", K, 100, {"MPC_B200_STREAM": "1"}, fast_math=fast, lanes=lanes)
    for rows, depth in ((4, 4), (8, 2), (8, 6), (12, 4), (16, 3)):
        run("acrobot", K, 100, {"MPC_B200_STREAM": "1", "MPC_B200_STREAM_ROWS": str(rows),
                                "MPC_B200_STREAM_DEPTH": str(depth)}, fast_math=1, lanes=2)
    for W in (8, 12):
        run("acrobot", K, 100, {"MPC_B200_STREAM": "1", "MPC_B200_WARPS": str(W)}, fast_math=1, lanes=2)
    for env in ("pendulum", "cartpole"):
        for st in ("0", "1"):
            run(env, K, 50, {"MPC_B200_STREAM": st}, fast_math=0, lanes=2)

#!/usr/bin/env python
"""Development probe (GPU box): host-side p50 latency of mpc_mppi_solve at small K with and without the nominal
rollout (want_traj), to split the fixed per-solve cost into launch + tile + merge and the serial trajectory."""
import os
import sys
import time
import statistics

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import tune_probe as tp  # noqa: E402

for env in ("pendulum", "cartpole", "acrobot"):
    H = 100 if env == "acrobot" else 50
    for K in [int(v) for v in os.environ.get("PROBE_KS", "64,4096,65536").split(",")]:
        s = tp.solver(env, K, H, dtype="f32", noise="tensor")
        s.fill_noise()
        x = tp.START[env] + 0.01
        U = np.zeros(H)
        res = {}
        for traj in (True, False):
            for _ in range(20):
                s.mppi_solve(x, None, U, want_traj=traj)
            lat = []
            for _ in range(200):
                t = time.perf_counter()
                s.mppi_solve(x, None, U, want_traj=traj)
                lat.append(time.perf_counter() - t)
            res[traj] = 1e3 * statistics.median(lat)
        s.set_profiling(True, rollout_events=True)
        s.upload_inputs(x, None, U)
        ks = []
        for _ in range(20):
            s.solve_enqueue()
            ks.append(s.rollout_ms())
        print("%-9s K=%6d H=%3d  p50 with traj %.4f ms  without %.4f ms  kernel (with traj) %.4f ms  geo=%s" % (
            env, K, H, res[True], res[False], statistics.median(ks), {k: s.geometry[k] for k in ("grid", "block")}), flush=True)
        s.close()

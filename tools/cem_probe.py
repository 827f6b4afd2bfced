#!/usr/bin/env python
"""Development probe (GPU box): BASELINE configs[1] — CEM CartPole K=4096 H=50 n_elite=409, 5 iterations, Philox —
p50 latency through the host-buffer call and device time per solve.   python tools/cem_probe.py [n_solves]"""
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from oracle import oracle as orc  # noqa: E402

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    for resample in (False, True):
        s = capi.Solver("CEM", 1, orc.default_params("cartpole"), K=4096, H=50, dtype="f32", noise="philox", sigma=1.0,
                        max_iter=5, n_elite=409, epsilon=1e-5, alpha=0.2, bounds=orc.BOUNDS["cartpole"], resample=resample)
        x = np.array([0.0, 0.0, np.pi, 0.0]) + 0.01
        for _ in range(min(20, n)):
            s.cem_solve(x, None, np.zeros(50))
        lat = []
        for _ in range(n):
            t = time.perf_counter()
            out = s.cem_solve(x, None, np.zeros(50))
            lat.append(time.perf_counter() - t)
        s.timer_start()
        for _ in range(n):
            s.solve_enqueue()
        ms = s.timer_stop()
        print("resample=%d p50 %.1f us  device %.1f us per solve  iters %d  u0 %.6f" % (
            resample, 1e6 * statistics.median(lat), 1e3 * ms / n, out[4], out[0][0]), flush=True)
        s.close()

#!/usr/bin/env python
"""Development probe (GPU box): CEM CartPole H=50, elite ratio 0.1, 5 iterations, Philox normals (BASELINE configs[1] at
K = 4096, and larger K) — p50 latency through the host-buffer call, device time per solve, launches per solve and
block 0's phase stamps of iteration 1 of the single-launch kernel.   python tools/cem_probe.py [n_solves] [K ...]"""
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from mpc_b200.models import ENV_NAMES, env_spec  # noqa: E402

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    Ks = [int(v) for v in sys.argv[2:]] or [4096]
    model = ENV_NAMES["cartpole"]()
    env_id, env_params = env_spec(model)
    for K in Ks:
        for fused in ("1", "0"):
            os.environ["MPC_B200_CEM_FUSED"] = fused
            s = capi.Solver("CEM", env_id, env_params, K=K, H=50, dtype="f32", noise="philox", sigma=1.0, max_iter=5,
                            n_elite=max(1, K // 10), epsilon=1e-5, alpha=0.2, bounds=(-1.0, 1.0))
            s.set_profiling(True)
            x = np.array([0.0, 0.0, np.pi, 0.0]) + 0.01
            nn = n if K <= 65536 else max(5, n // 20)
            for _ in range(min(20, nn)):
                s.cem_solve(x, None, np.zeros(50))
            lat = []
            for _ in range(nn):
                t = time.perf_counter()
                out = s.cem_solve(x, None, np.zeros(50))
                lat.append(time.perf_counter() - t)
            l0 = s.launch_count
            s.timer_start()
            for _ in range(nn):
                s.solve_enqueue(want_traj=False)
            ms = s.timer_stop()
            launches = (s.launch_count - l0) / nn
            tr = ""
            if fused == "1":
                t = s.trace()
                tr = "  ".join("%s %.1f" % (k, v / 1e3) for k, v in t.items() if v is not None and k != "start")
            print("K=%d fused=%s p50 %.1f us  device %.1f us per solve  launches %.0f  iters %d  u0 %.6f | %s" % (
                K, fused, 1e6 * statistics.median(lat), 1e3 * ms / nn, launches, out[4], out[0][0], tr), flush=True)
            s.close()

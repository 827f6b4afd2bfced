#!/usr/bin/env python
"""Development probe (GPU box): for every float32 rollout flavour (lanes 1|2 x accurate|fast trig) print the parity of
the per-sample costs against the CPU oracle on identical noise, and the rollout-kernel time at K = 2^20.
    python tools/variant_probe.py [--quick]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from oracle import oracle as orc  # noqa: E402

ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}
FLOPS = {"pendulum": 27, "cartpole": 56, "acrobot": 211}


def solver(env, K, H, **kw):
    return capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], **kw)


def parity():
    orc.build()
    print("== parity of per-sample costs vs oracle (K=4096): rel err p50 / p95 / p99 / max ; plan abs err")
    for env, H in (("pendulum", 50), ("cartpole", 50), ("acrobot", 100)):
        rng = np.random.RandomState(11)
        K, lam = 4096, 0.5
        x = START[env] + rng.normal(0, 0.2, len(START[env]))
        U = rng.uniform(-0.5, 0.5, H)
        eps = rng.normal(0, 1.0, (K, H))
        S_ref, U_ref, _ = orc.mppi_solve(orc.model(env), x, U, eps, lam, 1.0, nthreads=0)
        for lanes in (1, 2):
            for fast in (0, 1):
                s = solver(env, K, H, dtype="f32", lam=lam, sigma=1.0, lanes=lanes, fast_math=fast)
                s.set_noise(eps)
                u0, Un, xs, cost = s.mppi_solve(x, None, U)
                S = s.mppi_costs()[0]
                rel = np.abs(S - S_ref) / np.abs(S_ref)
                print("%-9s lanes=%d fast=%d  %.2e %.2e %.2e %.2e   plan %.2e" % (
                    env, lanes, fast, np.quantile(rel, .5), np.quantile(rel, .95), np.quantile(rel, .99), rel.max(),
                    np.abs(Un[0] - U_ref).max()))
                s.close()


def timing(K=1 << 20):
    peak = capi.measure_fma_peak(0, "f32")
    print("== rollout kernel at K=%d (FMA peak %.1f TFLOP/s): ms, steps/s, frac of FP32 roofline" % (K, peak))
    for env, H in (("pendulum", 50), ("cartpole", 50), ("acrobot", 100)):
        for noise in ("tensor", "philox"):
            for lanes in (1, 2):
                for fast in (0, 1):
                    s = solver(env, K, H, dtype="f32", noise=noise, lam=0.5, sigma=1.0, lanes=lanes, fast_math=fast)
                    if noise == "tensor":
                        s.fill_noise()
                    s.set_profiling(True, rollout_events=True)
                    s.upload_inputs(START[env] + 0.01, None, np.zeros(H))
                    for _ in range(3):
                        s.solve_enqueue()
                        s.flush_l2()
                    s.sync()
                    tot, roll = 0.0, 0.0
                    n = 5
                    for _ in range(n):
                        s.timer_start()
                        s.solve_enqueue()
                        tot += s.timer_stop()
                        roll += s.rollout_ms()
                        s.flush_l2()
                    s.sync()
                    sps = K * H / (roll / n * 1e-3)
                    print("%-9s %-6s lanes=%d fast=%d  rollout %.4f ms  solve %.4f ms  %.3e steps/s  frac %.3f  grid=%d smem=%d"
                          % (env, noise, lanes, fast, roll / n, tot / n, sps, sps * FLOPS[env] / (peak * 1e12),
                             s.geometry["grid"], s.geometry["smem"]))
                    s.close()


if __name__ == "__main__":
    mpc_b200.build_library()
    parity()
    timing()

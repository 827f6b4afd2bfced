#!/usr/bin/env python
"""Development probe (GPU box): phase stamps of the finalising block (mpc_get_trace) for the K = 2^20 workloads on
one GPU and, when two GPUs are visible, on two linked ranks driven by this one process (peer exchange in the kernel).
    python tools/trace_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import tune_probe as tp  # noqa: E402
from mpc_b200 import capi  # noqa: E402
import torch  # noqa: E402


def run(env, K, H, world):
    ss = [tp.solver(env, K, H, dtype="f32", noise="tensor", device=r, rank=r, world=world) for r in range(world)]
    if world > 1:
        capi.Solver.p2p_link(ss)
    for s in ss:
        s.fill_noise()
        s.set_profiling(True)
        s.upload_inputs(tp.START[env] + 0.01, None, np.zeros(H))
    acc = []
    for it in range(12):
        for s in ss:
            s.solve_enqueue()
        for s in ss:
            s.sync()
        if it >= 4:
            acc.append([s.trace() for s in ss])
    for r in range(world):
        keys = [k for k in acc[0][r] if acc[0][r][k] is not None]
        med = {k: float(np.median([a[r][k] for a in acc])) / 1e3 for k in keys}
        print("%-9s K=%8d world=%d rank=%d  us since kernel start: %s" % (
            env, K, world, r, "  ".join("%s %.1f" % (k, med[k]) for k in keys if k != "start")), flush=True)
    for s in ss:
        s.close()


if __name__ == "__main__":
    n = torch.cuda.device_count()
    for env, H in (("pendulum", 50), ("cartpole", 50), ("acrobot", 100)):
        for K in (4096, 1 << 20):
            run(env, K, H, 1)
            if n >= 2:
                run(env, K, H, 2)

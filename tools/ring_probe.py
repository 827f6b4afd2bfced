#!/usr/bin/env python
"""Development probe (GPU box): warps x ring stages of the whole-tile solve kernel at K = 2^20, H = 50 — kernel ms
(median of back-to-back solves, no logging rollout) and the warps' exit spread of the finalising block."""
import json
import os
import statistics
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402,F401
from mpc_b200 import capi  # noqa: E402
from mpc_b200.models import ENV_NAMES, env_spec  # noqa: E402

START = {"pendulum": [np.pi, 0.0], "cartpole": [0.0, 0.0, np.pi, 0.0]}
for env, trig in (("pendulum", 1), ("cartpole", 0)):
    for W in (12, 13, 14, 15, 16):
        for S in (16, 17):
            if S < W:
                continue
            os.environ["MPC_B200_WARPS"], os.environ["MPC_B200_STAGES"] = str(W), str(S)
            model = ENV_NAMES[env]()
            env_id, params = env_spec(model)
            s = capi.Solver("MPPI", env_id, params, K=1 << 20, H=50, dtype="f32", noise="tensor", lam=0.5, sigma=1.0,
                            bounds=(float(model.bounds_low[0]), float(model.bounds_high[0])), fast_math=trig, lanes=2)
            s.fill_noise()
            s.set_profiling(True)
            s.upload_inputs(np.array(START[env]) + 0.01, None, np.zeros(50))
            s.time_solves(3, want_traj=False)
            ms = statistics.median(s.time_solves(12, want_traj=False))
            msf = statistics.median(s.time_solves(12, want_traj=False, flush_l2=True))
            t = s.trace()
            print(json.dumps(dict(env=env, W=W, S=s.geometry["stages"], block=s.geometry["block"], ms=ms, ms_flushed=msf,
                                  first_warp_out=t["first_warp_out"] / 1e3, last_warp_out=t["last_warp_out"] / 1e3,
                                  end=t["end"] / 1e3)), flush=True)
            s.close()

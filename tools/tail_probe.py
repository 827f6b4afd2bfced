#!/usr/bin/env python
"""Development probe (GPU box): tail of the solve kernel (last warp out -> applied) with the two-level merge and with the
flat merge (MPC_B200_MERGE_FLAT=1), K = 2^20, H = 50."""
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
    for flat in ("0", "1"):
        os.environ["MPC_B200_MERGE_FLAT"] = flat
        model = ENV_NAMES[env]()
        env_id, params = env_spec(model)
        s = capi.Solver("MPPI", env_id, params, K=1 << 20, H=50, dtype="f32", noise="tensor", lam=0.5, sigma=1.0,
                        bounds=(float(model.bounds_low[0]), float(model.bounds_high[0])), fast_math=trig, lanes=2)
        s.fill_noise()
        s.set_profiling(True)
        s.upload_inputs(np.array(START[env]) + 0.01, None, np.zeros(50))
        s.time_solves(3, want_traj=False)
        msf = statistics.median(s.time_solves(20, want_traj=False, flush_l2=True))
        t = {k: round(v / 1e3, 1) for k, v in s.trace().items() if v is not None and k != "start"}
        print(json.dumps(dict(env=env, flat=flat, ms_flushed=msf, **t)), flush=True)
        s.close()

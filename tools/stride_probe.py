import json, os, statistics, sys
import numpy as np
sys.path.insert(0, "/root/repo")
import mpc_b200
from mpc_b200 import capi
from mpc_b200.models import ENV_NAMES, env_spec
START = {"pendulum": [np.pi, 0.0], "cartpole": [0.0, 0.0, np.pi, 0.0], "acrobot": [0, 0, 0, 0]}
for env, trig, H, lanes in (("pendulum", 1, 50, 2), ("cartpole", 0, 50, 2), ("acrobot", 1, 100, 1)):
    for K in (1 << 20, (1 << 20) + 128 * 7, (1 << 20) + 128 * 37, (1 << 20) - 128 * 53):
        model = ENV_NAMES[env]()
        env_id, params = env_spec(model)
        s = capi.Solver("MPPI", env_id, params, K=K, H=H, dtype="f32", noise="tensor", lam=0.5, sigma=1.0,
                        bounds=(float(model.bounds_low[0]), float(model.bounds_high[0])), fast_math=trig, lanes=lanes)
        s.fill_noise()
        s.upload_inputs(np.array(START[env]) + 0.01, None, np.zeros(H))
        s.time_solves(3, want_traj=False)
        ms = statistics.median(s.time_solves(12, want_traj=False))
        msf = statistics.median(s.time_solves(12, want_traj=False, flush_l2=True))
        print(env, K, "ms %.4f flushed %.4f  ns per 1k samples %.2f" % (ms, msf, 1e6 * msf / K * 1000), flush=True)
        s.close()

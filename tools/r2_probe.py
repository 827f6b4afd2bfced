#!/usr/bin/env python
"""Round-2 variant probe (run on the B200 box).  For every (workload, lanes, trig flavour, whole-tile | STREAM | WIDE)
combination: block shape, device time per solve (median of n back-to-back solves, own event pair each), and parity of
the per-sample costs / plan against the library's own float64 flavour on the identical (Philox-filled) noise at
K = 65536.  Output: one JSON line per configuration (profiles/r2*_variant_probe.jsonl)."""
import json
import os
import statistics
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402
from mpc_b200 import capi  # noqa: E402
from mpc_b200.models import ENV_NAMES, env_spec  # noqa: E402

START = {"pendulum": [np.pi, 0.0], "cartpole": [0.0, 0.0, np.pi, 0.0], "acrobot": [0.0, 0.0, 0.0, 0.0]}
FLOPS = {"pendulum": 27, "cartpole": 56, "acrobot": 211}
K_PAR = 65536
which = sys.argv[1] if len(sys.argv) > 1 else "all"
K_BIG = int(os.environ.get("PROBE_K", 1 << 20))


def solver(env, K, H, dtype, **kw):
    model = ENV_NAMES[env]()
    env_id, env_params = env_spec(model)
    lo, hi = float(model.bounds_low[0]), float(model.bounds_high[0])
    return capi.Solver("MPPI", env_id, env_params, K=K, H=H, dtype=dtype, noise="tensor", bounds=(lo, hi), lam=0.5,
                       sigma=1.0, seed=0, **kw)


def setenv(envs):
    for k in ("MPC_B200_STREAM", "MPC_B200_WIDE", "MPC_B200_WARPS", "MPC_B200_STAGES", "MPC_B200_STREAM_ROWS",
              "MPC_B200_STREAM_DEPTH"):
        os.environ.pop(k, None)
    os.environ.update(envs)


ref_cache = {}


def reference(env, H):
    if (env, H) not in ref_cache:
        setenv({})
        s = solver(env, K_PAR, H, "f64")
        s.fill_noise()
        x = np.array(START[env]) + 0.01
        u0, U, _, _ = s.mppi_solve(x, None, np.zeros(H), want_traj=False)
        ref_cache[(env, H)] = (s.mppi_costs()[0].copy(), U[0].copy())
        s.close()
    return ref_cache[(env, H)]


def run(env, H, lanes, trig, mode, extra_env=None, n=12):
    envs = {"whole": {"MPC_B200_STREAM": "0"}, "stream": {"MPC_B200_STREAM": "1"},
            "wide": {"MPC_B200_STREAM": "0", "MPC_B200_WIDE": "1"}}[mode]
    envs = dict(envs, **(extra_env or {}))
    rec = dict(env=env, H=H, K=K_BIG, lanes=lanes, trig=trig, mode=mode, knobs=extra_env or {})
    try:
        S64, U64 = reference(env, H)
        setenv(envs)
        x = np.array(START[env]) + 0.01
        # parity at K_PAR
        s = solver(env, K_PAR, H, "f32", lanes=lanes, fast_math=trig)
        s.fill_noise()
        u0, U, _, _ = s.mppi_solve(x, None, np.zeros(H), want_traj=False)
        S = s.mppi_costs()[0]
        rel = np.abs(S - S64) / np.abs(S64)
        rec["parity"] = dict(cost_rel_p50=float(np.quantile(rel, .5)), cost_rel_p99=float(np.quantile(rel, .99)),
                             cost_rel_max=float(rel.max()), plan_abs_max=float(np.abs(U[0] - U64).max()),
                             n_over_1e5=int((rel > 1e-5).sum()))
        s.close()
        # speed at K_BIG
        s = solver(env, K_BIG, H, "f32", lanes=lanes, fast_math=trig)
        s.fill_noise()
        s.upload_inputs(x, None, np.zeros(H))
        g = s.geometry
        rec["shape"] = dict(grid=g["grid"], block=g["block"], smem=g["smem"], stages=g["stages"], stream_rows=g["stream_rows"])
        s.time_solves(3, want_traj=False)
        ms0 = statistics.median(s.time_solves(n, want_traj=False))
        ms1 = statistics.median(s.time_solves(n, want_traj=True))
        rec["ms_no_traj"], rec["ms_traj"] = ms0, ms1
        rec["steps_per_s"] = K_BIG * H / (ms0 * 1e-3)
        rec["tflops_alg"] = rec["steps_per_s"] * FLOPS[env] / 1e12
        s.close()
    except Exception as e:  # keep going: one bad flavour must not cost the GPU call
        rec["error"] = repr(e)
    setenv({})
    print(json.dumps(rec), flush=True)


if which in ("all", "acrobot"):
    for lanes in (1, 2):
        for trig in (0, 1):
            run("acrobot", 100, lanes, trig, "whole")
        for trig in (0, 1, 2, 4, 5):
            run("acrobot", 100, lanes, trig, "stream")
    run("acrobot", 100, 2, 1, "stream", {"MPC_B200_WARPS": "16"})
    run("acrobot", 100, 2, 1, "stream", {"MPC_B200_WARPS": "20"})
    run("acrobot", 100, 1, 1, "stream", {"MPC_B200_WARPS": "24"})
    run("acrobot", 100, 1, 1, "stream", {"MPC_B200_STREAM_ROWS": "4", "MPC_B200_STREAM_DEPTH": "4"})
    run("acrobot", 100, 2, 1, "stream", {"MPC_B200_STREAM_ROWS": "4", "MPC_B200_STREAM_DEPTH": "4"})
if which in ("all", "h50"):
    for env in ("pendulum", "cartpole"):
        trig = 1 if env == "pendulum" else 0
        run(env, 50, 2, trig, "whole")
        run(env, 50, 2, trig, "stream")
        run(env, 50, 1, trig, "whole")
        run(env, 50, 1, trig, "wide")
        run(env, 50, 1, trig, "stream")
        run(env, 50, 2, trig, "whole", {"MPC_B200_WARPS": "12"})
        run(env, 50, 2, trig, "whole", {"MPC_B200_WARPS": "8"})
    run("pendulum", 50, 2, 0, "whole")
    run("acrobot", 50, 1, 1, "wide")
    run("acrobot", 50, 2, 1, "stream")
    run("acrobot", 50, 2, 1, "whole")

"""Pins the CPU oracle (oracle/mpc_oracle.c) to golden vectors produced by executing the unmodified reference
(oracle/gen_golden.py).  The reference has no tests of its own; these fixtures are the parity anchor.

Tolerances: the oracle follows numpy's operation order, so single steps agree to a few ulp (libm sin/cos/pow vs
numpy's SIMD kernels differ by <= 1-2 ulp).  Over a rollout that ulp noise is amplified by the dynamics
(a 1-ulp input change moves CartPole costs by 1.8e-13 rel), hence 2e-12 on rollout costs."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as orc

ENVS = ["pendulum", "cartpole", "acrobot"]


def _load(golden_dir, name):
    d = np.load(os.path.join(golden_dir, name))
    meta = json.loads(str(d["meta"])) if "meta" in d.files else None
    return d, meta


@pytest.mark.parametrize("env", ENVS)
def test_single_step_kat(golden_dir, env):
    d, _ = _load(golden_dir, "step_kat.npz")
    m = orc.model(env)
    xs, us, gs, cost, xn = (d[env + "_" + k] for k in ("x", "u", "g", "cost", "xn"))
    for i in range(xs.shape[0]):
        c, x2 = orc.step(m, xs[i], us[i, 0], gs[i])
        assert c == pytest.approx(cost[i], rel=1e-14, abs=1e-14), (env, i)
        np.testing.assert_allclose(x2, xn[i], rtol=1e-14, atol=1e-14, err_msg="%s %d" % (env, i))


def test_mppi_cases(golden_dir):
    d, meta = _load(golden_dir, "mppi_cases.npz")
    for ci, mt in enumerate(meta):
        p = "c%d_" % ci
        m = orc.model(mt["env"])
        U = d[p + "U_in"].reshape(-1).copy()
        for s in range(mt["n_solves"]):
            S, U, u0 = orc.mppi_solve(m, d[p + "x"], U, d[p + "delta_u"], mt["lam"], mt["std"], d[p + "goal"])
            np.testing.assert_allclose(S, d[p + "costs"][s], rtol=2e-12, err_msg="costs case %d solve %d" % (ci, s))
            np.testing.assert_allclose(U, d[p + "U_out"][s].reshape(-1), rtol=1e-10, atol=1e-12,
                                       err_msg="U case %d solve %d" % (ci, s))
            assert u0 == pytest.approx(float(d[p + "u0"][s][0]), rel=1e-10, abs=1e-12)


def test_mppi_threads_equal_serial(golden_dir):
    d, meta = _load(golden_dir, "mppi_cases.npz")
    ci = 3
    p, mt = "c%d_" % ci, meta[ci]
    m = orc.model(mt["env"])
    U = d[p + "U_in"].reshape(-1)
    S1 = orc.mppi_costs(m, d[p + "x"], U, d[p + "delta_u"], mt["lam"], mt["std"], d[p + "goal"], nthreads=1)
    S4 = orc.mppi_costs(m, d[p + "x"], U, d[p + "delta_u"], mt["lam"], mt["std"], d[p + "goal"], nthreads=4)
    assert np.array_equal(S1, S4)


def test_cem_cases(golden_dir):
    d, meta = _load(golden_dir, "cem_cases.npz")
    for ci, mt in enumerate(meta):
        p = "c%d_" % ci
        m = orc.model(mt["env"])
        z = d[p + "z"]
        # hand the oracle max_iter slabs of noise; only `iters` of them exist in the fixture (early break)
        zfull = np.zeros((mt["max_iter"],) + z.shape[1:])
        zfull[: z.shape[0]] = z
        r = orc.cem_solve(m, d[p + "x"], d[p + "U_in"], zfull, mt["n_elite"], mt["epsilon"], mt["alpha"], mt["std"])
        assert r["iters"] == mt["iters"], ci
        np.testing.assert_allclose(r["rewards"], d[p + "rewards"], rtol=2e-12, err_msg="rewards case %d" % ci)
        assert np.array_equal(r["elites"], d[p + "elites"]), "elite sets differ, case %d" % ci
        np.testing.assert_allclose(r["mu_hist"], d[p + "mu"][:, :, 0], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(r["sd_hist"], d[p + "sd"][:, :, 0], rtol=1e-11, atol=1e-14)
        assert r["u0"] == pytest.approx(float(d[p + "u0"][0]), rel=1e-12, abs=1e-14)


def test_simulate_matches_episode_fixture(golden_dir):
    d, meta = _load(golden_dir, "episode_cases.npz")
    for ci, mt in enumerate(meta):
        p = "c%d_" % ci
        m = orc.model(mt["env"])
        # planned_xs/costs were simulated from states[i] with the clipped, not-yet-shifted plan (agent.py:64-69);
        # U[i] is stored AFTER the shift, so rebuild the pre-shift plan from it where possible: rows 0..H-2 of the
        # shifted plan are rows 1..H-1 of the pre-shift plan; row 0 of the pre-shift plan is the returned action
        # only if it was not clipped differently -> use the xs themselves as the check instead.
        xs = d[p + "planned_xs"]
        H = xs.shape[1] - 1
        for i in range(xs.shape[0]):
            pre = np.empty(H)
            pre[1:] = d[p + "U"][i][: H - 1, 0]
            pre[0] = d[p + "actions"][i][0]
            xs_o, cost_o = orc.simulate(m, d[p + "states"][i], pre)
            np.testing.assert_allclose(xs_o, xs[i], rtol=1e-11, atol=1e-11)
            assert cost_o == pytest.approx(d[p + "planned_costs"][i], rel=1e-11)

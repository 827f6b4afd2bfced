"""Stress test of the solve kernel's synchronisation protocols (run on the B200 box: pytest -m gpu).

compute-sanitizer is closed on this pool, so the mbarrier ring (gen[] + phase parity), the self-cleaning arrival
tickets of the two-level merge and the STREAM chunk rings are exercised the blunt way: ~2000 back-to-back solves over
random (K, H, warps, stages, whole tiles | STREAM | WIDE, lanes, batch) shapes — K not a multiple of 32, H = 1,
H = MPPI_MAX_H, fewer tiles than warps, more solves than SMs — every solve's per-sample costs BIT-EQUAL to the same
solve in the library's default shape (per-sample arithmetic never depends on the tiling), the plan equal up to
summation order, every ticket back at zero, no flag time-out."""
import os

import numpy as np
import pytest

import mpc_b200  # noqa: F401
from mpc_b200 import capi
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}
KNOBS = ("MPC_B200_STREAM", "MPC_B200_WIDE", "MPC_B200_WARPS", "MPC_B200_STAGES", "MPC_B200_STREAM_ROWS",
         "MPC_B200_STREAM_DEPTH")


def _solver(env, K, H, dtype, batch, lanes, knobs):
    for k in KNOBS:
        os.environ.pop(k, None)
    os.environ.update(knobs)
    try:
        return capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], dtype=dtype,
                           noise="philox", lam=0.7, sigma=1.0, seed=11, batch=batch, lanes=lanes)
    finally:
        for k in KNOBS:
            os.environ.pop(k, None)


def test_two_thousand_solves_over_random_shapes():
    rng = np.random.RandomState(2024)
    n_solves, n_shapes, shapes_seen = 0, 0, set()
    fixed_H = [1, 2, 7, 8, 9, 31, 32, 33, 50, 100, 101, 255, 256]
    while n_solves < 2000:
        env = ("pendulum", "cartpole", "acrobot")[rng.randint(3)]
        dtype = "f32" if rng.rand() < 0.75 else "f64"
        H = int(fixed_H[rng.randint(len(fixed_H))]) if rng.rand() < 0.6 else int(rng.randint(1, 257))
        if env == "acrobot" and dtype == "f64":
            H = min(H, 64)  # the literal float64 Acrobot step is slow; the protocols do not depend on the env
        batch = 1 if rng.rand() < 0.7 else int(rng.choice([2, 7, 150, 300]))
        kmax = 40000 if batch == 1 else (3000 if batch < 100 else 400)
        K = int(rng.choice([1, 31, 32, 33, 63, 64, 65, 127, 129, 1000, 4096, 4100])) if rng.rand() < 0.5 else int(rng.randint(1, kmax))
        K = min(K, kmax)
        lanes = int(rng.choice([1, 2])) if dtype == "f32" else 0
        knobs = {}
        r = rng.rand()
        if r < 0.35:
            knobs["MPC_B200_STREAM"] = "1"
            if rng.rand() < 0.5:
                knobs["MPC_B200_STREAM_ROWS"] = str(int(rng.choice([4, 8, 16])))
                knobs["MPC_B200_STREAM_DEPTH"] = str(int(rng.choice([2, 3, 4, 6])))
        elif r < 0.7:
            knobs["MPC_B200_STREAM"] = "0"
            if lanes == 1 and rng.rand() < 0.3:
                knobs["MPC_B200_WIDE"] = "1"
        if rng.rand() < 0.6:
            knobs["MPC_B200_WARPS"] = str(int(rng.choice([1, 2, 3, 4, 5, 8, 12, 16, 20, 24, 32])))
        if rng.rand() < 0.3:
            knobs["MPC_B200_STAGES"] = str(int(rng.randint(1, 24)))
        try:
            s = _solver(env, K, H, dtype, batch, lanes, knobs)
        except capi.MpcError:
            continue  # a shape this H does not fit (whole tiles at H = 256, f64): refused loudly, nothing to stress
        ref = _solver(env, K, H, dtype, batch, lanes, {})
        shapes_seen.add((s.geometry["grid"], s.geometry["block"], s.geometry["stages"], s.geometry["stream_rows"]))
        x = START[env] + rng.normal(0, 0.2, (batch, orc.STATE_SIZE[env]))
        U = rng.uniform(-0.5, 0.5, (batch, H))
        reps = int(rng.randint(2, 9))
        for rep in range(reps):
            u0, Un, xs, cost = s.mppi_solve(x, None, U, want_traj=bool(rep & 1))
            S = s.mppi_costs()
            if rep == 0:
                u0r, Unr, _, _ = ref.mppi_solve(x, None, U, want_traj=False)
                Sr = ref.mppi_costs()
                U_first, S_first = Un.copy(), S.copy()
                # (long CartPole horizons with random plans overflow to inf / nan costs, exactly as the reference's do)
                assert np.array_equal(S, Sr, equal_nan=True), (env, K, H, dtype, batch, lanes, knobs, s.geometry)
                scale = max(1.0, float(np.abs(Unr).max()))
                np.testing.assert_allclose(Un, Unr, rtol=0, atol=(2e-6 if dtype == "f32" else 1e-12) * scale,
                                           err_msg=str((env, K, H, dtype, batch, lanes, knobs, s.geometry)))
            else:  # same inputs again: bit-identical (no ordering nondeterminism), rings restart from the left state
                assert np.array_equal(S, S_first, equal_nan=True) and np.array_equal(Un, U_first, equal_nan=True), (
                    rep, env, K, H, knobs, s.geometry)
            n_solves += 1
        assert s.nonzero_tickets() == 0 and ref.nonzero_tickets() == 0
        s.close()
        ref.close()
        n_shapes += 1
    assert n_shapes >= 200 and len(shapes_seen) >= 60, (n_shapes, len(shapes_seen))

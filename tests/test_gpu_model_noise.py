"""Model sensor noise inside the rollouts (Model.set_sensor_noise, models/model.py:20-24,42-43; enabled by
scripts/noise_tests.py:44 through Experiment's "model_noise") — `-m gpu`.

The reference draws this noise from numpy's global RNG inside forked pool workers, so its stream is not reproducible
even there.  Parity is therefore checked deterministically on the ARITHMETIC: the Philox normals the kernel used are
read back (mpc_philox_model_noise) and handed to the CPU oracle's noisy MPPI.f restatement; the costs must then agree
to the usual tolerances.  The distribution of the draws is checked separately."""
import numpy as np
import pytest

import mpc_b200
from mpc_b200 import capi
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0]),
         "acrobot": np.array([0.0, 0.0, 0.0, 0.0])}


@pytest.mark.parametrize("env", ["pendulum", "cartpole", "acrobot"])
@pytest.mark.parametrize("dtype,fast", [("f64", False), ("f32", False), ("f32", True)])
def test_noisy_model_costs_match_oracle_on_the_same_draws(env, dtype, fast):
    K, H, lam, sig_m = 3000, 30, 0.5, 0.05
    rng = np.random.RandomState(3)
    x = START[env] + rng.normal(0, 0.2, len(START[env]))
    U = rng.uniform(-0.5, 0.5, H)
    eps = rng.normal(0, 1.0, (K, H))
    m = orc.model(env)
    s = capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], dtype=dtype,
                    lam=lam, sigma=1.0, fast_math=fast, model_noise_std=sig_m, seed=11)
    assert s.geometry["stream_rows"] == 0
    s.set_noise(eps)
    for solve in range(2):  # fresh draws per solve
        z = s.philox_model_noise(solve)[:, :, :m.ns]
        s.mppi_solve(x, None, U)
        S = s.mppi_costs()[0]
        S_ref = orc.mppi_costs_noisy(m, x, U, eps, lam, 1.0, sig_m * z, nthreads=0)
        rel = np.abs(S - S_ref) / np.abs(S_ref)
        if dtype == "f64":
            assert rel.max() <= 1e-11, rel.max()  # the float32 normals are widened exactly; 30 chaotic steps
        else:
            assert np.quantile(rel, 0.95) <= 1e-5, np.quantile(rel, 0.95)
        if solve == 0:
            z0, S0 = z, S.copy()
    assert not np.array_equal(z0, z) and not np.array_equal(S0, S)
    # the noise matters at this std: without it the costs differ far beyond tolerance
    S_clean = orc.mppi_costs(m, x, U, eps, lam, 1.0, nthreads=0)
    assert np.median(np.abs(S0 - S_clean) / np.abs(S_clean)) > 1e-4
    s.close()


def test_model_noise_draws_are_standard_normal():
    s = capi.Solver("MPPI", 0, orc.default_params("pendulum"), K=20000, H=25, bounds=orc.BOUNDS["pendulum"],
                    dtype="f32", noise="philox", lam=0.5, sigma=1.0, model_noise_std=0.3, seed=5)
    z = s.philox_model_noise(0)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1.0) < 0.01
    assert abs(np.corrcoef(z[:, :, 0].ravel(), z[:, :, 1].ravel())[0, 1]) < 0.01
    assert abs(np.corrcoef(z[:, :-1, 0].ravel(), z[:, 1:, 0].ravel())[0, 1]) < 0.01
    s.close()


def test_agents_pick_up_a_noisy_model_and_cem_runs_with_it():
    """The drop-in path of noise_tests.py:41-49: Experiment(..., sensor_noise=s, model_noise=True)."""
    np.random.seed(0)
    for agent_cls, cfg in ((mpc_b200.MPPI, {"K": 50, "T": 10, "std": 1, "lam": 0.5}),
                           (mpc_b200.CEM, {"K": 50, "T": 10, "max_iter": 10, "n_elite": 5, "epsilon": 1e-5,
                                           "alpha": 0.2, "std": 1})):
        params = {"env": mpc_b200.PendulumEnv, "model": mpc_b200.PendulumModel, "agent": agent_cls,
                  "agent_config": cfg, "experiment_length": 8, "start_state": np.array([np.pi, 0.0]),
                  "sensor_noise": 0.5, "model_noise": True}
        res = mpc_b200.Experiment(params).run(warmstart=3)
        assert res["model_noise"] is True and res["env_states"].shape == (8, 2)
        assert np.isfinite(res["env_costs"]).all()
        assert "With Noise std=0.5" in res["model_name"] or res["model_name"]
    # the noisy and the clean model give different plans from the same agent noise
    np.random.seed(1)
    a = mpc_b200.MPPI(mpc_b200.PendulumModel(), {"K": 256, "T": 20, "std": 1.0, "lam": 0.5})
    np.random.seed(1)
    noisy = mpc_b200.PendulumModel()
    noisy.set_sensor_noise(0.5)
    b = mpc_b200.MPPI(noisy, {"K": 256, "T": 20, "std": 1.0, "lam": 0.5})
    x = np.array([np.pi, 0.0])
    ua, ub = a.predict_action(x, goal_state=np.zeros(3)), b.predict_action(x, goal_state=np.zeros(3))
    assert np.array_equal(a.delta_u, b.delta_u) and not np.array_equal(ua, ub)
    a.close()
    b.close()


def test_noise_sweep_on_the_device_fills_the_sink(tmp_path):
    """noise_tests.py as one call: actuation grid + sensor grid with the noisy model, episodes resident on the GPU."""
    sink = mpc_b200.ResultSink(str(tmp_path))
    res = mpc_b200.noise_sweep(mpc_b200.PendulumEnv, mpc_b200.PendulumModel, mpc_b200.MPPI,
                               {"K": 50, "T": 10, "std": 1, "lam": 0.5, "noise": "philox"}, np.array([np.pi, 0.0]),
                               stds=(0.1, 1.0), experiment_length=20, sink=sink, device=True, warmstart=5)
    assert len(res) == 4 and len(sink.find()) == 4
    noisy = sink.find({"model_noise": True})
    assert len(noisy) == 2 and all(r["sensor_noise"] for r in noisy)
    assert all(np.isfinite(r["env_costs"]).all() and r["env_states"].shape == (20, 2) for r in res)


def test_sample_and_temperature_sweeps_fill_the_sink(tmp_path):
    """The other two sweep drivers (scripts/sample_test.py:28-51, scripts/mppi_temperature.py:21-32): every grid point
    runs an episode (device-resident and through the reference-style host loop, `exp(warmstart)`), lands in the local
    sink with the reference's dictionary schema, and can be queried back by its agent_config."""
    sink = mpc_b200.ResultSink(str(tmp_path), "sweeps")
    start = np.array([np.pi, 0.0])
    base = {"std": 1.0, "lam": 0.5, "dtype": "f32"}
    res = mpc_b200.sample_sweep(mpc_b200.PendulumEnv, mpc_b200.PendulumModel, mpc_b200.MPPI, base, start,
                                samples=(64, 300), horizons=(5, 20), experiment_length=6, sink=sink, device=True)
    assert len(res) == 4 and {(r["agent_config"]["K"], r["agent_config"]["T"]) for r in res} == {(64, 5), (300, 5), (64, 20), (300, 20)}
    for r in res:
        assert r["env_states"].shape == (6, 2) and r["env_costs"].shape == (6,) and np.isfinite(r["env_costs"]).all()
        assert r["agent_planning_actions"].shape == (6 * r["agent_config"]["T"], 1)
    res_t = mpc_b200.temperature_sweep(mpc_b200.PendulumEnv, mpc_b200.PendulumModel, mpc_b200.MPPI,
                                       dict(base, K=128, T=10), start, lams=(0.01, 1.0), experiment_length=5, sink=sink,
                                       device=False, warmstart=2)
    assert [r["agent_config"]["lam"] for r in res_t] == [0.01, 1.0] and all(r["warmstart"] for r in res_t)
    assert len(sink.find()) == 6 and len(sink.find({"agent_config": {"K": 300, "T": 20}})) == 1
    hit = sink.find({"agent_config": {"lam": 0.01, "K": 128}})
    assert len(hit) == 1 and np.array_equal(hit[0]["env_states"], res_t[0]["env_states"])
    # Experiment.__call__ is what the reference scripts use
    exp = mpc_b200.Experiment({"env": mpc_b200.PendulumEnv, "model": mpc_b200.PendulumModel, "agent": mpc_b200.MPPI,
                               "agent_config": dict(base, K=64, T=8), "experiment_length": 4, "start_state": start})
    out = exp(3)
    assert out["env_states"].shape == (4, 2) and out["warmstart"] is True

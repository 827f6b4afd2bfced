"""Host-side pieces that need no GPU: result sink (the MongoDB stand-in), environment history frame, experiment
plumbing errors."""
import numpy as np
import pytest

import mpc_b200
from mpc_b200 import envs as envs_mod


def test_result_sink_round_trip(tmp_path):
    sink = mpc_b200.ResultSink(str(tmp_path), "runs")
    res = {"name": "MPPI solving PendulumEnvironment", "agent_name": "MPPI", "agent_config": {"K": 500, "T": 50},
           "env_states": np.arange(6.0).reshape(3, 2), "env_costs": np.array([1.0, 2.0, 3.0]), "passed_time": 0.25,
           "computation_time": [0.1, 0.15], "warmstart": True}
    sink.insert_one(res)
    sink.insert_one(dict(res, agent_config={"K": 50, "T": 50}, passed_time=0.05))
    enc = sink.encodeDict(res)
    assert type(enc["env_states"]) == bytes and "date" in enc  # DBTools.encodeDict: arrays pickled, date stamped
    hit = sink.find({"agent_name": "MPPI", "agent_config": {"K": 500}})
    assert len(hit) == 1 and np.array_equal(hit[0]["env_states"], res["env_states"])
    assert len(sink.find()) == 2 and len(sink.find({"agent_name": "CEM"})) == 0
    # table_generation.py-style aggregate: sum of passed_time over a query
    assert sum(r["passed_time"] for r in sink.find({"agent_name": "MPPI"})) == pytest.approx(0.30)
    lines = open(sink.index_path).read().strip().split("\n")
    assert len(lines) == 2


def test_history_frame_has_the_reference_columns():
    h = envs_mod.History()
    h = h.append({"state": np.zeros(2), "true_state": np.ones(2), "action": np.array([0.5]),
                  "true_action": np.array([0.5]), "cost": 1.25}, ignore_index=True)
    h = h.append({"state": np.ones(2), "true_state": np.ones(2), "action": np.array([0.1]),
                  "true_action": np.array([0.1]), "cost": 0.5}, ignore_index=True)
    states = np.array([x for x in h["state"].to_numpy()])  # Experiment.py:98
    assert states.shape == (2, 2) and len(h) == 2
    assert [c for c in h["cost"]] == [1.25, 0.5]


def test_environment_surface():
    env = mpc_b200.PendulumEnv()
    assert env.name == "PendulumEnvironment" and env.model.state_size == 2
    env.set_actuation_noise(0.1)
    env.set_sensor_noise(0.2)
    assert env.actuation_noise and env.sensor_noise_std == 0.2
    env.seed(3)
    s = env.reset()
    assert s.shape == (2,)
    s = env.reset(np.array([np.pi, 0.0]))
    assert np.array_equal(env.state, [np.pi, 0.0]) and len(env.history) == 0
    assert mpc_b200.CartPoleSwingUpEnv().name == "CartpoleSwingupEnvironment"
    assert mpc_b200.AcrobotEnv().name == "AcrobotEnvironment"
    with pytest.raises(NotImplementedError):
        env.render()


def test_environment_reset_quirk_of_the_reference():
    """env.py:52-59: reset(state) sets true_state but leaves `state` on the uniform draw, which for the reference's
    bounds is degenerate ([2, 2] for the Pendulum).  Default here: state = start state (documented in envs.py);
    `reference_reset_quirk = True` reproduces the reference — checked against the reference class itself when the
    reference tree is present."""
    start = np.array([np.pi, 0.0])
    env = mpc_b200.PendulumEnv()
    env.reset(start)
    assert np.array_equal(env.state, start) and np.array_equal(env.true_state, start)
    env.reference_reset_quirk = True
    env.reset(start)
    assert np.array_equal(env.true_state, start)
    assert np.array_equal(env.state, [2.0, 2.0])  # uniform(low = -1 * bounds_low = +2, high = +2)
    cp = mpc_b200.CartPoleSwingUpEnv()
    cp.reference_reset_quirk = True
    cp.reset(np.array([0.0, 0.0, np.pi, 0.0]))
    assert np.array_equal(cp.state, np.ones(4))
    from oracle import shims
    if shims.reference_available():
        shims.load_reference()
        from MPCBenchmark.envs.pendulum_env import PendulumEnv as RefPendulumEnv
        ref = RefPendulumEnv()
        ref.reset(start)
        assert np.array_equal(ref.state, env.state) and np.array_equal(ref.true_state, env.true_state)


class _FakeSolver:
    """Stands in for capi.Solver so that the agent classes run without a GPU: a deterministic `solve`."""
    tensor_noise, is_mppi = True, True

    def __init__(self, kind, env, env_params, K, H, **kw):
        self.H, self.calls, self.noise = H, 0, None

    def set_noise(self, d):
        self.noise = np.array(d)

    def mppi_solve(self, x, g_z, U, want_traj=True):
        self.calls += 1
        Un = np.asarray(U).reshape(1, self.H) * 0.5 + 3.0 * np.sin(np.arange(self.H) + self.calls + np.sum(x))
        xs = np.tile(np.asarray(x, dtype=float), (1, self.H + 1, 1)) + self.calls
        return Un[:, 0].copy(), Un, xs, np.array([float(self.calls)])

    def close(self):
        pass


def _drive(agent_cls, model):
    np.random.seed(0)
    agent = agent_cls(model, {"K": 8, "T": 6, "std": 1.0, "lam": 0.5})
    goal = np.zeros(3)
    agent.warm_start(np.array([0.1, 0.2]), 3, goal_state=goal)
    acts = [agent.predict_action(np.array([0.1 * i, -0.2]), goal_state=goal) for i in range(4)]
    out = (np.array(acts), agent.planned_us.copy(), agent.planned_xs.copy(), float(agent.planned_costs),
           [u.copy() for _, u in agent.planning_action_history], agent.step_iteration_variable,
           len(agent.warmstart_trajectories))
    agent.reset()
    assert agent.step_iteration_variable == 0 and agent.planning_state_history == []
    return out


def test_agents_subclass_the_reference_agent_when_it_is_importable(monkeypatch):
    """with MPCBenchmark importable the drop-in MPPI / CEM derive from the REFERENCE's Agent, so
    predict_action / warm_start / reset / _log_iteration are literally the reference's code; without it they derive
    from the behavioural mirror.  Both are driven here through the same fake solver and must behave identically
    (clip, logging, receding-horizon shift, return value)."""
    from mpc_b200 import agents
    monkeypatch.setattr(agents.capi, "Solver", _FakeSolver)
    assert agents.agent_base() is agents.Agent  # nothing importable by default
    mirror = _drive(agents.MPPI, mpc_b200.PendulumModel())
    from oracle import shims
    if not shims.reference_available():
        pytest.skip("reference tree not present")
    shims.load_reference()
    try:
        base = agents.rebind_agent_base()
        from MPCBenchmark.agents.agent import Agent as RefAgent
        assert base is RefAgent and issubclass(agents.MPPI, RefAgent) and issubclass(agents.CEM, RefAgent)
        assert mpc_b200.MPPI is agents.MPPI
        for name in ("predict_action", "warm_start", "reset", "_log_iteration"):
            assert getattr(agents.MPPI, name) is getattr(RefAgent, name), name     # the reference's own functions
            assert name not in vars(agents.MPPI) and name not in vars(agents.MPPI.__mro__[1])
        ref = _drive(agents.MPPI, mpc_b200.PendulumModel())
    finally:
        agents.rebind_agent_base(prefer_reference=False)
    assert agents.agent_base() is agents.Agent
    for a, b in zip(mirror, ref):
        if isinstance(a, list):
            assert all(np.array_equal(x, y) for x, y in zip(a, b))
        else:
            assert np.array_equal(a, b)


def test_experiment_is_callable_like_the_reference():
    """Experiment.py:143-158: every reference script runs `exp(50)`."""
    assert callable(mpc_b200.Experiment)
    calls = {}
    exp = mpc_b200.Experiment({"env": None, "model": None, "agent": None, "agent_config": {}, "experiment_length": 1,
                               "start_state": np.zeros(2)})
    exp.run = lambda warmstart=None: calls.setdefault("warm", warmstart) or {"ok": True}
    assert exp(50) == 50 or calls["warm"] == 50

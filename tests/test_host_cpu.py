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

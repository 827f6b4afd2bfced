"""Closed-loop experiments: the mirror of MPCBenchmark/ExperimentCore/Experiment.py plus a local-file result sink in
place of the MongoDB one (ExperimentCore/DBTools.py, scripts/mppi_temperature.py:31-32: there is no network here).

    exp = Experiment({"env": PendulumEnv, "model": PendulumModel, "agent": MPPI,
                      "agent_config": {"K": 4096, "T": 50, "std": 1.0, "lam": 0.5},
                      "experiment_length": 200, "start_state": np.array([np.pi, 0.0])})
    res = exp.run(warmstart=50)          # the reference's loop, one predict_action + env.step per host iteration
    res = exp.run_device(warmstart=50)   # the same episode resident on the GPU: one host synchronisation in total
    ResultSink("results/").insert_one(res)

`run` keeps the reference's control flow (Experiment.py:52-139) and its result dictionary key for key.
`run_device` produces the same dictionary from mpc_episode_run (include/mpc_b200.h): every solve, the plan clip / log /
shift of Agent.predict_action and the plant step of Environment.step are kernels enqueued back to back.
"""
import datetime
import json
import os
import pickle
import time

import numpy as np


class Experiment:
    def __init__(self, params):  # Experiment.py:9-33
        self.Environment = params["env"]
        self.Model = params["model"]
        self.Agent = params["agent"]
        self.agent_config = params["agent_config"]
        self.experiment_length = params["experiment_length"]
        self.start_state = params["start_state"]
        self.actuation_noise_std = params.get("actuation_noise", 0)
        self.actuation_noise = self.actuation_noise_std > 0
        self.sensor_noise_std = params.get("sensor_noise", 0)
        self.sensor_noise = self.sensor_noise_std > 0
        self.model_is_noised = params.get("model_noise", False)
        self.experiment_results = None
        self.verbose = params.get("verbose", False)  # the reference prints five lines per step (Experiment.py:91-97)

    # ------------------------------------------------------------------------------------------ shared pieces
    def _make(self):
        self.env = env = self.Environment()
        if self.sensor_noise:
            env.set_sensor_noise(self.sensor_noise_std)
        if self.actuation_noise:
            env.set_actuation_noise(self.actuation_noise_std)
        self.model = model = self.Model()
        if self.model_is_noised:
            self.model.set_sensor_noise(self.sensor_noise_std)
        agent = self.Agent(model, self.agent_config, cores=12)
        return env, model, agent

    def _pack(self, agent, env_name, model, passedtime, computation_time, states, actions, costs, warm_trajs,
              plan_states, plan_actions, plan_costs, true_states=None):
        """The dictionary of Experiment.py:109-139, key for key."""
        exp_name = agent.name + " solving " + env_name + " with " + model.name + "\n Starting from " + \
            str(self.start_state) + " Time: " + str(passedtime)
        res = {
            "name": exp_name,
            "computation_time": computation_time,
            "passed_time": passedtime,
            "warmstart": agent.warmstart,
            "warmstart_trajectories": warm_trajs,
            "env_name": env_name,
            "model_name": model.name,
            "agent_name": agent.name,
            "agent_config": self.agent_config,
            "env_start_state": np.asarray(self.start_state).tolist(),
            "env_states": states,
            "env_actions": actions,
            "env_costs": costs,
            "actuation_noise": self.actuation_noise,
            "sensor_noise": self.sensor_noise,
            "model_noise": self.model_is_noised,
            "agent_planning_states": plan_states,
            "agent_planning_actions": plan_actions,
            "agent_planning_costs": plan_costs,
        }
        if self.sensor_noise:
            # Experiment.py:130-134 stores env.history["state"] under "env_true_states" (sic); the device path has
            # the real true states and stores those
            res["env_true_states"] = states if true_states is None else true_states
            res["sensor_noise_std"] = self.sensor_noise_std
        if self.actuation_noise:
            res["env_true_actions"] = actions
            res["actuation_noise_std"] = self.actuation_noise_std
        return res

    # ------------------------------------------------------------------------------------------ the reference's loop
    def run(self, warmstart=None):  # Experiment.py:35-139
        env, model, agent = self._make()
        env.reset(self.start_state)
        agent.reset()
        goal = np.zeros(model.state_size + model.action_size)
        if warmstart is not None:
            agent.warm_start(env.true_state, warmstart, goal_state=goal)
        passedtime = 0
        computation_time = []
        for _ in range(self.experiment_length):
            t0 = time.time()
            action = agent.predict_action(env.state, goal_state=goal)
            iterationtime = time.time() - t0
            passedtime += iterationtime
            computation_time.append(np.around(iterationtime, decimals=3))
            _, r, done, _ = env.step(action)
            if self.verbose:
                print("Solver:", agent.name, "Config:", self.agent_config.get("T"), self.agent_config.get("K"),
                      self.agent_config.get("max_iter"), "Time Passed:", passedtime, "Cost", -r)
        passedtime = np.around(passedtime, decimals=3)
        states = np.array([x for x in env.history["state"].to_numpy()])
        actions = np.array([x for x in env.history["action"].to_numpy()])
        costs = np.array([x for x in env.history["cost"].to_numpy()])
        # Experiment.py:113,126-127 concatenates lists of (step, array) tuples, which numpy >= 1.24 rejects (ragged);
        # the arrays themselves are concatenated here, step-major
        warm = np.concatenate(agent.warmstart_trajectories, axis=0) if agent.warmstart_trajectories else np.zeros((0, 1))
        plan_states = np.concatenate([xs for _, xs in agent.planning_state_history], axis=0)
        plan_actions = np.concatenate([us for _, us in agent.planning_action_history], axis=0)
        res = self._pack(agent, env.name, model, passedtime, computation_time, states, actions, costs, warm,
                         plan_states, plan_actions, agent.planning_costs_history)
        agent.close()
        self.experiment_results = res
        return res

    def __call__(self, warmstart=None):  # Experiment.py:143-158: every reference script runs `exp(50)`
        """Runs `run` (the reference's host loop) and returns its result dictionary."""
        return self.run(warmstart=warmstart)

    # ------------------------------------------------------------------------------------------ resident on the GPU
    def run_device(self, warmstart=None, seed=0):
        """Same episode, same result dictionary, one host synchronisation for the whole control loop.  Environment
        noise is drawn on the device (Philox, key `seed`): reproducible, unlike the reference's global numpy RNG."""
        from .models import env_spec
        env, model, agent = self._make()
        if env_spec(env.model) != env_spec(model):
            raise ValueError("the environment's plant and the agent's model differ; run_device steps the plant with "
                             "the handle's own model parameters")
        solver = getattr(agent, "_solver", None)
        if solver is None:
            raise TypeError("run_device needs one of this package's GPU agents (MPPI / CEM)")
        if not solver.is_mppi and solver.tensor_noise:
            raise ValueError('a device-resident CEM episode needs agent_config["noise"] = "philox" (host numpy draws '
                             "cannot be resident on the device)")
        env.reset(self.start_state)
        agent.reset()
        goal = np.zeros(model.state_size + model.action_size)
        if warmstart is not None:  # through the agent: keeps warmstart_trajectories exactly as the reference's
            agent.warm_start(env.true_state, warmstart, goal_state=goal)
        n, H, ns = self.experiment_length, agent.horizon_length, model.state_size
        t0 = time.time()
        ep = solver.episode_run(env.true_state, n, U_in=np.asarray(agent.planned_us, dtype=np.float64).reshape(-1),
                                shift=agent.logging, actuation_std=self.actuation_noise_std if self.actuation_noise else 0.0,
                                sensor_std=self.sensor_noise_std if self.sensor_noise else 0.0, seed=seed)
        passedtime = np.around(time.time() - t0, decimals=3)
        # what the host loop would have left behind on the agent and the environment
        agent.planned_us = ep["U_final"].reshape(H, 1)
        agent.planned_xs, agent.planned_costs = ep["plan_xs"][-1], ep["plan_costs"][-1]
        for i in range(n):
            agent.planning_state_history.append((i, ep["plan_xs"][i]))
            agent.planning_action_history.append((i, ep["plan_us"][i].reshape(H, 1)))
            agent.planning_costs_history.append(ep["plan_costs"][i])
            env.history.append({"state": ep["states"][i + 1], "true_state": ep["true_states"][i + 1],
                                "action": ep["actions"][i:i + 1], "true_action": ep["actions"][i:i + 1],
                                "cost": ep["costs"][i]})
        agent.step_iteration_variable = n
        env.true_state, env.state = ep["true_states"][-1], ep["states"][-1]
        warm = np.concatenate(agent.warmstart_trajectories, axis=0) if agent.warmstart_trajectories else np.zeros((0, 1))
        res = self._pack(agent, env.name, model, passedtime, [float(passedtime) / n] * n, ep["states"][1:],
                         ep["actions"].reshape(n, 1), ep["costs"], warm, ep["plan_xs"].reshape(n * (H + 1), ns),
                         ep["plan_us"].reshape(n * H, 1), list(ep["plan_costs"]), true_states=ep["true_states"][1:])
        agent.close()
        self.experiment_results = res
        return res


# ================================================================================================ result sink
class ResultSink:
    """Local-file stand-in for the MongoDB collection of the reference's scripts
    (`collection.insert_one(DBTools.encodeDict(result))`, scripts/mppi_temperature.py:31-32, sample_test.py:49-51):
    one record per experiment, same dictionary schema (Experiment.py:109-139), arrays pickled with protocol 2 exactly
    as DBTools.encodeDict does (without the bson.Binary wrapper) and a "date" stamp added.  Records are appended to
    `<dir>/<collection>.pkl` (a stream of pickles) and indexed in `<dir>/<collection>.index.jsonl` (scalar fields
    only) so that table_generation-style queries do not have to unpickle everything."""

    def __init__(self, directory, collection="experiments"):
        self.directory = directory
        os.makedirs(directory, exist_ok=True)
        self.path = os.path.join(directory, collection + ".pkl")
        self.index_path = os.path.join(directory, collection + ".index.jsonl")

    @staticmethod
    def encodeDict(d):  # DBTools.py:7-15
        prepared = {}
        for key, value in d.items():
            prepared[key] = pickle.dumps(value, protocol=2) if type(value) == np.ndarray else value
        prepared["date"] = datetime.datetime.now()
        return prepared

    @staticmethod
    def decodeDict(d):  # DBTools.py:18-25
        return {k: (pickle.loads(v) if type(v) == bytes else v) for k, v in d.items()}

    def insert_one(self, result):
        rec = self.encodeDict(result)
        with open(self.path, "ab") as f:
            offset = f.tell()
            pickle.dump(rec, f, protocol=2)
        scalars = {k: v for k, v in result.items() if isinstance(v, (str, int, float, bool)) or k == "agent_config"}
        scalars.update(_offset=offset, date=rec["date"].isoformat(), passed_time=float(result.get("passed_time", 0.0)))
        with open(self.index_path, "a") as f:
            f.write(json.dumps(scalars, default=str) + "\n")
        return offset

    def find(self, query=None):
        """Records whose fields match `query` (nested dicts match key-wise: {"agent_config": {"K": 500}}), decoded."""
        def match(rec, q):
            for k, v in q.items():
                if k not in rec:
                    return False
                if isinstance(v, dict):
                    if not isinstance(rec[k], dict) or not match(rec[k], v):
                        return False
                elif rec[k] != v:
                    return False
            return True
        out = []
        if not os.path.exists(self.path):
            return out
        with open(self.path, "rb") as f:
            while True:
                try:
                    rec = pickle.load(f)
                except EOFError:
                    break
                if not query or match(rec, query):
                    out.append(self.decodeDict(rec))
        return out


# ================================================================================================ sweep drivers
def _sweep(configs, sink, device, warmstart, repeats):
    results = []
    for params in configs:
        for _ in range(repeats):
            exp = Experiment(params)
            res = exp.run_device(warmstart=warmstart) if device else exp.run(warmstart=warmstart)
            if sink is not None:
                sink.insert_one(res)
            results.append(res)
    return results


def sample_sweep(env, model, agent, base_config, start_state, samples=(10, 50, 100, 200, 500), horizons=(5, 10, 25, 50),
                 experiment_length=100, sink=None, device=True, warmstart=None, repeats=1):
    """scripts/sample_test.py:28-51: K x T grid for one agent / environment."""
    cfgs = [{"env": env, "model": model, "agent": agent, "agent_config": dict(base_config, K=K, T=T),
             "experiment_length": experiment_length, "start_state": start_state} for T in horizons for K in samples]
    return _sweep(cfgs, sink, device, warmstart, repeats)


def temperature_sweep(env, model, agent, base_config, start_state, lams=(0.001, 0.01, 0.1, 0.5, 1, 10),
                      experiment_length=100, sink=None, device=True, warmstart=None, repeats=1):
    """scripts/mppi_temperature.py:21-32: MPPI temperature grid."""
    cfgs = [{"env": env, "model": model, "agent": agent, "agent_config": dict(base_config, lam=lam),
             "experiment_length": experiment_length, "start_state": start_state} for lam in lams]
    return _sweep(cfgs, sink, device, warmstart, repeats)


def noise_sweep(env, model, agent, base_config, start_state, stds=(0.1, 0.5, 1.0), kinds=("actuation_noise", "sensor_noise"),
                experiment_length=100, sink=None, device=True, warmstart=None, repeats=1):
    """scripts/noise_tests.py:24-49: actuation noise grid, and sensor noise grid with the noisy model (:44)."""
    cfgs = [dict({"env": env, "model": model, "agent": agent, "agent_config": dict(base_config),
                  "experiment_length": experiment_length, "start_state": start_state, kind: std},
                 **({"model_noise": True} if kind == "sensor_noise" else {}))
            for kind in kinds for std in stds]
    return _sweep(cfgs, sink, device, warmstart, repeats)

"""Host-side mirror of the reference's environments (MPCBenchmark/envs/env.py, pendulum_env.py,
cartpole_swingup_env.py, acrobot_env.py): the plant the closed loop is run against.  Same class names, same
`reset / step / set_actuation_noise / set_sensor_noise / seed / history` surface; the plant step itself is
`Model.predict`, i.e. the GPU (mpc_model_step) — there is no numpy dynamics here.

Differences from the reference, on purpose:
  * `history` is a list-of-rows frame with the reference's columns (env.py:47-48 uses DataFrame.append, removed in
    pandas 2); `history["state"].to_numpy()` — the one access pattern Experiment.run uses (Experiment.py:98-100) — works.
  * `step` does not print (env.py:36,40) and does not mutate the caller's action array in place (env.py:39 `u +=`).
  * rendering (pyglet / OpenGL) is out of scope: unavailable on this box.
  * `reset(state)`: the reference sets `true_state = state.copy()` but leaves `self.state` — what `Experiment.run`
    hands to the FIRST `predict_action` (Experiment.py:81) — on the degenerate uniform draw of env.py:53-54
    (`low = -ones * bounds_low = +2`, `high = +2` for the Pendulum, so `[2, 2]`; `[1, 1, 1, 1]` for the other two), so
    the first action of every reference episode is planned from the wrong state.  Here `state` is
    set to the start state.  `Environment.reference_reset_quirk = True` (class or instance attribute) reproduces the
    reference bit for bit; tests/test_host_cpu.py covers both behaviours.
"""
import numpy as np

from .models import AcrobotModel, CartPoleSwingUpModel, PendulumModel


class _Column:
    def __init__(self, values):
        self._v = values

    def to_numpy(self):
        arr = np.empty(len(self._v), dtype=object)
        for i, v in enumerate(self._v):
            arr[i] = v
        return arr

    def __iter__(self):
        return iter(self._v)

    def __len__(self):
        return len(self._v)


class History:
    """The five columns env.py:47-48 appends per step."""
    COLUMNS = ("state", "true_state", "action", "true_action", "cost")

    def __init__(self):
        self.rows = []

    def append(self, row, ignore_index=True):
        self.rows.append({k: row[k] for k in self.COLUMNS})
        return self

    def __getitem__(self, col):
        return _Column([r[col] for r in self.rows])

    def __len__(self):
        return len(self.rows)


class Environment:
    """envs/env.py:7-69."""
    reference_reset_quirk = False  # True: reset(state) leaves `self.state` on the uniform draw, as env.py:52-59 does

    def __init__(self, name=None):
        self.name = "BaseEnvironment" if name is None else name
        self.true_state = None
        self.state = None
        self.observation = None
        self.model = None
        self.history = History()
        self.actuation_noise = False
        self.sensor_noise = False
        self.actuation_noise_std = 0
        self.sensor_noise_std = 0
        self.last_u = None

    def __str__(self):
        return "Name: " + self.name + "\n State: " + str(self.true_state)

    def set_actuation_noise(self, actuation_std):  # env.py:24-26
        self.actuation_noise = True
        self.actuation_noise_std = actuation_std

    def set_sensor_noise(self, sensor_std):  # env.py:28-30
        self.sensor_noise = True
        self.sensor_noise_std = sensor_std

    def seed(self, seed=None):  # env.py:32-33
        self.model.seed(seed)

    def step(self, u):  # env.py:35-50
        u = np.array(u, dtype=np.float64).reshape(-1)
        self.last_u = true_u = u
        if self.actuation_noise:
            u = u + np.random.normal(0, self.actuation_noise_std, self.model.action_size)
            true_u = u  # the reference's `u +=` aliases both names to the noisy action (env.py:36-39)
        self.true_state = self.model.predict(self.true_state, u)
        self.state = self.true_state + (np.random.normal(0, self.sensor_noise_std, self.model.state_size)
                                        if self.sensor_noise else 0)
        costs = -self.model.get_reward()
        self.history = self.history.append({"state": self.state, "true_state": self.true_state, "action": u,
                                            "true_action": true_u, "cost": costs}, ignore_index=True)
        return self.true_state, -costs, self._done(), {}

    def reset(self, state=None):  # env.py:52-60
        lo = -np.ones(self.model.state_size) * self.model.bounds_low
        hi = np.ones(self.model.state_size) * self.model.bounds_high
        self.true_state = self.state = self.model.np_random.uniform(low=lo, high=hi)
        if state is not None:
            self.true_state = np.array(state, dtype=np.float64).copy()
            if not self.reference_reset_quirk:
                self.state = self.true_state  # Experiment.py:81 reads env.state before the first step
        self.last_u = None
        self.history = History()
        return self.true_state

    def render(self, *a, **k):
        raise NotImplementedError("rendering needs a graphics stack: unavailable on this box")

    def close(self):
        pass

    def _done(self):
        raise NotImplementedError


class PendulumEnv(Environment):
    """envs/pendulum_env.py:8-55."""

    def __init__(self, g=10.0):
        super().__init__("PendulumEnvironment")
        self.model = PendulumModel(g)

    def _done(self):
        return False


class CartPoleSwingUpEnv(Environment):
    """envs/cartpole_swingup_env.py:19-146."""

    def __init__(self):
        super().__init__("CartpoleSwingupEnvironment")
        self.model = CartPoleSwingUpModel()
        self.x_threshold = 2.4

    def _done(self):  # cartpole_swingup_env.py:140-146
        x = self.state[0]
        return bool(x < -self.x_threshold or x > self.x_threshold)


class AcrobotEnv(Environment):
    """envs/acrobot_env.py:6-62."""

    def __init__(self):
        super().__init__("AcrobotEnvironment")
        self.model = AcrobotModel()

    def _done(self):  # acrobot_env.py:60-62
        s = self.state
        return bool(-np.cos(s[0]) - np.cos(s[1] + s[0]) > 1.)


ENV_FOR_MODEL = {"pendulum": PendulumEnv, "cartpole": CartPoleSwingUpEnv, "acrobot": AcrobotEnv}

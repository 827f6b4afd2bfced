"""Host-side mirror of the reference's model objects (MPCBenchmark/models/*.py) — parameter holders with the same
class names, attributes and `predict` / `get_reward` / `batch_predict` surface.  The arithmetic itself lives in
csrc/envs.cuh; `predict` runs it on the GPU through mpc_model_step (there is no numpy implementation here).

The agents accept either these classes or instances of the reference's own PendulumModel /
CartPoleSwingUpModel / AcrobotModel: `env_spec(model)` reads the parameters off the instance by attribute name.
"""
import numpy as np

from . import capi


class Model:
    """models/model.py:5-80 (interface only)."""
    bounds_low = 0
    bounds_high = 0
    state_size = -1
    action_size = -1

    def __init__(self, name="Base Model"):
        self.name = name
        self.last_reward = 0
        self.last_observation = None
        self.last_u = None
        self.sensor_noise = False
        self.sensor_std = 0
        self._stepper = None
        self._device = 0

    # model.py:46-48 (gym.utils.seeding.np_random(seed) of the gym the reference pins returns a RandomState-like
    # generator; only .uniform is used, by Environment.reset)
    def seed(self, seed=None):
        self.np_random = np.random.RandomState(seed)
        return [seed]

    @property
    def np_random(self):
        if getattr(self, "_np_random", None) is None:
            self._np_random = np.random.RandomState()
        return self._np_random

    @np_random.setter
    def np_random(self, rng):
        self._np_random = rng

    # model.py:20-24
    def set_sensor_noise(self, sensor_std):
        self.sensor_noise = True
        self.sensor_std = sensor_std
        self.name = self.name.replace("No Noise", "With Noise std=" + str(sensor_std))

    def _solver(self):
        # rebuilt when model.W was replaced: the handle folds the cost matrix at creation
        Wkey = np.asarray(self.W, dtype=np.float64).tobytes() if hasattr(self, "W") else None
        if self._stepper is not None and Wkey != getattr(self, "_stepper_W", None):
            self.close()
        if self._stepper is None:
            env, params = env_spec(self)
            self._stepper = capi.Solver("MPPI", env, params, K=1, H=1, dtype="f64", noise="philox",
                                        bounds=(float(np.ravel(self.bounds_low)[0]), float(np.ravel(self.bounds_high)[0])),
                                        device=self._device, W=cost_matrix(self))
            self._stepper_W = Wkey
        return self._stepper

    # model.py:26-44: cost of (x, u) against goal row 0, then the step
    def predict(self, current_state, action, goal=None):
        x = np.asarray(current_state, dtype=np.float64).reshape(1, -1)
        u = np.asarray(action, dtype=np.float64).reshape(1, -1)
        g = None if goal is None else np.asarray(goal, dtype=np.float64).reshape(-1, self.state_size + self.action_size)[0:1]
        cost, xn = self._solver().model_step(x, u[:, 0], g)
        self.last_reward = -cost[0]
        self.last_observation = xn[0]
        self.last_u = u[0]
        newstate = xn[0]
        if self.sensor_noise:
            newstate = newstate + np.random.normal(0, self.sensor_std, self.state_size)
        return newstate

    # model.py:50-60
    def batch_predict(self, current_state, action, goal=None):
        x = np.asarray(current_state, dtype=np.float64).reshape(-1, self.state_size)
        u = np.asarray(action, dtype=np.float64).reshape(x.shape[0], -1)
        g = None
        if goal is not None:
            g = np.broadcast_to(np.asarray(goal, dtype=np.float64), (x.shape[0], self.state_size + self.action_size))
        cost, xn = self._solver().model_step(x, u[:, 0], g)
        self.last_reward = -cost
        self.last_observation = xn
        self.last_u = u
        return xn

    def get_reward(self):
        return self.last_reward

    def get_observation(self):
        return self.last_observation

    def close(self):
        if self._stepper is not None:
            self._stepper.close()
            self._stepper = None


class PendulumModel(Model):
    """models/pendulum_model.py:7-22"""
    max_torque = 2.
    bounds_low = -np.array([max_torque])
    bounds_high = np.array([max_torque])
    action_size = 1
    state_size = 2

    def __init__(self, g=10.0):
        super().__init__("Pendulum Model (No Noise)")
        self.max_speed = 8
        self.dt = .05
        self.g = g
        self.m = 1.
        self.l = 1.
        self.W = np.diag([1.0, .1, .001])
        self.W_t = np.diag([1.0, 2.0, 0.0])


class CartPoleSwingUpModel(Model):
    """models/cartpole_swingup_model.py:15-46"""
    bounds_low = np.array([-1])
    bounds_high = np.array([1])
    state_size = 4
    action_size = 1

    def __init__(self):
        super().__init__("CartpoleSwingUp Model (No Noise)")
        self.g = 9.82
        self.m_c = 0.5
        self.m_p = 0.5
        self.total_m = (self.m_p + self.m_c)
        self.l = 0.6
        self.m_p_l = (self.m_p * self.l)
        self.force_mag = 10.0
        self.dt = 0.05
        self.b = 0.1
        self.x_threshold = 2.4
        self.W = -np.diag([1, 0, 5, 0, 0])
        self.W_t = -np.diag([5, 0, 10, 0, 0])


class AcrobotModel(Model):
    """models/acrobot_model.py:19-97"""
    dt = .2
    LINK_LENGTH_1 = 1.
    LINK_LENGTH_2 = 1.
    LINK_MASS_1 = 1.
    LINK_MASS_2 = 1.
    LINK_COM_POS_1 = 0.5
    LINK_COM_POS_2 = 0.5
    LINK_MOI = 1.
    MAX_VEL_1 = 4 * np.pi
    MAX_VEL_2 = 9 * np.pi
    torque_noise_max = 0.
    book_or_nips = "book"
    bounds_low = np.array([-1])
    bounds_high = np.array([1])
    state_size = 4
    action_size = 1

    def __init__(self, torque_noise_max=0):
        super().__init__("Acrobot Model (No Noise)")
        self.torque_noise_max = torque_noise_max
        self.W = np.diag([1, 0, 0, 0, 0])
        self.W_t = np.diag([1, 0, 0, 0, 0])


def _diag(W, n, what):
    """The diagonal of model.W for the parameter vector.  A non-diagonal W (or, in float32, a negative weight) is
    carried by `cost_matrix` instead and selects the generic-cost flavour of the kernels."""
    W = np.asarray(W, dtype=np.float64)
    if W.shape != (n, n):
        raise ValueError("%s.W must be %dx%d" % (what, n, n))
    return [float(v) for v in np.diag(W)]


def cost_matrix(model):
    """None when model.W is one of the reference's own cost matrices in kind (diagonal: the fast, folded cost of the
    kernels applies), else the full nz x nz matrix to hand to capi.Solver(W=...) — einsum("bi,ij,bj->b") over any W
    (pendulum_model.py:46-50, cartpole_swingup_model.py:95-104, acrobot_model.py:172-176)."""
    W = np.asarray(model.W, dtype=np.float64)
    sign = -1.0 if "CartPoleSwingUpModel" in {c.__name__ for c in type(model).__mro__} else 1.0  # its cost is -zd W zd
    if np.count_nonzero(W - np.diag(np.diag(W))) == 0 and np.all(sign * np.diag(W) >= 0):
        return None  # (a negative weight cannot be folded as sqrt(W_i) by the float32 kernels either)
    return W


def env_spec(model):
    """(env id, parameter vector in the layout of include/mpc_b200.h) for one of our models OR a reference model
    instance (matched by class name, parameters read by attribute)."""
    names = {c.__name__ for c in type(model).__mro__}
    if "PendulumModel" in names:
        p = [model.g, model.m, model.l, model.dt, model.max_speed, model.max_torque] + _diag(model.W, 3, "PendulumModel")
        return capi.ENV_PENDULUM, [float(v) for v in p]
    if "CartPoleSwingUpModel" in names:
        p = [model.g, model.m_c, model.m_p, model.l, model.force_mag, model.dt, model.b, model.x_threshold] + \
            _diag(model.W, 5, "CartPoleSwingUpModel")
        return capi.ENV_CARTPOLE, [float(v) for v in p]
    if "AcrobotModel" in names:
        if getattr(model, "torque_noise_max", 0) > 0:
            raise ValueError("AcrobotModel.torque_noise_max > 0 is not supported: the reference draws it from the "
                             "model's own RNG inside forked workers (acrobot_model.py:138-140), which no run "
                             "reproduces; default 0, set by no reference script")
        kind = getattr(model, "book_or_nips", "book")
        if kind not in ("book", "nips"):
            raise ValueError("AcrobotModel.book_or_nips must be 'book' or 'nips'")
        # "nips" (acrobot_model.py:124-128) is implemented in float64 mode; a float32 handle refuses it in mpc_create
        p = [model.dt, model.LINK_MASS_1, model.LINK_MASS_2, model.LINK_LENGTH_1, model.LINK_COM_POS_1,
             model.LINK_COM_POS_2, model.LINK_MOI, model.LINK_MOI, 9.8, model.MAX_VEL_1, model.MAX_VEL_2] + \
            _diag(model.W, 5, "AcrobotModel") + [1.0 if kind == "nips" else 0.0]
        return capi.ENV_ACROBOT, [float(v) for v in p]
    raise TypeError("unsupported model %r: the CUDA path implements PendulumModel, CartPoleSwingUpModel and "
                    "AcrobotModel" % type(model).__name__)


ENV_NAMES = {"pendulum": PendulumModel, "cartpole": CartPoleSwingUpModel, "acrobot": AcrobotModel}

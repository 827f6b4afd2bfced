"""Drop-in MPPI / CEM agents: the reference's Agent API (MPCBenchmark/agents/agent.py, mppi.py, cem.py) with the
sampling inner loop executed by libmpcb200.so on a B200.

    agent = MPPI(model, {"K": 4096, "T": 50, "std": 1.0, "lam": 0.5})        # same ctor as mppi.py:10
    u = agent.predict_action(state, goal_state=np.zeros(ns + du))            # same call as Experiment.py:81-82

`model` may be one of this package's models or an instance of the reference's own model classes.
Extra, optional params keys (defaults reproduce the reference): "dtype" ("f64" | "f32"), "noise" ("tensor": host
numpy noise exactly as the reference draws it | "philox": generated in-kernel), "fast_math" (f32: MUFU trig),
"lanes" (f32 MPPI: samples per thread, 1 | 2), "seed", "resample", "device", "per_step_goal" (sample rollouts use goal
row t at step t — an extension; the reference's rollouts only see row 0),
"devices" (list: ONE process drives several GPUs, K sharded over them — no torchrun needed),
"rank"/"world"/"comm_id" (K-sharding across GPUs, one process per GPU).

What stays on the host is only what the reference's Agent base class does around `_calc_action`
(agent.py:30-92): goal plumbing, clipping, logging lists and the receding-horizon shift.

Base class.  When the reference package is importable (`MPCBenchmark.agents.agent`), MPPI and CEM SUBCLASS THE
REFERENCE'S OWN `Agent`: `predict_action`, `warm_start`, `reset` and `_log_iteration` are then literally the reference's
code and only `_calc_action` / `simulate_trajectory` are replaced.  Where it is not importable (the GPU
box has no reference tree) the behavioural mirror `Agent` below is the base.  `agent_base()` reports which one is in
use; `rebind_agent_base()` re-evaluates the choice (tests install the import shims first).
"""
import numpy as np

from . import capi
from .models import cost_matrix, env_spec


class Agent:
    """Behavioural mirror of agents/agent.py:5-108, used when the reference package cannot be imported."""

    def __init__(self, name, model, logging=True):
        self.bounds_low = model.bounds_low
        self.bounds_high = model.bounds_high
        self.horizon_length = -1
        if model.state_size == -1 or model.action_size == -1:
            raise Exception("Given Model not fully implemented Error")
        self.state_size = model.state_size
        self.action_size = model.action_size
        self.model = model
        self.name = name
        self.logging = logging
        self.planned_xs = None
        self.planned_us = None
        self.planned_costs = None
        self.planning_state_history = []
        self.planning_action_history = []
        self.planning_costs_history = []
        self.step_iteration_variable = 0
        self.warmstart = False
        self.warmstart_trajectories = []

    def _goal_trajectory(self, g_z, goal_state):
        # agent.py:32-41 / 52-60: note np.array([None]) is not None, so a missing goal only fails later, as upstream
        goal_state = np.array([goal_state])
        if g_z is None:
            if goal_state is None:
                raise AttributeError("goal_state can't be null if no target trajectory g_z is given!")
            g_z = np.repeat(goal_state, self.horizon_length, axis=0)
        elif len(np.array(g_z).shape) <= 1:
            raise AttributeError("g_z can't be 1-Dimensional")
        return np.array(g_z)

    def warm_start(self, state, no_iterations=50, g_z=None, goal_state=None):  # agent.py:30-48
        self.warmstart = True
        g_z = self._goal_trajectory(g_z, goal_state)
        for _ in range(no_iterations):
            self._calc_action(state, g_z)
            self.planned_us = np.clip(self.planned_us, self.bounds_low, self.bounds_high)
            self.warmstart_trajectories.append(self.planned_us)

    def predict_action(self, state, g_z=None, goal_state=None):  # agent.py:50-76
        g_z = self._goal_trajectory(g_z, goal_state)
        u0 = self._calc_action(state, g_z)
        self.planned_us = np.clip(self.planned_us, self.bounds_low, self.bounds_high)
        if self.logging:
            self.planned_xs, self.planned_costs = self.simulate_trajectory(state, self.planned_us, g_z)
            self._log_iteration(self.planned_xs, self.planned_us, self.planned_costs)
            self.planned_us = np.roll(self.planned_us, -1)
            self.planned_us[-1] = np.zeros(self.action_size)
        self.step_iteration_variable += 1
        return np.clip(u0, self.bounds_low, self.bounds_high)

    def _calc_action(self, x, g_z=None):
        raise NotImplementedError

    def _log_iteration(self, planned_xs, planned_us, planned_costs):  # agent.py:81-87
        self.planning_state_history.append((self.step_iteration_variable, planned_xs.copy()))
        self.planning_action_history.append((self.step_iteration_variable, planned_us.copy()))
        self.planning_costs_history.append(planned_costs.copy())

    def reset(self):  # agent.py:89-92
        self.planning_state_history = []
        self.planning_action_history = []
        self.step_iteration_variable = 0

    def simulate_trajectory(self, x, us, g_z):  # agent.py:97-108, on the GPU
        raise NotImplementedError


def _reference_agent_class():
    """The reference's Agent class, or None when MPCBenchmark cannot be imported here."""
    try:
        from MPCBenchmark.agents.agent import Agent as RefAgent  # noqa: PLC0415
        return RefAgent
    except Exception:  # not installed / its own dependencies (gym ...) missing
        return None


def _make_agent_classes(base):
    """MPPI / CEM on top of `base` (the reference's Agent, or the mirror above)."""

    class _GpuAgent(base):
        """Shared handle management for the two sampling agents."""

        def _make_solver(self, kind, params, **kw):
            env, env_params = env_spec(self.model)
            self._env_id = env
            devices = params.get("devices")
            if devices is not None and len(devices) > 1:
                # one process, several GPUs: K sharded over `devices`, exchange inside the kernels over NVLink
                self._solver = capi.SolverGroup(
                    kind, env, env_params, self.K, self.horizon_length, devices,
                    dtype=params.get("dtype", "f64"), noise=params.get("noise", "tensor"),
                    bounds=(float(np.ravel(self.bounds_low)[0]), float(np.ravel(self.bounds_high)[0])),
                    seed=params.get("seed", 0), resample=params.get("resample", False),
                    fast_math=params.get("fast_math", False), lanes=params.get("lanes", 0),
                    model_noise_std=float(getattr(self.model, "sensor_std", 0.0)) if getattr(self.model, "sensor_noise", False) else 0.0,
                    W=cost_matrix(self.model), per_step_goal=params.get("per_step_goal", False), **kw)
                self._traj = None
                return
            if devices is not None and len(devices) == 1:
                params = dict(params, device=int(devices[0]))
            self._solver = capi.Solver(
                kind, env, env_params, K=self.K, H=self.horizon_length,
                dtype=params.get("dtype", "f64"), noise=params.get("noise", "tensor"),
                bounds=(float(np.ravel(self.bounds_low)[0]), float(np.ravel(self.bounds_high)[0])),
                device=params.get("device", 0), rank=params.get("rank", 0), world=params.get("world", 1),
                seed=params.get("seed", 0), resample=params.get("resample", False),
                fast_math=params.get("fast_math", False), lanes=params.get("lanes", 0),
                # a model with sensor noise adds it in every predict of the rollouts (model.py:42-43, noise_tests.py:44)
                model_noise_std=float(getattr(self.model, "sensor_std", 0.0)) if getattr(self.model, "sensor_noise", False) else 0.0,
                # a non-diagonal model.W, or params["per_step_goal"] (sample rollouts follow the goal TRAJECTORY: an
                # opt-in extension, the reference's rollouts use row 0), select the generic-cost flavour of the kernels
                W=cost_matrix(self.model), per_step_goal=params.get("per_step_goal", False), **kw)
            if params.get("world", 1) > 1:
                if "comm_id" in params:   # NCCL all-gather + merge kernel (MPPI)
                    self._solver.comm_init(params["comm_id"])
                else:                     # exchange inside the kernels over NVLink peer memory (MPPI and CEM)
                    from . import parallel
                    parallel.connect_p2p(self._solver)
            self._traj = None  # (x, clipped plan, g_z) -> (xs, cost) produced by the last solve

        def _remember_trajectory(self, x, U, g_z, xs, cost):
            Uc = np.clip(U.reshape(self.horizon_length, self.action_size), self.bounds_low, self.bounds_high)
            self._traj = (np.array(x, dtype=np.float64).reshape(-1), Uc, np.array(g_z, dtype=np.float64),
                          xs.reshape(self.horizon_length + 1, self.state_size), np.float64(cost))

        def simulate_trajectory(self, x, us, g_z):
            """agent.py:97-108.  The solve kernel already rolled out the clipped plan it produced; reuse that result when
            predict_action asks for exactly that trajectory, otherwise run the one-thread rollout kernel."""
            us = np.asarray(us, dtype=np.float64).reshape(self.horizon_length, self.action_size)
            g_z = np.asarray(g_z, dtype=np.float64)
            x = np.asarray(x, dtype=np.float64).reshape(-1)
            t = self._traj
            if t is not None and np.array_equal(t[0], x) and np.array_equal(t[1], us) and t[2].shape == g_z.shape \
                    and np.array_equal(t[2], g_z):
                return t[3].copy(), t[4]
            xs, cost = self._solver.simulate(x, us[:, 0], g_z)
            return xs, np.float64(cost)

        def close(self):  # mppi.py:31-33 / cem.py:38-40
            if getattr(self, "_solver", None) is not None:
                self._solver.close()
                self._solver = None

        def __del__(self):
            try:
                self.close()
            except Exception:
                pass


    class MPPI(_GpuAgent):
        """agents/mppi.py:7-75 on the GPU."""
        name = "MPPI"

        def __init__(self, model, params, cores=8):
            super().__init__("MPPI", model)
            self.K = params["K"]
            self.horizon_length = params["T"]
            self.planned_us = np.zeros((self.horizon_length, self.action_size), dtype=np.float64)
            self.std = np.ones((1, 1)) * params["std"]
            self.lam = params["lam"]
            self.x = np.zeros((self.horizon_length + 1, self.state_size))
            self.cores = cores  # accepted for signature compatibility; the GPU grid replaces the process pool
            if self.action_size != 1:
                raise ValueError("the CUDA path implements action_size == 1 (all three reference models)")
            self._make_solver("MPPI", params, lam=self.lam, sigma=float(params["std"]))
            self._delta_u = None
            if self._solver.tensor_noise:
                if params.get("device_noise", False):
                    # synthetic noise drawn on the device (Philox), for sizes where a host draw + upload is wasteful
                    self._solver.fill_noise()
                else:
                    # mppi.py:20-21: ONE draw in the constructor, reused by every solve; same global-RNG consumption
                    self.delta_u = np.random.normal(0, self.std, (self.K, self.horizon_length, self.action_size))
            self.sample_costs = np.zeros(self.K)

        @property
        def delta_u(self):
            return self._delta_u

        @delta_u.setter
        def delta_u(self, value):
            """Rebinding agent.delta_u re-uploads the tensor ([K][H][du] -> [H][K] tiles on the device).  After an
            in-place edit call sync_noise()."""
            self._delta_u = np.asarray(value, dtype=np.float64)
            self.sync_noise()

        def sync_noise(self):
            if self._delta_u is not None:
                self._solver.set_noise(self._delta_u.reshape(self.K, self.horizon_length))

        def _calc_action(self, x, g_z):  # mppi.py:54-75
            x = np.asarray(x, dtype=np.float64).reshape(-1)
            u0, U, xs, cost = self._solver.mppi_solve(x, g_z, self.planned_us.reshape(-1), want_traj=self.logging)
            self.planned_us[:] = U.reshape(self.horizon_length, self.action_size)
            if self.logging:
                self._remember_trajectory(x, U, g_z, xs[0], cost[0])
            self.sample_costs[:] = 0  # mppi.py:73; use last_sample_costs() to read them
            return self.planned_us[0]

        def last_sample_costs(self):
            """Per-sample rollout costs of the last solve (what mppi.py:57 held before line 73 zeroed it)."""
            return self._solver.mppi_costs()[0]


    class CEM(_GpuAgent):
        """agents/cem.py:8-78 on the GPU."""
        name = "CEM"

        def __init__(self, model, params, cores=8):
            super().__init__("CEM", model)
            self.K = params["K"]
            self.horizon_length = params["T"]
            self.max_iter = params["max_iter"]
            self.n_elite = params["n_elite"]
            self.epsilon = params["epsilon"]
            self.alpha = params["alpha"]
            self.std = np.ones((self.horizon_length, self.action_size)) * params["std"]
            self.planned_us = np.zeros((self.horizon_length, self.action_size))
            self.cores = cores
            if self.action_size != 1:
                raise ValueError("the CUDA path implements action_size == 1 (all three reference models)")
            self._make_solver("CEM", params, sigma=float(params["std"]), max_iter=self.max_iter, n_elite=self.n_elite,
                              epsilon=self.epsilon, alpha=self.alpha)
            self.last_iterations = 0

        def _calc_action(self, x, g_z=None):  # cem.py:54-78
            x = np.asarray(x, dtype=np.float64).reshape(-1)
            s = self._solver
            if s.tensor_noise:
                # Host RNG contract of cem.py:58-59: one draw of K*H*du normals per executed iteration.
                # np.random.normal(mean, std, size) == mean + std * standard_normal(size) bitwise for the legacy RNG.
                s.cem_begin(x, g_z, self.planned_us.reshape(-1))
                iters = 0
                for _ in range(self.max_iter):
                    z = np.random.standard_normal((self.K, self.horizon_length, self.action_size))
                    iters += 1
                    if s.cem_iterate(z.reshape(self.K, self.horizon_length)):
                        break
                u0, U, xs, cost = s.cem_end(want_traj=self.logging)
            else:
                u0, U, xs, cost, iters = s.cem_solve(x, g_z, self.planned_us.reshape(-1), want_traj=self.logging)
            self.last_iterations = iters
            self.planned_us = U.reshape(self.horizon_length, self.action_size).copy()
            if self.logging:
                self._remember_trajectory(x, U, g_z, xs, cost[0])
            return self.planned_us[0]


    return _GpuAgent, MPPI, CEM


_BASE = None


def agent_base():
    """The class MPPI and CEM currently derive from."""
    return _BASE


def rebind_agent_base(prefer_reference=True):
    """(Re)build MPPI / CEM on the reference's Agent if it is importable (and wanted), else on the mirror."""
    global _BASE, _GpuAgent, MPPI, CEM, MPPIB200, CEMB200
    ref = _reference_agent_class() if prefer_reference else None
    _BASE = ref if ref is not None else Agent
    _GpuAgent, MPPI, CEM = _make_agent_classes(_BASE)
    for cls in (_GpuAgent, MPPI, CEM):
        cls.__qualname__ = cls.__name__
    MPPIB200, CEMB200 = MPPI, CEM  # explicit names for code that imports both the reference agents and these
    import sys
    pkg = sys.modules.get(__name__.rpartition(".")[0])
    if pkg is not None and hasattr(pkg, "MPPI"):  # keep the package-level names on the current classes
        pkg.MPPI, pkg.CEM, pkg.MPPIB200, pkg.CEMB200 = MPPI, CEM, MPPI, CEM
    return _BASE


rebind_agent_base()

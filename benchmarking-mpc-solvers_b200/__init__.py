"""benchmarking-mpc-solvers_b200 — B200-native (sm_100a) sampling-MPC inner loop for MPCBenchmark.

The directory name is not a Python identifier; import it as `mpc_b200` (the shim module at the repo root loads this
directory under that name):

    import mpc_b200
    agent = mpc_b200.MPPI(mpc_b200.PendulumModel(), {"K": 4096, "T": 50, "std": 1.0, "lam": 0.5})
"""
from . import capi
from .agents import CEM, CEMB200, MPPI, MPPIB200, Agent
from .build import build as build_library
from .models import ENV_NAMES, AcrobotModel, CartPoleSwingUpModel, Model, PendulumModel, env_spec
from .envs import AcrobotEnv, CartPoleSwingUpEnv, Environment, PendulumEnv
from .experiment import Experiment, ResultSink, noise_sweep, sample_sweep, temperature_sweep

__all__ = ["capi", "Agent", "MPPI", "CEM", "MPPIB200", "CEMB200", "Model", "PendulumModel", "CartPoleSwingUpModel",
           "AcrobotModel", "env_spec", "ENV_NAMES", "build_library", "Environment", "PendulumEnv", "CartPoleSwingUpEnv",
           "AcrobotEnv", "Experiment", "ResultSink", "sample_sweep", "temperature_sweep", "noise_sweep"]

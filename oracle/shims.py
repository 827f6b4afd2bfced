"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

Shim loader that lets the UNMODIFIED reference package (`/root/reference/MPCBenchmark`) be imported in
the build container, where `gym`, `numdifftools`, `matplotlib` are absent and numpy is 2.x.  It is used only by
`oracle/gen_golden.py` (to produce `tests/golden/*.npz`) and by the container-only cross-checks in
`tests/test_oracle_vs_reference.py`.  `/root/reference` does not exist on the GPU box, so nothing that runs
there may call `load_reference()`.

What the reference imports that is missing here (file:line in /root/reference/MPCBenchmark):
  gym.utils.seeding.np_random   models/model.py:2,16,47 ; models/acrobot_model.py:6,94
  gym.utils.EzPickle            agents/cem.py:4 (imported, unused)
  gym.core / gym.spaces         models/acrobot_model.py:5 (imported, unused)
  numdifftools                  agents/ilqr.py:2 (pulled in by agents/__init__.py)
  matplotlib.pyplot             agents/ilqr.py:5, ExperimentCore/Plot.py
  np.float_                     models/acrobot_model.py:264,266 (removed in numpy 2)
"""
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("MPC_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "MPCBenchmark"))


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_shims():
    if not hasattr(np, "float_"):
        np.float_ = np.float64

    if "gym" not in sys.modules:
        def np_random(seed=None):
            return np.random.RandomState(seed), seed

        class EzPickle:
            def __init__(self, *a, **k):
                pass

        class Env:
            pass

        seeding = _module("gym.utils.seeding", np_random=np_random)
        utils = _module("gym.utils", seeding=seeding, EzPickle=EzPickle)
        core = _module("gym.core", Env=Env)
        spaces = _module("gym.spaces")
        rendering = _module("gym.envs.classic_control.rendering")
        cc = _module("gym.envs.classic_control", rendering=rendering)
        envs = _module("gym.envs", classic_control=cc)
        _module("gym", utils=utils, core=core, spaces=spaces, envs=envs, Env=Env,
                make=lambda *a, **k: (_ for _ in ()).throw(RuntimeError("gym shim: no envs")))

    if "numdifftools" not in sys.modules:
        class _NoDiff:
            def __init__(self, *a, **k):
                raise RuntimeError("numdifftools shim: iLQR is out of scope")
        _module("numdifftools", Jacobian=_NoDiff, Hessian=_NoDiff, Gradient=_NoDiff)

    if "matplotlib" not in sys.modules:
        plt = _module("matplotlib.pyplot", rcParams={})
        _module("matplotlib", pyplot=plt, rcParams={}, use=lambda *a, **k: None)


def load_reference():
    """Import and return the reference package (MPCBenchmark) unmodified."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    install_shims()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import MPCBenchmark.agents  # noqa: F401
    import MPCBenchmark.models  # noqa: F401
    import MPCBenchmark
    return MPCBenchmark

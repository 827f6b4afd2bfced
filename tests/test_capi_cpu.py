"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol the header declares, the
config struct layouts agree, and the product fails loudly (no fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

import mpc_b200
from mpc_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "mpc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mpc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    mpc_b200.build_library()
    lib = ctypes.CDLL(capi.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), "libmpcb200.so does not export %s" % name
    assert sorted(capi.SYMBOLS) == declared, "capi.SYMBOLS out of sync with include/mpc_b200.h"
    assert capi.load().mpc_abi_version() == 5


def test_config_struct_size_is_checked():
    L = capi.load()
    cfg = capi.MpcConfig()
    cfg.struct_size = 8
    h = ctypes.c_void_p()
    assert L.mpc_create(ctypes.byref(cfg), ctypes.byref(h)) == capi.MPC_EINVAL
    assert b"size mismatch" in L.mpc_last_error(None)


def test_env_spec_reads_parameters():
    env, p = mpc_b200.env_spec(mpc_b200.PendulumModel(g=9.81))
    assert env == capi.ENV_PENDULUM and p[0] == 9.81 and p[3] == 0.05 and p[6:9] == [1.0, 0.1, 0.001]
    env, p = mpc_b200.env_spec(mpc_b200.CartPoleSwingUpModel())
    assert env == capi.ENV_CARTPOLE and p[:8] == [9.82, 0.5, 0.5, 0.6, 10.0, 0.05, 0.1, 2.4] and p[8] == -1 and p[10] == -5
    env, p = mpc_b200.env_spec(mpc_b200.AcrobotModel())
    assert env == capi.ENV_ACROBOT and p[0] == 0.2 and p[8] == 9.8 and p[9] == pytest.approx(4 * np.pi)
    from oracle import oracle as orc
    for name, cls in mpc_b200.ENV_NAMES.items():
        assert list(mpc_b200.env_spec(cls())[1]) == [float(v) for v in orc.default_params(name)]

    class Other:
        pass
    with pytest.raises(TypeError):
        mpc_b200.env_spec(Other())
    # a full (non-diagonal) cost matrix is carried next to the parameter vector and selects the generic-cost kernels
    from mpc_b200.models import cost_matrix
    m = mpc_b200.PendulumModel()
    assert cost_matrix(m) is None
    m.W = np.ones((3, 3))
    assert mpc_b200.env_spec(m)[1][6:9] == [1.0, 1.0, 1.0] and np.array_equal(cost_matrix(m), np.ones((3, 3)))
    c = mpc_b200.CartPoleSwingUpModel()
    assert cost_matrix(c) is None          # W = -diag(...): its cost is -zd W zd, the effective weights are >= 0
    c.W = -c.W
    assert cost_matrix(c) is not None      # negative effective weights cannot be folded as sqrt(W_i)
    m.W = np.ones((2, 2))
    with pytest.raises(ValueError):
        mpc_b200.env_spec(m)


def test_no_silent_cpu_fallback():
    """Without a GPU the agents must raise, not compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.MpcError) as e:
        mpc_b200.MPPI(mpc_b200.PendulumModel(), {"K": 8, "T": 4, "std": 1.0, "lam": 1.0})
    assert e.value.code == capi.MPC_ECUDA
    with pytest.raises(capi.MpcError):
        mpc_b200.CEM(mpc_b200.PendulumModel(), {"K": 8, "T": 4, "std": 1.0, "max_iter": 2, "n_elite": 2,
                                                 "epsilon": 1e-5, "alpha": 0.2})


def test_product_never_imports_oracle():
    """The checker must never be on the product path: no import, no dlopen, no call of anything under oracle/."""
    pat = re.compile(r"import\s+oracle|from\s+oracle|from\s+\.+\s*oracle|liboracle|\borc_[a-z]+\s*\(|oracle[/.]")
    pkg = os.path.join(ROOT, "benchmarking-mpc-solvers_b200")
    checked = 0
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                checked += 1
                assert not pat.search(text), "%s reaches into oracle/: the product must not use the checker" % f
    assert checked >= 8

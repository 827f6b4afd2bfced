"""mpc_mppi_update_static — the allocation-free entry point (include/mpc_b200.h; no counterpart in the reference)
— against the existing entry point mpc_mppi_solve, through the C ABI.

Named test_zz_* so that pytest collects it last: a failure here must not hide any
other test.  It must pass alone in a fresh process and after every other test (the scratch is a static device array that
keeps whatever the last call left).

The existing entry point solves in float64 on a caller-supplied noise tensor; its per-sample costs (mpc_mppi_get_costs)
and its updated plan are the reference values.  The new entry point receives those costs, the same noise and the same
input plan as DEVICE arrays (torch supplies the device memory: the entry point itself allocates nothing) and must
return the same plan — the float64 bar of the parity tests (1e-12; the two kernels sum in different orders)."""
import numpy as np
import pytest

import mpc_b200  # noqa: F401
from mpc_b200 import capi
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": np.array([np.pi, 0.0]), "cartpole": np.array([0.0, 0.0, np.pi, 0.0])}
TOL = 1e-12


def _existing(env, K, H, lam, seed):
    """One float64 solve through mpc_mppi_solve: (costs S [K], noise [K][H], plan in, plan out)."""
    rng = np.random.RandomState(seed)
    x = START[env] + rng.normal(0, 0.2, orc.STATE_SIZE[env])
    U = rng.uniform(-0.5, 0.5, H)
    eps = rng.normal(0, 1.0, (K, H))
    s = capi.Solver("MPPI", ENV_ID[env], orc.default_params(env), K=K, H=H, bounds=orc.BOUNDS[env], dtype="f64",
                    lam=lam, sigma=1.0)
    try:
        s.set_noise(eps)
        _, U_out, _, _ = s.mppi_solve(x, None, U)
        S = s.mppi_costs()[0].copy()
        return S, eps, U, np.array(U_out[0], dtype=np.float64).copy()
    finally:
        s.close()


def _static(torch, S, eps, U, lam, stream=None):
    dev = torch.device("cuda", 0)
    K, H = eps.shape
    dS = torch.from_numpy(S).to(dev)
    de = torch.from_numpy(np.ascontiguousarray(eps)).to(dev)
    dU = torch.from_numpy(U).to(dev)
    dO = torch.full((H,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    capi.mppi_update_static(0, K, H, lam, dS.data_ptr(), de.data_ptr(), dU.data_ptr(), dO.data_ptr(), stream)
    torch.cuda.synchronize()
    return dO.cpu().numpy()


def test_static_update_matches_mppi_solve_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available(), "marked gpu: needs the B200"
    # two different problems back to back: the second call finds the first call's records (more of them: 296 blocks
    # against 17) and its ticket history in the shared static array
    cases = [("cartpole", 100000, 50, 0.5, 5), ("pendulum", 4100, 37, 1.0, 6)]
    for env, K, H, lam, seed in cases:
        S, eps, U, U_ref = _existing(env, K, H, lam, seed)
        assert np.abs(U_ref - U).max() > 1e-3  # the update moved the plan: the comparison is not vacuous
        for rep in range(2):  # and the same call twice: identical bits (fixed summation order, ticket cleared)
            U_new = _static(torch, S, eps, U, lam)
            assert np.all(np.isfinite(U_new))
            scale = max(1.0, float(np.abs(U_ref).max()))
            assert np.abs(U_new - U_ref).max() <= TOL * scale, (env, rep, np.abs(U_new - U_ref).max())
            if rep:
                assert np.array_equal(U_new, first)
            first = U_new


def test_static_update_on_a_caller_stream_and_in_place():
    torch = pytest.importorskip("torch")
    S, eps, U, U_ref = _existing("pendulum", 8192, 50, 0.5, 9)
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    dS, de = torch.from_numpy(S).to(dev), torch.from_numpy(eps).to(dev)
    dU = torch.from_numpy(U).to(dev)
    torch.cuda.synchronize()
    # U_out aliases U_in; two updates enqueued on one stream are ordered, so the second starts from the first's plan
    for _ in range(2):
        capi.mppi_update_static(0, 8192, 50, 0.5, dS.data_ptr(), de.data_ptr(), dU.data_ptr(), dU.data_ptr(),
                                st.cuda_stream)
    st.synchronize()
    twice = U + 2.0 * (U_ref - U)
    assert np.abs(dU.cpu().numpy() - twice).max() <= 4 * TOL * max(1.0, float(np.abs(twice).max()))


def test_static_update_rejects_sizes_beyond_the_static_array():
    torch = pytest.importorskip("torch")
    d = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    p = d.data_ptr()
    with pytest.raises(capi.MpcError) as e:
        capi.mppi_update_static(0, 8, 257, 1.0, p, p, p, p)
    assert e.value.code == -1 and "MPC_STATIC_MAX_H" in str(e.value)
    with pytest.raises(capi.MpcError) as e:
        capi.mppi_update_static(0, (1 << 24) + 1, 1, 1.0, p, p, p, p)
    assert e.value.code == -1 and "MPC_STATIC_MAX_K" in str(e.value)
    with pytest.raises(capi.MpcError):
        capi.mppi_update_static(0, 8, 1, 0.0, p, p, p, p)

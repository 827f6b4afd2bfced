"""Worker of tests/test_gpu_multiproc.py — launched by `python -m torch.distributed.run --nproc-per-node 2`.
One process per GPU: the ranks' exchange buffers are shared through CUDA IPC (mpc_p2p_export / mpc_p2p_import, the
path bench.py and torchrun users take — not the one-process mpc_p2p_link), K is sharded, and every solve's all-gather
happens inside the solve kernels over NVLink peer memory.  Checks, on every rank: the plan equals the single-GPU plan,
all ranks hold the same bits, also when one rank runs ahead of the other (double-buffered exchange slots)."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mpc_b200  # noqa: E402,F401
from mpc_b200 import capi, parallel  # noqa: E402
from mpc_b200.models import ENV_NAMES, env_spec  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo")  # the handles travel through any backend
    env_id, params = env_spec(ENV_NAMES["cartpole"]())
    K, H, lam = 6000 + 37, 40, 0.5
    rng = np.random.RandomState(3)
    x = np.array([0.0, 0.0, np.pi, 0.0]) + rng.normal(0, 0.2, 4)
    U = rng.uniform(-0.5, 0.5, H)
    kw = dict(K=K, H=H, dtype="f64", noise="philox", lam=lam, sigma=1.0, bounds=(-1.0, 1.0), seed=5)
    s = capi.Solver("MPPI", env_id, params, device=rank, rank=rank, world=world, **kw)
    parallel.connect_p2p(s)
    one = capi.Solver("MPPI", env_id, params, device=rank, **kw)
    # (1) a chain of solves, each refining the last plan: sharded == single GPU
    Us, U1 = U.copy(), U.copy()
    for i in range(4):
        _, Un, xs, cost = s.mppi_solve(x, None, Us)
        _, Un1, xs1, cost1 = one.mppi_solve(x, None, U1)
        np.testing.assert_allclose(Un, Un1, rtol=1e-11, atol=1e-12)
        np.testing.assert_allclose(xs, xs1, rtol=1e-9, atol=1e-9)
        Us, U1 = Un[0], Un1[0]
    # (2) one rank runs ahead: rank 0 enqueues six solves back to back, rank 1 sleeps between its solves, so rank 0's
    # solve e + 1 publishes while rank 1 is still inside (or before) solve e — different parity slots
    s.upload_inputs(x, None, U)
    for i in range(6):
        if rank == 1:
            time.sleep(0.003)
        s.solve_enqueue(want_traj=False)
    _, Ud, _, _ = s.download_outputs()
    one.upload_inputs(x, None, U)
    for i in range(6):
        one.solve_enqueue(want_traj=False)
    _, Ud1, _, _ = one.download_outputs()
    np.testing.assert_allclose(Ud, Ud1, rtol=1e-10, atol=1e-12)
    t = torch.from_numpy(np.ascontiguousarray(Ud).view(np.int64).copy())
    parts = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    assert all(torch.equal(parts[0], p) for p in parts), "ranks disagree on the plan"
    # (3) CEM over the same transport: candidates all-gathered inside the select kernel
    ck = dict(K=1000, H=20, dtype="f64", noise="philox", sigma=1.0, max_iter=3, n_elite=100, epsilon=1e-5, alpha=0.2,
              bounds=(-1.0, 1.0), seed=9)
    c = capi.Solver("CEM", env_id, params, device=rank, rank=rank, world=world, **ck)
    parallel.connect_p2p(c)
    c1 = capi.Solver("CEM", env_id, params, device=rank, **ck)
    c.upload_inputs(x, None, np.zeros(20))
    c.solve_enqueue()
    _, Uc, _, _ = c.download_outputs()
    _, Uc1, _, _, _ = c1.cem_solve(x, None, np.zeros(20))
    np.testing.assert_allclose(np.ravel(Uc), np.ravel(Uc1), rtol=1e-10, atol=1e-12)
    for h in (s, one, c, c1):
        h.close()
    dist.barrier()
    dist.destroy_process_group()
    print("rank %d ok" % rank, flush=True)


if __name__ == "__main__":
    main()

"""world_size-2 `gloo` test of the multi-GPU exchange protocol on CPU: each rank computes the partial record of its
K shard (here with the oracle standing in for the rank-local CUDA rollout), the records are all-gathered, and the
merge rule of benchmarking-mpc-solvers_b200/parallel.py (the specification of mppi_finalize_kernel stage 2) must
reproduce the single-process solve."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, K, H, lam, env, out_q):
    import sys
    sys.path.insert(0, ROOT)
    import mpc_b200  # noqa: F401
    from mpc_b200 import parallel
    from oracle import oracle as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.RandomState(0)          # same inputs on every rank
    x = rng.normal(0, 0.3, orc.STATE_SIZE[env])
    U = rng.uniform(-0.5, 0.5, H)
    eps = rng.normal(0, 1, (K, H))
    k0, k1 = parallel.shard_range(K, rank, world)
    S = orc.mppi_costs(orc.model(env), x, U, eps[k0:k1], lam, 1.0)
    beta = S.min()
    w = np.exp(-(1 / lam) * (S - beta))
    rec = np.concatenate([[beta, w.sum()], w @ eps[k0:k1]])
    gathered = [torch.zeros(2 + H, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(rec))
    _, _, erg = parallel.merge_partials(torch.stack(gathered).numpy(), lam)
    # comm-id broadcast helper works on any backend
    ident = parallel.broadcast_comm_id(lambda: bytes(range(128)))
    assert ident == bytes(range(128))
    out_q.put((rank, k0, k1, U + erg))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("env,K,H,lam", [("pendulum", 1001, 20, 0.5), ("cartpole", 64, 10, 0.01)])
def test_two_rank_exchange_reproduces_single_solve(env, K, H, lam):
    from oracle import oracle as orc
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, K, H, lam, env, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.RandomState(0)
    x = rng.normal(0, 0.3, orc.STATE_SIZE[env])
    U = rng.uniform(-0.5, 0.5, H)
    eps = rng.normal(0, 1, (K, H))
    _, U_ref, _ = orc.mppi_solve(orc.model(env), x, U, eps, lam, 1.0)
    results.sort()
    assert results[0][1] == 0 and results[0][2] == results[1][1] and results[1][2] == K   # shards tile [0, K)
    for _, _, _, U_rank in results:
        np.testing.assert_allclose(U_rank, U_ref, rtol=1e-10, atol=1e-12)
    np.testing.assert_array_equal(results[0][3], results[1][3])   # every rank ends with the same plan


def _cem_worker(rank, world, port, K, n_elite, out_q):
    import sys
    sys.path.insert(0, ROOT)
    import mpc_b200  # noqa: F401
    from mpc_b200 import parallel
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.RandomState(1)
    rewards = np.round(rng.normal(0, 1, K), 1)          # rounded: plenty of ties, also at the threshold
    k0, k1 = parallel.shard_range(K, rank, world)
    r, idx = parallel.local_candidates(rewards[k0:k1], k0, n_elite)
    n_cand = min(n_elite, -(-K // world))                # equal-sized contributions: pad like the kernel does
    rp = np.full(n_cand, -np.inf)
    ip = np.full(n_cand, np.iinfo(np.int32).max, dtype=np.int64)
    rp[:len(r)], ip[:len(idx)] = r, idx
    gr = [torch.zeros(n_cand, dtype=torch.float64) for _ in range(world)]
    gi = [torch.zeros(n_cand, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gr, torch.from_numpy(rp))
    dist.all_gather(gi, torch.from_numpy(ip))
    elites = parallel.select_global_elites([t.numpy() for t in gr], [t.numpy() for t in gi], n_elite)
    out_q.put((rank, elites))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("K,n_elite", [(1001, 100), (64, 64), (300, 200)])
def test_two_rank_cem_candidates_reproduce_global_topk(K, n_elite):
    """The CEM exchange protocol (local candidates -> all-gather -> identical global select) picks exactly the
    n_elite best samples of the whole population with ties broken by lowest index, on every rank."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cem_worker, args=(r, world, port, K, n_elite, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.RandomState(1)
    rewards = np.round(rng.normal(0, 1, K), 1)
    want = np.sort(np.lexsort((np.arange(K), -rewards))[:n_elite])
    for _, elites in results:
        np.testing.assert_array_equal(elites, want)

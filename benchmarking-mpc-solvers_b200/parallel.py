"""K-sharding across GPUs — host side.  One process per GPU; only the sample dimension shards (the reference's
Pool.map fan-out over samples, mppi.py:55-57, becomes rank-local rollouts plus ONE small exchange per solve).

Exchange record per rank and solve (float64, length 2 + H):
    rec[0]   beta_r = min_k S_k over the rank's shard
    rec[1]   eta_r  = sum_k exp(-(S_k - beta_r) / lam)
    rec[2+t] E_r[t] = sum_k exp(-(S_k - beta_r) / lam) * eps[k][t]
`merge_partials` is the specification of what the device does with the gathered records.  Transports, fastest first:
  * peer memory (default): the solve kernel itself stores the record into every rank's exchange buffer over NVLink and
    waits on flags — `connect_p2p` exchanges the buffers' IPC handles once (mpc_p2p_export / mpc_p2p_import);
  * NCCL all-gather on the solver's stream + a merge kernel (mpc_comm_init);
  * anything the caller has (mpc_mppi_rollout_partial / mpc_mppi_apply_partials).
"""
import numpy as np


def shard_range(K, rank, world):
    """Contiguous shard [k_begin, k_end) of rank `rank` — identical to the geometry computed in mpc_create."""
    return K * rank // world, K * (rank + 1) // world


def merge_partials(recs, lam):
    """recs [world][2+H] -> (beta, eta, erg[H]) with erg = sum_k w_k eps[k] for the normalised weights of mppi.py:59-62."""
    recs = np.asarray(recs, dtype=np.float64)
    beta = recs[:, 0].min()
    scale = np.exp(-(1 / lam) * (recs[:, 0] - beta))
    eta = float(np.sum(scale * recs[:, 1]))
    erg = (1 / eta) * (scale[:, None] * recs[:, 2:]).sum(axis=0)
    return beta, eta, erg


def local_candidates(rewards, k_begin, n_elite):
    """CEM over shards, step 1 (specification of cem_select_kernel's first half): a rank's min(n_elite, K_local) best
    samples as (reward, global index) pairs, ascending in global index; ties at the threshold go to the lowest index."""
    rewards = np.asarray(rewards, dtype=np.float64)
    n = min(n_elite, len(rewards))
    order = np.lexsort((np.arange(len(rewards)), -rewards))[:n]   # by reward descending, then index ascending
    order.sort()
    return rewards[order], order + k_begin


def select_global_elites(cand_rewards, cand_index, n_elite):
    """Step 2, identical on every rank after the all-gather (rank-major concatenation): the n_elite best candidates,
    ties by lowest global index; returns their global indices in ascending order."""
    r = np.concatenate([np.asarray(c, dtype=np.float64) for c in cand_rewards])
    idx = np.concatenate([np.asarray(c, dtype=np.int64) for c in cand_index])
    order = np.lexsort((idx, -r))[:n_elite]
    return np.sort(idx[order])


def broadcast_comm_id(create_id, group=None, src=0):
    """Rank `src` creates the communicator id (bytes), everybody receives it through torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    if dist.get_rank(group) == src:
        buf = torch.tensor(list(create_id()), dtype=torch.uint8)
    else:
        buf = torch.zeros(128, dtype=torch.uint8)
    dev = None
    if dist.get_backend(group) == "nccl":
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = buf.to(dev)
    dist.broadcast(buf, src=src, group=group)
    return bytes(buf.cpu().tolist())


def connect_p2p(solver, group=None):
    """One process per GPU: all-gather the 64-byte IPC handles of the ranks' exchange buffers through
    torch.distributed (any backend) and map them (mpc_p2p_import).  After this every solve is ONE kernel launch per
    rank; the all-gather of the rank records happens inside it over NVLink peer memory."""
    import torch
    import torch.distributed as dist
    mine = torch.tensor(list(solver.p2p_export()), dtype=torch.uint8)
    world = dist.get_world_size(group)
    if dist.get_backend(group) == "nccl":
        mine = mine.to(torch.device("cuda", torch.cuda.current_device()))
    out = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    solver.p2p_import([bytes(t.cpu().tolist()) for t in out])
    dist.barrier(group)

#!/usr/bin/env python
"""bench.py — rollout-steps/s of the sampling-MPC hot path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME] [--no-extra]

One "step" = one MPPI solve = MPPI._calc_action (mppi.py:54-75): rollout of K samples x H steps + softmin update.
Headline workload (BASELINE.json configs[2], largest K): MPPI on Acrobot, H = 100, K = 2^20, lam = 0.5, std = 1,
float32 rollout arithmetic with float64 accumulation, resident noise tensor (the reference's delta_u semantics,
mppi.py:20-21).  Only the sample dimension shards: each of N ranks rolls out K/N samples and one all-gather of
(min cost, sum w, sum w*eps[H]) per rank combines them => "strong" scaling (total work fixed).

value   device-timed (CUDA events on the solver's stream, inputs resident in HBM, max over ranks): MPPI._calc_action
e2e     same metric through the host-buffer C-ABI call (H2D of state+plan, D2H of action+plan+trajectory inside):
        what Agent.predict_action costs, logging rollout of the clipped plan included
The reference arm (--impl reference) times the CPU oracle port of the reference algorithm on all host threads
(the reference itself is pure Python and is not on the GPU box).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_ID = {"pendulum": 0, "cartpole": 1, "acrobot": 2}
START = {"pendulum": [np.pi, 0.0], "cartpole": [0.0, 0.0, np.pi, 0.0], "acrobot": [0.0, 0.0, 0.0, 0.0]}
# algorithmic flops per rollout step (SURVEY.md §8d convention: + - * / min max cmp = 1, FMA = 2, sin/cos/mod/exp = 1)
ALG_FLOPS = {"pendulum": 27, "cartpole": 56, "acrobot": 211}
WORKLOADS = {
    # name: (agent, env, K, H, params)
    "mppi_acrobot_h100_k1m": ("MPPI", "acrobot", 1 << 20, 100, dict(lam=0.5, sigma=1.0)),
    "mppi_pendulum_h50_k1m": ("MPPI", "pendulum", 1 << 20, 50, dict(lam=0.5, sigma=1.0)),
    "mppi_cartpole_h50_k1m": ("MPPI", "cartpole", 1 << 20, 50, dict(lam=0.5, sigma=1.0)),
    "mppi_pendulum_h50_k65536": ("MPPI", "pendulum", 65536, 50, dict(lam=0.5, sigma=1.0)),
}
DEFAULT_WORKLOAD = "mppi_acrobot_h100_k1m"
# float32 rollout flavour quoted per env: (fast_math, lanes).  fast_math = MUFU.SIN/COS instead of the polynomial sincos.
# Measured per-sample cost error against the float64 oracle on identical noise (profiles/r2b_parity_probe.jsonl,
# tests/test_gpu_parity.py holds both flavours to the SAME bars): Pendulum max 4.5e-7 (either flavour); Acrobot H <= 50
# max 2.5e-7; Acrobot H = 100 p99 2.2e-7 polynomial / 4.3e-7 MUFU with a chaotic tail beyond 1e-5 in BOTH (5 / 7 of
# 65536 samples); CartPole p95 5e-6, ~1 % of the samples beyond 1e-5 in any float32 evaluation (MUFU doubles that, so
# CartPole keeps the polynomial).  Every bench line carries a `parity` block measured in the same run.
# lanes: Acrobot's MUFU flavour is XU-pipe bound and runs best one sample per thread; the others use packed f32x2.
FLAVOUR = {"acrobot": (1, 1), "pendulum": (1, 2), "cartpole": (0, 2)}
L2_BYTES = 126 << 20


def ncu_evidence(workload, fast_math, lanes):
    """Per-launch DRAM traffic and pipe utilisation of the solve kernel from the committed ncu --set full captures
    (profiles/ncu_evidence.json, written from the raw pages by tools/ncu_summary.py); None when this flavour has none."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_evidence.json")))
        return table.get("%s|fast%d|lanes%d" % (workload, int(fast_math), int(lanes)))
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons during the timed region (NVML; nvidia-smi as a fallback)."""
    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
    NOTE = {"sw_power_cap": 0x4, "hw_power_brake": 0x80, "sync_boost": 0x10}

    def __init__(self, device):
        self.device, self.samples, self.reasons, self.max_mhz = device, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[device]) if vis and vis.split(",")[device].isdigit() else device
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                try:
                    r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in list(self.BAD.items()) + list(self.NOTE.items()):
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        if not self.samples:
            return self._smi_once()
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}

    def _smi_once(self):
        try:
            import subprocess
            out = subprocess.check_output(["nvidia-smi", "-i", str(self.device), "--query-gpu=clocks.sm,clocks.max.sm",
                                           "--format=csv,noheader,nounits"], timeout=10).decode().split(",")
            return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]), "reasons": [], "samples": 1,
                    "note": "single idle nvidia-smi sample (NVML unavailable)"}
        except Exception as e:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "note": "unavailable: %s" % e}


# --------------------------------------------------------------------------------------------- distributed plumbing
class Dist:
    def __init__(self, gpus):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.torch = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            self.torch, self.dist = torch, dist
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        if gpus != self.world:
            if self.rank == 0:
                print("bench.py: --gpus %d but WORLD_SIZE=%d; launch with torch.distributed.run for N>1" %
                      (gpus, self.world), file=sys.stderr)
            if gpus > 1 and self.world == 1:
                sys.exit(2)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def comm_id(self, create):
        from mpc_b200 import parallel
        return parallel.broadcast_comm_id(create) if self.world > 1 else None

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- our arm
class One:
    """A single-GPU view of this rank (extras that do not shard)."""
    world, rank = 1, 0

    def __init__(self, dist):
        self.local_rank = dist.local_rank

    def barrier(self):
        pass

    def max(self, v):
        return v


def make_solver(capi, agent, env, K, H, prm, dist, dtype="f32", noise="tensor", **kw):
    from mpc_b200.models import ENV_NAMES, env_spec
    if dtype == "f32" and agent == "MPPI":
        if kw.get("fast_math", -1) is None or kw.get("fast_math", -1) < 0:
            kw["fast_math"] = FLAVOUR[env][0]
        if not kw.get("lanes"):
            kw["lanes"] = FLAVOUR[env][1]
    else:
        kw["fast_math"] = max(0, kw.get("fast_math", 0) or 0)
        kw.pop("lanes", None)
    model = ENV_NAMES[env]()
    env_id, env_params = env_spec(model)
    lo, hi = float(model.bounds_low[0]), float(model.bounds_high[0])
    s = capi.Solver(agent, env_id, env_params, K=K, H=H, dtype=dtype, noise=noise, bounds=(lo, hi),
                    device=dist.local_rank, rank=dist.rank if dist.world > 1 else 0, world=dist.world, seed=0,
                    **prm, **kw)
    if dist.world > 1:
        if os.environ.get("MPC_B200_EXCHANGE", "p2p") == "p2p" or agent == "CEM":
            from mpc_b200 import parallel
            parallel.connect_p2p(s)  # in-kernel all-gather over NVLink peer memory: one launch per solve
        else:
            s.comm_init(dist.comm_id(capi.Solver.comm_create_id))
    if agent == "MPPI" and noise == "tensor":
        s.fill_noise()  # synthetic N(0, sigma) tensor drawn on the device (Philox); resident like agent.delta_u
    return s


def time_device(s, steps, warmup, dist, flush, want_traj=False):
    """K solves enqueued back to back on the solver's stream between a barrier + synchronise on both sides.  With the
    untimed L2 flush between solves every solve has its own pair of CUDA events (the flush sits between two pairs) and
    the step time is the sum of the K intervals; without a flush ONE pair brackets all K solves.  Max over ranks.
    For MPPI the solve is ONE kernel, so time / K is also the solve kernel's duration (launch gap included).
    K sharded over several GPUs: one untimed solve goes first after the barrier — its in-kernel exchange lines the
    ranks' device timelines up (a host barrier leaves them tens of microseconds apart).
    Returns (total ms, launches in the timed region, wall seconds of the bracket, per-solve ms of this rank or None)."""
    s.time_solves(max(1, warmup), want_traj=want_traj, flush_l2=flush)
    s.sync()
    dist.barrier()
    l0 = s.launch_count
    t_wall = time.perf_counter()
    each = None
    extra_launches = 0
    if flush:
        each = s.time_solves(steps, want_traj=want_traj, flush_l2=True)
        total = sum(each)
        extra_launches = steps + (2 if dist.world > 1 else 0)  # the flushes (+ the untimed aligning solve and its flush)
    else:
        if dist.world > 1:
            s.solve_enqueue(want_traj=want_traj)  # untimed: its in-kernel exchange aligns the ranks' device timelines
            extra_launches = 1
        s.timer_start()
        for _ in range(steps):
            s.solve_enqueue(want_traj=want_traj)
        total = s.timer_stop()
    s.sync()
    dist.barrier()
    wall = time.perf_counter() - t_wall
    launches = s.launch_count - l0 - extra_launches
    return dist.max(total), launches, wall, each


def time_e2e(s, x, U, steps, warmup, dist):
    """Host-buffer call (the C-ABI call a user's agent makes): H2D of state + plan and D2H of action, plan,
    trajectory and cost inside every timed call."""
    for _ in range(warmup):
        s.mppi_solve(x, None, U, want_traj=True)
    dist.barrier()
    lat = []
    t0 = time.perf_counter()
    for _ in range(steps):
        t1 = time.perf_counter()
        s.mppi_solve(x, None, U, want_traj=True)
        lat.append(time.perf_counter() - t1)
    total = time.perf_counter() - t0
    dist.barrier()
    return dist.max(total), lat


_PAR_CACHE = {}


def parity_block(capi, dist, env, H, prm, K_par=65536, **flavour):
    """Per-sample costs and plan of the benchmarked float32 flavour against the library's own float64 flavour (pinned
    to the reference at 1e-12 by tests/, so `oracle/` stays out of this process) on the same Philox noise stream, for a
    K_par-sample solve from the bench's start state.  Outside every timed region; one GPU (this rank)."""
    one = One(dist)
    x = np.array(START[env]) + 0.01
    U = np.zeros(H)
    key = (env, H, prm["lam"], prm["sigma"], K_par)
    if key not in _PAR_CACHE:
        r = make_solver(capi, "MPPI", env, K_par, H, prm, one, dtype="f64")
        u0, Un, _, _ = r.mppi_solve(x, None, U, want_traj=False)
        _PAR_CACHE[key] = (r.mppi_costs()[0].copy(), Un[0].copy(), float(u0[0]))
        r.close()
    S64, U64, u064 = _PAR_CACHE[key]
    s = make_solver(capi, "MPPI", env, K_par, H, prm, one, dtype="f32", **flavour)
    u0, Un, _, _ = s.mppi_solve(x, None, U, want_traj=False)
    S = s.mppi_costs()[0]
    s.close()
    rel = np.abs(S - S64) / np.abs(S64)
    return {"against": "float64 flavour of this library (pinned to the reference at rel 1e-12), same Philox noise, "
                       "K=%d subset" % K_par,
            "cost_rel_p50": float(np.quantile(rel, .5)), "cost_rel_p99": float(np.quantile(rel, .99)),
            "cost_rel_max": float(rel.max()), "n_cost_rel_over_1e-5": int((rel > 1e-5).sum()),
            "plan_abs_max": float(np.abs(Un[0] - U64).max()),
            "u0_rel": float(abs(u0[0] - u064) / max(abs(u064), 1e-300))}


def plan_hash(U):
    b = np.ascontiguousarray(U, dtype=np.float64).view(np.uint64).ravel()
    h = np.uint64(1469598103934665603)
    with np.errstate(over="ignore"):
        for v in b:
            h = (h ^ v) * np.uint64(1099511628211)
    return int(h >> np.uint64(12))  # 52 bits: exact in a float64 all-reduce


def run_mppi_workload(capi, dist, name, steps, warmup, dtype="f32", noise="tensor", lanes=0, fast_math=-1,
                      with_parity=True, with_e2e=True, fma_peak=None, peaks=None, peaks_src=""):
    """One K-sharded MPPI workload on all ranks: device-timed value, e2e, both rooflines, phase stamps, parity of the
    flavour against the float64 flavour, and at N > 1 the sharded plan against the same solve on one GPU."""
    agent, env, K, H, prm = WORKLOADS[name]
    s = make_solver(capi, agent, env, K, H, prm, dist, dtype=dtype, noise=noise, lanes=lanes, fast_math=fast_math)
    s.set_profiling(True)
    esz = 4 if dtype == "f32" else 8
    shard_bytes = s.K_local * H * esz
    flush = noise == "philox" or shard_bytes < 2 * L2_BYTES
    x = np.array(START[env]) + 0.01
    U = np.zeros(H)
    s.upload_inputs(x, None, U)
    clocks = ClockSampler(dist.local_rank)
    clocks.start()
    total_ms, launches, wall, each = time_device(s, steps, warmup, dist, flush)
    clk = clocks.stop()
    steps_total = float(K) * H * steps
    value = steps_total / (total_ms * 1e-3)
    try:
        tr = s.trace()
        trace_us = {k: round(v / 1e3, 1) for k, v in tr.items() if v is not None and k != "start"}
    except Exception:
        trace_us = None
    out = {"value": value, "ms_per_step": total_ms / steps, "clocks": clk, "gpu_launches": launches,
           "wall_s_timed_loop": wall, "solve_phases_us": trace_us, "K_per_gpu": s.K_local,
           "step_ms_min_median_max": [min(each), statistics.median(each), max(each)] if each else None,
           "lanes": s.lanes, "fast_math": int(s.fast_math), "flush": flush, "shard_bytes": shard_bytes,
           "geometry": {k: s.geometry[k] for k in ("grid", "block", "smem", "stages", "stream_rows")}}
    if with_e2e:
        e2e_total, lat = time_e2e(s, x, U, steps, warmup, dist)
        out["e2e"] = {"value": steps_total / e2e_total, "unit": "rollout-steps/s", "h2d_bytes_per_step": (s.ns + H) * 8,
                      "d2h_bytes_per_step": (2 + H + (H + 1) * s.ns + 1) * 8,
                      "p50_solve_ms": round(1e3 * statistics.median(lat), 4)}
    # ---- rooflines of the solve kernel: algorithmic flops and bytes over its event-timed duration
    kern_s = total_ms * 1e-3 / steps
    local_steps = float(s.K_local) * H
    ach_tf = local_steps * ALG_FLOPS[env] / kern_s / 1e12
    ach_gb = local_steps * esz / kern_s / 1e9 if noise == "tensor" else 0.0
    pipe = {"bound": "fp32_pipe" if dtype == "f32" else "fp64_pipe", "kernel": "mppi_solve_kernel", "achieved": ach_tf,
            "peak": fma_peak, "unit": "TFLOP/s", "frac": ach_tf / fma_peak if fma_peak else None, "traffic": None,
            "peak_source": "FMA-chain microbenchmark of this dtype run live (mpc_measure_fma_peak); "
                           "MEASURED_PEAKS.json has no CUDA-core figure",
            "alg_flops_per_step": ALG_FLOPS[env], "kernel_ms": kern_s * 1e3}
    hbm = {"bound": "hbm", "kernel": "mppi_solve_kernel", "achieved": ach_gb, "peak": peaks["hbm_gbs"], "unit": "GB/s",
           "frac": ach_gb / peaks["hbm_gbs"], "traffic": None, "peak_source": peaks_src,
           "alg_bytes_per_step": esz if noise == "tensor" else 0, "kernel_ms": kern_s * 1e3}
    ev = ncu_evidence(name, s.fast_math, s.lanes) if (noise == "tensor" and dist.world == 1 and dtype == "f32") else None
    if ev:
        for r in (pipe, hbm):
            r["traffic"] = ev["dram_bytes_per_launch"]
            r["traffic_source"] = "committed ncu --set full capture of this flavour (profiles/%s)" % ev.get("source")
        pipe["traffic_over_algorithmic"] = ev["dram_bytes_per_launch"] / (local_steps * esz)
        pipe["ncu"] = {k: ev[k] for k in ev if k != "dram_bytes_per_launch"}
    # which roof binds: the one whose floor (work / peak) is the longer time
    t_pipe = local_steps * ALG_FLOPS[env] / (fma_peak * 1e12) if fma_peak else 0.0
    t_hbm = local_steps * esz / (peaks["hbm_gbs"] * 1e9) if noise == "tensor" else 0.0
    if t_hbm > t_pipe:
        out["roofline"], out["roofline_other"] = hbm, pipe
    else:
        out["roofline"], out["roofline_other"] = pipe, hbm
    out["roofline"]["binding"] = "floor %.1f us vs %.1f us for the other roof" % (
        1e6 * max(t_pipe, t_hbm), 1e6 * min(t_pipe, t_hbm))
    # ---- N > 1: the sharded plan against the same solve on ONE GPU, and every rank's plan bit-equal
    if dist.world > 1:
        s.upload_inputs(x, None, U)
        s.solve_enqueue(want_traj=False)
        _, Un, _, _ = s.download_outputs()
        h = plan_hash(Un)
        hmin, hmax = -dist.max(-float(h)), dist.max(float(h))
        sv = {"all_ranks_bit_equal": bool(hmin == hmax)}
        if dist.rank == 0:
            one = One(dist)
            r = make_solver(capi, agent, env, K, H, prm, one, dtype=dtype, noise=noise, lanes=s.lanes,
                            fast_math=s.fast_math)
            r.upload_inputs(x, None, U)
            r.solve_enqueue(want_traj=False)
            _, U1, _, _ = r.download_outputs()
            r.close()
            sv.update(plan_max_abs=float(np.abs(Un - U1).max()), u0_abs=float(abs(Un[0, 0] - U1[0, 0])),
                      note="same flavour, same Philox noise (indexed by global sample), K unsharded on rank 0's GPU; "
                           "differences are the summation order of the weighted sum")
        dist.barrier()
        out["sharded_vs_single"] = sv
    s.close()
    if with_parity and dtype == "f32" and dist.rank == 0:
        try:
            out["parity"] = parity_block(capi, dist, env, H, prm, lanes=out["lanes"], fast_math=out["fast_math"])
        except Exception as e:
            out["parity"] = {"error": repr(e)}
    return out


def latency_extras(capi, dist, reps=200, warm=20):
    """p50 solve latency at K = 4096 (north_star: sub-ms) through the host-buffer call, 1 GPU."""
    out = {}
    one = One(dist)
    for env in ("pendulum", "cartpole", "acrobot"):
        H = 50
        s = make_solver(capi, "MPPI", env, 4096, H, dict(lam=0.5, sigma=1.0), one)
        x = np.array(START[env]) + 0.01
        U = np.zeros(H)
        for _ in range(warm):
            s.mppi_solve(x, None, U)
        lat = []
        for _ in range(reps):
            t = time.perf_counter()
            s.mppi_solve(x, None, U)
            lat.append(time.perf_counter() - t)
        out["mppi_%s_k4096_h50_f32_p50_ms" % env] = round(1e3 * statistics.median(lat), 4)
        s.close()
    return out


def cem_extra(capi, dist, fma_peak, K=4096, reps=100, warm=10):
    """BASELINE configs[1]: CEM CartPole, H = 50, elite ratio 0.1, 5 refit iterations, in-kernel Philox normals.  p50 of
    the host-buffer call, device time per solve, the solve against the FP32 roofline, and float32 parity against the
    library's float64 flavour on the same normals (elite-set overlap per iteration, mean)."""
    one = One(dist)
    H, iters, n_elite = 50, 5, max(1, K // 10)
    prm = dict(sigma=1.0, max_iter=iters, n_elite=n_elite, epsilon=1e-5, alpha=0.2)
    x = np.array(START["cartpole"]) + 0.01
    tag = "cem_cartpole_k%d_h50_5it" % K
    out = {}
    # parity first: fresh handles, so both draw the normals of solve 0
    res = {}
    for dt in ("f32", "f64"):
        s = make_solver(capi, "CEM", "cartpole", K, H, prm, one, noise="philox", dtype=dt)
        u0, Un, _, _, n_it = s.cem_solve(x, None, np.zeros(H))
        res[dt] = (Un.copy(), [np.sort(s.cem_elites(i)) for i in range(n_it)], n_it)
        s.close()
    n_cmp = min(res["f32"][2], res["f64"][2])
    out[tag + "_parity"] = {
        "against": "float64 flavour of this library on the same Philox normals",
        "elite_overlap_min": min(np.intersect1d(res["f32"][1][i], res["f64"][1][i]).size / n_elite for i in range(n_cmp)),
        "mean_abs_max": float(np.abs(res["f32"][0] - res["f64"][0]).max()), "iterations": [res["f32"][2], res["f64"][2]]}
    s = make_solver(capi, "CEM", "cartpole", K, H, prm, one, noise="philox")
    for _ in range(warm):
        s.cem_solve(x, None, np.zeros(H))
    lat = []
    for _ in range(reps):
        t = time.perf_counter()
        *_, n_it = s.cem_solve(x, None, np.zeros(H))
        lat.append(time.perf_counter() - t)
    p50 = statistics.median(lat)
    s.upload_inputs(x, None, np.zeros(H))
    dev_ms = statistics.median(s.time_solves(20, want_traj=False, flush_l2=False))
    l0 = s.launch_count
    s.solve_enqueue(want_traj=False)
    s.sync()
    launches = s.launch_count - l0
    s.close()
    out.update({tag + "_f32_philox_p50_ms": round(1e3 * p50, 4), tag + "_steps_per_s": K * H * n_it / p50,
                tag + "_device_ms_per_solve": dev_ms, tag + "_launches_per_solve": launches})
    if K == 4096:
        # the same solve called from plain C (tools/cabi_latency.c, built by build.py next to the library): the C-ABI
        # call with host buffers without the Python wrapper's ctypes / numpy time; with the logging trajectory (what
        # Agent.predict_action needs) and without it (CEM._calc_action alone, cem.py:54-78)
        exe = os.path.join(os.path.dirname(capi.LIB_PATH), "build", "cabi_latency")
        if os.path.exists(exe):
            import subprocess
            for traj, key in ((1, "_c_abi_p50_ms"), (0, "_c_abi_calc_action_only_p50_ms")):
                try:
                    r = subprocess.run([exe, "500", str(traj)], capture_output=True, text=True, timeout=60)
                    out[tag + key] = round(json.loads(r.stdout.strip().splitlines()[-1])["p50_us"] / 1e3, 4)
                except Exception as e:  # a probe, not the product: report and carry on
                    out[tag + key] = "unavailable: %r" % (e,)
    ach = K * H * n_it * ALG_FLOPS["cartpole"] / (dev_ms * 1e-3) / 1e12
    out[tag + "_roofline"] = {"bound": "fp32_pipe", "kernel": "cem_solve_kernel (the whole solve: one persistent launch)"
                              if launches == 1 else "all kernels of one CEM solve", "achieved": ach,
                              "peak": fma_peak, "unit": "TFLOP/s", "frac": ach / fma_peak if fma_peak else None,
                              "kernel_ms": dev_ms, "alg_flops_per_step": ALG_FLOPS["cartpole"],
                              "note": "K*H*iters = %d rollout-steps per solve; at K = 4096 the solve is a chain of "
                                      "latencies (one warp of rollouts per SM, two grid barriers per iteration), not "
                                      "a throughput problem" % (K * H * n_it)}
    return out


def batch_extra(capi, dist):
    """BASELINE configs[4]: 4096 independent CartPole MPPI solves (K = 1024, H = 50 each).  No data-path collective:
    rank r owns solves [4096 r / N, 4096 (r + 1) / N) (replicas), one launch per rank; value = all solves / max time."""
    Btot, K, H = 4096, 1024, 50
    b0, b1 = Btot * dist.rank // dist.world, Btot * (dist.rank + 1) // dist.world
    B = b1 - b0
    one = One(dist)
    s = make_solver(capi, "MPPI", "cartpole", K, H, dict(lam=0.5, sigma=1.0), one, batch=B)
    rng = np.random.RandomState(0)
    xs_all = np.array(START["cartpole"]) + rng.normal(0, 0.2, (Btot, 4))  # cartpole_swingup_model.py:50-51
    xs = xs_all[b0:b1]
    s.upload_inputs(xs, None, np.zeros((B, H)))
    s.time_solves(3, want_traj=False, flush_l2=True)
    dist.barrier()
    each = s.time_solves(10, want_traj=False, flush_l2=True)
    ms = dist.max(sum(each)) / 10
    dist.barrier()
    U0 = np.zeros((B, H))
    for _ in range(2):
        s.mppi_solve(xs, None, U0, want_traj=True, copy=False)
    dist.barrier()
    t = time.perf_counter()
    for _ in range(5):  # copy=False: results land in the solver's pinned output arrays (valid until the next solve)
        u0, Un, _, _ = s.mppi_solve(xs, None, U0, want_traj=True, copy=False)
    e2e_s = dist.max(time.perf_counter() - t) / 5
    out = {"batch4096_cartpole_k1024_h50_f32_steps_per_s": Btot * K * H / (ms * 1e-3),
           "batch4096_cartpole_k1024_h50_f32_ms_per_batch": ms,
           "batch4096_cartpole_k1024_h50_f32_e2e_steps_per_s": Btot * K * H / e2e_s,
           "batch4096_cartpole_k1024_h50_solves_per_rank": B,
           "batch4096_cartpole_k1024_h50_scaling": "replicas: %d solves per rank, no collective" % B}
    if dist.rank == 0:
        # parity of the config's own size and dtype: three of the batch's solves against the float64 flavour
        pick = [0, B // 2, B - 1]
        r = make_solver(capi, "MPPI", "cartpole", K, H, dict(lam=0.5, sigma=1.0), one, dtype="f64", batch=len(pick))
        _, U64, _, _ = r.mppi_solve(xs[pick], None, np.zeros((len(pick), H)), want_traj=False)
        S64 = r.mppi_costs()
        r.close()
        S32 = s.mppi_costs()[pick]
        rel = np.abs(S32 - S64) / np.abs(S64)
        out["batch4096_cartpole_k1024_h50_parity"] = {
            "against": "float64 flavour, same Philox noise, solves %s of this rank" % pick,
            "cost_rel_p50": float(np.quantile(rel, .5)), "cost_rel_p99": float(np.quantile(rel, .99)),
            "cost_rel_max": float(rel.max()), "plan_abs_max": float(np.abs(Un[pick] - U64).max())}
    s.close()
    return out


def temperature_extra(capi, dist):
    """BASELINE configs[3]: MPPI Pendulum K = 65536, H = 50 over the temperature grid of scripts/mppi_temperature.py:23.
    Latency of the host call per lambda, and what the small temperatures do to the float32 plan (the action
    is a softmin of the costs, its error is the cost error times max|S| / lambda)."""
    out = {}
    one = One(dist)
    for lam in (0.001, 0.01, 0.1, 0.5, 1.0, 10.0):
        prm = dict(lam=lam, sigma=1.0)
        s = make_solver(capi, "MPPI", "pendulum", 65536, 50, prm, one)
        x, U = np.array(START["pendulum"]), np.zeros(50)
        for _ in range(10):
            s.mppi_solve(x, None, U)
        lat = []
        for _ in range(100):
            t = time.perf_counter()
            s.mppi_solve(x, None, U)
            lat.append(time.perf_counter() - t)
        s.close()
        out["mppi_pendulum_k65536_h50_lam%g_p50_ms" % lam] = round(1e3 * statistics.median(lat), 4)
        p = parity_block(capi, dist, "pendulum", 50, prm, lanes=FLAVOUR["pendulum"][1], fast_math=FLAVOUR["pendulum"][0])
        out["mppi_pendulum_k65536_h50_lam%g_parity" % lam] = {k: p[k] for k in ("cost_rel_p99", "cost_rel_max",
                                                                                 "plan_abs_max", "u0_rel")}
    return out


def episode_extra(capi, dist):
    """configs[0]-style closed loop: a 200-step Pendulum episode resident on the device (mpc_episode_run) next to the
    same episode driven from the host, one predict_action-sized call per step."""
    out = {}
    one = One(dist)
    n = 200
    s = make_solver(capi, "MPPI", "pendulum", 4096, 50, dict(lam=0.5, sigma=1.0), one)
    x0 = np.array(START["pendulum"])
    s.episode_run(x0, 20, U_in=np.zeros(50))
    t = time.perf_counter()
    ep = s.episode_run(x0, n, U_in=np.zeros(50), actuation_std=0.1, sensor_std=0.1, seed=1)
    dt_dev = time.perf_counter() - t
    out["episode_pendulum_k4096_h50_device_us_per_control_step"] = round(1e6 * dt_dev / n, 2)
    out["episode_pendulum_k4096_h50_device_steps_per_s"] = 4096 * 50 * n / dt_dev
    out["episode_pendulum_k4096_h50_final_cost"] = float(ep["costs"][-1])
    U = np.zeros(50)
    x = x0.copy()
    t = time.perf_counter()
    for _ in range(n):  # the host loop: solve (H2D + launch + D2H) then one plant step through mpc_model_step
        u0, Un, xs_, c_ = s.mppi_solve(x, None, U, want_traj=True)
        Uc = np.clip(Un[0], -2.0, 2.0)
        U = np.append(Uc[1:], 0.0)
        _, xn = s.model_step(x[None, :], np.clip(u0, -2.0, 2.0))
        x = xn[0]
    out["episode_pendulum_k4096_h50_hostloop_us_per_control_step"] = round(1e6 * (time.perf_counter() - t) / n, 2)
    s.close()
    return out


def cpu_baseline_sample(env, K_full, H, prm, budget_s=12.0):
    """The oracle port (oracle/mpc_oracle.c, pthreads over samples) timed on this host: a bounded sample of the same
    workload (same env, H, lam, std; fewer samples)."""
    from oracle import oracle as orc
    m = orc.model(env)
    cores = orc.max_threads()
    rng = np.random.RandomState(0)
    x = np.array(START[env]) + 0.01
    U = np.zeros(H)

    def run(K):
        eps = rng.standard_normal((K, H)) * prm["sigma"]
        t = time.perf_counter()
        S = orc.mppi_costs(m, x, U, eps, prm["lam"], prm["sigma"], nthreads=cores)
        orc.mppi_update(S, eps, prm["lam"], U)
        return time.perf_counter() - t

    k_probe = min(K_full, 16384)
    run(k_probe)
    t_probe = run(k_probe)
    K = int(min(K_full, max(k_probe, k_probe * budget_s / max(t_probe, 1e-6))))
    K = min(K, 1 << 21)
    t = run(K)
    return {"value": K * H / t, "unit": "rollout-steps/s", "cores": cores, "kind": "port",
            "sample": "MPPI %s H=%d, K=%d of %d samples, one solve (rollout + update), %.1f s; C port of the "
                      "reference algorithm (oracle/mpc_oracle.c), %d pthreads" % (env, H, K, K_full, t, cores)}, K, t


def run_reference_arm(args, dist):
    agent, env, K, H, prm = WORKLOADS[args.workload]
    if dist.rank != 0:
        return
    from oracle import oracle as orc
    orc.build()
    m = orc.model(env)
    cores = orc.max_threads()
    rng = np.random.RandomState(0)
    x = np.array(START[env]) + 0.01
    U = np.zeros(H)
    # bounded sample per step: about 2 s of CPU work
    k_probe = 8192
    eps = rng.standard_normal((k_probe, H))
    t = time.perf_counter()
    orc.mppi_costs(m, x, U, eps, prm["lam"], prm["sigma"], nthreads=cores)
    rate = k_probe * H / (time.perf_counter() - t)
    Ks = int(min(K, max(k_probe, 2.0 * rate / H)))
    eps = rng.standard_normal((Ks, H)) * prm["sigma"]

    def step():
        S = orc.mppi_costs(m, x, U, eps, prm["lam"], prm["sigma"], nthreads=cores)
        orc.mppi_update(S, eps, prm["lam"], U)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = Ks * H * args.steps / dt
    sample = "MPPI %s H=%d: each step = one solve of K=%d of the %d samples on %d host threads" % (env, H, Ks, K, cores)
    line = {
        "impl": "reference", "metric": "rollout_steps_per_sec", "value": value, "unit": "rollout-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "agent": agent, "env": env, "K": K, "H": H, **prm,
                   "note": "CPU port of the reference algorithm; the Python reference is not on this box"},
        "cpu_baseline": {"value": value, "unit": "rollout-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "rollout-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_ours(args, dist):
    import mpc_b200
    from mpc_b200 import capi
    mpc_b200.build_library()
    agent, env, K, H, prm = WORKLOADS[args.workload]
    peaks, peaks_src = measured_peaks()
    fma_peak = {dt: capi.measure_fma_peak(dist.local_rank, dt) for dt in ("f32", "f64")}
    common = dict(peaks=peaks, peaks_src=peaks_src)
    r = run_mppi_workload(capi, dist, args.workload, args.steps, args.warmup, dtype=args.dtype, noise=args.noise,
                          lanes=args.lanes, fast_math=args.fast_math, fma_peak=fma_peak[args.dtype], **common)
    line = {
        "metric": "rollout_steps_per_sec", "value": r["value"], "unit": "rollout-steps/s", "n_gpus": dist.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": args.workload, "agent": agent, "env": env, "K": K, "H": H, **prm,
                   "noise": args.noise, "lanes_per_thread": r["lanes"], "fast_math": bool(r["fast_math"]),
                   "K_per_gpu": r["K_per_gpu"],
                   "timed_step": "one MPPI._calc_action: rollout of K x H + softmin update of the plan, one kernel; "
                                 "K solves enqueued back to back, one CUDA-event pair each (e2e adds the host "
                                 "buffers and the logging rollout of Agent.predict_action)",
                   "cache": ("L2 flushed (256 MiB write, untimed) between timed steps" if r["flush"] else
                             "noise shard %.0f MB per GPU > L2 (126 MB): no flush" % (r["shard_bytes"] / 1e6)),
                   "parallelism": "K-shard x%d, one all-gather of (beta, eta, E[H]) per solve (%s)" % (
                       dist.world, "single GPU" if dist.world == 1 else
                       ("inside the solve kernel over NVLink peer memory"
                        if os.environ.get("MPC_B200_EXCHANGE", "p2p") == "p2p" else "NCCL + merge kernel"))},
        "clocks": r["clocks"], "gpu_launches": r["gpu_launches"], "e2e": r["e2e"],
        "roofline": r["roofline"], "roofline_other": r["roofline_other"], "wall_s_timed_loop": r["wall_s_timed_loop"],
        "solve_phases_us": r["solve_phases_us"], "geometry": r["geometry"],
        "step_ms_min_median_max": r["step_ms_min_median_max"],
    }
    for k in ("parity", "sharded_vs_single"):
        if k in r:
            line[k] = r[k]
    if not args.no_extra:
        extra = {}

        def brief(w):
            keep = {k: w[k] for k in ("value", "ms_per_step", "step_ms_min_median_max", "e2e", "solve_phases_us", "parity",
                                      "sharded_vs_single", "K_per_gpu", "lanes", "fast_math") if k in w}
            keep["roofline"] = {k: w["roofline"][k] for k in ("bound", "achieved", "peak", "unit", "frac", "kernel_ms")}
            keep["roofline_other"] = {k: w["roofline_other"][k] for k in ("bound", "achieved", "peak", "unit", "frac")}
            return keep
        try:
            # north_star's own target configuration (K = 2^20, H = 50), K-sharded over all ranks like the headline
            for wl in ("mppi_pendulum_h50_k1m", "mppi_cartpole_h50_k1m"):
                if wl != args.workload:
                    extra[wl] = brief(run_mppi_workload(capi, dist, wl, 10, 3, fma_peak=fma_peak["f32"], **common))
            extra.update(batch_extra(capi, dist))  # configs[4]: replicas over the ranks
            if dist.world == 1:
                # float64 (the reference's own arithmetic, parity unconditional) against the measured DFMA peak
                for wl in (args.workload, "mppi_pendulum_h50_k1m", "mppi_cartpole_h50_k1m"):
                    w = run_mppi_workload(capi, dist, wl, 5, 3, dtype="f64", with_parity=False, with_e2e=False,
                                          fma_peak=fma_peak["f64"], **common)
                    extra[wl + "_f64"] = brief(w)
                for wl in ("mppi_pendulum_h50_k1m", "mppi_cartpole_h50_k1m", args.workload):
                    w = run_mppi_workload(capi, dist, wl, 5, 3, noise="philox", with_parity=False, with_e2e=False,
                                          fma_peak=fma_peak["f32"], **common)
                    extra[wl + "_philox"] = brief(w)
                if r["fast_math"]:  # the agents' default flavour (polynomial sincos) of the headline workload
                    w = run_mppi_workload(capi, dist, args.workload, 5, 3, fast_math=0, lanes=2, with_e2e=False,
                                          fma_peak=fma_peak["f32"], **common)
                    extra[args.workload + "_polynomial_trig"] = brief(w)
                extra.update(latency_extras(capi, dist))
                extra.update(cem_extra(capi, dist, fma_peak["f32"]))                      # BASELINE configs[1]
                extra.update(cem_extra(capi, dist, fma_peak["f32"], K=65536, reps=30, warm=5))
                extra.update(cem_extra(capi, dist, fma_peak["f32"], K=1 << 20, reps=6, warm=2))
                extra.update(temperature_extra(capi, dist))
                extra.update(episode_extra(capi, dist))
        except Exception as e:  # extras must never take the headline line down
            import traceback
            extra["error"] = repr(e) + " | " + traceback.format_exc().splitlines()[-2].strip()
        line["extra"] = extra
    if dist.rank == 0:
        if dist.world == 1:
            cb, _, _ = cpu_baseline_sample(env, K, H, prm)
            line["cpu_baseline"] = cb
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--noise", default="tensor", choices=["tensor", "philox"])
    ap.add_argument("--lanes", type=int, default=0, choices=[0, 1, 2], help="f32 samples per thread (0 = default 2)")
    ap.add_argument("--fast-math", type=int, default=-1, choices=[-1, 0, 1],
                    help="f32: MUFU trig instead of polynomial (-1: per-env flavour that meets the 1e-5 parity bar)")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        class R0:
            world = int(os.environ.get("WORLD_SIZE", "1"))
            rank = int(os.environ.get("RANK", "0"))
        run_reference_arm(args, R0())
        return
    dist = Dist(args.gpus)
    try:
        run_ours(args, dist)
    finally:
        dist.close()


if __name__ == "__main__":
    main()

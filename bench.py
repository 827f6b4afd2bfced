#!/usr/bin/env python
"""bench.py — rollout-steps/s of the sampling-MPC hot path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME] [--no-extra]

One "step" = one MPPI solve (rollout of K samples x H steps + softmin update + nominal rollout).
Headline workload (BASELINE.json configs[2], largest K): MPPI on Acrobot, H = 100, K = 2^20, lam = 0.5, std = 1,
float32 rollout arithmetic with float64 accumulation, resident noise tensor (the reference's delta_u semantics,
mppi.py:20-21).  Only the sample dimension shards: each of N ranks rolls out K/N samples and one all-gather of
(min cost, sum w, sum w*eps[H]) per rank combines them => "strong" scaling (total work fixed).

value   device-timed (CUDA events on the solver's stream, inputs resident in HBM, max over ranks)
e2e     same metric through the host-buffer C-ABI call (H2D of state+plan, D2H of action+plan+trajectory inside)
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
# float32 rollout flavour quoted per env: (fast_math, lanes).  fast_math = MUFU.SIN/COS instead of the polynomial sincos;
# it is used where EVERY per-sample cost stays within the north_star's rel 1e-5 of the float64 oracle
# (tests/test_gpu_parity.py::test_mppi_mufu_trig_flavour_meets_the_float32_bar, tools/variant_probe.py): Pendulum and
# Acrobot.  CartPole keeps the polynomial (its chaotic tail leaves no margin).  lanes: Acrobot's MUFU flavour is XU-pipe
# bound and runs best one sample per thread with 16 warps; the others use packed f32x2.
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
def make_solver(capi, orc_params, agent, env, K, H, prm, dist, dtype="f32", noise="tensor", **kw):
    from mpc_b200.models import ENV_NAMES, env_spec
    if dtype == "f32" and agent == "MPPI":
        if kw.get("fast_math", -1) < 0:
            kw["fast_math"] = FLAVOUR[env][0]
        if not kw.get("lanes"):
            kw["lanes"] = FLAVOUR[env][1]
    else:
        kw["fast_math"] = max(0, kw.get("fast_math", 0))
    model = ENV_NAMES[env]()
    env_id, env_params = env_spec(model)
    lo, hi = float(model.bounds_low[0]), float(model.bounds_high[0])
    s = capi.Solver(agent, env_id, env_params, K=K, H=H, dtype=dtype, noise=noise, bounds=(lo, hi),
                    device=dist.local_rank, rank=dist.rank if dist.world > 1 else 0, world=dist.world, seed=0,
                    **prm, **kw)
    if dist.world > 1:
        if agent == "MPPI" and os.environ.get("MPC_B200_EXCHANGE", "p2p") == "p2p":
            from mpc_b200 import parallel
            parallel.connect_p2p(s)  # in-kernel all-gather over NVLink peer memory: one launch per solve
        else:
            s.comm_init(dist.comm_id(capi.Solver.comm_create_id))
    if agent == "MPPI" and noise == "tensor":
        s.fill_noise()  # synthetic N(0, sigma) tensor drawn on the device (Philox); resident like agent.delta_u
    return s


def time_device(s, steps, warmup, dist, flush):
    """Per-step CUDA-event timing on the solver's stream; L2 flushed (untimed) between steps when asked."""
    for _ in range(warmup):
        s.solve_enqueue()
        if flush:
            s.flush_l2()
    s.sync()
    dist.barrier()
    total_ms, roll_ms = 0.0, 0.0
    l0 = s.launch_count
    n_flush = 0
    t_wall = time.perf_counter()
    for _ in range(steps):
        s.timer_start()
        s.solve_enqueue()
        total_ms += s.timer_stop()
        if s.is_mppi:
            roll_ms += s.rollout_ms()
        if flush:
            s.flush_l2()
            n_flush += 1
    s.sync()
    dist.barrier()
    wall = time.perf_counter() - t_wall
    launches = s.launch_count - l0 - n_flush
    return dist.max(total_ms), dist.max(roll_ms), launches, wall


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


def latency_extras(capi, dist, reps=200, warm=20):
    """p50 solve latency at K = 4096 (north_star: sub-ms) through the host-buffer call, 1 GPU."""
    out = {}

    class One:  # single-GPU view
        world, rank, local_rank = 1, 0, dist.local_rank
    one = One()
    for env in ("pendulum", "cartpole", "acrobot"):
        H = 50
        s = make_solver(capi, None, "MPPI", env, 4096, H, dict(lam=0.5, sigma=1.0), one)
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
    # BASELINE configs[1]: CEM CartPole K=4096 H=50 n_elite=409 5 iterations
    s = make_solver(capi, None, "CEM", "cartpole", 4096, 50, dict(sigma=1.0, max_iter=5, n_elite=409, epsilon=1e-5,
                                                                 alpha=0.2), one, noise="philox")
    x = np.array(START["cartpole"]) + 0.01
    for _ in range(warm):
        s.cem_solve(x, None, np.zeros(50))
    lat = []
    for _ in range(reps):
        t = time.perf_counter()
        *_, iters = s.cem_solve(x, None, np.zeros(50))
        lat.append(time.perf_counter() - t)
    p50 = statistics.median(lat)
    out["cem_cartpole_k4096_h50_5it_f32_philox_p50_ms"] = round(1e3 * p50, 4)
    out["cem_cartpole_k4096_h50_5it_steps_per_s"] = 4096 * 50 * iters / p50
    s.close()
    return out


def config_extras(capi, dist):
    """The other BASELINE.json configs on one GPU (reported in `extra`, not the headline):
    configs[4]: batch of 4096 independent CartPole MPPI solves (K = 1024, H = 50) in ONE launch;
    configs[3]: MPPI Pendulum K = 65536, H = 50 over the temperature grid of scripts/mppi_temperature.py:23;
    configs[0]-style closed loop: a 200-step Pendulum episode resident on the device (mpc_episode_run) next to the
    same episode driven from the host one predict_action-sized call per step."""
    out = {}
    one = dist  # extras run on world == 1 only
    # ---- configs[4]
    B, K, H = 4096, 1024, 50
    s = make_solver(capi, None, "MPPI", "cartpole", K, H, dict(lam=0.5, sigma=1.0), one, batch=B)
    s.set_profiling(True)
    rng = np.random.RandomState(0)
    xs = np.array(START["cartpole"]) + rng.normal(0, 0.2, (B, 4))  # cartpole_swingup_model.py:50-51
    s.upload_inputs(xs, None, np.zeros((B, H)))
    ms, _, _, _ = time_device(s, 10, 3, one, flush=True)
    out["batch4096_cartpole_k1024_h50_f32_steps_per_s"] = B * K * H * 10 / (ms * 1e-3)
    out["batch4096_cartpole_k1024_h50_f32_ms_per_batch"] = ms / 10
    t = time.perf_counter()
    for _ in range(5):
        s.mppi_solve(xs, None, np.zeros((B, H)), want_traj=True)
    out["batch4096_cartpole_k1024_h50_f32_e2e_steps_per_s"] = B * K * H * 5 / (time.perf_counter() - t)
    s.close()
    # ---- configs[3]
    for lam in (0.001, 0.01, 0.1, 0.5, 1.0, 10.0):
        s = make_solver(capi, None, "MPPI", "pendulum", 65536, 50, dict(lam=lam, sigma=1.0), one)
        x, U = np.array(START["pendulum"]), np.zeros(50)
        for _ in range(10):
            s.mppi_solve(x, None, U)
        lat = []
        for _ in range(100):
            t = time.perf_counter()
            s.mppi_solve(x, None, U)
            lat.append(time.perf_counter() - t)
        out["mppi_pendulum_k65536_h50_lam%g_p50_ms" % lam] = round(1e3 * statistics.median(lat), 4)
        s.close()
    # ---- closed loop
    n = 200
    s = make_solver(capi, None, "MPPI", "pendulum", 4096, 50, dict(lam=0.5, sigma=1.0), one)
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
    s = make_solver(capi, None, agent, env, K, H, prm, dist, dtype=args.dtype, noise=args.noise, lanes=args.lanes,
                    fast_math=args.fast_math)
    s.set_profiling(True)
    esz = 4 if args.dtype == "f32" else 8
    shard_bytes = s.K_local * H * esz
    flush = args.noise == "tensor" and shard_bytes < 2 * L2_BYTES or args.noise == "philox"
    x = np.array(START[env]) + 0.01
    U = np.zeros(H)
    s.upload_inputs(x, None, U)

    clocks = ClockSampler(dist.local_rank)
    clocks.start()
    total_ms, roll_ms, launches, wall = time_device(s, args.steps, args.warmup, dist, flush)
    clk = clocks.stop()
    steps_total = float(K) * H * args.steps
    value = steps_total / (total_ms * 1e-3)

    # phase stamps of the last timed solve's finalising block on rank 0 (us since its kernel start): where a solve's
    # time goes when the shard is small (tiles_done = rollouts finished; peers_in = every rank's record has arrived)
    try:
        tr = s.trace()
        trace_us = {k: round(v / 1e3, 1) for k, v in tr.items() if v is not None and k != "start"}
    except Exception:
        trace_us = None
    e2e_total, lat = time_e2e(s, x, U, args.steps, args.warmup, dist)
    e2e_value = steps_total / e2e_total
    h2d = (s.ns + H) * 8
    d2h = (2 + H + (H + 1) * s.ns + 1) * 8

    # roofline of the dominant kernel (mppi_rollout_kernel): algorithmic flops / its own event-timed duration
    roll_s = roll_ms * 1e-3 / args.steps
    local_steps = float(s.K_local) * H
    fma_peak = capi.measure_fma_peak(dist.local_rank, args.dtype) if dist.rank == 0 else 0.0
    ach_tflops = local_steps * ALG_FLOPS[env] / roll_s / 1e12
    roofline = {"bound": "fp32_pipe" if args.dtype == "f32" else "fp64_pipe", "kernel": "mppi_solve_kernel",
                "achieved": ach_tflops, "peak": fma_peak, "unit": "TFLOP/s",
                "frac": ach_tflops / fma_peak if fma_peak else None, "traffic": None,
                "peak_source": "FMA-chain microbenchmark run live (mpc_measure_fma_peak); MEASURED_PEAKS.json has no "
                               "CUDA-core figure", "alg_flops_per_step": ALG_FLOPS[env],
                "kernel_ms": roll_ms / args.steps}
    ev = ncu_evidence(args.workload, s.fast_math, s.lanes) if (args.noise == "tensor" and dist.world == 1) else None
    if ev:
        roofline["traffic"] = ev["dram_bytes_per_launch"]
        roofline["traffic_over_algorithmic"] = ev["dram_bytes_per_launch"] / (local_steps * esz)
        roofline["ncu"] = {k: ev[k] for k in ev if k != "dram_bytes_per_launch"}
    hbm_ach = local_steps * esz / roll_s / 1e9 if args.noise == "tensor" else 0.0
    roofline_hbm = {"bound": "hbm", "achieved": hbm_ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": hbm_ach / peaks["hbm_gbs"], "traffic": ev["dram_bytes_per_launch"] if ev else None,
                    "peak_source": peaks_src,
                    "alg_bytes_per_step": esz if args.noise == "tensor" else 0}

    line = {
        "metric": "rollout_steps_per_sec", "value": value, "unit": "rollout-steps/s", "n_gpus": dist.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": args.workload, "agent": agent, "env": env, "K": K, "H": H, **prm,
                   "noise": args.noise, "lanes_per_thread": s.lanes, "fast_math": bool(s.fast_math), "K_per_gpu": s.K_local,
                   "cache": ("L2 flushed (256 MiB write) between timed steps" if flush else
                             "noise shard %.0f MB per GPU > L2 (126 MB): no flush" % (shard_bytes / 1e6)),
                   "parallelism": "K-shard x%d, one all-gather of (beta, eta, E[H]) per solve (%s)" % (
                       dist.world, "single GPU" if dist.world == 1 else
                       ("inside the solve kernel over NVLink peer memory"
                        if os.environ.get("MPC_B200_EXCHANGE", "p2p") == "p2p" else "NCCL + merge kernel"))},
        "clocks": clk, "gpu_launches": launches,
        "e2e": {"value": e2e_value, "unit": "rollout-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "p50_solve_ms": round(1e3 * statistics.median(lat), 4)},
        "roofline": roofline, "roofline_hbm": roofline_hbm, "wall_s_timed_loop": wall, "solve_phases_us": trace_us,
    }
    s.close()
    if dist.rank == 0:
        if dist.world == 1:
            cb, _, _ = cpu_baseline_sample(env, K, H, prm)
            line["cpu_baseline"] = cb
            if not args.no_extra:
                extra = {}
                try:
                    extra.update(latency_extras(capi, dist))
                    extra.update(config_extras(capi, dist))
                    one = dist
                    for wl in ("mppi_pendulum_h50_k1m", "mppi_cartpole_h50_k1m"):
                        a2, e2, K2, H2, p2 = WORKLOADS[wl]
                        for noise in ("tensor", "philox"):
                            s2 = make_solver(capi, None, a2, e2, K2, H2, p2, one, dtype="f32", noise=noise)
                            s2.set_profiling(True)
                            s2.upload_inputs(np.array(START[e2]) + 0.01, None, np.zeros(H2))
                            ms, rms, _, _ = time_device(s2, 10, 3, one, flush=(noise == "philox"))
                            extra["%s_%s_f32_steps_per_s" % (wl, noise)] = K2 * H2 * 10 / (ms * 1e-3)
                            extra["%s_%s_f32_rollout_kernel_ms" % (wl, noise)] = rms / 10
                            s2.close()
                    s3 = make_solver(capi, None, agent, env, K, H, prm, one, dtype="f32", noise="philox")
                    s3.set_profiling(True)
                    s3.upload_inputs(x, None, U)
                    ms, rms, _, _ = time_device(s3, 5, 3, one, flush=True)
                    extra["%s_philox_f32_steps_per_s" % args.workload] = K * H * 5 / (ms * 1e-3)
                    s3.close()
                except Exception as e:  # extras must never take the headline line down
                    extra["error"] = repr(e)
                line["extra"] = extra
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

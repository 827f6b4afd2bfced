"""ctypes binding of libmpcb200.so — the repo's own user of the C ABI in include/mpc_b200.h.

There is no CPU fallback: if the library is missing, cannot be loaded, or no CUDA device is present, the calls raise.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmpcb200.so")

MPC_OK, MPC_EINVAL, MPC_ECUDA, MPC_ENCCL, MPC_ENOMEM, MPC_ESTATE = 0, -1, -2, -3, -4, -5
AGENT_MPPI, AGENT_CEM = 0, 1
ENV_PENDULUM, ENV_CARTPOLE, ENV_ACROBOT = 0, 1, 2
F64, F32 = 0, 1
NOISE_TENSOR, NOISE_PHILOX = 0, 1
COMM_ID_BYTES = 128
MAX_ENV_PARAMS = 24

P2P_HANDLE_BYTES = 64
DTYPES = {"f64": F64, "float64": F64, "fp64": F64, "f32": F32, "float32": F32, "fp32": F32}
NOISES = {"tensor": NOISE_TENSOR, "numpy": NOISE_TENSOR, "philox": NOISE_PHILOX}

# every symbol include/mpc_b200.h declares
SYMBOLS = [
    "mpc_create", "mpc_destroy", "mpc_last_error", "mpc_abi_version", "mpc_set_noise", "mpc_fill_noise",
    "mpc_philox_noise", "mpc_mppi_solve", "mpc_mppi_get_costs", "mpc_cem_begin", "mpc_cem_iterate", "mpc_cem_end",
    "mpc_cem_solve", "mpc_cem_get_rewards", "mpc_cem_get_elites", "mpc_cem_get_std", "mpc_simulate", "mpc_model_step",
    "mpc_comm_create_id", "mpc_comm_init", "mpc_mppi_rollout_partial", "mpc_mppi_apply_partials", "mpc_upload_inputs", "mpc_solve_enqueue", "mpc_solve_enqueue_ex", "mpc_time_solves", "mpc_download_outputs",
    "mpc_timer_start", "mpc_timer_stop", "mpc_sync", "mpc_set_profiling", "mpc_get_rollout_ms", "mpc_launch_count", "mpc_get_geometry", "mpc_measure_fma_peak",
    "mpc_flush_l2", "mpc_debug_tickets", "mpc_pin_host_buffer", "mpc_unpin_host_buffer", "mpc_p2p_export", "mpc_p2p_import", "mpc_p2p_link", "mpc_get_trace", "mpc_episode_run", "mpc_philox_model_noise", "mpc_cem_get_iterations",
    "mpc_mppi_update_static",
]


class MpcConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("agent", C.c_int32), ("env", C.c_int32), ("dtype", C.c_int32),
        ("noise", C.c_int32), ("device", C.c_int32), ("K", C.c_int32), ("H", C.c_int32), ("batch", C.c_int32),
        ("rank", C.c_int32), ("world", C.c_int32), ("max_iter", C.c_int32), ("n_elite", C.c_int32),
        ("resample", C.c_int32), ("fast_math", C.c_int32), ("lanes", C.c_int32),
        ("lam", C.c_double), ("sigma", C.c_double), ("epsilon", C.c_double), ("alpha", C.c_double),
        ("bounds_low", C.c_double), ("bounds_high", C.c_double), ("seed", C.c_uint64),
        ("env_params", C.c_double * MAX_ENV_PARAMS), ("model_noise_std", C.c_double),
        ("full_W", C.c_int32), ("per_step_goal", C.c_int32), ("W", C.c_double * 25),
    ]


class MpcEpisodeOpts(C.Structure):  # include/mpc_b200.h: mpc_episode_opts
    _fields_ = [("struct_size", C.c_uint32), ("n_steps", C.c_int32), ("warm_start_iters", C.c_int32),
                ("shift", C.c_int32), ("actuation_std", C.c_double), ("sensor_std", C.c_double), ("seed", C.c_uint64)]


class MpcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libmpcb200 error %d: %s" % (code, msg))
        self.code = code


_lib = None
_dp = C.POINTER(C.c_double)


def load():
    """Load libmpcb200.so.  Raises if it has not been built — there is deliberately no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: build it with `python %s` (nvcc, sm_100a). "
                           "This package has no CPU or PyTorch fallback." % (LIB_PATH, os.path.join(HERE, "build.py")))
    L = C.CDLL(LIB_PATH)
    hp = C.c_void_p
    L.mpc_create.argtypes = [C.POINTER(MpcConfig), C.POINTER(hp)]
    L.mpc_destroy.argtypes = [hp]
    L.mpc_last_error.argtypes = [hp]
    L.mpc_last_error.restype = C.c_char_p
    L.mpc_abi_version.argtypes = []
    L.mpc_set_noise.argtypes = [hp, _dp]
    L.mpc_fill_noise.argtypes = [hp]
    L.mpc_philox_noise.argtypes = [hp, C.c_uint64, _dp]
    L.mpc_mppi_solve.argtypes = [hp] + [C.c_void_p] * 7  # plain addresses: the wrapper caches them (hot call)
    L.mpc_mppi_get_costs.argtypes = [hp, _dp]
    L.mpc_cem_begin.argtypes = [hp, _dp, _dp, _dp]
    L.mpc_cem_iterate.argtypes = [hp, _dp, C.POINTER(C.c_int32)]
    L.mpc_cem_end.argtypes = [hp, _dp, _dp, _dp, _dp]
    L.mpc_cem_solve.argtypes = [hp] + [C.c_void_p] * 7 + [C.POINTER(C.c_int32)]
    L.mpc_cem_get_rewards.argtypes = [hp, C.c_int32, _dp]
    L.mpc_cem_get_elites.argtypes = [hp, C.c_int32, C.POINTER(C.c_int32)]
    L.mpc_cem_get_std.argtypes = [hp, _dp]
    L.mpc_simulate.argtypes = [hp, _dp, _dp, _dp, _dp, _dp]
    L.mpc_model_step.argtypes = [hp, C.c_int32, _dp, _dp, _dp, _dp, _dp]
    L.mpc_comm_create_id.argtypes = [C.c_void_p]
    L.mpc_comm_init.argtypes = [hp, C.c_void_p]
    L.mpc_p2p_export.argtypes = [hp, C.c_void_p]
    L.mpc_p2p_import.argtypes = [hp, C.c_void_p]
    L.mpc_p2p_link.argtypes = [C.POINTER(hp), C.c_int32]
    L.mpc_get_trace.argtypes = [hp, C.POINTER(C.c_uint64)]
    L.mpc_cem_get_iterations.argtypes = [hp, C.POINTER(C.c_int32)]
    L.mpc_philox_model_noise.argtypes = [hp, C.c_uint64, _dp]
    L.mpc_episode_run.argtypes = [hp, _dp, _dp, C.POINTER(MpcEpisodeOpts)] + [_dp] * 8
    L.mpc_mppi_rollout_partial.argtypes = [hp, _dp, _dp, _dp, _dp]
    L.mpc_mppi_apply_partials.argtypes = [hp, _dp, _dp, _dp, _dp, _dp]
    L.mpc_upload_inputs.argtypes = [hp, _dp, _dp, _dp]
    L.mpc_solve_enqueue.argtypes = [hp]
    L.mpc_solve_enqueue_ex.argtypes = [hp, C.c_int32]
    L.mpc_time_solves.argtypes = [hp, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_float)]
    L.mpc_download_outputs.argtypes = [hp, _dp, _dp, _dp, _dp]
    L.mpc_timer_start.argtypes = [hp]
    L.mpc_timer_stop.argtypes = [hp, C.POINTER(C.c_float)]
    L.mpc_sync.argtypes = [hp]
    L.mpc_set_profiling.argtypes = [hp, C.c_int32]
    L.mpc_get_rollout_ms.argtypes = [hp, C.POINTER(C.c_float)]
    L.mpc_launch_count.argtypes = [hp]
    L.mpc_launch_count.restype = C.c_int64
    L.mpc_get_geometry.argtypes = [hp, C.POINTER(C.c_int32)]
    L.mpc_measure_fma_peak.argtypes = [C.c_int32, C.c_int32, _dp]
    L.mpc_flush_l2.argtypes = [hp]
    L.mpc_debug_tickets.argtypes = [hp, C.POINTER(C.c_int32)]
    L.mpc_pin_host_buffer.argtypes = [C.c_void_p, C.c_size_t]
    L.mpc_unpin_host_buffer.argtypes = [C.c_void_p]
    L.mpc_mppi_update_static.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_double] + [C.c_void_p] * 5 + \
        [C.c_char_p, C.c_size_t]
    _lib = L
    return L


def mppi_update_static(device, K, H, lam, S_dev, eps_dev, U_in_dev, U_out_dev, stream=None):
    """mpc_mppi_update_static (include/mpc_b200.h): the MPPI control update on DEVICE addresses (ints), its scratch in a
    static device array — no allocation.  Enqueues on `stream` (a cudaStream_t as int; None = default) and returns."""
    err = C.create_string_buffer(512)
    rc = load().mpc_mppi_update_static(device, K, H, lam, S_dev, eps_dev, U_in_dev, U_out_dev, stream, err, len(err))
    if rc != 0:
        raise MpcError(rc, err.value.decode())


def _in(a, shape=None):
    """float64 C-contiguous view/copy + pointer (None -> NULL)."""
    if a is None:
        return None, None
    arr = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None and arr.size != int(np.prod(shape)):
        raise ValueError("expected %s elements, got array of shape %s" % (shape, arr.shape))
    return arr, arr.ctypes.data_as(_dp)


def _out(shape):
    arr = np.empty(shape, dtype=np.float64)
    return arr, arr.ctypes.data_as(_dp)


class Solver:
    """Thin object wrapper around one mpc_handle."""

    def __init__(self, agent, env, env_params, K, H, *, dtype="f64", noise="tensor", lam=1.0, sigma=1.0,
                 max_iter=1, n_elite=1, epsilon=0.0, alpha=0.0, bounds=(-1.0, 1.0), batch=1, device=0, rank=0,
                 world=1, seed=0, resample=False, fast_math=False, lanes=0, model_noise_std=0.0, W=None,
                 per_step_goal=False):
        self._h = None
        L = load()
        cfg = MpcConfig()
        cfg.struct_size = C.sizeof(MpcConfig)
        cfg.agent = AGENT_MPPI if str(agent).upper() == "MPPI" else AGENT_CEM
        cfg.env = int(env)
        cfg.dtype = DTYPES[dtype] if isinstance(dtype, str) else int(dtype)
        cfg.noise = NOISES[noise] if isinstance(noise, str) else int(noise)
        cfg.device, cfg.K, cfg.H, cfg.batch = int(device), int(K), int(H), int(batch)
        cfg.rank, cfg.world = int(rank), int(world)
        cfg.max_iter, cfg.n_elite, cfg.resample = int(max_iter), int(n_elite), int(bool(resample))
        cfg.fast_math, cfg.lanes = int(fast_math), int(lanes)  # fast_math: 0 / 1, Acrobot also 2, 4, 5 (mpc_b200.h)
        cfg.model_noise_std = float(model_noise_std)
        cfg.per_step_goal = int(bool(per_step_goal))
        if W is not None:  # the model's whole cost matrix (generic-cost flavour: any W, not only diagonal ones)
            Wm = np.asarray(W, dtype=np.float64)
            nz = 3 if int(env) == ENV_PENDULUM else 5
            if Wm.shape != (nz, nz):
                raise ValueError("W must be %dx%d for this env" % (nz, nz))
            cfg.full_W = 1
            for i, v in enumerate(Wm.reshape(-1)):
                cfg.W[i] = float(v)
        cfg.lam, cfg.sigma, cfg.epsilon, cfg.alpha = float(lam), float(sigma), float(epsilon), float(alpha)
        cfg.bounds_low, cfg.bounds_high = float(bounds[0]), float(bounds[1])
        cfg.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        if len(env_params) > MAX_ENV_PARAMS:
            raise ValueError("too many env params")
        for i, v in enumerate(env_params):
            cfg.env_params[i] = float(v)
        self.cfg = cfg
        self.L = L
        self.K, self.H, self.B = int(K), int(H), int(batch)
        self.rank, self.world = int(rank), int(world)
        self.ns = 2 if cfg.env == ENV_PENDULUM else 4
        self.nz = self.ns + 1
        self.is_mppi = cfg.agent == AGENT_MPPI
        self.tensor_noise = cfg.noise == NOISE_TENSOR
        self.fast_math = int(cfg.fast_math)
        self.lanes = int(lanes) if lanes else (2 if (dtype == "f32" and self.is_mppi) else 1)
        h = C.c_void_p()
        rc = L.mpc_create(C.byref(cfg), C.byref(h))
        if rc != MPC_OK:
            raise MpcError(rc, (L.mpc_last_error(None) or b"").decode())
        self._h = h
        geo = (C.c_int32 * 8)()
        self._ck(L.mpc_get_geometry(self._h, geo))
        self.geometry = dict(grid=geo[0], block=geo[1], smem=geo[2], tiles=geo[3], K_local=geo[4], k_begin=geo[5],
                             stages=geo[6], stream_rows=geo[7])
        self.K_local, self.k_begin = geo[4], geo[5]

    # -- plumbing
    def _ck(self, rc):
        if rc != MPC_OK:
            raise MpcError(rc, (self.L.mpc_last_error(self._h) or b"").decode())

    def close(self):
        if self._h is not None:
            for a in getattr(self, "_pinned", []):
                self.L.mpc_unpin_host_buffer(a.ctypes.data)
            self._pinned = []
            self.L.mpc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _gz(self, g_z):
        """None or an all-zero goal is sent as NULL (zero goal resident on the device)."""
        if g_z is None:
            return None, None
        g = np.asarray(g_z, dtype=np.float64)
        if not g.any():
            return None, None
        return _in(g, (self.B, self.H, self.nz))

    # -- noise
    def set_noise(self, delta_u):
        a, p = _in(delta_u, (self.K, self.H))
        self._ck(self.L.mpc_set_noise(self._h, p))

    def fill_noise(self):
        self._ck(self.L.mpc_fill_noise(self._h))

    def philox_noise(self, draw=0):
        out, p = _out((self.K_local, self.H))
        self._ck(self.L.mpc_philox_noise(self._h, int(draw), p))
        return out

    # -- MPPI
    # The solve calls are the hot host path (p50 latency at K = 4096 is ~30 us of kernel): inputs are staged into
    # persistent numpy buffers whose addresses were taken once, outputs land in persistent buffers and are returned
    # as copies — no per-call ctypes pointer objects (those cost more than the kernel launch).
    def _hot_buffers(self):
        hb = getattr(self, "_hb", None)
        if hb is None:
            B, H, ns = self.B, self.H, self.ns
            arrs = dict(x=np.empty((B, ns)), U=np.empty((B, H)), g=np.empty((B, H, self.nz)), u0=np.empty(B),
                        Uo=np.empty((B, H)), xs=np.empty((B if B == 1 else 0, H + 1, ns)), cost=np.empty(B))
            hb = self._hb = (arrs, {k: v.ctypes.data for k, v in arrs.items()})
        return hb

    def _stage_hot(self, x, g_z, U_in):
        arrs, ptr = self._hot_buffers()
        arrs["x"][...] = np.reshape(x, arrs["x"].shape)
        up = gp = None
        if U_in is not None:
            arrs["U"][...] = np.reshape(U_in, arrs["U"].shape)
            up = ptr["U"]
        if g_z is not None:
            g = np.asarray(g_z, dtype=np.float64)
            if g.any():  # an all-zero goal is sent as NULL (zero goal resident on the device)
                arrs["g"][...] = np.reshape(g, arrs["g"].shape)
                gp = ptr["g"]
        return arrs, ptr, gp, up

    def mppi_solve(self, x, g_z=None, U_in=None, want_traj=True, copy=True):
        arrs, ptr, gp, up = self._stage_hot(x, g_z, U_in)
        if self.B > 1:
            # batches: the outputs are megabytes.  The library copies every field straight from the device into these
            # arrays (pinned once: direct DMA); copy=False hands out the solver-owned arrays themselves, valid until
            # the next solve of this handle
            bo = getattr(self, "_batch_out", None)
            if bo is None:
                bo = self._batch_out = dict(u0=np.empty(self.B), U=np.empty((self.B, self.H)),
                                            xs=np.empty((self.B, self.H + 1, self.ns)), cost=np.empty(self.B))
                self._pinned = getattr(self, "_pinned", [])
                for a in bo.values():
                    if self.L.mpc_pin_host_buffer(a.ctypes.data, a.nbytes) == MPC_OK:
                        self._pinned.append(a)
            self._ck(self.L.mpc_mppi_solve(self._h, ptr["x"], gp, up, bo["u0"].ctypes.data, bo["U"].ctypes.data,
                                           bo["xs"].ctypes.data if want_traj else None,
                                           bo["cost"].ctypes.data if want_traj else None))
            if copy:
                return (bo["u0"].copy(), bo["U"].copy(), bo["xs"].copy() if want_traj else None,
                        bo["cost"].copy() if want_traj else None)
            return bo["u0"], bo["U"], (bo["xs"] if want_traj else None), (bo["cost"] if want_traj else None)
        if want_traj:
            self._ck(self.L.mpc_mppi_solve(self._h, ptr["x"], gp, up, ptr["u0"], ptr["Uo"], ptr["xs"], ptr["cost"]))
            return arrs["u0"].copy(), arrs["Uo"].copy(), arrs["xs"].copy(), arrs["cost"].copy()
        self._ck(self.L.mpc_mppi_solve(self._h, ptr["x"], gp, up, ptr["u0"], ptr["Uo"], None, None))
        return arrs["u0"].copy(), arrs["Uo"].copy(), None, None

    def mppi_costs(self):
        out, p = _out((self.B, self.K_local))
        self._ck(self.L.mpc_mppi_get_costs(self._h, p))
        return out

    # -- CEM
    def cem_begin(self, x, g_z=None, U_in=None):
        xa, xp = _in(x, (self.ns,))
        ga, gp = self._gz(g_z)
        ua, up = _in(U_in, (self.H,))
        self._ck(self.L.mpc_cem_begin(self._h, xp, gp, up))

    def cem_iterate(self, z=None, wait=True):
        za, zp = _in(z, (self.K, self.H))
        if not wait:  # enqueue only (several ranks driven by one thread)
            self._ck(self.L.mpc_cem_iterate(self._h, zp, None))
            return None
        done = C.c_int32(0)
        self._ck(self.L.mpc_cem_iterate(self._h, zp, C.byref(done)))
        return bool(done.value)

    def cem_end(self, want_traj=True):
        u0, u0p = _out((1,))
        U, Up = _out((self.H,))
        if want_traj:
            xs, xsp = _out((self.H + 1, self.ns))
            cost, cp = _out((1,))
        else:
            xs = cost = xsp = cp = None
        self._ck(self.L.mpc_cem_end(self._h, u0p, Up, xsp, cp))
        return u0, U, xs, cost

    def cem_solve(self, x, g_z=None, U_in=None, want_traj=True):
        arrs, ptr, gp, up = self._stage_hot(x, g_z, U_in)
        iters = C.c_int32(0)
        self._ck(self.L.mpc_cem_solve(self._h, ptr["x"], gp, up, ptr["u0"], ptr["Uo"], ptr["xs"] if want_traj else None,
                                      ptr["cost"] if want_traj else None, C.byref(iters)))
        if want_traj:
            return arrs["u0"].copy(), arrs["Uo"][0].copy(), arrs["xs"][0].copy(), arrs["cost"].copy(), iters.value
        return arrs["u0"].copy(), arrs["Uo"][0].copy(), None, None, iters.value

    def cem_iterations(self):
        n = C.c_int32(0)
        self._ck(self.L.mpc_cem_get_iterations(self._h, C.byref(n)))
        return n.value

    def cem_rewards(self, it):
        out, p = _out((self.K_local,))
        self._ck(self.L.mpc_cem_get_rewards(self._h, int(it), p))
        return out

    def cem_elites(self, it):
        out = np.empty(self.cfg.n_elite, dtype=np.int32)
        self._ck(self.L.mpc_cem_get_elites(self._h, int(it), out.ctypes.data_as(C.POINTER(C.c_int32))))
        return out

    def cem_std(self):
        out, p = _out((self.H,))
        self._ck(self.L.mpc_cem_get_std(self._h, p))
        return out

    def philox_model_noise(self, solve_index):
        out, p = _out((self.K_local, self.H, 4))
        self._ck(self.L.mpc_philox_model_noise(self._h, int(solve_index), p))
        return out

    # -- nominal rollout / single steps
    def simulate(self, x, us, g_z=None):
        xa, xp = _in(x, (self.ns,))
        ua, up = _in(us, (self.H,))
        ga, gp = _in(g_z, (self.H, self.nz)) if g_z is not None else (None, None)
        xs, xsp = _out((self.H + 1, self.ns))
        cost = C.c_double()
        self._ck(self.L.mpc_simulate(self._h, xp, up, gp, xsp, C.byref(cost)))
        return xs, cost.value

    # -- closed loop resident on the device (Experiment.py:67-90 without the host in the loop)
    def episode_run(self, x0, n_steps, U_in=None, warm_start=0, shift=True, actuation_std=0.0, sensor_std=0.0, seed=0,
                    want_plans=True):
        """Returns a dict: states [n+1][ns] (observations), true_states, actions [n], costs [n], U_final [H] and, when
        want_plans, plan_us [n][H], plan_xs [n][H+1][ns], plan_costs [n] (the agent's planning histories)."""
        n = int(n_steps)
        opts = MpcEpisodeOpts(C.sizeof(MpcEpisodeOpts), n, int(warm_start), int(bool(shift)), float(actuation_std),
                              float(sensor_std), int(seed))
        xa, xp = _in(x0, (self.ns,))
        ua, up = _in(U_in, (self.H,)) if U_in is not None else (None, None)
        states, sp = _out((n + 1, self.ns))
        true_states, tp = _out((n + 1, self.ns))
        actions, ap = _out((n,))
        costs, cp = _out((n,))
        Uf, ufp = _out((self.H,))
        pus = pxs = pcs = None
        pup = pxp = pcp = None
        if want_plans:
            pus, pup = _out((n, self.H))
            pxs, pxp = _out((n, self.H + 1, self.ns))
            pcs, pcp = _out((n,))
        self._ck(self.L.mpc_episode_run(self._h, xp, up, C.byref(opts), sp, tp, ap, cp, pup, pxp, pcp, ufp))
        return dict(states=states, true_states=true_states, actions=actions, costs=costs, U_final=Uf, plan_us=pus,
                    plan_xs=pxs, plan_costs=pcs)

    def model_step(self, x, u, g=None):
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, self.ns)
        n = x.shape[0]
        xa, xp = _in(x)
        ua, up = _in(u, (n,))
        ga, gp = _in(g, (n, self.nz)) if g is not None else (None, None)
        cost, cp = _out((n,))
        xn, xnp = _out((n, self.ns))
        self._ck(self.L.mpc_model_step(self._h, n, xp, up, gp, cp, xnp))
        return cost, xn

    # -- multi-GPU
    @staticmethod
    def comm_create_id():
        buf = C.create_string_buffer(COMM_ID_BYTES)
        rc = load().mpc_comm_create_id(buf)
        if rc != MPC_OK:
            raise MpcError(rc, (load().mpc_last_error(None) or b"").decode())
        return bytes(buf.raw)

    def comm_init(self, id_bytes):
        buf = C.create_string_buffer(bytes(id_bytes), COMM_ID_BYTES)
        self._ck(self.L.mpc_comm_init(self._h, buf))

    def p2p_export(self):
        """64-byte IPC handle of this rank's exchange buffer (in-kernel all-gather over NVLink peer memory)."""
        buf = C.create_string_buffer(P2P_HANDLE_BYTES)
        self._ck(self.L.mpc_p2p_export(self._h, buf))
        return bytes(buf.raw)

    def p2p_import(self, handles):
        """handles: the ranks' exported handles, rank-major (bytes of length world * 64, or a list of bytes)."""
        raw = b"".join(handles) if isinstance(handles, (list, tuple)) else bytes(handles)
        if len(raw) != self.world * P2P_HANDLE_BYTES:
            raise ValueError("expected %d handle bytes, got %d" % (self.world * P2P_HANDLE_BYTES, len(raw)))
        buf = C.create_string_buffer(raw, len(raw))
        self._ck(self.L.mpc_p2p_import(self._h, buf))

    @staticmethod
    def p2p_link(solvers):
        """One process driving several GPUs: wire the rank-ordered solvers' exchange buffers to each other."""
        L = load()
        arr = (C.c_void_p * len(solvers))(*[s._h.value for s in solvers])
        rc = L.mpc_p2p_link(arr, len(solvers))
        if rc != MPC_OK:
            msgs = [(L.mpc_last_error(s._h) or b"").decode() for s in solvers]
            raise MpcError(rc, "; ".join(m for m in msgs if m))

    def mppi_rollout_partial(self, x, g_z=None, U_in=None):
        xa, xp = _in(x, (self.B, self.ns))
        ga, gp = self._gz(g_z)
        ua, up = _in(U_in, (self.B, self.H))
        rec, rp = _out((self.B, 2 + self.H))
        self._ck(self.L.mpc_mppi_rollout_partial(self._h, xp, gp, up, rp))
        return rec

    def mppi_apply_partials(self, recs, want_traj=True):
        ra, rp = _in(recs, (self.cfg.world, self.B, 2 + self.H))
        u0, u0p = _out((self.B,))
        U, Up = _out((self.B, self.H))
        if want_traj:
            xs, xsp = _out((self.B, self.H + 1, self.ns))
            cost, cp = _out((self.B,))
        else:
            xs = cost = xsp = cp = None
        self._ck(self.L.mpc_mppi_apply_partials(self._h, rp, u0p, Up, xsp, cp))
        return u0, U, xs, cost

    # -- measurement
    def upload_inputs(self, x, g_z=None, U_in=None):
        xa, xp = _in(x, (self.B, self.ns))
        ga, gp = self._gz(g_z)
        ua, up = _in(U_in, (self.B, self.H))
        self._ck(self.L.mpc_upload_inputs(self._h, xp, gp, up))

    def solve_enqueue(self, want_traj=True):
        self._ck(self.L.mpc_solve_enqueue_ex(self._h, int(bool(want_traj))))

    def time_solves(self, n, want_traj=True, flush_l2=False):
        """Device time (ms) of each of n back-to-back device-resident solves (own event pair each, one sync at the end)."""
        ms = (C.c_float * int(n))()
        self._ck(self.L.mpc_time_solves(self._h, int(n), int(bool(want_traj)), int(bool(flush_l2)), ms))
        return [float(v) for v in ms]

    def download_outputs(self):
        u0, u0p = _out((self.B,))
        U, Up = _out((self.B, self.H))
        xs, xsp = _out((self.B, self.H + 1, self.ns))
        cost, cp = _out((self.B,))
        self._ck(self.L.mpc_download_outputs(self._h, u0p, Up, xsp, cp))
        return u0, U, xs, cost

    def timer_start(self):
        self._ck(self.L.mpc_timer_start(self._h))

    def timer_stop(self):
        ms = C.c_float()
        self._ck(self.L.mpc_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def set_profiling(self, on=True, store_costs=True, rollout_events=False):
        """on: phase stamps of the finalising block (trace()); store_costs=False: skip the per-sample cost store;
        rollout_events: an event pair around every solve kernel (rollout_ms()) — two more stream operations per solve."""
        self._ck(self.L.mpc_set_profiling(self._h, int(bool(on)) | (0 if store_costs else 2) | (4 if rollout_events else 0)))

    def rollout_ms(self):
        ms = C.c_float()
        self._ck(self.L.mpc_get_rollout_ms(self._h, C.byref(ms)))
        return ms.value

    def trace(self):
        """Phase stamps (ns) of the last solve's finalising block, relative to the kernel start (profiling on)."""
        buf = (C.c_uint64 * 16)()
        self._ck(self.L.mpc_get_trace(self._h, buf))
        if self.is_mppi:
            names = ["start", "tiles_done", "ticket", "merged", "published", "peers_in", "merged2", "applied", "end",
                     "recs_in_smem", "rec_written", "fenced", "first_warp_out", "last_warp_out", "recs_stored"]
        else:  # single-launch CEM: block 0's phases of iteration 1, then the tail
            names = ["start", "it1_begin", "it1_rollout_done", "it1_keys_barrier", "it1_threshold", "it1_partial_sums",
                     "it1_sums_barrier", "it1_updated", "iterations_done", "end", "it1_samples_sorted", "it1_keys_located"]
        t0 = buf[0]
        return {n: (int(buf[i]) - int(t0)) if buf[i] else None for i, n in enumerate(names)}

    def nonzero_tickets(self):
        n = C.c_int32(0)
        self._ck(self.L.mpc_debug_tickets(self._h, C.byref(n)))
        return n.value

    def sync(self):
        self._ck(self.L.mpc_sync(self._h))

    def flush_l2(self):
        self._ck(self.L.mpc_flush_l2(self._h))

    @property
    def launch_count(self):
        return int(self.L.mpc_launch_count(self._h))


def measure_fma_peak(device=0, dtype="f32"):
    out = C.c_double()
    rc = load().mpc_measure_fma_peak(int(device), DTYPES[dtype], C.byref(out))
    if rc != MPC_OK:
        raise MpcError(rc, (load().mpc_last_error(None) or b"").decode())
    return out.value


class SolverGroup:
    """One process driving several GPUs ("no torchrun needed for a drop-in class"): one handle per
    device, K sharded across them, exchange buffers wired with mpc_p2p_link so that the all-gather of the rank records
    (MPPI) / elite candidates (CEM) happens inside the kernels over NVLink.  Presents the part of Solver's surface the
    agents use; every rank ends with the identical plan, rank 0's copy is returned."""

    def __init__(self, agent, env, env_params, K, H, devices, **kw):
        kw.pop("device", None), kw.pop("rank", None), kw.pop("world", None)
        if kw.get("batch", 1) != 1:
            raise ValueError("a device group shards ONE solve; use one handle per device for batches")
        self.ranks = []
        try:
            for r, d in enumerate(devices):
                self.ranks.append(Solver(agent, env, env_params, K, H, device=int(d), rank=r, world=len(devices), **kw))
            Solver.p2p_link(self.ranks)
        except Exception:
            self.close()
            raise
        s0 = self.ranks[0]
        self.is_mppi, self.tensor_noise = s0.is_mppi, s0.tensor_noise
        self.K, self.H, self.B, self.ns, self.nz = s0.K, s0.H, s0.B, s0.ns, s0.nz
        self.fast_math, self.lanes, self.geometry = s0.fast_math, s0.lanes, s0.geometry

    def close(self):
        for s in self.ranks:
            s.close()
        self.ranks = []

    # noise: every handle takes its own shard of the GLOBAL tensor
    def set_noise(self, delta_u):
        for s in self.ranks:
            s.set_noise(delta_u)

    def fill_noise(self):
        for s in self.ranks:
            s.fill_noise()

    def _solve_all(self, x, g_z, U_in):
        for s in self.ranks:
            s.upload_inputs(x, g_z, U_in)
        for s in self.ranks:  # enqueue everywhere before waiting anywhere: the kernels wait for each other's records
            s.solve_enqueue()
        outs = [s.download_outputs() for s in self.ranks]
        return outs[0]

    def mppi_solve(self, x, g_z=None, U_in=None, want_traj=True):
        u0, U, xs, cost = self._solve_all(x, g_z, U_in)
        return (u0, U, xs, cost) if want_traj else (u0, U, None, None)

    def mppi_costs(self):
        return np.concatenate([s.mppi_costs() for s in self.ranks], axis=1)

    def cem_solve(self, x, g_z=None, U_in=None, want_traj=True):
        u0, U, xs, cost = self._solve_all(x, g_z, U_in)
        return u0, U[0], (xs[0] if want_traj else None), (cost if want_traj else None), self.ranks[0].cem_iterations()

    def cem_begin(self, x, g_z=None, U_in=None):
        for s in self.ranks:
            s.cem_begin(x, g_z, U_in)

    def cem_iterate(self, z=None):
        for s in self.ranks[:-1]:
            s.cem_iterate(z, wait=False)
        return self.ranks[-1].cem_iterate(z, wait=True)  # every rank reaches the same verdict

    def cem_end(self, want_traj=True):
        outs = [s.cem_end(want_traj) for s in self.ranks]
        return outs[0]

    def cem_elites(self, it):
        return self.ranks[0].cem_elites(it)

    def simulate(self, x, us, g_z=None):
        return self.ranks[0].simulate(x, us, g_z)

    def model_step(self, x, u, g=None):
        return self.ranks[0].model_step(x, u, g)

    def episode_run(self, *a, **k):
        raise MpcError(MPC_ESTATE, "device-resident episodes run on one GPU (world = 1); use Experiment.run with a device group")

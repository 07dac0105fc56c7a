"""mpc_cbf_b200 — Python host glue over the C ABI of libmpccbf.so (include/mpccbf.h).

This module is plumbing for tests and bench.py: ctypes prototypes, a `Config` mirror of
`mpccbf_config`, and a `Handle` wrapper. All compute happens in the sm_100a kernels behind the
C ABI; there is NO fallback of any kind — if the library is missing or no B200-class device is
usable, calls raise.

The directory name contains a hyphen (`mpc-cbf_b200/`), so import it with `load()` from
`__graft_entry__` / `tests/conftest.py`, or via importlib:

    spec = importlib.util.spec_from_file_location("mpc_cbf_b200", ".../mpc-cbf_b200/__init__.py")
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPCCBF_LIB") or os.path.join(_HERE, "libmpccbf.so")  # override: kernel-variant experiments

OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_ALLOC = 0, -1, -2, -3, -4
STATUS_OPTIMAL, STATUS_FEASIBLE, STATUS_UNBOUNDED, STATUS_INFEASIBLE = 0, 1, 2, 3
STATUS_ERROR, STATUS_UNKNOWN, STATUS_INFEASIBLEORUNBOUNDED, STATUS_SKIPPED = 4, 5, 6, 7
SOLVER_AUTO, SOLVER_REDUCED, SOLVER_DENSE = 0, 1, 2

# every symbol include/mpccbf.h declares (tests assert the library exports all of them)
EXPORTED_SYMBOLS = [
    "mpccbf_config_default", "mpccbf_config_validate", "mpccbf_config_from_json", "mpccbf_create",
    "mpccbf_destroy", "mpccbf_last_error", "mpccbf_version", "mpccbf_num_decision_vars",
    "mpccbf_active_solver", "mpccbf_step_batch", "mpccbf_step_batch_host", "mpccbf_optimize_one",
    "mpccbf_solve_qp_dense", "mpccbf_cbf_rows", "mpccbf_rollout", "mpccbf_rollout_agents", "mpccbf_step_agents",
    "mpccbf_last_iteration_counts", "mpccbf_copy_iteration_counts", "mpccbf_measure_fp64_tflops", "mpccbf_get_table",
    "mpccbf_linear_cost", "mpccbf_get_counters", "mpccbf_host_alloc", "mpccbf_host_free",
    "mpccbf_cbf_filter_batch", "mpccbf_cbf_filter_batch_host", "mpccbf_conn_rows", "mpccbf_run_episode",
    "mpccbf_run_episode_ex", "mpccbf_step_batch_host_async", "mpccbf_synchronize", "mpccbf_slack_fallback_count",
    "mpccbf_step_batch_compact", "mpccbf_step_batch_host_compact", "mpccbf_step_batch_host_compact_async",
    "mpccbf_expand_ctrl_pts", "mpccbf_eval_compact", "mpccbf_episode_opts_from_json",
    "mpccbf_follow_reset", "mpccbf_follow_step", "mpccbf_set_kernel_variant", "mpccbf_auto_kernel_variant",
    "mpccbf_multi_create", "mpccbf_multi_destroy", "mpccbf_multi_num_devices", "mpccbf_multi_comm_ranks",
    "mpccbf_multi_handle", "mpccbf_multi_step_batch_host", "mpccbf_multi_step_batch_host_compact",
    "mpccbf_multi_run_sharded", "mpccbf_solve_qp_dense_noalloc",
]


class Config(C.Structure):
    """Mirror of `mpccbf_config` (include/mpccbf.h); field order must match the header."""

    _fields_ = [
        ("h", C.c_double), ("Ts", C.c_double),
        ("k_hor", C.c_int32), ("num_pieces", C.c_int32), ("num_ctrl_pts", C.c_int32),
        ("continuity_degree", C.c_int32),
        ("piece_max_param", C.c_double), ("w_pos_err", C.c_double), ("w_u_eff", C.c_double),
        ("spd_f", C.c_int32), ("cbf_horizon", C.c_int32),
        ("v_min", C.c_double * 3), ("v_max", C.c_double * 3),
        ("a_min", C.c_double * 3), ("a_max", C.c_double * 3),
        ("d_min", C.c_double), ("d_max", C.c_double), ("cbf_gamma", C.c_double),
        ("impc_iter", C.c_int32), ("slack_mode", C.c_int32),
        ("slack_cost", C.c_double), ("slack_decay", C.c_double),
        ("num_agents", C.c_int32), ("max_scenarios", C.c_int32),
        ("device", C.c_int32), ("solver", C.c_int32),
        ("feas_tol", C.c_double),
        ("admm_max_iter", C.c_int32), ("admm_check_every", C.c_int32),
        ("admm_eps", C.c_double), ("admm_rho", C.c_double), ("admm_sigma", C.c_double),
        ("conn_mode", C.c_int32),
    ]

    def copy(self, **kw):
        c = Config.from_buffer_copy(self)
        for k, v in kw.items():
            if isinstance(v, (list, tuple, np.ndarray)):
                getattr(c, k)[:] = list(v)
            else:
                setattr(c, k, v)
        return c

    @property
    def n(self):
        return self.num_pieces * 3 * self.num_ctrl_pts


class EpisodeOpts(C.Structure):
    """Mirror of `mpccbf_episode_opts` (include/mpccbf.h)."""
    _fields_ = [("goal_radius", C.c_double), ("collision_radius", C.c_double), ("order", C.c_int32),
                ("pos_std", C.c_double), ("vel_std", C.c_double), ("seed", C.c_uint64),
                ("collision_shape", C.c_int32), ("box_half", C.c_double * 2)]


ORDER_JACOBI, ORDER_GAUSS_SEIDEL = 0, 1
SHAPE_CIRCLE, SHAPE_BOX = 0, 1


def episode_opts_from_json(path):
    """collision shape of a reference experiment file with the metric's precedence (aligned_box first)."""
    o = EpisodeOpts()
    rc = lib().mpccbf_episode_opts_from_json(os.fsencode(path), C.byref(o))
    if rc != OK:
        raise MpcCbfError(rc, lib().mpccbf_last_error().decode())
    return o


class MpcCbfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"mpccbf error {code}: {msg}")
        self.code = code


def build(force=False):
    """Compiles libmpccbf.so for sm_100a with nvcc (in-tree, next to this file)."""
    src = os.path.join(_HERE, "csrc")
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    subprocess.run(["make", "-C", src], check=True, stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib(path=None):
    """Loads the C-ABI library (raises if it has not been built: there is no fallback)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise FileNotFoundError(f"{p} is missing: build it with __graft_entry__.build() "
                                "(make -C mpc-cbf_b200/csrc); there is no CPU fallback")
    L = C.CDLL(p)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    cfgp = C.POINTER(Config)
    L.mpccbf_config_default.argtypes = [cfgp]
    L.mpccbf_config_default.restype = None
    L.mpccbf_config_validate.argtypes = [cfgp]
    L.mpccbf_config_from_json.argtypes = [C.c_char_p, cfgp, vp, vp, i32]
    L.mpccbf_create.argtypes = [cfgp, C.POINTER(vp)]
    L.mpccbf_destroy.argtypes = [vp]
    L.mpccbf_last_error.restype = C.c_char_p
    L.mpccbf_version.restype = C.c_char_p
    L.mpccbf_num_decision_vars.argtypes = [vp]
    L.mpccbf_active_solver.argtypes = [vp]
    L.mpccbf_slack_fallback_count.argtypes = [vp]
    L.mpccbf_slack_fallback_count.restype = C.c_int64
    L.mpccbf_step_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp]
    L.mpccbf_step_batch_host.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_step_batch_host_async.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_synchronize.argtypes = [vp]
    L.mpccbf_episode_opts_from_json.argtypes = [C.c_char_p, vp]
    L.mpccbf_follow_reset.argtypes = [vp, i32, vp]
    L.mpccbf_follow_step.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp]
    L.mpccbf_set_kernel_variant.argtypes = [vp, i32]
    L.mpccbf_auto_kernel_variant.argtypes = [vp, i64]
    L.mpccbf_multi_create.argtypes = [cfgp, i32, vp, C.POINTER(vp)]
    L.mpccbf_multi_destroy.argtypes = [vp]
    L.mpccbf_multi_num_devices.argtypes = [vp]
    L.mpccbf_multi_comm_ranks.argtypes = [vp]
    L.mpccbf_multi_handle.argtypes = [vp, i32]
    L.mpccbf_multi_handle.restype = vp
    L.mpccbf_multi_step_batch_host.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_multi_step_batch_host_compact.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_multi_run_sharded.argtypes = [vp, i32, vp, vp, i32, vp]
    L.mpccbf_step_batch_compact.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp]
    L.mpccbf_step_batch_host_compact.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_step_batch_host_compact_async.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_expand_ctrl_pts.argtypes = [vp, i64, vp, vp, vp, vp]
    L.mpccbf_eval_compact.argtypes = [vp, i64, vp, vp, dbl, i32, vp]
    L.mpccbf_step_agents.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    L.mpccbf_optimize_one.argtypes = [vp, vp, i32, vp, vp, vp, vp]
    L.mpccbf_solve_qp_dense.argtypes = [vp, i32, i32, i32] + [vp] * 11
    L.mpccbf_solve_qp_dense_noalloc.argtypes = [C.POINTER(Config), i32, i32, i32] + [vp] * 10
    L.mpccbf_cbf_rows.argtypes = [vp, i32, vp, vp, vp]
    L.mpccbf_rollout.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.mpccbf_cbf_filter_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    L.mpccbf_cbf_filter_batch_host.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.mpccbf_conn_rows.argtypes = [vp, i32, vp, vp, vp, vp]
    L.mpccbf_run_episode.argtypes = [vp, i32, vp, vp, i32, dbl, dbl, vp, vp, vp, vp, vp, vp]
    L.mpccbf_run_episode_ex.argtypes = [vp, i32, vp, vp, i32, C.POINTER(EpisodeOpts), vp, vp, vp, vp, vp, vp]
    L.mpccbf_rollout_agents.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp]
    L.mpccbf_last_iteration_counts.argtypes = [vp]
    L.mpccbf_last_iteration_counts.restype = vp
    L.mpccbf_copy_iteration_counts.argtypes = [vp, vp, i64]
    L.mpccbf_measure_fp64_tflops.argtypes = [vp]
    L.mpccbf_measure_fp64_tflops.restype = dbl
    L.mpccbf_get_table.argtypes = [vp, C.c_char_p, vp, i64]
    L.mpccbf_get_table.restype = i64
    L.mpccbf_linear_cost.argtypes = [vp, vp, vp, vp]
    L.mpccbf_get_counters.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.mpccbf_host_alloc.argtypes = [C.c_size_t]
    L.mpccbf_host_alloc.restype = vp
    L.mpccbf_host_free.argtypes = [vp]
    L.mpccbf_host_free.restype = None
    if path is None:
        _lib = L
    return L


def default_config(**kw):
    c = Config()
    lib().mpccbf_config_default(C.byref(c))
    return c.copy(**kw) if kw else c


def config_from_json(path, cap_agents=1024, **kw):
    c = default_config(**kw)
    so = np.zeros((cap_agents, 3))
    sf = np.zeros((cap_agents, 3))
    rc = lib().mpccbf_config_from_json(os.fsencode(path), C.byref(c), so.ctypes.data, sf.ctypes.data, cap_agents)
    if rc != OK:
        raise MpcCbfError(rc, lib().mpccbf_last_error().decode())
    return c, so[:c.num_agents].copy(), sf[:c.num_agents].copy()


def validate(cfg):
    """Returns (code, message)."""
    L = lib()
    rc = L.mpccbf_config_validate(C.byref(cfg))
    return rc, (L.mpccbf_last_error().decode() if rc != OK else "")


def _f64(x):
    return np.ascontiguousarray(x, dtype=np.float64)


def _ptr(t):
    """Device/host pointer of a numpy array or torch tensor (None -> NULL)."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


NOALLOC_SCRATCH_BYTES = 8 << 20   # MPCCBF_NOALLOC_SCRATCH_BYTES (include/mpccbf.h)


def solve_qp_dense_noalloc(cfg, P, q, A, l, u, x, objective, status, iters, stream=None, library=None):
    """mpccbf_solve_qp_dense_noalloc: DEVICE tensors in and out (current device), asynchronous on `stream`,
    scratch in the library's fixed device array. Raises MpcCbfError (e.g. scratch limit exceeded)."""
    L = library if library is not None else lib()
    batch, n = q.shape
    m = l.shape[1] if l is not None else 0
    rc = L.mpccbf_solve_qp_dense_noalloc(C.byref(cfg) if cfg is not None else None, batch, n, m, _ptr(P), _ptr(q),
                                         _ptr(A) if m else None, _ptr(l) if m else None, _ptr(u) if m else None,
                                         _ptr(x), _ptr(objective), _ptr(status), _ptr(iters), stream)
    if rc != OK:
        raise MpcCbfError(rc, L.mpccbf_last_error().decode())


class Handle:
    """RAII wrapper of `mpccbf_handle`. `device=-1` gives a host-only handle (tables only)."""

    def __init__(self, cfg, library=None):
        self._L = library or lib()
        self.cfg = Config.from_buffer_copy(cfg)
        h = C.c_void_p()
        rc = self._L.mpccbf_create(C.byref(self.cfg), C.byref(h))
        if rc != OK:
            raise MpcCbfError(rc, self._L.mpccbf_last_error().decode())
        self._h = h
        self.n = self._L.mpccbf_num_decision_vars(h)
        self.solver = self._L.mpccbf_active_solver(h)

    def close(self):
        if getattr(self, "_h", None):
            self._L.mpccbf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != OK:
            raise MpcCbfError(rc, self._L.mpccbf_last_error().decode())

    # ---- tables -------------------------------------------------------------------------------
    def table(self, name):
        cnt = self._L.mpccbf_get_table(self._h, name.encode(), None, 0)
        if cnt < 0:
            raise MpcCbfError(cnt, self._L.mpccbf_last_error().decode())
        out = np.zeros(cnt)
        self._L.mpccbf_get_table(self._h, name.encode(), out.ctypes.data, cnt)
        K, n = self.cfg.k_hor, self.n
        cols = {"U_basis": n, "V_basis": n, "P": n, "E": n, "A_fix": n, "Y": 6}
        if name in cols:
            return out.reshape(-1, cols[name])
        if name == "Z":
            return out.reshape(n, -1)
        if name == "H_red":
            r = int(round(np.sqrt(cnt)))
            return out.reshape(r, r)
        return out

    def linear_cost(self, x0, ref):
        x0, ref = _f64(x0), _f64(ref)
        q = np.zeros(self.n)
        self._check(self._L.mpccbf_linear_cost(self._h, x0.ctypes.data, ref.ctypes.data, q.ctypes.data))
        return q

    # ---- host-pointer entry points (numpy in / numpy out) --------------------------------------
    def step_batch_host(self, states, targets, out=None):
        states, targets = _f64(states), _f64(targets)
        S, N = states.shape[0], states.shape[1]
        it = self.cfg.impc_iter
        if out is None:
            out = dict(ctrl_pts=np.zeros((S, N, self.n)), status=np.zeros((S, N, it), dtype=np.int32),
                       n_curves=np.zeros((S, N), dtype=np.int32), objective=np.zeros((S, N)))
        self._check(self._L.mpccbf_step_batch_host(self._h, S, states.ctypes.data, targets.ctypes.data,
                                                   _ptr(out["ctrl_pts"]), _ptr(out["status"]),
                                                   _ptr(out["n_curves"]), _ptr(out["objective"])))
        return out

    def step_batch_host_async(self, states, targets, out):
        """Enqueue only (buffers must be contiguous float64 / int32 arrays that outlive synchronize())."""
        S = states.shape[0]
        self._check(self._L.mpccbf_step_batch_host_async(self._h, S, states.ctypes.data, targets.ctypes.data,
                                                         _ptr(out["ctrl_pts"]), _ptr(out["status"]),
                                                         _ptr(out["n_curves"]), _ptr(out["objective"])))

    def synchronize(self):
        self._check(self._L.mpccbf_synchronize(self._h))

    # ---- compact result form: v_free [S][N][6] instead of ctrl_pts [S][N][n] ---------------------
    def step_batch_host_compact(self, states, targets, out=None):
        states, targets = _f64(states), _f64(targets)
        S, N = states.shape[0], states.shape[1]
        it = self.cfg.impc_iter
        if out is None:
            out = dict(v_free=np.zeros((S, N, 6)), status=np.zeros((S, N, it), dtype=np.int32),
                       n_curves=np.zeros((S, N), dtype=np.int32), objective=np.zeros((S, N)))
        self._check(self._L.mpccbf_step_batch_host_compact(self._h, S, states.ctypes.data, targets.ctypes.data,
                                                           _ptr(out["v_free"]), _ptr(out["status"]),
                                                           _ptr(out["n_curves"]), _ptr(out["objective"])))
        return out

    def step_batch_host_compact_async(self, states, targets, out):
        S = states.shape[0]
        self._check(self._L.mpccbf_step_batch_host_compact_async(self._h, S, states.ctypes.data, targets.ctypes.data,
                                                                 _ptr(out["v_free"]), _ptr(out["status"]),
                                                                 _ptr(out["n_curves"]), _ptr(out["objective"])))

    def step_batch_compact(self, S, states, targets, v_free, status, n_curves=None, objective=None, stream=None):
        """Device pointers (torch tensors or ints), asynchronous on `stream`."""
        self._check(self._L.mpccbf_step_batch_compact(self._h, S, _ptr(states), _ptr(targets), _ptr(v_free), _ptr(status),
                                                      _ptr(n_curves), _ptr(objective), stream))

    def expand_ctrl_pts(self, states, v_free, n_curves=None):
        states, v_free = _f64(states).reshape(-1, 6), _f64(v_free).reshape(-1, 6)
        cnt = states.shape[0]
        nc = np.ascontiguousarray(n_curves, dtype=np.int32).ravel() if n_curves is not None else None
        out = np.zeros((cnt, self.n))
        self._check(self._L.mpccbf_expand_ctrl_pts(self._h, cnt, states.ctypes.data, v_free.ctypes.data,
                                                   nc.ctypes.data if nc is not None else None, out.ctypes.data))
        return out

    def eval_compact(self, states, v_free, t, deriv=0):
        states, v_free = _f64(states).reshape(-1, 6), _f64(v_free).reshape(-1, 6)
        out = np.zeros((states.shape[0], 3))
        self._check(self._L.mpccbf_eval_compact(self._h, states.shape[0], states.ctypes.data, v_free.ctypes.data,
                                                float(t), int(deriv), out.ctypes.data))
        return out

    def cbf_filter_batch_host(self, states, desired_u):
        """cbf::ConnectivityControl::optimize for every agent: dict(u [S][N][3], status [S][N], lambda2 [S])."""
        states, desired_u = _f64(states), _f64(desired_u)
        S, N = states.shape[0], states.shape[1]
        out = dict(u=np.zeros((S, N, 3)), status=np.zeros((S, N), dtype=np.int32), lambda2=np.zeros(S))
        self._check(self._L.mpccbf_cbf_filter_batch_host(self._h, S, states.ctypes.data, desired_u.ctypes.data,
                                                         _ptr(out["u"]), _ptr(out["status"]), _ptr(out["lambda2"])))
        return out

    def optimize_one(self, states, self_idx, ref):
        states, ref = _f64(states), _f64(ref)
        it = self.cfg.impc_iter
        cp = np.zeros((it, self.n))
        status = np.zeros(it, dtype=np.int32)
        nc = C.c_int32()
        self._check(self._L.mpccbf_optimize_one(self._h, states.ctypes.data, self_idx, ref.ctypes.data,
                                                cp.ctypes.data, C.addressof(nc), status.ctypes.data))
        return dict(curves=cp[:nc.value], status=status, n_curves=nc.value, success=bool(nc.value == it))

    def solve_qp_dense(self, P, q, A, l, u, lb=None, ub=None):
        P, q, A, l, u = map(_f64, (P, q, A, l, u))
        batch, n = q.shape
        m = l.shape[1]
        x = np.zeros((batch, n))
        obj = np.zeros(batch)
        status = np.zeros(batch, dtype=np.int32)
        iters = np.zeros(batch, dtype=np.int32)
        lb = _f64(lb) if lb is not None else None
        ub = _f64(ub) if ub is not None else None
        self._check(self._L.mpccbf_solve_qp_dense(self._h, batch, n, m, P.ctypes.data, q.ctypes.data,
                                                  A.ctypes.data if m else None, l.ctypes.data if m else None,
                                                  u.ctypes.data if m else None, _ptr(lb), _ptr(ub),
                                                  x.ctypes.data, obj.ctypes.data, status.ctypes.data,
                                                  iters.ctypes.data))
        return dict(x=x, objective=obj, status=status, iters=iters)

    # ---- device-pointer entry points (torch CUDA tensors; stream = raw cudaStream_t or None) ----
    def step_batch(self, S, states, targets, ctrl_pts, status, n_curves=None, objective=None, stream=None):
        self._check(self._L.mpccbf_step_batch(self._h, S, _ptr(states), _ptr(targets), _ptr(ctrl_pts),
                                              _ptr(status), _ptr(n_curves), _ptr(objective), stream))

    def step_agents(self, S, agent_begin, agent_count, states, targets, ctrl_pts, status, n_curves=None,
                    objective=None, stream=None):
        self._check(self._L.mpccbf_step_agents(self._h, S, agent_begin, agent_count, _ptr(states), _ptr(targets),
                                               _ptr(ctrl_pts), _ptr(status), _ptr(n_curves), _ptr(objective),
                                               stream))

    def cbf_rows(self, S, states, rows, stream=None):
        self._check(self._L.mpccbf_cbf_rows(self._h, S, _ptr(states), _ptr(rows), stream))

    def cbf_filter_batch(self, S, states, desired_u, u, status, lambda2=None, stream=None):
        self._check(self._L.mpccbf_cbf_filter_batch(self._h, S, _ptr(states), _ptr(desired_u), _ptr(u), _ptr(status),
                                                    _ptr(lambda2), stream))

    def run_episode(self, S, states, targets, steps, goal_radius, collision_radius, success=None, finish_step=None,
                    min_dist=None, failed_calls=None, traj=None, stream=None, order=ORDER_JACOBI, pos_std=0.0,
                    vel_std=0.0, seed=0, box_half=None):
        """Closed-loop episode on the device (states is updated in place); see include/mpccbf.h.
        box_half = (cx, cy): the metric's box collision shape instead of the circle."""
        if order == ORDER_JACOBI and pos_std == 0.0 and vel_std == 0.0 and box_half is None:
            self._check(self._L.mpccbf_run_episode(self._h, S, _ptr(states), _ptr(targets), steps, goal_radius,
                                                   collision_radius, _ptr(success), _ptr(finish_step), _ptr(min_dist),
                                                   _ptr(failed_calls), _ptr(traj), stream))
            return
        o = EpisodeOpts(goal_radius, collision_radius, order, pos_std, vel_std, seed)
        if box_half is not None:
            o.collision_shape = SHAPE_BOX
            o.box_half[:] = [float(box_half[0]), float(box_half[1])]
        self._check(self._L.mpccbf_run_episode_ex(self._h, S, _ptr(states), _ptr(targets), steps, C.byref(o),
                                                  _ptr(success), _ptr(finish_step), _ptr(min_dist),
                                                  _ptr(failed_calls), _ptr(traj), stream))

    def conn_rows(self, S, states, rows, lambda2=None, stream=None):
        self._check(self._L.mpccbf_conn_rows(self._h, S, _ptr(states), _ptr(rows), _ptr(lambda2), stream))

    def rollout(self, S, states, ctrl_pts, n_curves, next_states, stream=None):
        self._check(self._L.mpccbf_rollout(self._h, S, _ptr(states), _ptr(ctrl_pts), _ptr(n_curves),
                                           _ptr(next_states), stream))

    def rollout_agents(self, S, agent_count, states, ctrl_pts, n_curves, next_states, stream=None):
        self._check(self._L.mpccbf_rollout_agents(self._h, S, agent_count, _ptr(states), _ptr(ctrl_pts),
                                                  _ptr(n_curves), _ptr(next_states), stream))

    def follow_reset(self, S, stream=None):
        self._check(self._L.mpccbf_follow_reset(self._h, S, stream))

    def follow_step(self, S, agent_begin, agent_count, states, ctrl_new, n_curves, stream=None):
        """Reference-driver curve following of an agent shard; `states` [S][N][6] is updated in place."""
        self._check(self._L.mpccbf_follow_step(self._h, S, agent_begin, agent_count, _ptr(states), _ptr(ctrl_new),
                                               _ptr(n_curves), stream))

    def set_kernel_variant(self, lanes):
        self._check(self._L.mpccbf_set_kernel_variant(self._h, lanes))

    def auto_kernel_variant(self, num_agent_qps):
        return int(self._L.mpccbf_auto_kernel_variant(self._h, num_agent_qps))

    def slack_fallback_count(self):
        """Always 0 since round 2 (the slack caps live in the active-set kernel; kept for round-1 callers)."""
        return int(self._L.mpccbf_slack_fallback_count(self._h))

    def last_iteration_counts_ptr(self):
        return self._L.mpccbf_last_iteration_counts(self._h)

    def last_iteration_counts(self, count):
        """Active-set iteration counts [count] (int32) of the last controller step, on the host."""
        out = np.zeros(count, dtype=np.int32)
        self._check(self._L.mpccbf_copy_iteration_counts(self._h, out.ctypes.data, count))
        return out

    def measure_fp64_tflops(self):
        v = self._L.mpccbf_measure_fp64_tflops(self._h)
        if v < 0:
            raise MpcCbfError(int(v), self._L.mpccbf_last_error().decode())
        return v

    def counters(self):
        a, b = C.c_int64(), C.c_int64()
        self._L.mpccbf_get_counters(self._h, C.byref(a), C.byref(b))
        return a.value, b.value


class Multi:
    """RAII wrapper of `mpccbf_multi`: one process driving several GPUs through the C ABI."""

    def __init__(self, cfg, num_devices, devices=None):
        self._L = lib()
        self.cfg = cfg
        self._m = C.c_void_p()
        dv = None
        if devices is not None:
            dv = (C.c_int32 * num_devices)(*devices)
        rc = self._L.mpccbf_multi_create(C.byref(cfg), num_devices, dv, C.byref(self._m))
        if rc != OK:
            raise MpcCbfError(rc, self._L.mpccbf_last_error().decode())
        self.num_devices = int(self._L.mpccbf_multi_num_devices(self._m))
        self.comm_ranks = int(self._L.mpccbf_multi_comm_ranks(self._m))
        self.n = cfg.n

    def _check(self, rc):
        if rc != OK:
            raise MpcCbfError(rc, self._L.mpccbf_last_error().decode())

    def step_batch_host(self, states, targets, out=None, compact=False):
        states, targets = _f64(states), _f64(targets)
        S, N = states.shape[0], states.shape[1]
        it = self.cfg.impc_iter
        key = "v_free" if compact else "ctrl_pts"
        if out is None:
            out = {key: np.zeros((S, N, 6 if compact else self.n)), "status": np.zeros((S, N, it), dtype=np.int32),
                   "n_curves": np.zeros((S, N), dtype=np.int32), "objective": np.zeros((S, N))}
        fn = self._L.mpccbf_multi_step_batch_host_compact if compact else self._L.mpccbf_multi_step_batch_host
        self._check(fn(self._m, S, states.ctypes.data, targets.ctypes.data, _ptr(out[key]), _ptr(out["status"]),
                       _ptr(out["n_curves"]), _ptr(out["objective"])))
        return out

    def run_sharded(self, states, targets, steps):
        """Agent-sharded closed loop; returns (final states [S][N][6], device ms per control step)."""
        st, tg = _f64(states).copy(), _f64(targets)
        ms = C.c_double()
        self._check(self._L.mpccbf_multi_run_sharded(self._m, st.shape[0], st.ctypes.data, tg.ctypes.data, steps, C.byref(ms)))
        return st, ms.value

    def close(self):
        if self._m:
            self._L.mpccbf_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

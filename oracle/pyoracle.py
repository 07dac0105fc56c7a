"""TEST INFRASTRUCTURE — ctypes loader for the CPU oracle (oracle/liboracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The product package (mpc-cbf_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class Config(C.Structure):
    """Mirror of `mpccbf_config` (include/mpccbf.h). Field order must match the header."""

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


def base_config(**kw):
    """workspace/experiments/config/base_config.json (effective parameters after preprocess.py)."""
    c = Config()
    c.h, c.Ts, c.k_hor = 0.1, 0.01, 16
    c.num_pieces, c.num_ctrl_pts, c.continuity_degree, c.piece_max_param = 3, 4, 3, 0.5
    c.w_pos_err, c.w_u_eff, c.spd_f = 10.0, 10.0, 8
    c.cbf_horizon, c.impc_iter = 2, 2
    c.v_min[:] = [-2.0, -2.0, -2.6179938779914944]
    c.v_max[:] = [2.0, 2.0, 2.6179938779914944]
    c.a_min[:] = [-5.0, -5.0, -3.141592653589793]
    c.a_max[:] = [5.0, 5.0, 3.141592653589793]
    c.d_min, c.d_max, c.cbf_gamma = 2.0, 4.0, 5.0
    c.slack_mode, c.slack_cost, c.slack_decay = 0, 50000.0, 0.1
    c.num_agents, c.max_scenarios, c.device, c.solver = 2, 1, 0, 0
    c.feas_tol = 1e-9
    for k, v in kw.items():
        if isinstance(v, (list, tuple, np.ndarray)):
            getattr(c, k)[:] = list(v)
        else:
            setattr(c, k, v)
    return c


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".hpp"))]
    srcs.append(os.path.join(_HERE, "..", "include", "mpccbf.h"))
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "all"], check=True, stdout=subprocess.DEVNULL)
    return so


_lib = None
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_fac.restype = C.c_uint64
        _lib.oracle_fac.argtypes = [C.c_uint64]
        _lib.oracle_comb.restype = C.c_uint64
        _lib.oracle_comb.argtypes = [C.c_uint64, C.c_uint64]
        _lib.oracle_perm.restype = C.c_uint64
        _lib.oracle_perm.argtypes = [C.c_uint64, C.c_uint64]
        _lib.oracle_last_error.restype = C.c_char_p
        _lib.oracle_get_table.restype = C.c_int64
        _lib.oracle_get_table.argtypes = [C.POINTER(Config), C.c_char_p, C.c_void_p, C.c_int64, C.c_int]
    return _lib


def ref_lib():
    """The REAL reference's math::fac/comb/perm/pow (oracle/_ref), or None when not built."""
    p = os.path.join(_HERE, "_ref", "libref_combinatorics.so")
    if not os.path.exists(p):
        return None
    L = C.CDLL(p)
    L.ref_fac.restype = C.c_uint64
    L.ref_fac.argtypes = [C.c_uint64, C.POINTER(C.c_int)]
    L.ref_comb.restype = C.c_uint64
    L.ref_comb.argtypes = [C.c_uint64, C.c_uint64]
    L.ref_perm.restype = C.c_uint64
    L.ref_perm.argtypes = [C.c_uint64, C.c_uint64]
    L.ref_pow.restype = C.c_double
    L.ref_pow.argtypes = [C.c_double, C.c_uint64]
    return L


def _arr(x):
    return np.ascontiguousarray(x, dtype=np.float64)


def bernstein_basis(deg, T, t, d, prec=0):
    out = np.zeros(deg + 1)
    rc = lib().oracle_bernstein_basis(C.c_int(deg), C.c_double(T), C.c_double(t), C.c_int(d),
                                      out.ctypes.data_as(C.c_void_p), C.c_int(prec))
    if rc != 0:
        raise RuntimeError("bernsteinBasis:: given parameter is outside of [0, max_parameter].")
    return out


def bernstein_coefficient_matrix(deg, T, d):
    out = np.zeros((deg + 1, deg + 1))
    lib().oracle_bernstein_coefficient_matrix(C.c_int(deg), C.c_double(T), C.c_int(d),
                                              out.ctypes.data_as(C.c_void_p))
    return out


def double_integrator_apply(ts, x, u):
    out = np.zeros(6)
    x, u = _arr(x), _arr(u)
    lib().oracle_double_integrator_apply(C.c_double(ts), x.ctypes.data_as(C.c_void_p),
                                         u.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out


def safety_cbf_row(ego, nbr, dmin, gamma=5.0, prec=0):
    ego, nbr = _arr(ego), _arr(nbr)
    a = np.zeros(3)
    b = C.c_double(0)
    lib().oracle_safety_cbf_row(ego.ctypes.data_as(C.c_void_p), nbr.ctypes.data_as(C.c_void_p),
                                C.c_double(dmin), C.c_double(gamma), a.ctypes.data_as(C.c_void_p),
                                C.byref(b), C.c_int(prec))
    return a, b.value


def get_table(cfg, name, prec=0):
    L = lib()
    cnt = L.oracle_get_table(C.byref(cfg), name.encode(), None, 0, prec)
    if cnt < 0:
        raise KeyError(name)
    out = np.zeros(cnt)
    L.oracle_get_table(C.byref(cfg), name.encode(), out.ctypes.data_as(C.c_void_p), cnt, prec)
    K, n = cfg.k_hor, cfg.num_pieces * 3 * cfg.num_ctrl_pts
    shapes = {"U_basis": (3 * K, n), "V_basis": (3 * K, n), "A0_pos": (3 * K, 6), "A0_vel": (3 * K, 6),
              "Lambda_pos": (3 * K, 3 * K), "Lambda_vel": (3 * K, 3 * K), "h_samples": (K,)}
    return out.reshape(shapes[name])


def build_qp(cfg, states, self_idx, ref, it=0, pred_states=None, slack_weights=None):
    """The QP of IMPC iteration `it` as the reference generator assembles it (dict of arrays)."""
    L = lib()
    states, ref = _arr(states), _arr(ref)
    N = states.shape[0]
    n, m, ne, nc = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    pred = _arr(pred_states) if pred_states is not None else np.zeros((cfg.cbf_horizon, 6))
    sw = _arr(slack_weights) if slack_weights is not None else None
    args = [C.byref(cfg), states.ctypes.data_as(C.c_void_p), C.c_int(N), C.c_int(self_idx),
            ref.ctypes.data_as(C.c_void_p), C.c_int(it), pred.ctypes.data_as(C.c_void_p),
            sw.ctypes.data_as(C.c_void_p) if sw is not None else None,
            C.byref(n), C.byref(m), C.byref(ne), C.byref(nc)]
    rc = L.oracle_build_qp(*args, None, None, None, None, None, None, None)
    if rc != 0:
        raise RuntimeError("oracle_build_qp failed: " + lib().oracle_last_error().decode())
    P = np.zeros((n.value, n.value)); q = np.zeros(n.value)
    A = np.zeros((m.value, n.value)); l = np.zeros(m.value); u = np.zeros(m.value)
    lb = np.zeros(n.value); ub = np.zeros(n.value)
    rc = L.oracle_build_qp(*args, *[a.ctypes.data_as(C.c_void_p) for a in (P, q, A, l, u, lb, ub)])
    if rc != 0:
        raise RuntimeError("oracle_build_qp failed: " + lib().oracle_last_error().decode())
    return dict(P=P, q=q, A=A, l=l, u=u, lb=lb, ub=ub, n_eq=ne.value, n_cbf=nc.value)


def solve_qp(P, q, A, l, u, lb=None, ub=None, feas_tol=1e-9, prec=1):
    """min x'Px + q'x s.t. l <= Ax <= u, lb <= x <= ub. Returns dict(x, objective, status, ...)."""
    P, q, A, l, u = map(_arr, (P, q, A, l, u))
    n, m = q.size, l.size
    A = A.reshape(m, n) if m else np.zeros((0, n))
    x = np.zeros(n); y = np.zeros(max(m, 1)); kkt = np.zeros(4)
    obj = C.c_double(); st = C.c_int32(); it = C.c_int32()
    lbp = _arr(lb).ctypes.data_as(C.c_void_p) if lb is not None else None
    ubp = _arr(ub).ctypes.data_as(C.c_void_p) if ub is not None else None
    _keep = (lb, ub)
    rc = lib().oracle_solve_qp(C.c_int(n), C.c_int(m), P.ctypes.data_as(C.c_void_p),
                               q.ctypes.data_as(C.c_void_p), A.ctypes.data_as(C.c_void_p),
                               l.ctypes.data_as(C.c_void_p), u.ctypes.data_as(C.c_void_p), lbp, ubp,
                               C.c_double(feas_tol), C.c_int(prec), x.ctypes.data_as(C.c_void_p),
                               C.byref(obj), C.byref(st), C.byref(it), kkt.ctypes.data_as(C.c_void_p),
                               y.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError("oracle_solve_qp failed")
    return dict(x=x, objective=obj.value, status=st.value, iters=it.value, kkt=kkt, y=y[:m])


def optimize(cfg, states, self_idx, ref, feas_tol=1e-9, prec=1):
    """ConnectivityIMPCCBF::optimize restated. Returns dict(curves [n_curves][n], status, ...)."""
    states, ref = _arr(states), _arr(ref)
    N = states.shape[0]
    n = cfg.num_pieces * 3 * cfg.num_ctrl_pts
    it = cfg.impc_iter
    cp = np.zeros((it, n)); status = np.zeros(it, dtype=np.int32); obj = np.zeros(it)
    kkt = np.zeros((it, 4)); mrows = np.zeros(it, dtype=np.int32); nc = C.c_int32()
    rc = lib().oracle_optimize(C.byref(cfg), states.ctypes.data_as(C.c_void_p), C.c_int(N),
                               C.c_int(self_idx), ref.ctypes.data_as(C.c_void_p), C.c_double(feas_tol),
                               C.c_int(prec), cp.ctypes.data_as(C.c_void_p),
                               status.ctypes.data_as(C.c_void_p), C.byref(nc),
                               obj.ctypes.data_as(C.c_void_p), kkt.ctypes.data_as(C.c_void_p),
                               mrows.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError("oracle_optimize failed: " + lib().oracle_last_error().decode())
    return dict(curves=cp[:nc.value], status=status, n_curves=nc.value, objective=obj, kkt=kkt,
                m_rows=mrows, success=bool(nc.value == it))


def step_batch(cfg, states, targets, threads=1, feas_tol=1e-9, prec=0):
    """Jacobi control step for states [S][N][6], targets [S][N][3] (CPU baseline entry)."""
    states, targets = _arr(states), _arr(targets)
    S, N = states.shape[0], states.shape[1]
    n = cfg.num_pieces * 3 * cfg.num_ctrl_pts
    cp = np.zeros((S, N, n)); status = np.zeros((S, N, cfg.impc_iter), dtype=np.int32)
    ncv = np.zeros((S, N), dtype=np.int32); obj = np.zeros((S, N))
    c2 = Config.from_buffer_copy(cfg)
    c2.num_agents = N
    rc = lib().oracle_step_batch(C.byref(c2), C.c_int(S), states.ctypes.data_as(C.c_void_p),
                                 targets.ctypes.data_as(C.c_void_p), cp.ctypes.data_as(C.c_void_p),
                                 status.ctypes.data_as(C.c_void_p), ncv.ctypes.data_as(C.c_void_p),
                                 obj.ctypes.data_as(C.c_void_p), C.c_int(threads), C.c_double(feas_tol),
                                 C.c_int(prec))
    if rc != 0:
        raise RuntimeError("oracle_step_batch failed")
    return dict(ctrl_pts=cp, status=status, n_curves=ncv, objective=obj)


def eval_curve(cfg, ctrl_pts, t, d):
    out = np.zeros(3)
    cp = _arr(ctrl_pts)
    rc = lib().oracle_eval_curve(C.byref(cfg), cp.ctypes.data_as(C.c_void_p), C.c_double(t), C.c_int(d),
                                 out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise ValueError("eval_curve: parameter out of range")
    return out


def hardware_threads():
    return int(lib().oracle_hardware_threads())


# ---- CBF-only filter (cbf::ConnectivityControl) ------------------------------------------------------
def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def lambda2(states, dmax, prec=1):
    """ConnectivityCBF::getLambda2 for one scenario: (lambda2, normalised Fiedler vector)."""
    states = _arr(states)
    N = states.shape[0]
    fv = np.zeros(N)
    f = lib().oracle_lambda2
    f.restype = C.c_double
    l2 = f(_vp(states), C.c_int(N), C.c_double(dmax), C.c_int(prec), _vp(fv))
    return l2, fv


def conn_cbf_row(states, self_idx, dmax, gamma=5.0, prec=1):
    """Connectivity HOCBF row of one agent: (Ac [3], Bc, lambda2) with Ac.u + Bc >= 0."""
    states = _arr(states)
    Ac = np.zeros(3); Bc = C.c_double(); l2 = C.c_double()
    lib().oracle_conn_cbf_row(_vp(states), C.c_int(states.shape[0]), C.c_int(self_idx), C.c_double(dmax),
                              C.c_double(gamma), C.c_int(prec), _vp(Ac), C.byref(Bc), C.byref(l2))
    return Ac, Bc.value, l2.value


def clf_row(ego, nbr):
    """Distance CLF row: (Ac [3], Bc) with Ac.u <= -Bc."""
    ego, nbr = _arr(ego), _arr(nbr)
    Ac = np.zeros(3); Bc = C.c_double()
    lib().oracle_clf_row(_vp(ego), _vp(nbr), _vp(Ac), C.byref(Bc))
    return Ac, Bc.value


def cbf_filter_batch(cfg, states, desired_u, feas_tol=1e-9, prec=1):
    """ConnectivityControl::optimize for every agent: states [S][N][6], desired_u [S][N][3]."""
    states, desired_u = _arr(states), _arr(desired_u)
    S, N = states.shape[:2]
    c2 = Config.from_buffer_copy(cfg)
    c2.num_agents = N
    u = np.zeros((S, N, 3)); status = np.zeros((S, N), dtype=np.int32); l2 = np.zeros(S)
    rc = lib().oracle_cbf_filter_batch(C.byref(c2), C.c_int(S), _vp(states), _vp(desired_u), C.c_double(feas_tol),
                                       C.c_int(prec), _vp(u), _vp(status), _vp(l2))
    if rc != 0:
        raise RuntimeError("oracle_cbf_filter_batch failed: " + lib().oracle_last_error().decode())
    return dict(u=u, status=status, lambda2=l2)

"""ctypes loader of the CPU oracle (oracle/_build/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Allowed importers: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference legs).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from ilqr_planner_b200 import cabi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
# same source, other arithmetic (oracle/Makefile): "fma" = -ffp-contract=fast, "ld" = long double internally
VARIANTS = {None: LIB_PATH, "fma": os.path.join(_HERE, "_build", "liboracle_fma.so"), "ld": os.path.join(_HERE, "_build", "liboracle_ld.so")}
_libs = {}


def build(force=False):
    src = os.path.join(_HERE, "ilqr_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "ilqrb.h")
    newest = max(os.path.getmtime(src), os.path.getmtime(hdr), os.path.getmtime(os.path.join(_HERE, "Makefile")))
    if (not force and all(os.path.exists(p) and os.path.getmtime(p) >= newest for p in VARIANTS.values())):
        return LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-B", "-j3", "all"], stdout=subprocess.DEVNULL)
    return LIB_PATH


def load(variant=None):
    """variant None: the oracle proper (FP64, -ffp-contract=off); "fma" / "ld": see VARIANTS"""
    if variant in _libs:
        return _libs[variant]
    build()
    lib = C.CDLL(VARIANTS[variant])
    P = C.POINTER
    lib.oracle_last_error.restype = C.c_char_p
    lib.oracle_get_dims.argtypes = [P(cabi.ProblemDesc), P(cabi.Dims)]
    lib.oracle_solve_recursive.argtypes = [P(cabi.ProblemDesc), cabi.c_double_p, P(cabi.SolveOpts), P(cabi.Result), C.c_int]
    lib.oracle_solve_al.argtypes = [P(cabi.ProblemDesc), cabi.c_double_p, P(cabi.ALDesc), P(cabi.SolveOpts), P(cabi.Result), C.c_int]
    lib.oracle_solve_batch.argtypes = [P(cabi.ProblemDesc), cabi.c_double_p, P(cabi.BatchDesc), P(cabi.SolveOpts), P(cabi.Result), C.c_int]
    lib.oracle_rollout.argtypes = [P(cabi.ProblemDesc), cabi.c_double_p, cabi.c_double_p, cabi.c_double_p, cabi.c_double_p, cabi.c_double_p, C.c_int]
    lib.oracle_kinematics.argtypes = [P(cabi.ProblemDesc), C.c_int, C.c_int] + [cabi.c_double_p] * 7
    lib.oracle_batch_step_truth.argtypes = [P(cabi.ProblemDesc), cabi.c_double_p, P(cabi.BatchDesc), cabi.c_double_p, cabi.c_double_p, C.c_int]
    lib.oracle_set_margin_out.argtypes = [cabi.c_double_p]
    lib.oracle_eval_point.argtypes = [P(cabi.ProblemDesc), C.c_int, cabi.c_int32_p, cabi.c_int32_p] + [cabi.c_double_p] * 9
    lib.oracle_sd.argtypes = [C.c_int, C.c_int] + [cabi.c_double_p] * 4
    lib.oracle_lqt.argtypes = [C.c_int] * 5 + [cabi.c_double_p] * 4 + [C.c_double] + [cabi.c_double_p] * 6 + [C.c_int]
    lib.oracle_set_margin_out.restype = None
    _libs[variant] = lib
    return lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("oracle error %d: %s" % (rc, load().oracle_last_error().decode()))


def hardware_threads():
    return int(load().oracle_hardware_threads())


class _margins:
    """attach a [B][nb_iter] `margin_hist` (line-search decision margins, a test aid) to the result of a solve"""

    def __init__(self, out, pb, nb_iter, variant=None):
        self.out = out
        self.variant = variant
        out.margin_hist = np.full((pb.batch, max(nb_iter, 1)), np.nan)

    def __enter__(self):
        load(self.variant).oracle_set_margin_out(cabi.dptr(self.out.margin_hist))

    def __exit__(self, *a):
        load(self.variant).oracle_set_margin_out(cabi.dptr(None))


def solve_recursive(pb, U0, nb_iter, line_search=True, early_stop=False, threads=1, variant=None):
    from ilqr_planner_b200.problem import ResultBuffers
    d = pb.to_desc()
    out = ResultBuffers(pb, nb_iter, "recursive")
    opts = cabi.SolveOpts(nb_iter, int(line_search), int(early_stop), 0)
    U0 = None if U0 is None else np.ascontiguousarray(np.asarray(U0, dtype=np.float64).reshape(pb.batch, pb.horizon - 1, pb.m))
    with _margins(out, pb, nb_iter, variant):
        _check(load(variant).oracle_solve_recursive(C.byref(d), cabi.dptr(U0), C.byref(opts), C.byref(out.res), threads))
    return out


def make_al_desc(pb, A, b, lambda0, lag_update_step, penalty, scaling_factor):
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64))
    b = np.ascontiguousarray(np.asarray(b, dtype=np.float64))
    lam = np.ascontiguousarray(np.asarray(lambda0, dtype=np.float64))
    per_step = 1 if A.ndim == 3 else 0
    nc = A.shape[-2]
    al = cabi.ALDesc(nc, per_step, cabi.dptr(A), cabi.dptr(b), cabi.dptr(lam), int(lag_update_step), 0, float(penalty), float(scaling_factor))
    al._keep = (A, b, lam)
    return al


def solve_al(pb, U0, A, b, lambda0, nb_iter, lag_update_step, penalty, scaling_factor, line_search=True, early_stop=False, threads=1, variant=None):
    from ilqr_planner_b200.problem import ResultBuffers
    d = pb.to_desc()
    al = make_al_desc(pb, A, b, lambda0, lag_update_step, penalty, scaling_factor)
    out = ResultBuffers(pb, nb_iter, "al", nc=al.nc)
    opts = cabi.SolveOpts(nb_iter, int(line_search), int(early_stop), 0)
    U0 = None if U0 is None else np.ascontiguousarray(np.asarray(U0, dtype=np.float64).reshape(pb.batch, pb.horizon - 1, pb.m))
    with _margins(out, pb, nb_iter, variant):
        _check(load(variant).oracle_solve_al(C.byref(d), cabi.dptr(U0), C.byref(al), C.byref(opts), C.byref(out.res), threads))
    return out


def make_batch_desc(Q=None, psi=None):
    Q = None if Q is None else np.ascontiguousarray(np.asarray(Q, dtype=np.float64))
    psi = None if psi is None else np.ascontiguousarray(np.asarray(psi, dtype=np.float64))
    bd = cabi.BatchDesc(cabi.dptr(Q), cabi.dptr(psi), 0 if psi is None else psi.shape[1], 0)
    bd._keep = (Q, psi)
    return bd


def solve_batch(pb, u0, nb_iter, early_stop=False, Q=None, psi=None, threads=1, variant=None):
    from ilqr_planner_b200.problem import ResultBuffers
    d = pb.to_desc()
    bd = make_batch_desc(Q, psi)
    out = ResultBuffers(pb, nb_iter, "batch")
    opts = cabi.SolveOpts(nb_iter, 1, int(early_stop), 0)
    u0 = None if u0 is None else np.ascontiguousarray(np.asarray(u0, dtype=np.float64).reshape(pb.batch, -1))
    with _margins(out, pb, nb_iter, variant):
        _check(load(variant).oracle_solve_batch(C.byref(d), cabi.dptr(u0), C.byref(bd), C.byref(opts), C.byref(out.res), threads))
    return out


def batch_step_truth(pb, u0, Q=None, psi=None, threads=1):
    """first-iteration step of BatchILQR / BatchILQRCP: (du as the reference computes it, du solved in __float128 from the same inputs)"""
    d = pb.to_desc()
    bd = make_batch_desc(Q, psi)
    N = (pb.horizon - 1) * pb.m
    u0 = None if u0 is None else np.ascontiguousarray(np.asarray(u0, dtype=np.float64).reshape(pb.batch, -1))
    du_ref, du_q = np.empty((pb.batch, N)), np.empty((pb.batch, N))
    _check(load().oracle_batch_step_truth(C.byref(d), cabi.dptr(u0), C.byref(bd), cabi.dptr(du_ref), cabi.dptr(du_q), threads))
    return du_ref, du_q


def rollout(pb, U, threads=1):
    d = pb.to_desc()
    B, H = pb.batch, pb.horizon
    U = np.ascontiguousarray(np.asarray(U, dtype=np.float64).reshape(B, H - 1, pb.m))
    X = np.empty((B, H, pb.n))
    fX = np.empty((B, H, pb.nt))
    J = np.empty((B, H, pb.r, pb.n))
    cost = np.empty(B)
    _check(load().oracle_rollout(C.byref(d), cabi.dptr(U), cabi.dptr(X), cabi.dptr(fX), cabi.dptr(J), cabi.dptr(cost), threads))
    return X, fX, J, cost


def kinematics(pb, robot, q, dq=None):
    d = pb.to_desc()
    dof = pb.robots[robot].dof
    q = np.ascontiguousarray(np.asarray(q, dtype=np.float64).reshape(-1, dof))
    nq = q.shape[0]
    dq = np.zeros_like(q) if dq is None else np.ascontiguousarray(np.asarray(dq, dtype=np.float64).reshape(-1, dof))
    x, quat, J, dx, w = np.zeros((nq, 3)), np.zeros((nq, 4)), np.zeros((nq, 6, dof)), np.zeros((nq, 3)), np.zeros((nq, 3))
    _check(load().oracle_kinematics(C.byref(d), robot, nq, cabi.dptr(q), cabi.dptr(dq), cabi.dptr(x), cabi.dptr(quat), cabi.dptr(J), cabi.dptr(dx), cabi.dptr(w)))
    return x, quat, J, dx, w


def eval_point(pb, x, k=None, u=None, problem=None, fx_in=None):
    """twin of solver.Handle.eval_point"""
    d = pb.to_desc()
    x = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1, pb.n))
    nq = x.shape[0]
    ia = lambda v: None if v is None else np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.int32), (nq,)))
    da = lambda v, w: None if v is None else np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(nq, w))
    k, problem, u, fx_in = ia(k), ia(problem), da(u, pb.m), da(fx_in, pb.nt)
    o = dict(fx=np.empty((nq, pb.nt)), J=np.empty((nq, pb.r, pb.n)), diff=np.empty((nq, pb.r)), cost=np.empty(nq), cost_x=np.empty((nq, pb.n)),
             cost_xx=np.empty((nq, pb.n, pb.n)))
    _check(load().oracle_eval_point(C.byref(d), nq, cabi.iptr(problem), cabi.iptr(k), cabi.dptr(x), cabi.dptr(u), cabi.dptr(fx_in), cabi.dptr(o["fx"]),
                                    cabi.dptr(o["J"]), cabi.dptr(o["diff"]), cabi.dptr(o["cost"]), cabi.dptr(o["cost_x"]), cabi.dptr(o["cost_xx"])))
    return o


def sd(op, a, b=None, c=None):
    """twin of solver.sd"""
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1, 4))
    nq = a.shape[0]
    cv = lambda v: None if v is None else np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(nq, 4))
    b, c = cv(b), cv(c)
    out = np.empty(nq) if op == cabi.SD_DISTANCE else np.empty((nq, 3, 4)) if op == cabi.SD_DQUAT_TO_W_JAC else np.empty((nq, 4))
    _check(load().oracle_sd(op, nq, cabi.dptr(a), cabi.dptr(b), cabi.dptr(c), cabi.dptr(out)))
    return out


def lqt(A, B, Qs, mu, r, dense=True, variant=None):
    """twin of solver.lqt_solve: solver::LQT restated (lqt.cpp:16-126).  A [n,n] / B [n,m] shared or [nq,n,n] / [nq,n,m]; Qs [nq,T,n,n]; mu [nq,T*n];
    r = pow(rfactor, nb_deriv).  dense=False skips solveLinAl's (T-1)m-sized inverse (u, xhat stay NaN)."""
    Qs = np.ascontiguousarray(Qs, dtype=np.float64)
    nq, T, n = Qs.shape[0], Qs.shape[1], Qs.shape[2]
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    m = B.shape[-1]
    per = int(A.ndim == 3)
    mu = np.ascontiguousarray(np.asarray(mu, dtype=np.float64).reshape(nq, T * n))
    o = {"Ps": np.empty((nq, T, n, n)), "ds": np.empty((nq, T, n)), "K": np.empty((nq, T - 1, m, n)), "f": np.empty((nq, T - 1, m)),
         "u": np.full((nq, (T - 1) * m), np.nan), "xhat": np.full((nq, T * n), np.nan)}
    _check(load(variant).oracle_lqt(nq, n, m, T, per, cabi.dptr(A), cabi.dptr(B), cabi.dptr(Qs), cabi.dptr(mu), float(r), cabi.dptr(o["Ps"]), cabi.dptr(o["ds"]),
                                    cabi.dptr(o["K"]), cabi.dptr(o["f"]), cabi.dptr(o["u"]), cabi.dptr(o["xhat"]), int(dense)))
    return o

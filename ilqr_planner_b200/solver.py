"""Thin Python front of the C ABI: one `Handle` = one batch of problems resident on one GPU."""
import ctypes as C

import numpy as np

from . import cabi
from .problem import ProblemBatch, ResultBuffers


def make_al_desc(A, b, lambda0, lag_update_step, penalty, scaling_factor):
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64))
    b = np.ascontiguousarray(np.asarray(b, dtype=np.float64))
    lam = np.ascontiguousarray(np.asarray(lambda0, dtype=np.float64))
    per_step = 1 if A.ndim == 3 else 0
    al = cabi.ALDesc(A.shape[-2], per_step, cabi.dptr(A), cabi.dptr(b), cabi.dptr(lam), int(lag_update_step), 0, float(penalty), float(scaling_factor))
    al._keep = (A, b, lam)
    return al


def make_batch_desc(Q=None, psi=None, woodbury=False):
    Q = None if Q is None else np.ascontiguousarray(np.asarray(Q, dtype=np.float64))
    psi = None if psi is None else np.ascontiguousarray(np.asarray(psi, dtype=np.float64))
    bd = cabi.BatchDesc(cabi.dptr(Q), cabi.dptr(psi), 0 if psi is None else psi.shape[1], cabi.BATCH_WOODBURY if woodbury else 0)
    bd._keep = (Q, psi)
    return bd


def opts(nb_iter, line_search=True, early_stop=False):
    return cabi.SolveOpts(int(nb_iter), int(bool(line_search)), int(bool(early_stop)), 0)


class Handle:
    def __init__(self, pb: ProblemBatch, device=0, stream=None):
        self.pb = pb
        self.lib = cabi.load()
        self.desc = pb.to_desc()
        self.h = C.c_void_p()
        cabi.check(self.lib.ilqrb_create(C.byref(self.desc), int(device), C.byref(self.h)))
        if stream is not None:
            cabi.check(self.lib.ilqrb_set_stream(self.h, C.c_void_p(stream)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.ilqrb_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _u(self, U0):
        if U0 is None:
            return None
        return np.ascontiguousarray(np.asarray(U0, dtype=np.float64).reshape(self.pb.batch, (self.pb.horizon - 1) * self.pb.m))

    # ---- solvers: host buffers in, host buffers out
    def solve_recursive(self, U0, nb_iter, line_search=True, early_stop=False, out=None):
        out = out or ResultBuffers(self.pb, nb_iter, "recursive")
        U0 = self._u(U0)
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_solve_recursive(self.h, cabi.dptr(U0), C.byref(o), C.byref(out.res)))
        return out

    def solve_recursive_async(self, U0, nb_iter, line_search=True, early_stop=False, out=None):
        """queues H2D + solve + D2H and returns; U0 and `out` (pinned) must stay untouched until synchronize()"""
        out = out or ResultBuffers(self.pb, nb_iter, "recursive", pinned=True)
        self._u0_keep = self._u(U0)
        self._out_keep = out
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_solve_recursive_async(self.h, cabi.dptr(self._u0_keep), C.byref(o), C.byref(out.res)))
        return out

    def solve_al(self, U0, A, b, lambda0, nb_iter, lag_update_step, penalty, scaling_factor, line_search=True, early_stop=False, out=None):
        al = make_al_desc(A, b, lambda0, lag_update_step, penalty, scaling_factor)
        out = out or ResultBuffers(self.pb, nb_iter, "al", nc=al.nc)
        U0 = self._u(U0)
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_solve_al(self.h, cabi.dptr(U0), C.byref(al), C.byref(o), C.byref(out.res)))
        return out

    def solve_batch(self, u0, nb_iter, early_stop=False, Q=None, psi=None, out=None, woodbury=False):
        """woodbury=True: the optional reduced solve of ilqrb.h ILQRB_BATCH_WOODBURY (BatchILQR only; not the reference's algorithm)"""
        bd = make_batch_desc(Q, psi, woodbury)
        out = out or ResultBuffers(self.pb, nb_iter, "batch")
        u0 = self._u(u0)
        o = opts(nb_iter, True, early_stop)
        cabi.check(self.lib.ilqrb_solve_batch(self.h, cabi.dptr(u0), C.byref(bd), C.byref(o), C.byref(out.res)))
        return out

    # ---- split entry points (upload once / run async / download)
    def upload_u0(self, U0):
        self._u0_keep = self._u(U0)
        cabi.check(self.lib.ilqrb_upload_u0(self.h, cabi.dptr(self._u0_keep)))

    def upload_u0_device(self, ptr):
        """U0 [B][H-1][m] already on a GPU: `ptr` = its device address (e.g. torch_tensor.data_ptr())"""
        cabi.check(self.lib.ilqrb_upload_u0(self.h, C.cast(C.c_void_p(int(ptr)), cabi.c_double_p)))

    def run_recursive(self, nb_iter, line_search=True, early_stop=False):
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_run_recursive(self.h, C.byref(o)))

    def run_al(self, al, nb_iter, line_search=True, early_stop=False):
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_run_al(self.h, C.byref(al), C.byref(o)))

    def run_batch(self, bd, nb_iter, early_stop=False):
        o = opts(nb_iter, True, early_stop)
        cabi.check(self.lib.ilqrb_run_batch(self.h, C.byref(bd), C.byref(o)))

    def download(self, out):
        cabi.check(self.lib.ilqrb_download(self.h, C.byref(out.res)))
        return out

    def synchronize(self):
        cabi.check(self.lib.ilqrb_synchronize(self.h))

    def chain_after_compute(self, prev):
        """whatever this handle queues next starts when `prev`'s last solve has left the SMs (its result copies may still run): ilqrb_chain_after_compute"""
        cabi.check(self.lib.ilqrb_chain_after_compute(self.h, prev.h))

    # ---- test hooks
    def rollout(self, U):
        pb = self.pb
        B, H = pb.batch, pb.horizon
        U = np.ascontiguousarray(np.asarray(U, dtype=np.float64).reshape(B, H - 1, pb.m))
        X, fX, J, cost = np.empty((B, H, pb.n)), np.empty((B, H, pb.nt)), np.empty((B, H, pb.r, pb.n)), np.empty(B)
        cabi.check(self.lib.ilqrb_rollout(self.h, cabi.dptr(U), cabi.dptr(X), cabi.dptr(fX), cabi.dptr(J), cabi.dptr(cost)))
        return X, fX, J, cost

    def kinematics(self, robot, q, dq=None):
        dof = self.pb.robots[robot].dof
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float64).reshape(-1, dof))
        nq = q.shape[0]
        dq = np.zeros_like(q) if dq is None else np.ascontiguousarray(np.asarray(dq, dtype=np.float64).reshape(-1, dof))
        x, quat, J, dx, w = np.zeros((nq, 3)), np.zeros((nq, 4)), np.zeros((nq, 6, dof)), np.zeros((nq, 3)), np.zeros((nq, 3))
        cabi.check(self.lib.ilqrb_kinematics(self.h, robot, nq, cabi.dptr(q), cabi.dptr(dq), cabi.dptr(x), cabi.dptr(quat), cabi.dptr(J),
                                             cabi.dptr(dx), cabi.dptr(w)))
        return x, quat, J, dx, w

    def eval_point(self, x, k=None, u=None, problem=None, fx_in=None):
        """System::getFxJac(xk), diff, cost, cost_x, cost_xx at nq points (ilqrb_eval_point): x [nq][n], k [nq] timesteps,
        u [nq][m] (None = zeros), problem [nq] (None = problem 0), fx_in [nq][nt] (None = diff of f(x)).
        Returns a dict of fx, J, diff, cost, cost_x, cost_xx."""
        pb = self.pb
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1, pb.n))
        nq = x.shape[0]
        ia = lambda v: None if v is None else np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.int32), (nq,)))
        da = lambda v, w: None if v is None else np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(nq, w))
        k, problem, u, fx_in = ia(k), ia(problem), da(u, pb.m), da(fx_in, pb.nt)
        o = dict(fx=np.empty((nq, pb.nt)), J=np.empty((nq, pb.r, pb.n)), diff=np.empty((nq, pb.r)), cost=np.empty(nq), cost_x=np.empty((nq, pb.n)),
                 cost_xx=np.empty((nq, pb.n, pb.n)))
        cabi.check(self.lib.ilqrb_eval_point(self.h, nq, cabi.iptr(problem), cabi.iptr(k), cabi.dptr(x), cabi.dptr(u), cabi.dptr(fx_in), cabi.dptr(o["fx"]),
                                             cabi.dptr(o["J"]), cabi.dptr(o["diff"]), cabi.dptr(o["cost"]), cabi.dptr(o["cost_x"]), cabi.dptr(o["cost_xx"])))
        return o

    # ---- profiling
    def set_profiling(self, on):
        cabi.check(self.lib.ilqrb_set_profiling(self.h, int(on)))

    def reset_profile(self):
        cabi.check(self.lib.ilqrb_reset_profile(self.h))

    def profile(self):
        p = cabi.Profile()
        cabi.check(self.lib.ilqrb_get_profile(self.h, C.byref(p)))
        return {p.name[i].decode(): (p.ms[i], p.launches[i]) for i in range(cabi.PROF_SLOTS) if p.launches[i] > 0}

    def launch_count(self):
        return int(self.lib.ilqrb_launch_count(self.h))


def chain_from_urdf(urdf, base, tip, transform_rpy=(0, 0, 0), transform_xyz=(0, 0, 0), is_path=False):
    """URDF -> (segment dicts, dof) via the library's host builder ilqrb_chain_from_urdf"""
    lib = cabi.load()
    segs = (cabi.Segment * cabi.MAX_SEGMENTS)()
    n, dof = C.c_int(0), C.c_int(0)
    rpy = np.ascontiguousarray(np.asarray(transform_rpy, dtype=np.float64))
    xyz = np.ascontiguousarray(np.asarray(transform_xyz, dtype=np.float64))
    cabi.check(lib.ilqrb_chain_from_urdf(urdf.encode(), int(is_path), base.encode(), tip.encode(), cabi.dptr(rpy), cabi.dptr(xyz), segs,
                                         cabi.MAX_SEGMENTS, C.byref(n), C.byref(dof)))
    out = []
    for i in range(n.value):
        s = segs[i]
        out.append(dict(joint_type=int(s.joint_type), axis=np.array(s.axis[:]), origin=np.array(s.origin[:]), tip_R=np.array(s.tip_R[:]),
                        tip_p=np.array(s.tip_p[:])))
    return out, dof.value


def build_psi(kind, dim, K):
    lib = cabi.load()
    cols = 2 * K if kind == "linear" else K
    out = np.zeros((dim, cols))
    cabi.check(lib.ilqrb_build_psi(kind.encode(), int(dim), int(K), cabi.dptr(out)))
    return out


def fp64_peaks(device=0):
    """(DFMA, DMMA m8n8k4) TFLOP/s of `device`, measured in place by the library (ilqrb_fp64_peaks)"""
    lib = cabi.load()
    a, b = C.c_double(0), C.c_double(0)
    cabi.check(lib.ilqrb_fp64_peaks(int(device), C.cast(C.byref(a), cabi.c_double_p), C.cast(C.byref(b), cabi.c_double_p)))
    return a.value, b.value


def multi_plan(batch, n_slices):
    """contiguous slices of `batch` problems over n_slices devices (ilqrb_multi_plan; no GPU work)"""
    lib = cabi.load()
    lo, hi = (C.c_int * n_slices)(), (C.c_int * n_slices)()
    cabi.check(lib.ilqrb_multi_plan(int(batch), int(n_slices), lo, hi))
    return list(zip(lo[:], hi[:]))


class MultiHandle:
    """one batch sharded over several GPUs of one box (ilqrb_create_multi): same solve calls as Handle, whole-batch buffers"""

    def __init__(self, pb: ProblemBatch, devices):
        self.pb = pb
        self.lib = cabi.load()
        self.desc = pb.to_desc()
        self.h = C.c_void_p()
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        cabi.check(self.lib.ilqrb_create_multi(C.byref(self.desc), len(devices), devs, C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.ilqrb_multi_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    _u = Handle._u

    def slices(self):
        out = []
        for i in range(self.lib.ilqrb_multi_count(self.h)):
            d, lo, hi = C.c_int(), C.c_int(), C.c_int()
            cabi.check(self.lib.ilqrb_multi_slice(self.h, i, C.byref(d), C.byref(lo), C.byref(hi), None))
            out.append((d.value, lo.value, hi.value))
        return out

    def solve_recursive(self, U0, nb_iter, line_search=True, early_stop=False, out=None):
        out = out or ResultBuffers(self.pb, nb_iter, "recursive")
        U0 = self._u(U0)
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_multi_solve_recursive(self.h, cabi.dptr(U0), C.byref(o), C.byref(out.res)))
        return out

    def solve_al(self, U0, A, b, lambda0, nb_iter, lag_update_step, penalty, scaling_factor, line_search=True, early_stop=False, out=None):
        al = make_al_desc(A, b, lambda0, lag_update_step, penalty, scaling_factor)
        out = out or ResultBuffers(self.pb, nb_iter, "al", nc=al.nc)
        U0 = self._u(U0)
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_multi_solve_al(self.h, cabi.dptr(U0), C.byref(al), C.byref(o), C.byref(out.res)))
        return out

    def solve_batch(self, u0, nb_iter, early_stop=False, Q=None, psi=None, out=None):
        bd = make_batch_desc(Q, psi)
        out = out or ResultBuffers(self.pb, nb_iter, "batch")
        u0 = self._u(u0)
        o = opts(nb_iter, True, early_stop)
        cabi.check(self.lib.ilqrb_multi_solve_batch(self.h, cabi.dptr(u0), C.byref(bd), C.byref(o), C.byref(out.res)))
        return out

    def upload_u0(self, U0):
        self._u0_keep = self._u(U0)
        cabi.check(self.lib.ilqrb_multi_upload_u0(self.h, cabi.dptr(self._u0_keep)))

    def run_recursive(self, nb_iter, line_search=True, early_stop=False):
        o = opts(nb_iter, line_search, early_stop)
        cabi.check(self.lib.ilqrb_multi_run_recursive(self.h, C.byref(o)))

    def synchronize(self):
        cabi.check(self.lib.ilqrb_multi_synchronize(self.h))

    def download(self, out, nb_iter):
        cabi.check(self.lib.ilqrb_multi_download(self.h, C.byref(out.res), int(nb_iter)))
        return out


def keypoint_diff(sys_kind, task_kind, nb_deriv, dof, target, fx, deadzone=None, device=0):
    """Keypoint::diff without a system (ilqrb_keypoint_diff): target = one KP_STRIDE record, fx [nq][nt] -> [nq][r]"""
    lib = cabi.load()
    time = 1 if sys_kind == cabi.SYS_POS_ORN_TIME else 0
    if task_kind == cabi.TASK_JOINT:
        nt = r = nb_deriv * dof + time
    elif sys_kind == cabi.SYS_POSITION:
        nt = r = dof * nb_deriv
    else:
        nt, r = 7 * nb_deriv + time, 6 * nb_deriv + time
    target = np.ascontiguousarray(np.asarray(target, dtype=np.float64).reshape(cabi.KP_STRIDE))
    fx = np.ascontiguousarray(np.asarray(fx, dtype=np.float64).reshape(-1, nt))
    dz = None if deadzone is None else np.ascontiguousarray(np.asarray(deadzone, dtype=np.float64).reshape(4))
    out = np.empty((fx.shape[0], r))
    cabi.check(lib.ilqrb_keypoint_diff(device, sys_kind, task_kind, nb_deriv, dof, cabi.dptr(target), cabi.dptr(dz), fx.shape[0], cabi.dptr(fx), cabi.dptr(out)))
    return out


def sd(op, a, b=None, c=None, device=0):
    """utils::Sd on S^3 (ilqrb_sd): op = cabi.SD_*, operands [nq][4]"""
    lib = cabi.load()
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1, 4))
    nq = a.shape[0]
    cv = lambda v: None if v is None else np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(nq, 4))
    b, c = cv(b), cv(c)
    out = np.empty(nq) if op == cabi.SD_DISTANCE else np.empty((nq, 3, 4)) if op == cabi.SD_DQUAT_TO_W_JAC else np.empty((nq, 4))
    cabi.check(lib.ilqrb_sd(device, op, nq, cabi.dptr(a), cabi.dptr(b), cabi.dptr(c), cabi.dptr(out)))
    return out


def lqt_solve(A, B, Qs, mu, r, device=0):
    """solver::LQT for a batch (ilqrb_lqt_solve; lqt.cpp:16-126).  A [n,n] / B [n,m] shared by the problems, or [nq,n,n] / [nq,n,m]; Qs [nq,T,n,n];
    mu [nq,T*n]; r = pow(rfactor, nb_deriv).  Returns Ps, ds (solveDP, time order), K, f (getCommand(timestep, state)), u, xhat (solveLinAl), status."""
    lib = cabi.load()
    Qs = np.ascontiguousarray(Qs, dtype=np.float64)
    if Qs.ndim != 4 or Qs.shape[2] != Qs.shape[3]:
        raise ValueError("Qs: [nq, T, n, n] expected")
    nq, T, n = Qs.shape[0], Qs.shape[1], Qs.shape[2]
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    m = B.shape[-1]
    per = int(A.ndim == 3)
    if A.shape != ((nq, n, n) if per else (n, n)) or B.shape != ((nq, n, m) if per else (n, m)):
        raise ValueError("A / B: [n, n] and [n, m] (or one pair per problem) expected")
    mu = np.ascontiguousarray(np.asarray(mu, dtype=np.float64).reshape(nq, T * n))
    d = cabi.LqtDesc(nq, n, m, T, per, 0, cabi.dptr(A), cabi.dptr(B), cabi.dptr(Qs), cabi.dptr(mu), float(r))
    o = {"Ps": np.empty((nq, T, n, n)), "ds": np.empty((nq, T, n)), "K": np.empty((nq, T - 1, m, n)), "f": np.empty((nq, T - 1, m)),
         "u": np.empty((nq, (T - 1) * m)), "xhat": np.empty((nq, T * n)), "status": np.zeros(nq, dtype=np.int32)}
    cabi.check(lib.ilqrb_lqt_solve(device, C.byref(d), cabi.dptr(o["Ps"]), cabi.dptr(o["ds"]), cabi.dptr(o["K"]), cabi.dptr(o["f"]), cabi.dptr(o["u"]),
                                   cabi.dptr(o["xhat"]), cabi.iptr(o["status"])))
    return o


def lqt_command(K, f, mu, timestep, state, problem=None, device=0):
    """LQT::getCommand(timestep, curr_state) for a set of queries (ilqrb_lqt_command; lqt.cpp:105-124)"""
    lib = cabi.load()
    K = np.ascontiguousarray(K, dtype=np.float64)
    nq, Tm1, m, n = K.shape
    f = np.ascontiguousarray(f, dtype=np.float64)
    mu = np.ascontiguousarray(np.asarray(mu, dtype=np.float64).reshape(nq, (Tm1 + 1) * n))
    ts = np.ascontiguousarray(np.atleast_1d(timestep), dtype=np.int32)
    nc = ts.shape[0]
    st = np.ascontiguousarray(np.asarray(state, dtype=np.float64).reshape(nc, n))
    pr = None if problem is None else np.ascontiguousarray(np.atleast_1d(problem), dtype=np.int32)
    u = np.empty((nc, m))
    cabi.check(lib.ilqrb_lqt_command(device, nq, n, m, Tm1 + 1, cabi.dptr(K), cabi.dptr(f), cabi.dptr(mu), nc, cabi.iptr(pr), cabi.iptr(ts), cabi.dptr(st), cabi.dptr(u)))
    return u

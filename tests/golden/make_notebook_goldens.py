#!/usr/bin/env python
"""Generates tests/golden/notebook_goldens.json from the reference's tutorial notebooks.

Run HERE (the container that has /root/reference); the GPU box never needs the reference.

The reference has no solver tests; its only known-answer vectors are the stdout lines
"Iteration i, Cost: c, alpha= a[, time= t]" stored in pylqr_planner/Tutorials/*.ipynb.
This script executes each notebook's code cells against a *recording* stand-in for the `PyLQR` module:
constructors and solve() calls only record their arguments (robot kinematics come from the CPU oracle so that
cells computing object frames from FK keep working), and every solve() call is paired with the trace stored in
that cell's output.  Only DATA is written to the JSON (problem parameters + traces); no reference code is copied.
"""
import json
import math
import os
import re
import sys
import types
from unittest import mock

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from ilqr_planner_b200 import cabi  # noqa: E402
from ilqr_planner_b200.problem import ProblemBatch, RobotSpec, TaskSpec  # noqa: E402
from oracle import oracle, urdf_chain  # noqa: E402

TUT = "/root/reference/pylqr_planner/Tutorials"
NOTEBOOKS = ["POS_ORN_SYS", "POS_ORN_SYS_2ND", "POS_ORN_SYS_AL_ILQR", "POS_ORN_SYS_OBJ_FRAME", "POS_ORN_TIME_SYS",
             "POS_ORN_TIME_SYS_2ND", "POS_ORN_MULTI_SYS", "POS_ORN_MULTI_SYS_2ND", "POS_ORN_MULTI_SYS_TIME"]

CHAINS = {}
CASES = []
CURRENT = {"nb": None, "cell": None, "traces": None}


def _arr(a):
    return np.asarray(a, dtype=np.float64)


# ----------------------------------------------------------------------------------------------- sim
class SimulationInterface:
    pass


class KDLRobot(SimulationInterface):
    def __init__(self, urdf, base_frame, tip_frame, q, dq, transform_rpy=(0, 0, 0), transform_xyz=(0, 0, 0), is_path=True):
        text = open(os.path.join(TUT, urdf)).read() if is_path else urdf
        self.segs = urdf_chain.kdl_chain_from_urdf(text, base_frame, tip_frame, transform_rpy, transform_xyz)
        self.chain_key = "%s:%s->%s:%s:%s" % (os.path.basename(urdf) if is_path else "inline", base_frame, tip_frame,
                                              list(map(float, transform_rpy)), list(map(float, transform_xyz)))
        CHAINS[self.chain_key] = [dict(joint_type=int(s["joint_type"]), axis=list(map(float, s["axis"])), origin=list(map(float, s["origin"])),
                                       tip_R=list(map(float, s["tip_R"])), tip_p=list(map(float, s["tip_p"]))) for s in self.segs]
        self.dof = len(q)
        self.q, self.dq, self.t = _arr(q).copy(), _arr(dq).copy(), 0.0
        self.T = None

    def base(self):
        return self

    def _pb(self):
        return ProblemBatch(cabi.SYS_POS_ORN, 1, 2, 0.1, self.dof, [RobotSpec(cabi.ROBOT_KDL, self.dof, self.segs)], [], np.zeros(self.dof),
                            np.zeros((1, self.dof)), np.zeros((1, self.dof)))

    def _fk(self):
        b = self.base()
        x, quat, J, dx, w = oracle.kinematics(b._pb(), 0, b.q, b.dq)
        return x[0], quat[0], J[0], dx[0], w[0]

    # state
    def get_q(self):
        return self.base().q.copy()

    def get_dq(self):
        return self.base().dq.copy()

    def get_time(self):
        return self.base().t

    def set_time(self, t):
        self.base().t = t

    def set_conf(self, q, dq, reset_time=True):
        b = self.base()
        b.q, b.dq = _arr(q).copy(), _arr(dq).copy()
        if reset_time:
            b.t = 0.0

    def send_vel(self, dt, dq, updateKin=True):
        b = self.base()
        b.dq = _arr(dq).copy()
        b.q = b.q + dt * b.dq
        b.t += dt

    def send_acc(self, dt, ddq, updateKin=True):
        b = self.base()
        ddq = _arr(ddq)
        b.q = b.q + dt * b.dq + dt * dt / 2 * ddq
        b.dq = b.dq + dt * ddq
        b.t += dt

    def update_kinematics(self):
        pass

    def get_ee_pos(self):
        x = self._fk()[0]
        if self.T is not None:
            x = self.T[:3, :3].T @ (x - self.T[:3, 3])
        return x

    def get_ee_orn(self):
        quat = self._fk()[1]
        if self.T is None:
            return quat
        w, x, y, z = quat
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        M = self.T[:3, :3].T @ R
        tr = np.trace(M)
        # only used for plotting cells in the notebooks; branch-free is fine here
        qw = math.sqrt(max(tr + 1.0, 1e-300)) / 2
        return np.array([qw, (M[2, 1] - M[1, 2]) / (4 * qw), (M[0, 2] - M[2, 0]) / (4 * qw), (M[1, 0] - M[0, 1]) / (4 * qw)])

    def get_ee_vel(self):
        return self._fk()[3]

    def get_ee_ang_vel(self):
        return self._fk()[4]

    def get_ee_ang_vel_quat(self):
        return np.zeros(4)

    def J(self):
        return self._fk()[2]


class TransformedSimulationInterface(KDLRobot):
    def __init__(self, r, T):
        self.r = r
        self.T = _arr(T).copy()
        self.dof = r.dof

    def base(self):
        return self.r.base()


class Robot2D(SimulationInterface):
    def __init__(self, *a, **k):
        raise NotImplementedError


# -------------------------------------------------------------------------------------------- system
class PosOrnKeypoint:
    def __init__(self, *args):
        if len(args) == 4:
            pos, orn, prec, ts = args
            dpos, dorn, self.order = np.zeros(3), np.zeros(4), 1
        else:
            pos, dpos, orn, dorn, prec, ts = args
            self.order = 2
        self.target = np.zeros(cabi.KP_STRIDE)
        self.target[0:3], self.target[3:7], self.target[7:10], self.target[10:14] = _arr(pos), _arr(orn), _arr(dpos), _arr(dorn)
        self.precision = _arr(prec).copy()
        self.timestep = int(ts)


class SpacetimeKeypoint(PosOrnKeypoint):
    def __init__(self, *args):
        if len(args) == 5:
            pos, orn, prec, ct, ts = args
            super().__init__(pos, orn, prec, ts)
        else:
            pos, dpos, orn, dorn, prec, ct, ts = args
            super().__init__(pos, dpos, orn, dorn, prec, ts)
        self.target[14] = float(ct)


class _Sys:
    time = 0

    def _common(self, r, keypoints, RtDiag, horizon, nb_deriv, dt, limits):
        self.r, self.horizon, self.nb_deriv, self.dt = r, int(horizon), int(nb_deriv), float(dt)
        self.keypoints = sorted(keypoints, key=lambda k: k.timestep)
        self.r_diag = _arr(RtDiag).copy()
        self.q0, self.dq0 = r.get_q(), r.get_dq()  # localInit(): captured at construction
        dof = r.dof
        n = nb_deriv * dof + self.time
        self.limits = None
        if limits is not None:
            qMax, qMin, dqMax, dqMin = limits
            smax, smin = np.zeros(n), np.zeros(n)
            w = np.ones(n, dtype=np.int32)
            smax[:dof], smin[:dof] = _arr(qMax), _arr(qMin)
            if nb_deriv == 2:
                smax[dof:2 * dof], smin[dof:2 * dof] = _arr(dqMax), _arr(dqMin)
                a, b = _arr(dqMax), _arr(dqMin)
                if np.sum((a - b) ** 2) <= 1e-24 * min(np.sum(a * a), np.sum(b * b)):
                    w[dof:2 * dof] = 0
            if self.time:
                w[-1] = 0
            self.limits = (smax, smin, w)

    def task_json(self):
        return dict(T=None if self.r.T is None else self.r.T.reshape(-1).tolist(), kp_timestep=[k.timestep for k in self.keypoints],
                    kp_target=[k.target.tolist() for k in self.keypoints], kp_precision=[k.precision.tolist() for k in self.keypoints],
                    r_diag=self.r_diag.tolist(),
                    limits=None if self.limits is None else dict(state_max=self.limits[0].tolist(), state_min=self.limits[1].tolist(),
                                                                 weight=self.limits[2].tolist()))

    def spec(self):
        b = self.r.base()
        return dict(sys_kind=self.sys_kind, nb_deriv=self.nb_deriv, horizon=self.horizon, dt=self.dt, dof=b.dof, chain=b.chain_key,
                    q0=self.q0.tolist(), dq0=self.dq0.tolist(), r_diag=self.r_diag.tolist(), tasks=[self.task_json()])

    def get_mu_vector(self, *a):
        return None

    def get_Q_matrix(self, *a):
        return None


class PosOrnPlannerSys(_Sys):
    sys_kind = cabi.SYS_POS_ORN

    def __init__(self, r, keypoints, RtDiag, *rest):
        if len(rest) == 3:
            horizon, nb_deriv, dt = rest
            limits = None
        elif len(rest) == 5:
            qMax, qMin, horizon, nb_deriv, dt = rest
            limits = (qMax, qMin, np.zeros(r.dof), np.zeros(r.dof))
        else:
            qMax, qMin, dqMax, dqMin, horizon, nb_deriv, dt = rest
            limits = (qMax, qMin, dqMax, dqMin)
        self._common(r, keypoints, RtDiag, horizon, nb_deriv, dt, limits)


class PosOrnTimePlannerSys(_Sys):
    sys_kind = cabi.SYS_POS_ORN_TIME
    time = 1

    def __init__(self, r, keypoints, RtDiag, *rest):
        if len(rest) == 2:
            horizon, nb_deriv = rest
            limits = None
        elif len(rest) == 4:
            qMax, qMin, horizon, nb_deriv = rest
            limits = (qMax, qMin, np.zeros(r.dof), np.zeros(r.dof))
        else:
            qMax, qMin, dqMax, dqMin, horizon, nb_deriv = rest
            limits = (qMax, qMin, dqMax, dqMin)
        self._common(r, keypoints, RtDiag, horizon, nb_deriv, 0.0, limits)


class SequentialSystem(_Sys):
    def __init__(self, r, systems, RtDiag, horizon, nbDeriv):
        self.r, self.systems, self.r_diag = r, systems, _arr(RtDiag).copy()

    def spec(self):
        s0 = self.systems[0].spec()
        s0["tasks"] = [s.task_json() for s in self.systems]
        s0["r_diag"] = self.r_diag.tolist()
        return s0


class _Unsupported:
    def __init__(self, *a, **k):
        raise NotImplementedError("out of scope (SURVEY §2 rows 14/15)")


# -------------------------------------------------------------------------------------------- solver
def _next_trace():
    tr = CURRENT["traces"]
    if not tr:
        raise RuntimeError("solve() call without stored trace in %s cell %s" % (CURRENT["nb"], CURRENT["cell"]))
    return tr.pop(0)


def _compact_u(kw, m):
    for key in ("U0", "u0"):
        if key in kw:
            a = _arr(kw.pop(key)).reshape(-1, m)
            if np.all(a == a[0]):
                kw["u0_row"] = a[0].tolist()
            else:
                kw["u0_full"] = a.tolist()


def _record(solver, sysobj, **kw):
    sp = sysobj.spec()
    _compact_u(kw, sp["dof"] + (1 if sp["sys_kind"] == cabi.SYS_POS_ORN_TIME else 0))
    case = dict(name="%s/cell%d/%s" % (CURRENT["nb"], CURRENT["cell"], solver), notebook=CURRENT["nb"], cell=CURRENT["cell"], solver=solver,
                system=sysobj.spec(), trace=_next_trace())
    case.update(kw)
    CASES.append(case)
    return case


class ILQRRecursive:
    def __init__(self, s):
        self.s = s

    def solve(self, U0, nb_iter, line_search=True, early_stop=False, cb=None):
        U0 = _arr(U0)
        _record("recursive", self.s, U0=U0.tolist(), nb_iter=int(nb_iter), line_search=bool(line_search), early_stop=bool(early_stop))
        sp = self.s.spec()
        H, n = sp["horizon"], sp["nb_deriv"] * sp["dof"] + (1 if sp["sys_kind"] == cabi.SYS_POS_ORN_TIME else 0)
        nt = sum((7 * sp["nb_deriv"] + (1 if sp["sys_kind"] == cabi.SYS_POS_ORN_TIME else 0)) for _ in sp["tasks"])
        m = U0.shape[1]
        return ([np.zeros(n) for _ in range(H)], [np.zeros(nt) for _ in range(H)], [np.zeros(m) for _ in range(H - 1)],
                [np.zeros((m, n)) for _ in range(H - 1)], [np.zeros(m) for _ in range(H - 1)], 0.0)


class Constraint:
    def __init__(self):
        self.A, self.b = None, None


class AL_ILQR:
    def __init__(self, s, inequality, initLambda):
        self.s = s
        self.A = np.stack([_arr(c.A) for c in inequality])
        self.b = np.stack([_arr(c.b) for c in inequality])
        self.lam = np.stack([_arr(l) for l in initLambda])

    def solve(self, U0, nb_iter, lag_update_step, penalty, scaling_factor, line_search=True, early_stop=False, cb=None):
        U0 = _arr(U0)
        A, b, lam = self.A, self.b, self.lam
        if all(np.array_equal(A[0], A[k]) and np.array_equal(b[0], b[k]) and np.array_equal(lam[0], lam[k]) for k in range(len(A))):
            A, b, lam = A[0], b[0], lam[0]
        _record("al", self.s, U0=U0.tolist(), nb_iter=int(nb_iter), line_search=bool(line_search), early_stop=bool(early_stop), A=A.tolist(),
                b=b.tolist(), lambda0=lam.tolist(), lag_update_step=int(lag_update_step), penalty=float(penalty), scaling_factor=float(scaling_factor))
        sp = self.s.spec()
        H, n = sp["horizon"], sp["nb_deriv"] * sp["dof"]
        return [np.zeros(n) for _ in range(H)], [np.zeros(7 * sp["nb_deriv"]) for _ in range(H)], [np.zeros(U0.shape[1]) for _ in range(H - 1)]


class BatchILQR:
    def __init__(self, s, Q=None):
        self.s, self.Q = s, Q

    def solve(self, nb_iter, u0, early_stop=False, cb=None):
        u0 = _arr(u0)
        _record("batch", self.s, u0=u0.tolist(), nb_iter=int(nb_iter), early_stop=bool(early_stop), Q=None if self.Q is None else _arr(self.Q).tolist())
        return np.zeros_like(u0)


class BatchILQRCP:
    def __init__(self, s, *args):
        self.s = s
        if len(args) == 2:
            self.Q, self.psi = args
        else:
            self.Q, self.psi = None, args[0]

    def solve(self, nb_iter, u0, early_stop=False, cb=None):
        u0 = _arr(u0)
        psi = _arr(self.psi)
        # store psi compactly when it is kron(psi_small, I_m); otherwise dense
        small = PSI_LAST.get("small")
        if small is not None and psi.shape[1] % small.shape[1] == 0 and psi.shape[0] % small.shape[0] == 0 and \
                psi.shape[0] // small.shape[0] == psi.shape[1] // small.shape[1] and \
                np.array_equal(psi, np.kron(small, np.identity(psi.shape[1] // small.shape[1]))):
            extra = dict(psi_kron=small.tolist(), psi_kind=PSI_LAST["kind"], psi_K=PSI_LAST["K"])
        else:
            extra = dict(psi=psi.tolist())
        _record("batch_cp", self.s, u0=u0.tolist(), nb_iter=int(nb_iter), early_stop=bool(early_stop), Q=None if self.Q is None else _arr(self.Q).tolist(),
                **extra)
        return np.zeros_like(u0)


# --------------------------------------------------------------------------------------------- utils
PSI_LAST = {}


def _psi(kind):
    def f(dim, K):
        from tests.golden import psi_ref
        p = psi_ref.build_psi(kind, dim, K)
        PSI_LAST.update(small=p, kind=kind, K=int(K))
        return p
    return f


class PythonCallbackMessage:
    pass


def install_shim():
    root = types.ModuleType("PyLQR")
    sim = types.ModuleType("PyLQR.sim")
    system = types.ModuleType("PyLQR.system")
    solver = types.ModuleType("PyLQR.solver")
    utils = types.ModuleType("PyLQR.utils")
    primitives = types.ModuleType("PyLQR.utils.primitives")
    for k, v in dict(SimulationInterface=SimulationInterface, KDLRobot=KDLRobot, Robot2D=Robot2D,
                     TransformedSimulationInterface=TransformedSimulationInterface).items():
        setattr(sim, k, v)
    for k, v in dict(PosOrnKeypoint=PosOrnKeypoint, SpacetimeKeypoint=SpacetimeKeypoint, PosOrnPlannerSys=PosOrnPlannerSys,
                     PosOrnTimePlannerSys=PosOrnTimePlannerSys, SequentialSystem=SequentialSystem, JointSpacePlannerSys=_Unsupported,
                     JointSpaceTimePlannerSys=_Unsupported, AngularKeypoint=_Unsupported, AngularTimeKeypoint=_Unsupported).items():
        setattr(system, k, v)
    for k, v in dict(ILQRRecursive=ILQRRecursive, AL_ILQR=AL_ILQR, Constraint=Constraint, BatchILQR=BatchILQR, BatchILQRCP=BatchILQRCP).items():
        setattr(solver, k, v)
    for kind in ("rbf", "bernstein", "unitstep", "sawtooth", "linear"):
        setattr(primitives, "build_psi_" + kind, _psi(kind))
    utils.primitives = primitives
    utils.PythonCallbackMessage = PythonCallbackMessage
    root.sim, root.system, root.solver, root.utils = sim, system, solver, utils
    sys.modules.update({"PyLQR": root, "PyLQR.sim": sim, "PyLQR.system": system, "PyLQR.solver": solver, "PyLQR.utils": utils,
                        "PyLQR.utils.primitives": primitives})
    for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
        sys.modules[name] = mock.MagicMock()


LINE = re.compile(r"Iteration (\d+), Cost: (\S+), alpha= ([^,\s]+)")


def parse_traces(cell):
    """one trace per solve() call in the cell; the iteration counter restarting at 1 separates them"""
    text = "".join("".join(o.get("text", [])) for o in cell.get("outputs", []) if o.get("output_type") == "stream")
    traces, cur = [], None
    for line in text.splitlines():
        mm = LINE.match(line.strip())
        if not mm:
            continue
        if int(mm.group(1)) == 1:
            cur = []
            traces.append(cur)
        cur.append(dict(cost=mm.group(2), alpha=mm.group(3)))
    return traces


def run_notebook(name):
    nb = json.load(open(os.path.join(TUT, name + ".ipynb")))
    ns = {"__name__": "__notebook__"}
    for i, cell in enumerate(nb["cells"]):
        if cell["cell_type"] != "code":
            continue
        src = "".join(cell["source"])
        src = "\n".join(l for l in src.splitlines() if not l.lstrip().startswith("%"))
        CURRENT.update(nb=name, cell=i, traces=parse_traces(cell))
        exec(compile(src, "%s[cell %d]" % (name, i), "exec"), ns)
        for fn in ("set_axes_equal", "plot_pose_matrix"):  # plotting helpers defined by some notebooks
            if fn in ns:
                ns[fn] = lambda *a, **k: None


def main():
    install_shim()
    for name in NOTEBOOKS:
        n0 = len(CASES)
        run_notebook(name)
        print("%-26s %d solve calls recorded" % (name, len(CASES) - n0))
    out = dict(source="idiap/ilqr_planner @ eed29c7, pylqr_planner/Tutorials/*.ipynb stored outputs", chains=CHAINS, cases=CASES)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "notebook_goldens.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

"""Python-side builder of `ilqrb_problem_desc` batches (plumbing over the C ABI, include/ilqrb.h).

A `ProblemBatch` mirrors the reference's object graph for B problems of one structure:
  robots  ~ sim::KDLRobot / sim::Robot2D                      (KDLRobot.cpp, 2DRobot.cpp)
  tasks   ~ one sys::PosOrnPlannerSys / PosOrnTimePlannerSys per entry; >1 entries = sys::SequentialSystem,
            each optionally seen through a sim::TransformedSimulationInterface (T)
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import cabi


def _f64(a, shape=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if shape is not None:
        a = np.ascontiguousarray(a.reshape(shape))
    return a


@dataclass
class RobotSpec:
    kind: int = cabi.ROBOT_KDL
    dof: int = 7
    segments: Optional[list] = None  # list of dicts (joint_type, axis, origin, tip_R, tip_p)
    lengths: Optional[np.ndarray] = None

    def seg_array(self):
        n = len(self.segments)
        arr = (cabi.Segment * n)()
        for i, s in enumerate(self.segments):
            arr[i].joint_type = int(s["joint_type"])
            arr[i].axis[:] = [float(v) for v in s["axis"]]
            arr[i].origin[:] = [float(v) for v in s["origin"]]
            arr[i].tip_R[:] = [float(v) for v in np.asarray(s["tip_R"]).reshape(-1)]
            arr[i].tip_p[:] = [float(v) for v in s["tip_p"]]
        return arr


@dataclass
class TaskSpec:
    kp_timestep: List[int]
    kp_target: np.ndarray          # [B][Kp][15]
    kp_precision: np.ndarray       # [B or 1][Kp][r][r]
    r_diag: np.ndarray             # [m]
    robot: int = 0
    joint_offset: int = 0
    T: Optional[np.ndarray] = None  # 4x4
    state_max: Optional[np.ndarray] = None
    state_min: Optional[np.ndarray] = None
    limit_weight: Optional[np.ndarray] = None
    task_kind: int = cabi.TASK_POSE           # TASK_JOINT: JointSpace(Time)PlannerSys with Angular(Time)Keypoints
    kp_deadzone: Optional[np.ndarray] = None  # [Kp][4] pos_radius, orn_thresh[3] (PosOrnKeypointDistFunct); None = plain keypoints


@dataclass
class ProblemBatch:
    sys_kind: int
    nb_deriv: int
    horizon: int
    dt: float
    dof: int
    robots: List[RobotSpec]
    tasks: List[TaskSpec]
    r_diag: np.ndarray
    q0: np.ndarray   # [B][dof]
    dq0: np.ndarray  # [B][dof]
    _keep: list = field(default_factory=list, repr=False)

    @property
    def batch(self):
        return int(np.asarray(self.q0).reshape(-1, self.dof).shape[0])

    @property
    def time(self):
        return 1 if self.sys_kind == cabi.SYS_POS_ORN_TIME else 0

    @property
    def n(self):
        return self.nb_deriv * self.dof + self.time

    @property
    def m(self):
        return self.dof + self.time

    def task_dims(self, t):
        if self.tasks[t].task_kind == cabi.TASK_JOINT:
            d = self.nb_deriv * self.robots[self.tasks[t].robot].dof + self.time
            return d, d
        if self.sys_kind == cabi.SYS_POSITION:
            car = 2 if self.robots[self.tasks[t].robot].kind == cabi.ROBOT_2D else 3
            return car * self.nb_deriv, car * self.nb_deriv
        return 7 * self.nb_deriv + self.time, 6 * self.nb_deriv + self.time

    @property
    def nt(self):
        return sum(self.task_dims(t)[0] for t in range(len(self.tasks)))

    @property
    def r(self):
        return sum(self.task_dims(t)[1] for t in range(len(self.tasks)))

    @property
    def n_kp_total(self):
        return sum(len(t.kp_timestep) for t in self.tasks)

    def slice(self, lo, hi):
        """contiguous shard [lo, hi) of the batch (multi-GPU sharding)"""
        tasks = []
        for t in self.tasks:
            prec = np.asarray(t.kp_precision)
            tasks.append(TaskSpec(kp_timestep=list(t.kp_timestep), kp_target=np.asarray(t.kp_target)[lo:hi],
                                  kp_precision=prec[lo:hi] if prec.shape[0] > 1 else prec, r_diag=t.r_diag, robot=t.robot,
                                  joint_offset=t.joint_offset, T=t.T, state_max=t.state_max, state_min=t.state_min,
                                  limit_weight=t.limit_weight, task_kind=t.task_kind, kp_deadzone=t.kp_deadzone))
        return ProblemBatch(self.sys_kind, self.nb_deriv, self.horizon, self.dt, self.dof, self.robots, tasks, self.r_diag,
                            np.asarray(self.q0)[lo:hi], np.asarray(self.dq0)[lo:hi])

    def to_desc(self):
        """-> ctypes ProblemDesc (arrays kept alive on self._keep)"""
        keep = []
        B = self.batch
        n, m = self.n, self.m
        robots = (cabi.Robot * len(self.robots))()
        for i, rb in enumerate(self.robots):
            robots[i].kind = rb.kind
            robots[i].dof = rb.dof
            if rb.kind == cabi.ROBOT_KDL:
                sa = rb.seg_array()
                keep.append(sa)
                robots[i].n_segments = len(rb.segments)
                robots[i].segments = C.cast(sa, C.POINTER(cabi.Segment))
            else:
                ln = _f64(rb.lengths)
                keep.append(ln)
                robots[i].lengths = cabi.dptr(ln)
        tasks = (cabi.Task * len(self.tasks))()
        for i, t in enumerate(self.tasks):
            _, r_t = self.task_dims(i)
            Kp = len(t.kp_timestep)
            ts = np.ascontiguousarray(np.asarray(t.kp_timestep, dtype=np.int32))
            tg = _f64(t.kp_target, (B, Kp, cabi.KP_STRIDE))
            pr = _f64(t.kp_precision)
            per = 1 if pr.shape[0] == B and B > 1 else 0
            pr = _f64(pr, (B if per else 1, Kp, r_t, r_t))
            rd = _f64(t.r_diag, (m,))
            keep += [ts, tg, pr, rd]
            tasks[i].robot = t.robot
            tasks[i].joint_offset = t.joint_offset
            tasks[i].n_keypoints = Kp
            tasks[i].kp_timestep = cabi.iptr(ts)
            tasks[i].kp_target = cabi.dptr(tg)
            tasks[i].kp_precision = cabi.dptr(pr)
            tasks[i].precision_per_problem = per
            tasks[i].task_kind = int(t.task_kind)
            if t.kp_deadzone is not None:
                dz = _f64(t.kp_deadzone, (Kp, 4))
                keep.append(dz)
                tasks[i].kp_deadzone = cabi.dptr(dz)
            tasks[i].r_diag = cabi.dptr(rd)
            if t.T is not None:
                tasks[i].has_transform = 1
                tasks[i].T[:] = [float(v) for v in np.asarray(t.T, dtype=np.float64).reshape(-1)]
            if t.state_max is not None:
                smax, smin = _f64(t.state_max, (n,)), _f64(t.state_min, (n,))
                w = np.ascontiguousarray(np.asarray(t.limit_weight if t.limit_weight is not None else np.ones(n), dtype=np.int32))
                keep += [smax, smin, w]
                tasks[i].limits_set = 1
                tasks[i].state_max = cabi.dptr(smax)
                tasks[i].state_min = cabi.dptr(smin)
                tasks[i].limit_weight = cabi.iptr(w)
        rd = _f64(self.r_diag, (m,))
        q0 = _f64(self.q0, (B, self.dof))
        dq0 = _f64(self.dq0, (B, self.dof))
        keep += [robots, tasks, rd, q0, dq0]
        d = cabi.ProblemDesc()
        d.abi_version = cabi.ABI_VERSION
        d.sys_kind = self.sys_kind
        d.nb_deriv = self.nb_deriv
        d.horizon = self.horizon
        d.batch = B
        d.dof = self.dof
        d.n_robots = len(self.robots)
        d.n_tasks = len(self.tasks)
        d.dt = float(self.dt)
        d.robots = C.cast(robots, C.POINTER(cabi.Robot))
        d.tasks = C.cast(tasks, C.POINTER(cabi.Task))
        d.r_diag = cabi.dptr(rd)
        d.q0 = cabi.dptr(q0)
        d.dq0 = cabi.dptr(dq0)
        self._keep = keep
        return d


def limit_weights(nb_deriv, dof, time, dq_max=None, dq_min=None):
    """joint_limits_weight_ as System.cpp:49-60 + PosOrnTimePlannerSys.cpp:72-82 build it"""
    w = np.ones(nb_deriv * dof + time, dtype=np.int32)
    if nb_deriv == 2 and dq_max is not None:
        a, b = np.asarray(dq_max, dtype=np.float64), np.asarray(dq_min, dtype=np.float64)
        # Eigen isApprox(prec=1e-12): |a-b|^2 <= prec^2 * min(|a|^2, |b|^2)
        if np.sum((a - b) ** 2) <= 1e-24 * min(np.sum(a * a), np.sum(b * b)):
            w[dof:2 * dof] = 0
    if time:
        w[-1] = 0
    return w


class ResultBuffers:
    """numpy outputs + the ctypes ilqrb_result pointing at them"""

    def __init__(self, pb: ProblemBatch, nb_iter, solver="recursive", nc=0, pinned=False, gains=True):
        B, H, n, m, nt = pb.batch, pb.horizon, pb.n, pb.m, pb.nt
        self.res = cabi.Result()
        nan = np.nan

        def alloc(shape, dtype=np.float64):
            if pinned:
                import torch
                t = torch.empty(shape, dtype=torch.float64 if dtype == np.float64 else torch.int32).pin_memory()
                a = t.numpy()
                self._pins = getattr(self, "_pins", []) + [t]
            else:
                a = np.empty(shape, dtype=dtype)
            a[...] = nan if dtype == np.float64 else 0
            return a

        self.cost = alloc((B,))
        self.cost_hist = alloc((B, max(nb_iter, 1)))
        self.alpha_hist = alloc((B, max(nb_iter, 1)))
        self.iters = alloc((B,), np.int32)
        self.status = alloc((B,), np.int32)
        self.res.cost = cabi.dptr(self.cost)
        self.res.cost_hist = cabi.dptr(self.cost_hist)
        self.res.alpha_hist = cabi.dptr(self.alpha_hist)
        self.res.iters = cabi.iptr(self.iters)
        self.res.status = cabi.iptr(self.status)
        if solver in ("recursive", "al"):
            self.X = alloc((B, H, n))
            self.fX = alloc((B, H, nt))
            self.U = alloc((B, H - 1, m))
            self.res.X, self.res.fX, self.res.U = cabi.dptr(self.X), cabi.dptr(self.fX), cabi.dptr(self.U)
        if solver == "recursive" and gains:  # gains=False: out->K = out->d = NULL, the caller does not want Ks / ds back
            self.K = alloc((B, H - 1, m, n))
            self.d = alloc((B, H - 1, m))
            self.res.K, self.res.d = cabi.dptr(self.K), cabi.dptr(self.d)
        if solver == "al":
            self.lam = alloc((B, H - 1, nc))
            self.res.lambda_ = cabi.dptr(self.lam)
        if solver == "batch":
            self.U = alloc((B, (H - 1) * m))
            self.res.U = cabi.dptr(self.U)

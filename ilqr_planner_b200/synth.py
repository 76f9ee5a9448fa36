"""Deterministic synthetic problem batches for the BASELINE.json configs.

Host-side input generation only (numpy): the targets are made reachable by construction with a small numpy FK;
nothing here is on the solve path.  RNG: numpy.random.default_rng(20261104 + 1000*config + problem_index).
"""
import math
import os

import numpy as np

from . import cabi
from .problem import ProblemBatch, RobotSpec, TaskSpec

_HERE = os.path.dirname(os.path.abspath(__file__))
PANDA_URDF = os.path.join(_HERE, "data", "panda_arm.urdf")
PANDA_Q0 = np.array([0.62991112, -0.2329776, -0.01423721, -1.70254115, 0.06251303, 1.50592777, 0.71771416])
PANDA_LOWER = np.array([-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973])
PANDA_UPPER = np.array([2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973])
SEED0 = 20261104

_chain_cache = {}


def panda_chain():
    """segments of panda_link0 -> panda_tip from the library's own URDF builder (ilqrb_chain_from_urdf)"""
    if "panda" not in _chain_cache:
        from .solver import chain_from_urdf
        segs, dof = chain_from_urdf(PANDA_URDF, "panda_link0", "panda_tip", is_path=True)
        assert dof == 7
        _chain_cache["panda"] = segs
    return _chain_cache["panda"]


def _rodrigues(axis, ang):
    c, s = np.cos(ang), np.sin(ang)
    x, y, z = axis
    K = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
    return c[..., None, None] * np.eye(3) + s[..., None, None] * K + (1 - c)[..., None, None] * np.outer(axis, axis)


def fk_pose(segs, q):
    """numpy FK for target generation: q [B][dof] -> p [B][3], quat [B][4] (w,x,y,z; sign-free construction)"""
    q = np.asarray(q, dtype=np.float64)
    B = q.shape[0]
    R = np.tile(np.eye(3), (B, 1, 1))
    p = np.zeros((B, 3))
    j = 0
    for s in segs:
        tipR = np.asarray(s["tip_R"], dtype=np.float64).reshape(3, 3)
        tipp = np.asarray(s["tip_p"], dtype=np.float64)
        org = np.asarray(s["origin"], dtype=np.float64)
        if s["joint_type"] == cabi.JOINT_ROT:
            Rj = _rodrigues(np.asarray(s["axis"], dtype=np.float64), q[:, j])
            pj = np.tile(org, (B, 1))
            j += 1
        elif s["joint_type"] == cabi.JOINT_TRANS:
            Rj = np.tile(np.eye(3), (B, 1, 1))
            pj = org + q[:, j:j + 1] * np.asarray(s["axis"], dtype=np.float64)
            j += 1
        else:
            Rj = np.tile(np.eye(3), (B, 1, 1))
            pj = np.zeros((B, 3))
        ps = np.einsum("bij,j->bi", Rj, tipp) + pj
        p = np.einsum("bij,bj->bi", R, ps) + p
        R = R @ (Rj @ tipR)
    # quaternion (w >= 0 branch-free via eigen-like formula on the largest component)
    quat = np.zeros((B, 4))
    for b in range(B):
        M = R[b]
        tr = np.trace(M)
        if tr > 0:
            s4 = math.sqrt(tr + 1.0) * 2
            quat[b] = [0.25 * s4, (M[2, 1] - M[1, 2]) / s4, (M[0, 2] - M[2, 0]) / s4, (M[1, 0] - M[0, 1]) / s4]
        else:
            i = int(np.argmax(np.diag(M)))
            jx, kx = (i + 1) % 3, (i + 2) % 3
            s4 = math.sqrt(1.0 + M[i, i] - M[jx, jx] - M[kx, kx]) * 2
            v = np.zeros(3)
            v[i] = 0.25 * s4
            v[jx] = (M[jx, i] + M[i, jx]) / s4
            v[kx] = (M[kx, i] + M[i, kx]) / s4
            quat[b] = [(M[kx, jx] - M[jx, kx]) / s4, v[0], v[1], v[2]]
    return p, quat


def _draws(config, lo, hi, spec):
    """per-problem deterministic uniform draws; spec = list of (name, low, high, size)"""
    out = {name: np.zeros((hi - lo, size)) for name, _, _, size in spec}
    for b in range(lo, hi):
        rng = np.random.default_rng(SEED0 + 1000 * config + b)
        for name, a, c, size in spec:
            out[name][b - lo] = rng.uniform(a, c, size)
    return out


def _posorn_targets(segs, q0, dq_list):
    tg = np.zeros((q0.shape[0], len(dq_list), cabi.KP_STRIDE))
    for i, dq in enumerate(dq_list):
        p, quat = fk_pose(segs, q0 + dq)
        tg[:, i, 0:3] = p
        tg[:, i, 3:7] = quat
    return tg


def config2(batch=4096, horizon=100, dt=0.1, offset=0):
    """C2: 7-DoF chain, PosOrn, nb_deriv=1, limits +-10 pi (inactive), keypoints at H/2-1 and H-1"""
    segs = panda_chain()
    d = _draws(2, offset, offset + batch, [("q0", -0.5, 0.5, 7), ("t1", -0.6, 0.6, 7), ("t2", -0.6, 0.6, 7)])
    q0 = PANDA_Q0 + d["q0"]
    tg = _posorn_targets(segs, q0, [d["t1"], d["t2"]])
    Q = np.diag([1, 1, 1, .1, .1, .1])
    R = np.ones(7) * 1e-5
    lim = np.ones(7) * np.pi * 10
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=lim,
                    state_min=-lim, limit_weight=np.ones(7, dtype=np.int32))
    return ProblemBatch(cabi.SYS_POS_ORN, 1, horizon, dt, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [task], R, q0, np.zeros_like(q0))


def config3(batch=16384, horizon=100, dt=0.05, offset=0):
    """C3: same chain, nb_deriv=2; System limits kept inactive (+-10 pi), AL rows q <= qmax, -q <= -qmin (see al_config3)"""
    segs = panda_chain()
    d = _draws(3, offset, offset + batch, [("q0", -0.5, 0.5, 7), ("t1", -0.6, 0.6, 7), ("t2", -0.6, 0.6, 7)])
    q0 = np.clip(PANDA_Q0 + d["q0"], PANDA_LOWER + 0.05, PANDA_UPPER - 0.05)
    tg = _posorn_targets(segs, q0, [d["t1"], d["t2"]])
    Q = np.diag([1, 1, 1, .1, .1, .1, 1, 1, 1, .1, .1, .1])
    R = np.ones(7) * 1e-5
    smax = np.concatenate([np.ones(7) * np.pi * 10, np.ones(7) * 10])
    w = np.ones(14, dtype=np.int32)
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax,
                    state_min=-smax, limit_weight=w)
    return ProblemBatch(cabi.SYS_POS_ORN, 2, horizon, dt, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [task], R, q0, np.zeros_like(q0))


def al_config3():
    """AL constraint of C3: A = [[I7,0,0],[-I7,0,0]] (14 x 21), b = [qmax; -qmin], lambda0 = 0, lag 5, penalty .25, scaling 1.1"""
    A = np.zeros((14, 21))
    A[:7, :7] = np.eye(7)
    A[7:, :7] = -np.eye(7)
    b = np.concatenate([PANDA_UPPER, -PANDA_LOWER])
    return dict(A=A, b=b, lambda0=np.zeros(14), lag_update_step=5, penalty=0.25, scaling_factor=1.1)


def config4(batch=8192, horizon=200, offset=0):
    """C4: PosOrnTime nb_deriv=1, keypoint times 2 s and 5 s (Batch solvers)"""
    segs = panda_chain()
    d = _draws(4, offset, offset + batch, [("q0", -0.5, 0.5, 7), ("t1", -0.6, 0.6, 7), ("t2", -0.6, 0.6, 7)])
    q0 = PANDA_Q0 + d["q0"]
    tg = _posorn_targets(segs, q0, [d["t1"], d["t2"]])
    tg[:, 0, 14] = 2.0
    tg[:, 1, 14] = 5.0
    Q = np.diag([1, 1, 1, .1, .1, .1, .1])
    R = np.ones(8) * 1e-5
    smax = np.concatenate([np.ones(7) * np.pi * 10, [0.0]])
    w = np.array([1] * 7 + [0], dtype=np.int32)
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax,
                    state_min=-smax, limit_weight=w)
    return ProblemBatch(cabi.SYS_POS_ORN_TIME, 1, horizon, 0.0, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [task], R, q0, np.zeros_like(q0))


def config4_short(batch=16, horizon=60, dof=5, offset=0, tight_limits=True, two_tasks=False):
    """parity-test shape of the time-system fast path: the first `dof` joints of the chain, PosOrnTime nb_deriv=1, soft joint limits
    tight enough to be violated along the way; two_tasks: a second sub-system (its own keypoint, limits of its own) on the same chain"""
    segs = _truncated_chain(dof)
    d = _draws(11, offset, offset + batch, [("q0", -0.4, 0.4, dof), ("t1", -0.6, 0.6, dof), ("t2", -0.6, 0.6, dof), ("t3", -0.5, 0.5, dof)])
    q0 = PANDA_Q0[:dof] + d["q0"]
    tg = _posorn_targets(segs, q0, [d["t1"], d["t2"]])
    tg[:, 0, 14] = 1.5
    tg[:, 1, 14] = 3.5
    Q = np.diag([1, 1, 1, .1, .1, .1, .1])
    R = np.ones(dof + 1) * 1e-5
    span = 0.3 if tight_limits else 30.0
    smax = np.concatenate([PANDA_Q0[:dof] + 0.4 + span, [0.0]])
    smin = np.concatenate([PANDA_Q0[:dof] - 0.4 - span, [0.0]])
    w = np.array([1] * dof + [0], dtype=np.int32)
    tasks = [TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax, state_min=smin,
                      limit_weight=w)]
    if two_tasks:
        tg2 = _posorn_targets(segs, q0, [d["t3"]])
        tg2[:, 0, 14] = 0.8
        tasks.append(TaskSpec(kp_timestep=[horizon // 4], kp_target=tg2, kp_precision=np.array([[Q]]), r_diag=R, state_max=smax - 0.3, state_min=smin + 0.3,
                              limit_weight=w))
    return ProblemBatch(cabi.SYS_POS_ORN_TIME, 1, horizon, 0.0, dof, [RobotSpec(cabi.ROBOT_KDL, dof, segs)], tasks, R, q0, np.zeros_like(q0))


def u0_time(pb):
    """time systems start from s = 0.01 (POS_ORN_TIME_SYS cell 8)"""
    u = np.zeros((pb.batch, pb.horizon - 1, pb.m))
    u[:, :, -1] = 0.01
    return u


def config5(batch=65536, horizon=100, dt=0.1, offset=0):
    """C5: dual-arm composite (2 x 7-DoF chains over one 14-joint state), two PosOrn tasks, n = m = 14"""
    segs = panda_chain()
    d = _draws(5, offset, offset + batch, [("q0", -0.5, 0.5, 14), ("t1", -0.6, 0.6, 14)])
    q0 = np.tile(PANDA_Q0, 2) + d["q0"]
    Q = np.diag([1, 1, 1, .1, .1, .1])
    R = np.ones(14) * 1e-5
    tasks = []
    for a in range(2):
        sl = slice(7 * a, 7 * a + 7)
        # arm 0 has its keypoint mid-horizon, arm 1 at the end (distinct timesteps)
        tg = _posorn_targets(segs, q0[:, sl], [d["t1"][:, sl]])
        tasks.append(TaskSpec(kp_timestep=[horizon // 2 - 1 if a == 0 else horizon - 1], kp_target=tg, kp_precision=np.array([[Q]]), r_diag=R,
                              robot=a, joint_offset=7 * a))
    robots = [RobotSpec(cabi.ROBOT_KDL, 7, segs), RobotSpec(cabi.ROBOT_KDL, 7, segs)]
    return ProblemBatch(cabi.SYS_POS_ORN, 1, horizon, dt, 14, robots, tasks, R, q0, np.zeros_like(q0))


def _truncated_chain(dof):
    segs, nj = [], 0
    for sg in panda_chain():
        segs.append(sg)
        nj += int(sg["joint_type"] != cabi.JOINT_FIXED)
        if nj == dof:
            break
    return segs


def config5_uneven(batch=16, horizon=60, dt=0.1, offset=0, dofs=(5, 4), limits=True):
    """parity-test shape: composite of two truncated chains of different size (odd block offset, runtime block sizes of
    the block-diagonal backward pass), soft joint limits tight enough to be violated on the first arm"""
    n = sum(dofs)
    d = _draws(9, offset, offset + batch, [("q0", -0.4, 0.4, n), ("t1", -0.6, 0.6, n)])
    q0 = np.concatenate([PANDA_Q0[:k] for k in dofs]) + d["q0"]
    Q = np.diag([1, 1, 1, .1, .1, .1])
    R = np.ones(n) * 1e-5
    tasks, robots, at = [], [], 0
    for a, k in enumerate(dofs):
        segs = _truncated_chain(k)
        sl = slice(at, at + k)
        tg = _posorn_targets(segs, q0[:, sl], [d["t1"][:, sl]])
        kw = {}
        if limits and a == 0:
            center = np.concatenate([PANDA_Q0[:j] for j in dofs])
            kw = dict(state_max=center + 0.55, state_min=center - 0.55, limit_weight=np.ones(n, dtype=np.int32))
        tasks.append(TaskSpec(kp_timestep=[horizon // 2 - 1 if a == 0 else horizon - 1], kp_target=tg, kp_precision=np.array([[Q]]), r_diag=R,
                              robot=a, joint_offset=at, **kw))
        robots.append(RobotSpec(cabi.ROBOT_KDL, k, segs))
        at += k
    return ProblemBatch(cabi.SYS_POS_ORN, 1, horizon, dt, n, robots, tasks, R, q0, np.zeros_like(q0))


def config1(batch=1, horizon=50, dt=0.1, offset=0):
    """C1: Robot2D 3-link, position task, single integrator"""
    lengths = np.array([0.4, 0.3, 0.2])
    d = _draws(1, offset, offset + batch, [("q0", 0.0, np.pi, 3), ("t", -0.5, 0.5, 3)])
    q0 = d["q0"]
    qt = q0 + d["t"]
    tg = np.zeros((batch, 1, cabi.KP_STRIDE))
    tg[:, 0, 0] = (lengths * np.cos(qt)).sum(1)
    tg[:, 0, 1] = (lengths * np.sin(qt)).sum(1)
    Q = np.eye(2)
    R = np.ones(3) * 1e-5
    task = TaskSpec(kp_timestep=[horizon - 1], kp_target=tg, kp_precision=np.array([[Q]]), r_diag=R)
    return ProblemBatch(cabi.SYS_POSITION, 1, horizon, dt, 3, [RobotSpec(cabi.ROBOT_2D, 3, None, lengths)], [task], R, q0, np.zeros_like(q0))


def config3_short(batch=16, horizon=60, dt=0.05, dof=5, offset=0, tight_limits=True):
    """parity-test shape: the first `dof` joints of the chain, nb_deriv=2, soft joint limits tight enough to be violated
    along the way (exercises the runtime-size double-integrator kernels and the exact limit path of the line search)"""
    full = panda_chain()
    segs, nj = [], 0
    for sg in full:
        segs.append(sg)
        nj += int(sg["joint_type"] != cabi.JOINT_FIXED)
        if nj == dof:
            break
    d = _draws(7, offset, offset + batch, [("q0", -0.4, 0.4, dof), ("t1", -0.6, 0.6, dof), ("t2", -0.6, 0.6, dof)])
    q0 = PANDA_Q0[:dof] + d["q0"]
    tg = _posorn_targets(segs, q0, [d["t1"], d["t2"]])
    Q = np.diag([1, 1, 1, .1, .1, .1, 1, 1, 1, .1, .1, .1])
    R = np.ones(dof) * 1e-4
    span = 0.35 if tight_limits else 30.0
    smax = np.concatenate([PANDA_Q0[:dof] + 0.4 + span, np.ones(dof) * (0.6 if tight_limits else 50.0)])
    smin = np.concatenate([PANDA_Q0[:dof] - 0.4 - span, -np.ones(dof) * (0.6 if tight_limits else 50.0)])
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax,
                    state_min=smin, limit_weight=np.ones(2 * dof, dtype=np.int32))
    return ProblemBatch(cabi.SYS_POS_ORN, 2, horizon, dt, dof, [RobotSpec(cabi.ROBOT_KDL, dof, segs)], [task], R, q0, np.zeros_like(q0))


def config3_active(batch=64, horizon=100, dt=0.05, offset=0):
    """C3 shape (7-DoF, nb_deriv=2) whose AL constraints bind: starts within 0.1 rad of the tutorial posture, targets 0.4-0.6 rad
    away per joint, so the joint box / velocity / acceleration rows of `al_config3_active` are active over most of the solve."""
    segs = panda_chain()
    d = _draws(8, offset, offset + batch, [("q0", -0.1, 0.1, 7), ("t1", 0.4, 0.6, 7), ("s1", -1.0, 1.0, 7), ("t2", 0.4, 0.6, 7), ("s2", -1.0, 1.0, 7)])
    q0 = PANDA_Q0 + d["q0"]
    tg = _posorn_targets(segs, q0, [np.sign(d["s1"]) * d["t1"], np.sign(d["s2"]) * d["t2"]])
    Q = np.diag([1, 1, 1, .1, .1, .1, 1, 1, 1, .1, .1, .1])
    R = np.ones(7) * 1e-5
    smax = np.concatenate([np.ones(7) * np.pi * 10, np.ones(7) * 10])  # System limits inactive: the AL rows are what is exercised
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax,
                    state_min=-smax, limit_weight=np.ones(14, dtype=np.int32))
    return ProblemBatch(cabi.SYS_POS_ORN, 2, horizon, dt, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [task], R, q0, np.zeros_like(q0))


def al_config3_active(kind="qbox", horizon=100, lag_update_step=3):
    """AL rows for `config3_active` on s = [q(7); v(7); u(7)]:
      "qbox"   14 rows  |q - q_tutorial| <= 0.3                                   (shared by all steps; nc <= 16)
      "mixed"  29 rows  the box, v <= 0.25 (7 rows), |u| <= 0.3 on joints 0..3 (8 rows): rows on the state, the velocity and the control
      "steps"  16 rows per step: a box shrinking from 0.45 to 0.25 over the horizon and u_0 <= 0.3, -u_0 <= 0.3 (per_step = 1)"""
    def box(w):
        A = np.zeros((14, 21))
        A[:7, :7] = np.eye(7)
        A[7:, :7] = -np.eye(7)
        return A, np.concatenate([PANDA_Q0 + w, -(PANDA_Q0 - w)])
    if kind == "qbox":
        A, b = box(0.3)
        lam0 = np.zeros(14)
    elif kind == "mixed":
        Ab, bb = box(0.3)
        Av = np.zeros((7, 21))
        Av[:, 7:14] = np.eye(7)
        Au = np.zeros((8, 21))
        Au[:4, 14:18] = np.eye(4)
        Au[4:, 14:18] = -np.eye(4)
        A = np.concatenate([Ab, Av, Au])
        b = np.concatenate([bb, np.ones(7) * 0.25, np.ones(8) * 0.3])
        lam0 = np.zeros(29)
        lam0[3] = 0.01  # a multiplier that starts positive (AL-ILQR.cpp:27: such a row is active whatever its constraint value)
    elif kind == "steps":
        As, bs = [], []
        for k in range(horizon - 1):
            Ab, bb = box(0.45 - 0.2 * k / (horizon - 2))
            Au = np.zeros((2, 21))
            Au[0, 14] = 1.0
            Au[1, 14] = -1.0
            As.append(np.concatenate([Ab, Au]))
            bs.append(np.concatenate([bb, [0.3, 0.3]]))
        A, b = np.array(As), np.array(bs)
        lam0 = np.zeros((horizon - 1, 16))
    else:
        raise ValueError(kind)
    return dict(A=A, b=b, lambda0=lam0, lag_update_step=lag_update_step, penalty=0.25, scaling_factor=1.1)


# ---- joint-space systems, angular keypoints, dead-zone keypoints (the HYBRID_* / JOINT_SPACE_* notebook structures;
# the notebooks draw their joint targets with an unseeded np.random.uniform, so their stored traces cannot be reproduced) -------
def _joint_targets(q0, dq_list, times=None):
    tg = np.zeros((q0.shape[0], len(dq_list), cabi.KP_STRIDE))
    for i, dq in enumerate(dq_list):
        tg[:, i, 0:q0.shape[1]] = q0 + dq
        if times is not None:
            tg[:, i, 14] = times[i]
    return tg


def config_joint_space(batch=16, horizon=100, dt=0.1, offset=0, time=False):
    """JOINT_SPACE_SYS(_TIME): JointSpace(Time)PlannerSys, two Angular(Time)Keypoints at H/2-1 and H-1, Q = I (time entry .1),
    R = 1e-5, Panda joint limits as soft limits"""
    segs = panda_chain()
    d = _draws(10 + int(time), offset, offset + batch, [("q0", -0.3, 0.3, 7), ("t1", -0.8, 0.8, 7), ("t2", -0.8, 0.8, 7)])
    q0 = np.clip(PANDA_Q0 + d["q0"], PANDA_LOWER + 0.1, PANDA_UPPER - 0.1)
    tg = _joint_targets(q0, [d["t1"], d["t2"]], times=(2.5, 5.0) if time else None)
    nq = 7 + int(time)
    Q = np.identity(nq)
    if time:
        Q[-1, -1] = .1
    R = np.ones(nq) * 1e-5
    smax = np.concatenate([PANDA_UPPER, [0.0]]) if time else PANDA_UPPER
    smin = np.concatenate([PANDA_LOWER, [0.0]]) if time else PANDA_LOWER
    w = np.array([1] * 7 + ([0] if time else []), dtype=np.int32)
    task = TaskSpec(kp_timestep=[horizon // 2 - 1, horizon - 1], kp_target=tg, kp_precision=np.array([[Q, Q]]), r_diag=R, state_max=smax, state_min=smin,
                    limit_weight=w, task_kind=cabi.TASK_JOINT)
    return ProblemBatch(cabi.SYS_POS_ORN_TIME if time else cabi.SYS_POS_ORN, 1, horizon, 0.0 if time else dt, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [task], R, q0,
                        np.zeros_like(q0))


def config_hybrid(batch=16, horizon=120, dt=0.01, offset=0, time=False):
    """HYBRID_SYS(_TIME): SequentialSystem(JointSpace(Time)PlannerSys with one angular keypoint at H/2-1, PosOrn(Time)PlannerSys with one
    pose keypoint at H-1), sub-system R = 1e-3 (time: 1e-5), top-level R = 1e-6"""
    segs = panda_chain()
    d = _draws(12 + int(time), offset, offset + batch, [("q0", -0.3, 0.3, 7), ("t1", -0.5, 0.5, 7), ("t2", -0.5, 0.5, 7)])
    q0 = np.clip(PANDA_Q0 + d["q0"], PANDA_LOWER + 0.1, PANDA_UPPER - 0.1)
    nq = 7 + int(time)
    tgj = _joint_targets(q0, [d["t1"]], times=(2.5,) if time else None)
    tgp = _posorn_targets(segs, q0, [d["t2"]])
    Qj = np.identity(nq)
    Qp = np.diag([1, 1, 1, .1, .1, .1] + ([0.0] if time else []))
    if time:
        Qj[-1, -1] = 0.0
        tgp[:, 0, 14] = 5.0
    Rsub = np.ones(nq) * (1e-5 if time else 1e-3)
    smax = np.concatenate([PANDA_UPPER, [0.0]]) if time else PANDA_UPPER
    smin = np.concatenate([PANDA_LOWER, [0.0]]) if time else PANDA_LOWER
    w = np.array([1] * 7 + ([0] if time else []), dtype=np.int32)
    t1 = TaskSpec(kp_timestep=[horizon // 2 - 1], kp_target=tgj, kp_precision=np.array([[Qj]]), r_diag=Rsub, state_max=smax, state_min=smin, limit_weight=w,
                  task_kind=cabi.TASK_JOINT)
    t2 = TaskSpec(kp_timestep=[horizon - 1], kp_target=tgp, kp_precision=np.array([[Qp]]), r_diag=Rsub, state_max=smax, state_min=smin, limit_weight=w)
    R = np.ones(nq) * 1e-6
    return ProblemBatch(cabi.SYS_POS_ORN_TIME if time else cabi.SYS_POS_ORN, 1, horizon, 0.0 if time else dt, 7, [RobotSpec(cabi.ROBOT_KDL, 7, segs)], [t1, t2], R, q0,
                        np.zeros_like(q0))


def config2_deadzone(batch=16, horizon=100, dt=0.1, offset=0, nb_deriv=1):
    """C2 shape with PosOrnKeypointDistFunct keypoints: the first keypoint has a 5 cm / (0.05, 0.1, 2.0) rad dead zone, the second is a
    plain PosOrnKeypoint (negative radius)"""
    pb = config2(batch=batch, horizon=horizon, dt=dt, offset=offset) if nb_deriv == 1 else config3(batch=batch, horizon=horizon, dt=dt, offset=offset)
    pb.tasks[0].kp_deadzone = np.array([[0.05, 0.05, 0.1, 2.0], [-1.0, 0.0, 0.0, 0.0]])
    return pb

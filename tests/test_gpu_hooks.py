"""SURVEY 8f N2 on the GPU: the single-point hooks behind the reference's debugging API — System::getFxJac(xk), System::diff,
System::cost / cost_x / cost_xx (cost_F* = the same with u = 0 at k = H - 1), Keypoint::diff and utils::Sd — against the oracle's
restatement of the same functions at 1e-12 (element-wise, relative to the largest entry of the quantity)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _close(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    scale = max(1.0, float(np.max(np.abs(b))) if b.size else 1.0)
    err = float(np.max(np.abs(a - b))) / scale if a.size else 0.0
    assert err <= TOL, "%s: %.3e > %.0e" % (what, err, TOL)


CASES = {
    "c2": lambda s: s.config2(batch=4, horizon=40),
    "c3_limits_2nd_order": lambda s: s.config3_short(batch=4),
    "c4_time": lambda s: s.config4(batch=4, horizon=40),
    "c5_dual_arm": lambda s: s.config5(batch=4, horizon=40),
    "c1_robot2d": lambda s: s.config1(batch=2),
    "joint_space": lambda s: s.config_joint_space(batch=4, horizon=40),
    "joint_space_time": lambda s: s.config_joint_space(batch=4, horizon=40, time=True),
    "hybrid": lambda s: s.config_hybrid(batch=4, horizon=40),
    "deadzone": lambda s: s.config2_deadzone(batch=4, horizon=40),
}


def _points(pb, rng, nq):
    """states around the initial one (some beyond the joint limits), controls, every keypoint timestep and some others"""
    q0 = np.asarray(pb.q0, dtype=np.float64).reshape(pb.batch, -1)
    x = np.zeros((nq, pb.n))
    prob = rng.integers(0, pb.batch, size=nq).astype(np.int32)
    x[:, :q0.shape[1]] = q0[prob] + rng.uniform(-0.8, 0.8, size=(nq, q0.shape[1]))
    x[:, q0.shape[1]:] = rng.uniform(-1.5, 1.5, size=(nq, pb.n - q0.shape[1]))
    if pb.time:
        x[:, -1] = rng.uniform(0, 3, size=nq)
    u = rng.uniform(-1, 1, size=(nq, pb.m))
    ks = sorted({int(t) for tk in pb.tasks for t in np.asarray(tk.kp_timestep).ravel()})
    k = np.array([ks[i % len(ks)] if i % 3 else int(rng.integers(0, pb.horizon)) for i in range(nq)], dtype=np.int32)
    return x, u, k, prob


@pytest.mark.parametrize("case", list(CASES))
def test_eval_point_vs_oracle(oracle_lib, case):
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle
    from oracle import oracle
    pb = CASES[case](synth)
    rng = np.random.default_rng(11)
    x, u, k, prob = _points(pb, rng, 24)
    with Handle(pb) as h:
        g = h.eval_point(x, k=k, u=u, problem=prob)
        gF = h.eval_point(x, k=pb.horizon - 1, problem=prob)  # cost_F, cost_F_x, cost_F_xx
        fx_in = g["fx"] + rng.uniform(-0.05, 0.05, size=g["fx"].shape)
        gd = h.eval_point(x, k=k, problem=prob, fx_in=fx_in)  # System::diff on a caller's task vector
    o = oracle.eval_point(pb, x, k=k, u=u, problem=prob)
    oF = oracle.eval_point(pb, x, k=pb.horizon - 1, problem=prob)
    od = oracle.eval_point(pb, x, k=k, problem=prob, fx_in=fx_in)
    for key in ("fx", "J", "diff", "cost", "cost_x", "cost_xx"):
        _close(g[key], o[key], case + " " + key)
        _close(gF[key], oF[key], case + " final " + key)
    _close(gd["diff"], od["diff"], case + " diff(fx_in)")
    assert np.any(g["cost"] != 0) and np.any(g["diff"] != 0)


def test_eval_point_rejects_bad_indices():
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle
    pb = synth.config2(batch=2, horizon=20)
    with Handle(pb) as h:
        with pytest.raises(RuntimeError, match="outside the batch"):
            h.eval_point(np.zeros((1, pb.n)), k=0, problem=5)
        with pytest.raises(RuntimeError, match="outside the horizon"):
            h.eval_point(np.zeros((1, pb.n)), k=20)


def test_keypoint_diff_without_a_system(oracle_lib):
    """Keypoint::diff of each keypoint class from its record alone == System::diff of a system holding that keypoint"""
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import keypoint_diff
    from oracle import oracle
    rng = np.random.default_rng(5)
    for case, sys_kind, task_kind in (("c2", cabi.SYS_POS_ORN, cabi.TASK_POSE), ("c3_limits_2nd_order", cabi.SYS_POS_ORN, cabi.TASK_POSE),
                                      ("c4_time", cabi.SYS_POS_ORN_TIME, cabi.TASK_POSE), ("joint_space", cabi.SYS_POS_ORN, cabi.TASK_JOINT),
                                      ("joint_space_time", cabi.SYS_POS_ORN_TIME, cabi.TASK_JOINT), ("deadzone", cabi.SYS_POS_ORN, cabi.TASK_POSE),
                                      ("c1_robot2d", cabi.SYS_POSITION, cabi.TASK_POSE)):
        pb = CASES[case](synth)
        tk = pb.tasks[0]
        kp = 0
        k = int(np.asarray(tk.kp_timestep).ravel()[kp])
        tg = np.asarray(tk.kp_target, dtype=np.float64).reshape(pb.batch, -1, cabi.KP_STRIDE)[0, kp]
        dz = None if getattr(tk, "kp_deadzone", None) is None else np.asarray(tk.kp_deadzone, dtype=np.float64).reshape(-1, 4)[kp]
        nq = 16
        fx = rng.uniform(-1, 1, size=(nq, pb.nt))
        if task_kind == cabi.TASK_POSE and sys_kind != cabi.SYS_POSITION:
            nd = pb.nb_deriv
            fx[:, 3:7] /= np.linalg.norm(fx[:, 3:7], axis=1, keepdims=True)
            if case == "deadzone":
                fx[:8, :3] = tg[:3] + rng.uniform(-0.02, 0.02, size=(8, 3))  # inside and around the ball
            assert nd in (1, 2)
        fx[-1] = 0.0  # the all-zero task vector short-circuits (PosOrnKeypoint.cpp:27)
        dof = pb.robots[tk.robot].dof if task_kind == cabi.TASK_JOINT else (pb.nt // pb.nb_deriv if sys_kind == cabi.SYS_POSITION else 0)
        got = keypoint_diff(sys_kind, task_kind, pb.nb_deriv, dof, tg, fx, deadzone=dz)
        want = oracle.eval_point(pb, np.zeros((nq, pb.n)), k=k, fx_in=fx)["diff"][:, :got.shape[1]]
        _close(got, want, case + " Keypoint::diff")


def test_sd_functions_vs_oracle(oracle_lib):
    from ilqr_planner_b200 import cabi
    from ilqr_planner_b200.solver import sd
    from oracle import oracle
    rng = np.random.default_rng(3)
    nq = 64
    unit = lambda v: v / np.linalg.norm(v, axis=1, keepdims=True)
    a, b, c = unit(rng.normal(size=(nq, 4))), unit(rng.normal(size=(nq, 4))), unit(rng.normal(size=(nq, 4)))
    # edge cases of sd.h: identical points, antipodal points, a zero operand, unnormalised operands
    b[0] = a[0]
    b[1] = -a[1]
    b[2] = 0.0
    a[3] *= 2.5
    tang = rng.normal(size=(nq, 4)) * 0.7
    tang[4] = 0.0  # expMap of the zero tangent vector returns the base
    for op, args in ((cabi.SD_LOGMAP, (a, b)), (cabi.SD_EXPMAP, (a, tang)), (cabi.SD_DISTANCE, (a, b)), (cabi.SD_TRANSPORT, (tang, a, c)),
                     (cabi.SD_TRANSPORT, (tang, a, b)), (cabi.SD_DQUAT_TO_W_JAC, (a,))):
        _close(sd(op, *args), oracle.sd(op, *args), "Sd op %d" % op)
    # round trip: expMap(base, logMap(base, y)) == y on the hemisphere where distance() is the plain arc cosine
    y = np.where((np.sum(a * c, axis=1) < 0)[:, None], -c, c)
    base = unit(a)
    back = sd(cabi.SD_EXPMAP, base, sd(cabi.SD_LOGMAP, base, y))
    assert float(np.max(np.abs(back - y))) < 1e-9

"""The reference-facing host API: C++ classes of include/ilqr_planner_b200/ and the `PyLQR` module bound over them.

CPU part: headers compile stand-alone, PyLQR imports, constructors / size arithmetic / mu-Q layouts / error messages follow
the reference (System.cpp, SequentialSystem.cpp, PosOrn*PlannerSys.cpp).
GPU part: every notebook golden is rebuilt through the PyLQR API (the code path of the tutorials) and must print the
reference's stored "Iteration i, Cost: c, alpha= a" trace.
"""
import ast
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from . import _cases as tc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
URDF = os.path.join(ROOT, "ilqr_planner_b200", "data", "panda_arm.urdf")
sys.path.insert(0, ROOT)


def _pylqr():
    import PyLQR  # noqa: F401
    return PyLQR


def _keypoint(sp, rec, prec, step):
    from PyLQR.system import PosOrnKeypoint, SpacetimeKeypoint
    rec = np.asarray(rec, dtype=np.float64)
    p, e, dp, de, t = rec[0:3], rec[3:7], rec[7:10], rec[10:14], float(rec[14])
    prec = np.asarray(prec, dtype=np.float64)
    time = sp["sys_kind"] == 1
    if sp["nb_deriv"] == 1:
        return SpacetimeKeypoint(p, e, prec, t, step) if time else PosOrnKeypoint(p, e, prec, step)
    return SpacetimeKeypoint(p, dp, e, de, prec, t, step) if time else PosOrnKeypoint(p, dp, e, de, prec, step)


def system_from_spec(sp):
    """golden system spec -> PyLQR objects, the way the tutorials build them"""
    from PyLQR.sim import KDLRobot, TransformedSimulationInterface
    from PyLQR.system import PosOrnPlannerSys, PosOrnTimePlannerSys, SequentialSystem
    m = re.match(r"(.+?):(.+?)->(.+?):(\[.*?\]):(\[.*?\])$", sp["chain"])
    base, tip, rpy, xyz = m.group(2), m.group(3), ast.literal_eval(m.group(4)), ast.literal_eval(m.group(5))
    dof, nd, H = sp["dof"], sp["nb_deriv"], sp["horizon"]
    rbt = KDLRobot(URDF, base, tip, sp["q0"], sp["dq0"], rpy, xyz)
    subs = []
    for t in sp["tasks"]:
        r = rbt if t["T"] is None else TransformedSimulationInterface(rbt, np.asarray(t["T"], dtype=np.float64).reshape(4, 4))
        kps = [_keypoint(sp, rec, pr, k) for rec, pr, k in zip(t["kp_target"], t["kp_precision"], t["kp_timestep"])]
        args = [r, kps, t["r_diag"][:]]
        if t["limits"] is not None:
            smax, smin = np.asarray(t["limits"]["state_max"]), np.asarray(t["limits"]["state_min"])
            args += [smax[:dof], smin[:dof]]
            if nd == 2:
                args += [smax[dof:2 * dof], smin[dof:2 * dof]]
        if sp["sys_kind"] == 1:
            subs.append(PosOrnTimePlannerSys(*args, H, nd))
        else:
            subs.append(PosOrnPlannerSys(*args, H, nd, sp["dt"]))
    if len(subs) == 1:
        return rbt, subs[0]
    return rbt, SequentialSystem(rbt, subs, sp["r_diag"], H, nd)


def parse_trace(messages):
    out = []
    for msg in messages:
        m = re.match(r"Iteration (\d+), Cost: (\S+), alpha= ([^,\s]+)", msg)
        assert m, msg
        out.append(dict(cost=m.group(2).replace("-nan", "nan"), alpha=m.group(3)))
    return out


# ------------------------------------------------------------------------------------------------------- CPU
def test_headers_compile_standalone(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('#include <ilqr_planner_b200/ilqr_planner.hpp>\nusing namespace ilqr_planner;\n'
                   'int main() { solver::Constraint c; (void)c; std::shared_ptr<sys::System> s; solver::ILQRRecursive r(s); (void)r; return 0; }\n')
    r = subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_csv_helpers_round_trip(tmp_path):
    """EigenSerialize::save / load of utils.cpp:20-60 (4 significant digits, one vector per line, matrices row by row)"""
    src = tmp_path / "t.cpp"
    src.write_text('''#include <ilqr_planner_b200/serialize.hpp>
#include <cstdio>
using namespace ilqr_planner;
int main(int, char** argv) {
    std::string d = argv[1];
    std::vector<VectorXd> xs = {VectorXd({1.0, -2.5, 3.14159265}), VectorXd({1e-7, 12345.678, 0.0})};
    EigenSerialize::save(xs, d + "/v.csv");
    auto back = EigenSerialize::load<3>(d + "/v.csv");
    if (back.size() != 2 || back[0](1) != -2.5 || back[1](0) != 1e-7) return 1;
    MatrixXd A(2, 3);
    for (int i = 0; i < 2; i++) for (int j = 0; j < 3; j++) A(i, j) = i * 3 + j + 0.125;
    EigenSerialize::save(A, d + "/m.csv");
    EigenSerialize::save(std::vector<MatrixXd>{A, A}, d + "/ms.csv");
    EigenSerialize::save(xs[0], d + "/b.csv");
    auto v = eigenMatrixToStdVector(A);
    return (v.size() == 6 && v[1] == 3.125) ? 0 : 2;  // column-major like Eigen
}
''')
    exe = tmp_path / "t"
    r = subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert subprocess.run([str(exe), str(tmp_path)]).returncode == 0
    assert (tmp_path / "v.csv").read_text() == "1,-2.5,3.142\n1e-07,1.235e+04,0\n"
    assert (tmp_path / "m.csv").read_text() == "0.125,1.125,2.125\n3.125,4.125,5.125\n"
    assert (tmp_path / "b.csv").read_text() == "1,-2.5,3.142\n"
    assert (tmp_path / "ms.csv").read_text().count("=================================== \n") == 2


def test_pylqr_module_surface():
    P = _pylqr()
    for mod, names in (("sim", ["SimulationInterface", "KDLRobot", "Robot2D", "TransformedSimulationInterface"]),
                       ("system", ["Keypoint", "PosOrnKeypoint", "SpacetimeKeypoint", "System", "SequentialSystem", "PosOrnPlannerSys", "PosOrnTimePlannerSys",
                                   "PosOrnKeypointDistFunct", "AngularKeypoint", "AngularTimeKeypoint", "JointSpacePlannerSys", "JointSpaceTimePlannerSys"]),
                       ("solver", ["Constraint", "AL_ILQR", "BatchILQR", "BatchILQRCP", "ILQRRecursive"]),
                       ("utils", ["CallBackMessage", "PythonCallbackMessage", "primitives", "Sd"])):
        for n in names:
            assert hasattr(getattr(P, mod), n), "%s.%s" % (mod, n)
    for n in ("logMap", "expMap", "distance", "transport", "dquat_to_w_jac"):
        assert hasattr(P.utils.Sd, n)
    for cls in ("AL_ILQR", "BatchILQR", "BatchILQRCP", "ILQRRecursive"):
        assert hasattr(getattr(P.solver, cls), "solve_batch") and hasattr(getattr(P.solver, cls), "cache_stats")
    for n in ("diff", "diff_batch", "forward_pass_with_limits", "forward_pass_batch", "cost", "cost_x", "cost_xx", "cost_u", "cost_uu", "cost_xu", "cost_ux", "cost_F",
              "cost_F_x", "cost_F_xx", "get_fx_jac"):
        assert hasattr(P.system.System, n), n
    assert hasattr(P.system.Keypoint, "diff")
    from PyLQR.utils import primitives  # the tutorials' import form
    psi = primitives.build_psi_unitstep(99, 2)
    assert psi.shape == (99, 2)
    from .golden import psi_ref
    for kind, fn in (("rbf", primitives.build_psi_RBF), ("bernstein", primitives.build_psi_bernstein), ("unitstep", primitives.build_psi_unitstep),
                     ("sawtooth", primitives.build_psi_sawtooth), ("linear", primitives.build_psi_linear)):
        np.testing.assert_allclose(fn(20, 4), psi_ref.build_psi(kind, 20, 4), rtol=1e-13, atol=1e-15)


def test_sizes_layouts_and_messages_follow_the_reference():
    _pylqr()
    from PyLQR.sim import KDLRobot
    from PyLQR.system import PosOrnKeypoint, PosOrnPlannerSys, SpacetimeKeypoint
    c = tc.golden_case("POS_ORN_MULTI_SYS_2ND/cell23/recursive")
    rbt, s = system_from_spec(c["system"])
    sp = c["system"]
    assert (s.get_nb_state_var(), s.get_nb_ctrl_var(), s.get_nb_target_var(), s.get_horizon()) == (14, 7, 14 * len(sp["tasks"]), sp["horizon"])
    np.testing.assert_array_equal(s.get_init_state(), np.r_[sp["q0"], sp["dq0"]])
    # sparse mu: one slot per keypoint of the merged, time-sorted list; sub-system blocks inside; state order [p, dp, quat, dquat]
    steps = sorted(k for t in sp["tasks"] for k in t["kp_timestep"])
    assert s.get_kp_indexes() == steps
    mu = s.get_mu_vector(True).reshape(len(steps), -1)
    for i, k in enumerate(steps):
        o = 0
        for t in sp["tasks"]:
            blk = mu[i, o:o + 14]
            if k in t["kp_timestep"]:
                rec = np.asarray(t["kp_target"][t["kp_timestep"].index(k)])
                np.testing.assert_array_equal(blk, np.r_[rec[0:3], rec[7:10], rec[3:7], rec[10:14]])
            else:
                assert not blk.any()
            o += 14
    Q = s.get_Q_matrix(True)
    assert Q.shape == (12 * len(sp["tasks"]) * len(steps),) * 2
    assert s.get_Q_matrix(False).shape == (12 * len(sp["tasks"]) * sp["horizon"],) * 2
    # keypoint checks (System.cpp:363-374)
    q0 = sp["q0"]
    r = KDLRobot(URDF, "panda_link0", "panda_tip", q0, [0.0] * 7)
    k1 = PosOrnKeypoint(np.zeros(3), np.array([1.0, 0, 0, 0]), np.eye(6), 10)
    kt = SpacetimeKeypoint(np.zeros(3), np.array([1.0, 0, 0, 0]), np.eye(7), 1.0, 10)
    with pytest.raises(RuntimeError, match=r"Wrong keypoint order \(nb_deriv_\): Expecting 2 got 1"):
        PosOrnPlannerSys(r, [k1], [1e-5] * 7, 50, 2, 0.1)
    with pytest.raises(RuntimeError, match="Wrong keypoint type: got POS_ORN_TIME"):
        PosOrnPlannerSys(r, [kt], [1e-5] * 7, 50, 1, 0.1)
    with pytest.raises(RuntimeError, match=r"\[urdf_parsing\]"):
        KDLRobot("missing.urdf", "a", "b", q0, [0.0] * 7)
    # joint state bookkeeping happens on the host (SimulationInterface.cpp:19-31)
    r.send_vel(0.5, np.ones(7), False)
    np.testing.assert_allclose(r.get_q(), np.asarray(q0) + 0.5)
    assert r.get_time() == 0.5
    r.set_conf(q0, [0.0] * 7, True)
    assert r.get_time() == 0.0


def test_joint_space_classes_follow_the_reference():
    """sizes, state layouts, TAG checks of the joint-space classes (JointSpacePlannerSys.cpp:54-75, JointSpaceTimePlannerSys.cpp:41-59,
    AngularKeypoint.cpp:13-20, AngularTimeKeypoint.cpp:12-17)"""
    _pylqr()
    from PyLQR.sim import KDLRobot
    from PyLQR.system import (AngularKeypoint, AngularTimeKeypoint, JointSpacePlannerSys, JointSpaceTimePlannerSys, PosOrnKeypoint, PosOrnKeypointDistFunct, PosOrnPlannerSys,
                              SequentialSystem)
    q0 = [0.1 * i for i in range(7)]
    r = KDLRobot(URDF, "panda_link0", "panda_tip", q0, [0.0] * 7)
    qa, qb = np.linspace(-1, 1, 7), np.linspace(1, -1, 7)
    ka, kb = AngularKeypoint(qa, np.eye(7), 40), AngularKeypoint(qb, 2 * np.eye(7), 20)
    s = JointSpacePlannerSys(r, [ka, kb], [1e-3] * 7, 50, 1, 0.1)
    assert (s.get_nb_state_var(), s.get_nb_ctrl_var(), s.get_nb_target_var(), s.get_nb_Q_var(), s.get_horizon()) == (7, 7, 7, 7, 50)
    assert s.get_kp_indexes() == [20, 40]
    np.testing.assert_array_equal(ka.get_state(), qa)
    np.testing.assert_array_equal(s.get_mu_vector(True), np.r_[qb, qa])
    fx, J = s.get_fx_jac()
    np.testing.assert_array_equal(fx, q0)
    np.testing.assert_array_equal(J, np.eye(7))
    kt = AngularTimeKeypoint(qa, np.eye(8), 1.5, 30)
    np.testing.assert_array_equal(kt.get_state(), np.r_[qa, 1.5])
    assert kt.get_continuous_time() == 1.5
    st = JointSpaceTimePlannerSys(r, [kt], [1e-3] * 8, 50, 1)
    assert (st.get_nb_state_var(), st.get_nb_ctrl_var(), st.get_nb_target_var(), st.get_nb_Q_var()) == (8, 8, 8, 8)
    np.testing.assert_array_equal(st.get_init_state(), np.r_[q0, 0.0])
    k2 = AngularKeypoint(qa, qb, np.eye(14), 10)
    np.testing.assert_array_equal(k2.get_state(), np.r_[qa, qb])
    with pytest.raises(RuntimeError, match="Wrong keypoint type: got JNT_TIME"):
        JointSpacePlannerSys(r, [kt], [1e-3] * 7, 50, 1, 0.1)
    with pytest.raises(RuntimeError, match="Wrong keypoint type: got POS_ORN"):
        JointSpacePlannerSys(r, [PosOrnKeypoint(np.zeros(3), np.array([1.0, 0, 0, 0]), np.eye(6), 10)], [1e-3] * 7, 50, 1, 0.1)
    with pytest.raises(RuntimeError, match=r"Wrong keypoint order \(nb_deriv_\): Expecting 1 got 2"):
        JointSpacePlannerSys(r, [k2], [1e-3] * 7, 50, 1, 0.1)
    # a dead-zone keypoint is a PosOrnKeypoint for the planner (TAG POS_ORN); a hybrid system stacks both kinds of sub-system
    kd = PosOrnKeypointDistFunct(np.array([0.4, 0.1, 0.5]), np.array([0.0, 1.0, 0, 0]), np.eye(6), 0.05, np.array([0.1, 0.1, 0.1]), 49)
    se = PosOrnPlannerSys(r, [kd], [1e-3] * 7, 50, 1, 0.1)
    hy = SequentialSystem(r, [s, se], [1e-3] * 7, 50, 1)
    assert (hy.get_nb_target_var(), hy.get_nb_Q_var()) == (14, 13)
    with pytest.raises(RuntimeError, match="orn_thresh must have 3 entries"):
        PosOrnKeypointDistFunct(np.zeros(3), np.array([1.0, 0, 0, 0]), np.eye(6), 0.05, np.zeros(2), 10)


# ------------------------------------------------------------------------------------------------------- GPU
GOLD = [c["name"] for c in tc.goldens()["cases"]]


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLD)
def test_tutorial_solve_through_pylqr(name):
    _pylqr()
    from PyLQR.solver import AL_ILQR, BatchILQR, BatchILQRCP, Constraint, ILQRRecursive
    from PyLQR.utils import PythonCallbackMessage
    c = tc.golden_case(name)
    sp = c["system"]
    rbt, s = system_from_spec(sp)
    H, m = sp["horizon"], s.get_nb_ctrl_var()
    u0 = (np.tile(np.asarray(c["u0_row"], dtype=np.float64), (H - 1, 1)) if "u0_row" in c else np.asarray(c["u0_full"], dtype=np.float64).reshape(H - 1, m))
    cb = PythonCallbackMessage()
    exact = True
    if c["solver"] == "recursive":
        X, fX, U, K, d, cost = ILQRRecursive(s).solve(u0, c["nb_iter"], c["line_search"], c["early_stop"], cb)
        assert np.asarray(X).shape == (H, s.get_nb_state_var()) and np.asarray(K).shape == (H - 1, m, s.get_nb_state_var())
    elif c["solver"] == "al":
        A, b, lam = np.asarray(c["A"], dtype=np.float64), np.asarray(c["b"], dtype=np.float64), np.asarray(c["lambda0"], dtype=np.float64)
        cons = []
        for k in range(H - 1):
            ck = Constraint()
            ck.A, ck.b = (A[k], b[k]) if A.ndim == 3 else (A, b)
            cons.append(ck)
        lam0 = [lam[k] if lam.ndim == 2 else lam for k in range(H - 1)]
        X, fX, U = AL_ILQR(s, cons, lam0).solve(u0, c["nb_iter"], c["lag_update_step"], c["penalty"], c["scaling_factor"], c["line_search"], c["early_stop"], cb)
    else:
        exact = False
        if c["solver"] == "batch_cp":
            class _PB:  # case_psi only needs m
                pass
            pb = _PB()
            pb.m = m
            psi = tc.case_psi(c, pb)
            pl = BatchILQRCP(s, psi) if c["Q"] is None else BatchILQRCP(s, np.asarray(c["Q"]), psi)
        else:
            pl = BatchILQR(s) if c["Q"] is None else BatchILQR(s, np.asarray(c["Q"]))
        u = pl.solve(c["nb_iter"], u0.reshape(-1), c["early_stop"], cb)
        assert u.shape == ((H - 1) * m,)
    got, exp = parse_trace(cb.messages), tc.norm_trace(c["trace"])
    assert len(got) == len(exp), "iteration count %d != %d" % (len(got), len(exp))
    for i, (a, b) in enumerate(zip(got, exp)):
        assert a["alpha"] == b["alpha"], "iteration %d alpha %s != %s" % (i + 1, a["alpha"], b["alpha"])
        if exact:
            assert a["cost"] == b["cost"], "iteration %d cost %s != %s" % (i + 1, a["cost"], b["cost"])
        elif a["cost"] != b["cost"]:
            assert abs(float(a["cost"]) - float(b["cost"])) <= 2e-6 * abs(float(b["cost"])), "iteration %d cost %s != %s" % (i + 1, a["cost"], b["cost"])


@pytest.mark.gpu
def test_single_configuration_accessors_run_on_the_device(oracle_lib):
    """the tutorials replay U through rbt.send_vel / get_ee_pos (POS_ORN_SYS cell 14): same numbers as the oracle's FK"""
    _pylqr()
    from oracle import oracle
    from PyLQR.sim import TransformedSimulationInterface
    c = tc.golden_case("POS_ORN_SYS_OBJ_FRAME/cell19/recursive")
    sp = c["system"]
    rbt, s = system_from_spec(sp)
    pb = tc.problem_from_spec(sp, tc.goldens()["chains"])
    U = 0.05 * np.sin(np.arange((sp["horizon"] - 1) * 7, dtype=np.float64)).reshape(1, sp["horizon"] - 1, 7)
    X, fX, J, cost = oracle.rollout(pb, U)
    view = rbt if sp["tasks"][0]["T"] is None else TransformedSimulationInterface(rbt, np.asarray(sp["tasks"][0]["T"]).reshape(4, 4))
    rbt.set_conf(sp["q0"], sp["dq0"], True)
    for k in range(0, 12):
        np.testing.assert_allclose(np.r_[view.get_ee_pos(), view.get_ee_orn()], fX[0, k, :7], rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(view.J(), J[0, k], rtol=1e-11, atol=1e-13)
        fx, Jk = s.get_fx_jac()
        np.testing.assert_allclose(fx, fX[0, k], rtol=1e-12, atol=1e-13)
        rbt.send_vel(sp["dt"], U[0, k], True)
    np.testing.assert_allclose(rbt.get_q(), X[0, 12], rtol=1e-14)


def _close(a, b, what, tol=1e-12):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: shape %s != %s" % (what, a.shape, b.shape)
    scale = max(1.0, float(np.max(np.abs(b))) if b.size else 1.0)
    err = float(np.max(np.abs(a - b))) / scale if a.size else 0.0
    assert err <= tol, "%s: %.3e > %.0e" % (what, err, tol)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["POS_ORN_MULTI_SYS_2ND/cell23/recursive", "POS_ORN_SYS_OBJ_FRAME/cell19/recursive", "POS_ORN_TIME_SYS/cell16/recursive"])
def test_debug_api_through_pylqr(oracle_lib, name):
    """System.diff / diff_batch / cost* / cost_F* / get_fx_jac(xk) / forward_pass_batch / forward_pass_with_limits and Keypoint.diff of the
    reference's debugging API run on the GPU and agree with the oracle's restatement of the same functions at 1e-12"""
    _pylqr()
    from oracle import oracle
    names = [c["name"] for c in tc.goldens()["cases"]]
    if name not in names:
        name = next(n for n in names if n.startswith(name.split("/")[0]))
    c = tc.golden_case(name)
    sp = c["system"]
    rbt, s = system_from_spec(sp)
    pb = tc.problem_from_spec(sp, tc.goldens()["chains"])
    rng = np.random.default_rng(2)
    n, m, H = s.get_nb_state_var(), s.get_nb_ctrl_var(), sp["horizon"]
    ks = sorted(k for t in sp["tasks"] for k in t["kp_timestep"])
    for trial in range(4):
        xk = np.asarray(s.get_init_state()) + rng.uniform(-0.6, 0.6, size=n)
        if sp["sys_kind"] == 1:
            xk[-1] = rng.uniform(0, 2)
        uk = rng.uniform(-1, 1, size=m)
        k = ks[trial % len(ks)] if trial < 3 else (ks[0] + 1) % H
        o = oracle.eval_point(pb, xk, k=k, u=uk)
        fx, J = s.get_fx_jac(xk)
        _close(fx, o["fx"][0], "get_fx_jac fx")
        _close(J, o["J"][0], "get_fx_jac J")
        _close(s.cost(xk, uk, k), o["cost"], "cost")
        _close(s.cost_x(xk, uk, k), o["cost_x"][0], "cost_x")
        _close(s.cost_xx(xk, uk, k), o["cost_xx"][0], "cost_xx")
        fin = np.asarray(o["fx"][0]) + rng.uniform(-0.05, 0.05, size=o["fx"].shape[1])
        _close(s.diff(fin, k), oracle.eval_point(pb, xk, k=k, fx_in=fin)["diff"][0], "diff")
        oF = oracle.eval_point(pb, xk, k=H - 1)
        _close(s.cost_F(xk), oF["cost"], "cost_F")
        _close(s.cost_F_x(xk), oF["cost_x"][0], "cost_F_x")
        _close(s.cost_F_xx(xk), oF["cost_xx"][0], "cost_F_xx")
        _close(s.cost_u(xk, uk, k), np.asarray(s.get_Rt()) @ uk, "cost_u")
        assert s.cost_ux(xk, uk, k).shape == (m, n) and s.cost_xu(xk, uk, k).shape == (n, m) and not s.cost_ux(xk, uk, k).any()
    # the simulation is untouched by the point evaluations
    np.testing.assert_array_equal(s.get_state(), s.get_init_state())
    # forward_pass_batch: fX and the Jacobians of the rollout
    U = 0.05 * np.sin(np.arange((H - 1) * m, dtype=np.float64)).reshape(H - 1, m)
    if sp["sys_kind"] == 1:
        U[:, -1] = 0.1
    X, fXo, Jo, _ = oracle.rollout(pb, U[None])
    fX, qL, ABJL = s.forward_pass_batch(U.reshape(-1))
    _close(fX, fXo[0].reshape(-1), "forward_pass_batch fX")
    assert len(ABJL) == H and qL.shape == (H * n,)
    for kk in (0, 1, H // 2, H - 1):
        _close(ABJL[kk][2], Jo[0, kk], "forward_pass_batch J[%d]" % kk)
    A1, B1 = ABJL[1][0], ABJL[1][1]
    s.reset()
    x1, fx1, A, B, J1 = s.forward_pass(np.asarray(s.get_init_state()), U[0], 0)
    _close(A1, A, "A of step 1")
    _close(B1, B, "B of step 1")
    _close(x1, X[0, 1], "forward_pass x")
    s.reset()
    # diff_batch over the keypoints' own states is zero; over f(x0) it is the stacked diff
    mu = np.asarray(s.get_mu_vector(True))
    stack = np.tile(np.asarray(fXo[0, 0]), len(ks))
    want = np.concatenate([oracle.eval_point(pb, np.asarray(s.get_init_state()), k=k, fx_in=fXo[0, 0])["diff"][0] for k in ks])
    _close(s.diff_batch(stack), want, "diff_batch")
    assert mu.shape == stack.shape
    # Keypoint.diff of one sub-system's keypoint == that block of System.diff
    t0 = sp["tasks"][0]
    kp = _keypoint(sp, t0["kp_target"][0], t0["kp_precision"][0], t0["kp_timestep"][0])
    nt0 = len(np.asarray(kp.get_state()))
    blk = oracle.eval_point(pb, np.asarray(s.get_init_state()), k=t0["kp_timestep"][0], fx_in=fXo[0, 0])["diff"][0]
    _close(kp.diff(np.asarray(fXo[0, 0][:nt0])), blk[:len(np.asarray(kp.get_precision()))], "Keypoint.diff")


@pytest.mark.gpu
def test_sd_through_pylqr(oracle_lib):
    P = _pylqr()
    from ilqr_planner_b200 import cabi
    from oracle import oracle
    rng = np.random.default_rng(9)
    Sd = P.utils.Sd
    for _ in range(6):
        a, b, c = (v / np.linalg.norm(v) for v in rng.normal(size=(3, 4)))
        t = rng.normal(size=4) * 0.5
        _close(Sd.logMap(a, b), oracle.sd(cabi.SD_LOGMAP, a, b)[0], "logMap")
        _close(Sd.expMap(a, t), oracle.sd(cabi.SD_EXPMAP, a, t)[0], "expMap")
        _close(Sd.distance(a, b), oracle.sd(cabi.SD_DISTANCE, a, b)[0], "distance")
        _close(Sd.transport(t, a, c), oracle.sd(cabi.SD_TRANSPORT, t, a, c)[0], "transport")
        _close(Sd.dquat_to_w_jac(a), oracle.sd(cabi.SD_DQUAT_TO_W_JAC, a)[0], "dquat_to_w_jac")
    with pytest.raises(RuntimeError, match="size 4"):
        Sd.logMap(np.zeros(3), np.zeros(3))


def _joint_batch(rng, B):
    """B variations of one joint-space + end-effector hybrid built through PyLQR (the HYBRID_* notebooks' structure)"""
    _pylqr()
    from PyLQR.sim import KDLRobot
    from PyLQR.system import AngularKeypoint, JointSpacePlannerSys, PosOrnKeypoint, PosOrnPlannerSys, SequentialSystem
    q0 = np.array([0.0, -0.4, 0.0, -1.8, 0.0, 1.5, 0.7])
    r = KDLRobot(URDF, "panda_link0", "panda_tip", q0, [0.0] * 7)
    H = 60
    qmid = q0 + 0.3
    sj = JointSpacePlannerSys(r, [AngularKeypoint(qmid, 10 * np.eye(7), 30)], [1e-3] * 7, H, 1, 0.05)
    pe, oe = np.array([0.45, 0.1, 0.45]), np.array([0.0, 1.0, 0.0, 0.0])
    se = PosOrnPlannerSys(r, [PosOrnKeypoint(pe, oe, 100 * np.eye(6), H - 1)], [1e-3] * 7, H, 1, 0.05)
    s = SequentialSystem(r, [sj, se], [1e-3] * 7, H, 1)
    q0s = q0 + rng.uniform(-0.2, 0.2, size=(B, 7))
    kj = (qmid + rng.uniform(-0.2, 0.2, size=(B, 7)))[:, None, :]
    ke = np.concatenate([pe + rng.uniform(-0.08, 0.08, size=(B, 3)), np.tile(oe, (B, 1))], axis=1)[:, None, :]
    return s, (r, sj, se), q0s, kj, ke


@pytest.mark.gpu
def test_solve_batch_through_pylqr_matches_per_system_solves():
    """ILQRRecursive.solve_batch (arrays in, one GPU batch) == solve_systems (a list of system objects) == single solves, bit for bit;
    the handle is created once and reused; N3 classes (joint-space + end-effector hybrid) through the Python API"""
    _pylqr()
    from PyLQR.sim import KDLRobot
    from PyLQR.solver import ILQRRecursive
    from PyLQR.system import AngularKeypoint, JointSpacePlannerSys, PosOrnKeypoint, PosOrnPlannerSys, SequentialSystem
    rng = np.random.default_rng(4)
    B = 6
    s, _, q0s, kj, ke = _joint_batch(rng, B)
    H, m = s.get_horizon(), s.get_nb_ctrl_var()
    sol = ILQRRecursive(s)
    out = sol.solve_batch(q0=q0s, kp_states=[kj, ke], nb_iter=5)
    assert out["X"].shape == (B, H, 7) and out["K"].shape == (B, H - 1, m, 7) and out["fX"].shape == (B, H, 14) and out["cost_hist"].shape == (B, 5)
    first = {k: np.array(v) for k, v in out.items() if v is not None}
    assert np.all(first["iters"] == 5) and np.all(np.isfinite(first["cost"])) and np.all(first["cost_hist"][:, -1] < first["cost_hist"][:, 0])
    # per-system objects, one batch
    systems = []
    for b in range(B):
        r = KDLRobot(URDF, "panda_link0", "panda_tip", q0s[b], [0.0] * 7)
        sj = JointSpacePlannerSys(r, [AngularKeypoint(kj[b, 0], 10 * np.eye(7), 30)], [1e-3] * 7, H, 1, 0.05)
        se = PosOrnPlannerSys(r, [PosOrnKeypoint(ke[b, 0, :3], ke[b, 0, 3:], 100 * np.eye(6), H - 1)], [1e-3] * 7, H, 1, 0.05)
        systems.append(SequentialSystem(r, [sj, se], [1e-3] * 7, H, 1))
    lst = ILQRRecursive(systems[0]).solve_systems(systems, nb_iter=5)
    for key in ("X", "U", "K", "d", "cost", "cost_hist", "alpha_hist"):
        np.testing.assert_array_equal(first[key], lst[key], err_msg=key)
    # single solves of two of them
    for b in (0, B - 1):
        X, fX, U, K, d, cost = ILQRRecursive(systems[b]).solve(np.zeros((H - 1, m)), 5, True, False, None)
        np.testing.assert_array_equal(np.asarray(X), first["X"][b])
        np.testing.assert_array_equal(np.asarray(K), first["K"][b])
        assert cost == first["cost"][b]
    # second batch of the same structure: no new handle, new data, results follow the data
    out2 = sol.solve_batch(q0=q0s[::-1].copy(), kp_states=[kj[::-1].copy(), ke[::-1].copy()], nb_iter=5)
    assert sol.cache_stats() == (1, 1)
    np.testing.assert_array_equal(np.array(out2["X"]), first["X"][::-1])
    # gains left on the device
    out3 = sol.solve_batch(q0=q0s, kp_states=[kj, ke], nb_iter=5, gains=False)
    assert out3["K"] is None and out3["d"] is None
    np.testing.assert_array_equal(np.array(out3["U"]), first["U"])
    # another batch size: a new handle
    sol.solve_batch(q0=q0s[:3], kp_states=[kj[:3], ke[:3]], nb_iter=2)
    assert sol.cache_stats()[0] == 2
    with pytest.raises(ValueError, match="one array per sub-system"):
        sol.solve_batch(q0=q0s, kp_states=kj, nb_iter=2)
    with pytest.raises(ValueError, match="another batch size"):
        sol.solve_batch(q0=q0s, kp_states=[kj[:2], ke], nb_iter=2)


@pytest.mark.gpu
def test_batch_solvers_and_al_solve_batch_through_pylqr():
    _pylqr()
    from PyLQR.solver import AL_ILQR, BatchILQR, BatchILQRCP, Constraint
    from PyLQR.utils import primitives
    rng = np.random.default_rng(8)
    B = 3
    s, _, q0s, kj, ke = _joint_batch(rng, B)
    H, m, n = s.get_horizon(), s.get_nb_ctrl_var(), s.get_nb_state_var()
    u0 = np.zeros((B, (H - 1) * m))
    ob = BatchILQR(s).solve_batch(q0=q0s, kp_states=[kj, ke], u0=u0, nb_iter=3)
    assert ob["U"].shape == (B, (H - 1) * m) and np.all(np.array(ob["cost_hist"])[:, -1] < np.array(ob["cost_hist"])[:, 0])
    psi = np.kron(primitives.build_psi_RBF(H - 1, 8), np.eye(m))
    oc = BatchILQRCP(s, psi).solve_batch(q0=q0s, kp_states=[kj, ke], nb_iter=3)
    assert oc["U"].shape == (B, (H - 1) * m) and np.all(np.isfinite(np.array(oc["cost"])))
    cons = []
    for k in range(H - 1):
        c = Constraint()
        c.A = np.hstack([np.zeros((m, n)), np.eye(m)])  # u <= 0.6
        c.b = np.full(m, 0.6)
        cons.append(c)
    al = AL_ILQR(s, cons, [np.zeros(m)] * (H - 1))
    oa = al.solve_batch(q0=q0s, kp_states=[kj, ke], nb_iter=6, lag_update_step=2, penalty=1.0, scaling_factor=2.0)
    assert oa["lambda"].shape == (B, H - 1, m) and oa["X"].shape == (B, H, n) and np.all(np.isfinite(np.array(oa["cost"])))
    assert float(np.max(np.array(oa["U"]))) < 0.6 + 0.05


@pytest.mark.gpu
def test_solve_batch_keeps_up_with_the_ctypes_handle():
    """the PyLQR batched entry point must cost what the ctypes Handle costs — same problems, same pinned result memory,
    host buffers in and out; within 10 % once both are warm"""
    import time
    _pylqr()
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.problem import ResultBuffers
    from ilqr_planner_b200.solver import Handle
    from PyLQR.sim import KDLRobot
    from PyLQR.solver import ILQRRecursive
    from PyLQR.system import PosOrnKeypoint, PosOrnPlannerSys
    B, nit = 2048, 10
    pb = synth.config2(batch=B)
    H = pb.horizon
    tk = pb.tasks[0]
    tg = np.asarray(tk.kp_target).reshape(B, -1, 15)
    pr = np.asarray(tk.kp_precision).reshape(-1, tg.shape[1], 6, 6)[0]
    steps = [int(v) for v in np.asarray(tk.kp_timestep).ravel()]
    q0 = np.asarray(pb.q0).reshape(B, 7)
    r = KDLRobot(URDF, "panda_link0", "panda_tip", q0[0], [0.0] * 7)
    kps = [PosOrnKeypoint(tg[0, i, 0:3], tg[0, i, 3:7], pr[i], steps[i]) for i in range(len(steps))]
    s = PosOrnPlannerSys(r, kps, np.asarray(tk.r_diag), H, 1, pb.dt)
    sol = ILQRRecursive(s)
    kp_states = np.ascontiguousarray(tg[:, :, :7])
    u0 = np.zeros((B, H - 1, 7))

    def best(fn, reps=6):
        fn()
        fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        return min(ts)

    with Handle(pb) as h:
        out = ResultBuffers(pb, nit, "recursive", pinned=True)
        t_h = best(lambda: h.solve_recursive(u0, nit, True, False, out=out))
        ref = {k: np.array(getattr(out, k)) for k in ("X", "U", "K", "cost_hist")}
    res = {}
    t_p = best(lambda: res.update(sol.solve_batch(q0=q0, kp_states=kp_states, U0=u0, nb_iter=nit)))
    for k in ref:
        # the URDF chain and synth.panda_chain() are the same robot up to how the constants were typed in: same problems to rounding
        assert np.allclose(np.array(res[k]), ref[k], rtol=1e-6, atol=1e-9), k
    assert sol.cache_stats()[0] == 1
    print("ctypes Handle %.2f ms, PyLQR solve_batch %.2f ms (B=%d, %d iterations)" % (t_h * 1e3, t_p * 1e3, B, nit))
    assert t_p <= 1.10 * t_h, "PyLQR solve_batch %.2f ms vs ctypes Handle %.2f ms" % (t_p * 1e3, t_h * 1e3)

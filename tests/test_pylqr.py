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


def test_pylqr_module_surface():
    P = _pylqr()
    for mod, names in (("sim", ["SimulationInterface", "KDLRobot", "Robot2D", "TransformedSimulationInterface"]),
                       ("system", ["Keypoint", "PosOrnKeypoint", "SpacetimeKeypoint", "System", "SequentialSystem", "PosOrnPlannerSys", "PosOrnTimePlannerSys"]),
                       ("solver", ["Constraint", "AL_ILQR", "BatchILQR", "BatchILQRCP", "ILQRRecursive"]),
                       ("utils", ["CallBackMessage", "PythonCallbackMessage", "primitives"])):
        for n in names:
            assert hasattr(getattr(P, mod), n), "%s.%s" % (mod, n)
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

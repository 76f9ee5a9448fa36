"""Shared test helpers: golden cases -> ProblemBatch, synthetic BASELINE.json configs, comparison utilities."""
import json
import os

import numpy as np

from ilqr_planner_b200 import cabi
from ilqr_planner_b200.problem import ProblemBatch, RobotSpec, TaskSpec

HERE = os.path.dirname(os.path.abspath(__file__))
_GOLD = None


def goldens():
    global _GOLD
    if _GOLD is None:
        with open(os.path.join(HERE, "golden", "notebook_goldens.json")) as f:
            _GOLD = json.load(f)
    return _GOLD


def golden_case(name):
    for c in goldens()["cases"]:
        if c["name"] == name:
            return c
    raise KeyError(name)


def problem_from_spec(sp, chains, batch=1):
    """system spec (golden JSON) -> ProblemBatch with `batch` identical copies"""
    segs = chains[sp["chain"]]
    dof = sp["dof"]
    tasks = []
    for t in sp["tasks"]:
        tg = np.asarray(t["kp_target"], dtype=np.float64)[None].repeat(batch, 0)
        pr = np.asarray(t["kp_precision"], dtype=np.float64)[None]
        lim = t["limits"]
        tasks.append(TaskSpec(kp_timestep=t["kp_timestep"], kp_target=tg, kp_precision=pr, r_diag=np.asarray(t["r_diag"]),
                              T=None if t["T"] is None else np.asarray(t["T"]).reshape(4, 4),
                              state_max=None if lim is None else np.asarray(lim["state_max"]),
                              state_min=None if lim is None else np.asarray(lim["state_min"]),
                              limit_weight=None if lim is None else np.asarray(lim["weight"], dtype=np.int32)))
    q0 = np.asarray(sp["q0"], dtype=np.float64)[None].repeat(batch, 0)
    dq0 = np.asarray(sp["dq0"], dtype=np.float64)[None].repeat(batch, 0)
    return ProblemBatch(sp["sys_kind"], sp["nb_deriv"], sp["horizon"], sp["dt"], dof, [RobotSpec(cabi.ROBOT_KDL, dof, segs)], tasks,
                        np.asarray(sp["r_diag"]), q0, dq0)


def case_u0(case, pb):
    H, m = pb.horizon, pb.m
    if "u0_row" in case:
        u = np.tile(np.asarray(case["u0_row"], dtype=np.float64), (H - 1, 1))
    else:
        u = np.asarray(case["u0_full"], dtype=np.float64).reshape(H - 1, m)
    return u[None].repeat(pb.batch, 0)


def case_psi(case, pb):
    if "psi" in case:
        return np.asarray(case["psi"], dtype=np.float64)
    return np.kron(np.asarray(case["psi_kron"], dtype=np.float64), np.identity(pb.m))


def fmt_g(x):
    """std::ostream default formatting of a double (6 significant digits) as the reference prints it"""
    if np.isnan(x):
        return "nan"
    return "%g" % x


def trace_lines(cost_hist, alpha_hist, iters):
    return [dict(cost=fmt_g(cost_hist[i]), alpha=fmt_g(alpha_hist[i])) for i in range(iters)]


def norm_trace(tr):
    return [dict(cost=t["cost"].replace("-nan", "nan"), alpha=t["alpha"]) for t in tr]

"""GPU bring-up: stage-by-stage comparison of the CUDA path with the CPU oracle (prints, never asserts)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ilqr_planner_b200 import cabi, synth
from ilqr_planner_b200.solver import Handle
from oracle import oracle
from tests import _cases as tc

def rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    den = np.maximum(np.abs(b).max(), 1e-300)
    return float(np.nanmax(np.abs(a - b)) / den)

def stage(pb, name, iters=(1, 2, 3, 6), u0=None):
    print("=====", name, "B=%d n=%d m=%d H=%d" % (pb.batch, pb.n, pb.m, pb.horizon))
    with Handle(pb) as h:
        x, quat, J, dx, w = h.kinematics(0, pb.q0[:, :pb.robots[0].dof])
        xo, qo, Jo, dxo, wo = oracle.kinematics(pb, 0, pb.q0[:, :pb.robots[0].dof])
        print("kinematics rel: x %.2e quat %.2e J %.2e" % (rel(x, xo), rel(quat, qo), rel(J, Jo)))
        rng = np.random.default_rng(0)
        U = rng.normal(0, 0.05, (pb.batch, pb.horizon - 1, pb.m))
        if pb.time: U[:, :, -1] = np.abs(U[:, :, -1]) + 0.05
        X, fX, Jr, c = h.rollout(U)
        Xo, fXo, Jro, co = oracle.rollout(pb, U)
        print("rollout rel: X %.2e fX %.2e J %.2e cost %.2e" % (rel(X, Xo), rel(fX, fXo), rel(Jr, Jro), rel(c, co)))
        for it in iters:
            out = h.solve_recursive(u0, it, True, False)
            ref = oracle.solve_recursive(pb, u0, it, True, False)
            print("it=%d  cost %.2e  alpha_eq %s  X %.2e U %.2e K %.2e d %.2e fX %.2e  status %s" % (
                it, rel(out.cost_hist, ref.cost_hist), np.array_equal(out.alpha_hist, ref.alpha_hist), rel(out.X, ref.X), rel(out.U, ref.U),
                rel(out.K, ref.K), rel(out.d, ref.d), rel(out.fX, ref.fX), np.unique(out.status)))
        print("gpu cost_hist[0]", out.cost_hist[0]); print("ref cost_hist[0]", ref.cost_hist[0])
        print("gpu alpha[0]", out.alpha_hist[0]); print("ref alpha[0]", ref.alpha_hist[0])

def golden(name):
    g = tc.goldens(); c = tc.golden_case(name)
    pb = tc.problem_from_spec(c["system"], g["chains"]); u0 = tc.case_u0(c, pb)
    with Handle(pb) as h:
        if c["solver"] == "recursive":
            out = h.solve_recursive(u0, c["nb_iter"], c["line_search"], c["early_stop"])
        elif c["solver"] == "al":
            out = h.solve_al(u0, c["A"], c["b"], c["lambda0"], c["nb_iter"], c["lag_update_step"], c["penalty"], c["scaling_factor"], c["line_search"], c["early_stop"])
        else:
            return
    got = tc.trace_lines(out.cost_hist[0], out.alpha_hist[0], out.iters[0]); exp = tc.norm_trace(c["trace"])
    nm = sum(1 for a, b in zip(got, exp) if a == b)
    print("%-45s %s %d/%d (got %d) status %d" % (name, "OK  " if got == exp else "DIFF", nm, len(exp), len(got), out.status[0]))
    if got != exp:
        for i, (a, b) in enumerate(zip(got, exp)):
            if a != b:
                print("    first diff at it %d: got %s exp %s" % (i + 1, a, b)); break

if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "gold"]
    if "c2" in which: stage(synth.config2(batch=8), "config2")
    if "c3" in which: stage(synth.config3(batch=8), "config3 (recursive on nd=2)")
    if "c5" in which: stage(synth.config5(batch=8), "config5")
    if "c4" in which:
        pb = synth.config4(batch=4, horizon=60); stage(pb, "config4 (recursive on time system)", u0=synth.u0_time(pb))
    if "c1" in which: stage(synth.config1(batch=8), "config1")
    if "gold" in which:
        for c in tc.goldens()["cases"]:
            if c["solver"] in ("recursive", "al"): golden(c["name"])

"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI, against
  (1) the reference's stored notebook traces (tests/golden/notebook_goldens.json) and
  (2) the CPU oracle on seeded synthetic batches of every BASELINE.json config shape.

Tolerances (north_star: per iteration the same accepted alpha sequence; trajectories, gains and cost within 1e-9 relative),
see tests/_parity.py: element-wise relative differences, every iteration's cost against itself, bound max(1e-9, 4 x the
oracle's own noise floor under a 1-ulp change of q0); step sequences exact up to each problem's first tie, a tie being a
line-search margin below the measured instability threshold of the oracle itself.
  d (feed-forward)      : against max|U| + max|d|  (d -> 0 at convergence, so it is scaled by the control it corrects)
"""
import numpy as np
import pytest

from tests import _cases as tc

pytestmark = pytest.mark.gpu

from tests import _parity as pt
from tests._parity import RTOL, rel


def _handle(pb):
    from ilqr_planner_b200.solver import Handle
    return Handle(pb, device=0)


REC_FIELDS = {
    "cost_hist": (lambda r: r.cost_hist, dict(floor=0.0)),  # every iteration's cost against ITSELF
    "X": (lambda r: r.X, {}), "U": (lambda r: r.U, {}), "fX": (lambda r: r.fX, {}), "K": (lambda r: r.K, {}),
}


def _subprocess_case(code, env_extra, timeout=600):
    """the A/B switches of the library are read once per process: such cases run in a child process"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-c", "import sys; sys.path.insert(0, %r)\n" % root + code], env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


# ---------------------------------------------------------------------------------------------------------------
# (1) reference notebook traces, reproduced by the CUDA path
GOLD_GPU = [c["name"] for c in tc.goldens()["cases"] if c["solver"] in ("recursive", "al")]


@pytest.mark.parametrize("name", GOLD_GPU)
def test_notebook_trace_on_gpu(name):
    g = tc.goldens()
    c = tc.golden_case(name)
    pb = tc.problem_from_spec(c["system"], g["chains"])
    u0 = tc.case_u0(c, pb)
    with _handle(pb) as h:
        if c["solver"] == "recursive":
            out = h.solve_recursive(u0, c["nb_iter"], c["line_search"], c["early_stop"])
        else:
            out = h.solve_al(u0, c["A"], c["b"], c["lambda0"], c["nb_iter"], c["lag_update_step"], c["penalty"], c["scaling_factor"], c["line_search"],
                             c["early_stop"])
    got = tc.trace_lines(out.cost_hist[0], out.alpha_hist[0], out.iters[0])
    exp = tc.norm_trace(c["trace"])
    assert len(got) == len(exp), "iteration count %d != %d" % (len(got), len(exp))
    for i, (a, b) in enumerate(zip(got, exp)):
        assert a == b, "iteration %d: got %s, reference printed %s" % (i + 1, a, b)


# ---------------------------------------------------------------------------------------------------------------
# (2) oracle parity on synthetic batches
def _check_recursive(pb, u0, nb_iter, oracle, early_stop=False, min_tie_free=0.5):
    with _handle(pb) as h:
        out = h.solve_recursive(u0, nb_iter, True, early_stop)
    fields = dict(REC_FIELDS)
    dscale = None

    def solve(p, v=None):
        return oracle.solve_recursive(p, u0, nb_iter, True, early_stop, threads=8, variant=v)
    sens = pt.measure(solve, pb, fields)
    ref = sens.ref
    free = pt.check_sequences(out, sens)  # exact step sequences up to each problem's first tie
    assert free.sum() >= min_tie_free * pb.batch, "too few tie-free problems (%d of %d) for this test to mean anything: %s" % (free.sum(), pb.batch, sens.summary())
    # every iteration before a problem's first tie: the cost against itself
    ft = sens.first_tie()
    keep = np.arange(ref.cost_hist.shape[1])[None, :] < ft[:, None]
    err_c = rel(np.where(keep, out.cost_hist, 0.0), np.where(keep, ref.cost_hist, 0.0), floor=0.0) if keep.any() else 0.0
    assert err_c < sens.tol("cost_hist"), "cost trace %.2e (bound %.2e; %s)" % (err_c, sens.tol("cost_hist"), sens.summary())
    # tie-free problems: everything the reference returns
    assert np.array_equal(out.iters[free], ref.iters[free])
    for f in ("X", "U", "fX", "K"):
        e = sens.err(f, out, free)
        assert e < sens.tol(f), "%s %.2e (bound %.2e; %s)" % (f, e, sens.tol(f), sens.summary())
    dscale = float(np.abs(ref.U).max() + np.abs(ref.d).max())
    dn = max(rel(r.d[sens.stable], ref.d[sens.stable], scale=dscale) for r in sens.perturbed) if sens.stable.any() else 0.0
    assert rel(out.d[free], ref.d[free], scale=dscale) < max(RTOL, pt.NOISE_FACTOR * dn)
    assert rel(out.cost[free], ref.cost[free], floor=0.0) < sens.tol("cost_hist")
    sens.check_truth(out, free)
    return out, ref


@pytest.mark.parametrize("nb_iter", [1, 2, 3, 5])
def test_config2_per_iteration(oracle_lib, nb_iter):
    """per-iteration matching: run 1, 2, 3, 5 iterations and compare everything the reference returns"""
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config2(batch=24), None, nb_iter, oracle_lib)


def test_config2_full_20_iterations(oracle_lib):
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config2(batch=64), None, 20, oracle_lib)


def test_config2_early_stop_and_no_line_search(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config2(batch=16)
    out, ref = _check_recursive(pb, None, 20, oracle_lib, early_stop=True)
    assert (out.status & 4).any()  # some problems stop early
    with _handle(pb) as h:
        o2 = h.solve_recursive(None, 3, False, False)
    r2 = oracle_lib.solve_recursive(pb, None, 3, False, False)
    assert np.array_equal(o2.alpha_hist, r2.alpha_hist) and rel(o2.cost_hist, r2.cost_hist) < RTOL


def test_config1_robot2d_position(oracle_lib):
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config1(batch=32), None, 8, oracle_lib)


def test_config3_shape_recursive_double_integrator(oracle_lib):
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config3(batch=16), None, 6, oracle_lib)


def test_config5_dual_arm(oracle_lib):
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config5(batch=16), None, 6, oracle_lib)


@pytest.mark.parametrize("dof,batch", [(2, 9), (3, 16), (4, 5), (5, 11), (6, 16), (7, 13)])
def test_column_per_lane_backward_pass_sizes_and_ragged_batches(oracle_lib, dof, batch):
    """k_backward_n1_col<N> (four problems per warp, one column of the augmented matrices per lane) for every state size it is
    instantiated for, on batches that do not fill the last warp (idle lane groups shadow a live problem and must store nothing),
    without limits and with soft joint limits that the initial controls violate"""
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config5_uneven(batch=batch, dofs=(dof,), limits=False), None, 5, oracle_lib)
    pb = synth.config5_uneven(batch=batch, dofs=(dof,), limits=True)
    u0 = np.full((pb.batch, pb.horizon - 1, pb.m), 0.3)
    u0[:, :, ::2] *= -1
    _check_recursive(pb, u0, 2, oracle_lib)


def test_column_per_lane_and_warp_per_problem_backward_kernels_agree():
    """ILQRB_BWD=warp (read once per process, hence the subprocess) selects the warp-per-problem kernel: the two forms of the backward
    pass return the same solve up to rounding, on a batch that is not a multiple of 4"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from ilqr_planner_b200 import synth\n"
        "from ilqr_planner_b200.solver import Handle\n"
        "pb = synth.config2(batch=70, horizon=60)\n"
        "with Handle(pb) as h:\n"
        "    o = h.solve_recursive(None, 4, True, False)\n"
        "np.savez(sys.argv[1], X=o.X, U=o.U, K=o.K, cost=o.cost, alpha=o.alpha_hist)\n" % root)
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        outs = {}
        for form in ("col", "warp"):
            f = os.path.join(td, form + ".npz")
            env = dict(os.environ, ILQRB_BWD=form)
            subprocess.run([sys.executable, "-c", code, f], check=True, env=env, timeout=600)
            outs[form] = dict(np.load(f))
    a, b = outs["col"], outs["warp"]
    same = np.all((a["alpha"] == b["alpha"]) | (np.isnan(a["alpha"]) & np.isnan(b["alpha"])), axis=1)  # a numerical tie may part two sequences
    assert same.sum() >= 0.9 * same.size
    for k in ("X", "U", "K", "cost"):
        assert rel(a[k][same], b[k][same]) < 1e-8, (k, rel(a[k][same], b[k][same]))


def test_composite_uneven_blocks(oracle_lib):
    """block-diagonal backward pass (one warp per problem and per group of joints no task couples) with runtime block
    sizes and an odd block offset; then with soft joint limits that the initial controls violate over most of the horizon.
    The limit cases run 2 iterations: a quadratic limit penalty puts the next iterate on the limit itself, after which
    `x > limit` is decided by rounding in any implementation (the reference's too)."""
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config5_uneven(batch=16, limits=False), None, 6, oracle_lib)
    _check_recursive(synth.config5_uneven(batch=8, dofs=(3, 2, 4), limits=False), None, 5, oracle_lib)
    for pb in (synth.config5_uneven(batch=16, limits=True), synth.config5_uneven(batch=8, dofs=(5,), limits=True)):
        u0 = np.full((pb.batch, pb.horizon - 1, pb.m), 0.3)
        u0[:, :, ::2] *= -1
        for nit in (1, 2):
            out, ref = _check_recursive(pb, u0, nit, oracle_lib)
        X = ref.X.reshape(pb.batch, pb.horizon, -1)
        assert ((X > pb.tasks[0].state_max) | (X < pb.tasks[0].state_min)).any(), "the limits of this test must be active"


def test_time_system_recursive(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config4(batch=8, horizon=60)
    _check_recursive(pb, synth.u0_time(pb), 6, oracle_lib)


def test_warm_start_nonzero_u0_and_zero_iterations(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config2(batch=8)
    rng = np.random.default_rng(7)
    u0 = rng.normal(0, 0.02, (8, pb.horizon - 1, pb.m))
    _check_recursive(pb, u0, 3, oracle_lib)
    with _handle(pb) as h:
        o = h.solve_recursive(u0, 0, True, False)
    r = oracle_lib.solve_recursive(pb, u0, 0, True, False)
    assert rel(o.X, r.X) < RTOL and rel(o.cost, r.cost) < RTOL and np.all(o.K == 0) and np.all(o.iters == 0)


AL_FIELDS = {
    "cost_hist": (lambda r: r.cost_hist, dict(floor=0.0)),
    "X": (lambda r: r.X, {}), "U": (lambda r: r.U, {}), "lam": (lambda r: r.lam, {}),
}


def _check_al(pb, al, nb_iter, oracle, line_search=True, want_active=True):
    """AL-iLQR on the GPU against the oracle (AL-ILQR.cpp:21-44,116-134,189-208): cost trace, X, U and the multipliers within
    max(1e-9, 4 x the oracle's own 1-ulp noise), step sequences exact up to each problem's first tie"""
    with _handle(pb) as h:
        out = h.solve_al(None, al["A"], al["b"], al["lambda0"], nb_iter, al["lag_update_step"], al["penalty"], al["scaling_factor"], line_search, False)

    def solve(p, v=None):
        return oracle.solve_al(p, None, al["A"], al["b"], al["lambda0"], nb_iter, al["lag_update_step"], al["penalty"], al["scaling_factor"], line_search, False,
                               threads=8, variant=v)
    sens = pt.measure(solve, pb, AL_FIELDS)
    ref = sens.ref
    if want_active:
        # the constraints must really bind, or the AL terms of the backward pass and the multiplier update are not exercised
        A, b = np.asarray(al["A"]), np.asarray(al["b"])
        sv = np.concatenate([ref.X[:, :-1], ref.U], axis=2)
        c = (np.einsum("rj,bkj->bkr", A, sv) if A.ndim == 2 else np.einsum("krj,bkj->bkr", A, sv)) - b
        active = (c >= 0) | (ref.lam > 0)
        assert (ref.lam > 0).any(), "no multiplier ever became positive"
        assert active.any(axis=2).mean() >= 0.25, "active set non-empty in only %.0f %% of the steps" % (100 * active.any(axis=2).mean())
        assert nb_iter >= 3 * al["lag_update_step"], "the multiplier / penalty update must fire at least three times"
    free = pt.check_sequences(out, sens)
    assert free.sum() >= pb.batch // 2, sens.summary()
    ft = sens.first_tie()
    keep = np.arange(ref.cost_hist.shape[1])[None, :] < ft[:, None]
    err_c = rel(np.where(keep, out.cost_hist, 0.0), np.where(keep, ref.cost_hist, 0.0), floor=0.0)
    assert err_c < sens.tol("cost_hist"), "cost trace %.2e (bound %.2e; %s)" % (err_c, sens.tol("cost_hist"), sens.summary())
    for f in ("X", "U", "lam"):
        e = sens.err(f, out, free)
        assert e < sens.tol(f), "%s %.2e (bound %.2e; %s)" % (f, e, sens.tol(f), sens.summary())
    sens.check_truth(out, free)
    return out, ref


def test_config3_al_ilqr_inert_constraints(oracle_lib):
    """C3 as BASELINE.json words it (rows = the Panda joint limits, which these reachable targets never touch): the AL terms
    stay zero, the solve must still equal the oracle's.  The AL arithmetic itself is held by test_config3_al_active_rows."""
    from ilqr_planner_b200 import synth
    _check_al(synth.config3(batch=12), synth.al_config3(), 12, oracle_lib, want_active=False)


@pytest.mark.parametrize("kind,line_search,nb_iter,lag", [("qbox", True, 12, 3), ("mixed", True, 12, 3), ("steps", True, 12, 3), ("qbox", True, 10, 1),
                                                         ("qbox", False, 4, 1), ("mixed", False, 4, 1)])
def test_config3_al_active_rows(oracle_lib, kind, line_search, nb_iter, lag):
    """AL-iLQR on the double-integrator fast path (fused2.cuh k_backward_n2<.., AL=true>, k_accept's multiplier update) with
    constraints that bind over most of the solve: a joint box 0.3 rad around the start with targets 0.4-0.6 rad away ("qbox",
    14 rows shared by all steps), the box plus velocity and control rows ("mixed", 29 rows: c_u != 0, a multiplier that
    starts positive), and per-step rows ("steps").  Without line search every iteration takes the full AL step."""
    from ilqr_planner_b200 import synth
    pb = synth.config3_active(batch=64)
    al = synth.al_config3_active(kind, lag_update_step=lag)
    _check_al(pb, al, nb_iter, oracle_lib, line_search=line_search)


def test_al_fast_path_equals_generic_path():
    """the same active-row AL solve under ILQRB_GENERIC=1 (generic k_backward<.., AL> with element loops): both GPU paths must
    agree with each other as closely as each agrees with the oracle"""
    code = (
        "import numpy as np\n"
        "from ilqr_planner_b200 import synth\n"
        "from ilqr_planner_b200.solver import Handle\n"
        "pb = synth.config3_active(batch=32)\n"
        "al = synth.al_config3_active('mixed', lag_update_step=2)\n"
        "with Handle(pb, device=0) as h: out = h.solve_al(None, al['A'], al['b'], al['lambda0'], 6, 2, al['penalty'], al['scaling_factor'], True, False)\n"
        "np.savez(%r, cost_hist=out.cost_hist, alpha_hist=out.alpha_hist, X=out.X, U=out.U, lam=out.lam)\n")
    import os
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        fa, fb = os.path.join(td, "fast.npz"), os.path.join(td, "generic.npz")
        _subprocess_case(code % fa, {})
        _subprocess_case(code % fb, {"ILQRB_GENERIC": "1"})
        a, b = np.load(fa), np.load(fb)
        same = pt.same_sequence(a["alpha_hist"], b["alpha_hist"])
        assert same.sum() >= 24, "the two GPU paths choose different steps in %d of 32 problems" % (~same).sum()
        assert (b["lam"] > 0).any()
        assert rel(a["cost_hist"][same], b["cost_hist"][same], floor=0.0) < 1e-8
        assert rel(a["X"][same], b["X"][same]) < 1e-8 and rel(a["lam"][same], b["lam"][same]) < 1e-7


def test_generic_recursive_path_vs_oracle(oracle_lib):
    """ILQRB_GENERIC=1: the all-in-shared-memory kernels (k_backward<..>, k_linesearch, k_accept) on a system the fast path would
    take, against the oracle (the switch is read once per process, hence the child process)"""
    code = (
        "import numpy as np\n"
        "from ilqr_planner_b200 import synth\n"
        "from ilqr_planner_b200.solver import Handle\n"
        "from oracle import oracle\n"
        "from tests import _parity as pt\n"
        "for pb, nit in ((synth.config2(batch=16), 6), (synth.config3(batch=8), 4)):\n"
        "    with Handle(pb, device=0) as h: out = h.solve_recursive(None, nit, True, False)\n"
        "    F = {'cost_hist': (lambda r: r.cost_hist, dict(floor=0.0)), 'X': (lambda r: r.X, {}), 'U': (lambda r: r.U, {}), 'K': (lambda r: r.K, {})}\n"
        "    sens = pt.measure(lambda p, v=None: oracle.solve_recursive(p, None, nit, True, False, threads=8, variant=v), pb, F)\n"
        "    free = pt.check_sequences(out, sens)\n"
        "    assert free.sum() >= pb.batch // 2\n"
        "    for f in F:\n"
        "        e = sens.err(f, out, free)\n"
        "        assert e < sens.tol(f), (f, e, sens.tol(f), sens.summary())\n"
        "print('generic ok')\n")
    assert "generic ok" in _subprocess_case(code, {"ILQRB_GENERIC": "1"})


# ---------------------------------------------------------------------------------------------------------------
# kernel-level hooks and size-independent properties
def test_kinematics_and_rollout_hooks(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config3(batch=16)
    rng = np.random.default_rng(3)
    q = rng.uniform(-2, 2, (64, 7))
    dq = rng.uniform(-1, 1, (64, 7))
    U = rng.normal(0, 0.5, (16, pb.horizon - 1, pb.m))
    with _handle(pb) as h:
        got = h.kinematics(0, q, dq)
        X, fX, J, cost = h.rollout(U)
    exp = oracle_lib.kinematics(pb, 0, q, dq)
    for a, b in zip(got, exp):
        assert rel(a, b) < 1e-12
    Xo, fXo, Jo, co = oracle_lib.rollout(pb, U)
    assert np.array_equal(X, Xo)  # the integrator is evaluated without FMA contraction: bit-exact
    assert rel(fX, fXo) < 1e-12 and rel(J, Jo) < 1e-12 and rel(cost, co) < 1e-11


def test_full_size_properties():
    """BASELINE size (B=4096, T=100, 20 iterations): properties that need no oracle run:
    monotone cost under the line-search rule, duplicated problems give bit-identical answers, batch-order independence."""
    from ilqr_planner_b200 import synth
    pb = synth.config2(batch=4096)
    with _handle(pb) as h:
        out = h.solve_recursive(None, 20, True, False)
    ch = out.cost_hist
    assert not np.isnan(ch).any()
    # accepted cost never increases except through the unconditional alpha = 2^-10 acceptance (Q2)
    inc = ch[:, 1:] > ch[:, :-1]
    assert np.all(out.alpha_hist[:, 1:][inc] == 2.0 ** -10)
    # reversed batch order -> same per-problem results, bit for bit (problems are independent)
    idx = np.arange(4096)[::-1].copy()
    pr = synth.config2(batch=4096)
    pr.q0 = pr.q0[idx].copy()
    pr.dq0 = pr.dq0[idx].copy()
    pr.tasks[0].kp_target = pr.tasks[0].kp_target[idx].copy()
    with _handle(pr) as h:
        o2 = h.solve_recursive(None, 20, True, False)
    assert np.array_equal(o2.cost_hist[idx], out.cost_hist)
    assert np.array_equal(o2.U[idx], out.U)


def test_errors_are_loud():
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import Handle
    pb = synth.config2(batch=2)
    pb.tasks[0].kp_timestep = [10, 10]
    with pytest.raises(RuntimeError, match="strictly ascending"):
        Handle(pb)
    pb = synth.config2(batch=2)
    with pytest.raises(RuntimeError, match="device index"):
        Handle(pb, device=99)


# ---------------------------------------------------------------------------------------------------------------
# (3) Batch / control-primitive Batch solvers (BatchILQR.cpp, BatchILQRCP.cpp): single-integrator systems on the GPU
GOLD_BATCH = [c["name"] for c in tc.goldens()["cases"] if c["solver"] in ("batch", "batch_cp")]


@pytest.mark.parametrize("name", GOLD_BATCH)
def test_batch_notebook_trace_on_gpu(name):
    g = tc.goldens()
    c = tc.golden_case(name)
    pb = tc.problem_from_spec(c["system"], g["chains"])
    u0 = tc.case_u0(c, pb)
    psi = tc.case_psi(c, pb) if c["solver"] == "batch_cp" else None
    with _handle(pb) as h:
        out = h.solve_batch(u0, c["nb_iter"], c["early_stop"], Q=c["Q"], psi=psi)
    got = tc.trace_lines(out.cost_hist[0], out.alpha_hist[0], out.iters[0])
    exp = tc.norm_trace(c["trace"])
    assert len(got) == len(exp), "iteration count %d != %d" % (len(got), len(exp))
    for i, (a, b) in enumerate(zip(got, exp)):
        assert a["alpha"] == b["alpha"], "iteration %d alpha %s != %s" % (i + 1, a["alpha"], b["alpha"])
        if a["cost"] != b["cost"]:  # one unit of the 6th printed digit (dense N x N solve: Cholesky here, LU inverse there)
            assert abs(float(a["cost"]) - float(b["cost"])) <= 2e-6 * abs(float(b["cost"])), "iteration %d cost %s != %s" % (i + 1, a["cost"], b["cost"])


BATCH_FIELDS = {"cost_hist": (lambda r: r.cost_hist, dict(floor=0.0)), "U": (lambda r: r.U, {})}


def _check_batch(pb, u0, nb_iter, oracle, psi=None, early_stop=False):
    """The step comes out of a dense solve whose matrix R + Su^T W Su has condition number 1e5 (R = 1e-5 against O(1) task
    curvature) to 1e7 (time systems: the time column of B is large), so every implementation's answer carries cond * eps of
    noise that compounds over the iterations.  How much is MEASURED: the bound is max(1e-9, 4 x what the oracle's own outputs
    move by when q0 changes by one or two ulps) — on the time systems that noise alone is 1e-6 (cost) to 1e-4 (u), on the
    double-integrator and dual-arm systems 1e-10.  test_batch_step_against_quad_truth holds the solve itself to a
    __float128 solution."""
    with _handle(pb) as h:
        out = h.solve_batch(u0, nb_iter, early_stop, psi=psi)
    sens = pt.measure(lambda p, v=None: oracle.solve_batch(p, u0, nb_iter, early_stop, psi=psi, threads=8, variant=v), pb, BATCH_FIELDS)
    ref = sens.ref
    free = pt.check_sequences(out, sens)
    assert free.sum() >= pb.batch // 2, sens.summary()
    ft = sens.first_tie()
    keep = np.arange(ref.cost_hist.shape[1])[None, :] <= ft[:, None]  # the batch solvers print the PRE-step cost: the tie iteration's own is still common
    keep &= ~np.isnan(ref.cost_hist)
    err_c = rel(np.where(keep, out.cost_hist, 0.0), np.where(keep, ref.cost_hist, 0.0), floor=0.0)
    assert err_c < sens.tol("cost_hist"), "cost trace %.2e (bound %.2e; %s)" % (err_c, sens.tol("cost_hist"), sens.summary())
    e = sens.err("U", out, free)
    assert e < sens.tol("U"), "u %.2e (bound %.2e; %s)" % (e, sens.tol("U"), sens.summary())
    return out, ref


@pytest.mark.parametrize("case", ["full_horizon", "early_stop", "no_line_search", "five_joints_tight_limits", "two_sub_systems"])
def test_time_system_fast_path_vs_oracle(oracle_lib, case):
    """fused_time.cuh (k_backward_time / k_forward_time): compile-time and runtime joint counts, joint limits that are violated along
    the way (limit terms of the line search and of the backward pass), two limit-carrying sub-systems (the slow limit loops)"""
    from ilqr_planner_b200 import synth
    if case == "five_joints_tight_limits":
        pb = synth.config4_short(batch=16)
    elif case == "two_sub_systems":
        pb = synth.config4_short(batch=12, dof=6, two_tasks=True)
    else:
        pb = synth.config4(batch=24, horizon=100)
    u0 = synth.u0_time(pb)
    if case == "early_stop":
        _check_recursive(pb, u0, 25, oracle_lib, early_stop=True, min_tie_free=0.25)
    elif case == "no_line_search":
        from ilqr_planner_b200.solver import Handle
        from oracle import oracle
        with Handle(pb) as h:
            out = h.solve_recursive(u0, 3, False, False)
        ref = oracle.solve_recursive(pb, u0, 3, False, False, threads=8)
        assert np.all(out.alpha_hist == 1.0) and np.all(ref.alpha_hist == 1.0)
        ok = np.isfinite(ref.cost_hist[:, -1])
        assert rel(out.cost_hist[ok], ref.cost_hist[ok], floor=0.0) < 1e-6 and rel(out.X[ok], ref.X[ok]) < 1e-6
    else:
        _check_recursive(pb, u0, 8, oracle_lib, min_tie_free=0.25)


@pytest.mark.parametrize("case", ["single_integrator", "double_integrator", "time_system"])
def test_indefinite_quu_takes_the_pivoted_inverse(oracle_lib, case):
    """a keypoint with a NEGATIVE precision repels: the control Hessian Quu + 1e-6 I turns indefinite, the unpivoted elimination of the
    backward kernels meets a non-positive pivot, and the step is redone with the partial-pivot LU inverse the reference uses
    (ILQRRecursive.cpp:89; wmat.cuh wm_inverse_pivoted).  The status word says so, and the result follows the oracle's LU."""
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import Handle
    from oracle import oracle
    if case == "single_integrator":
        pb, u0 = synth.config2(batch=16), None
    elif case == "double_integrator":
        pb, u0 = synth.config3_short(batch=16, tight_limits=False), None
    else:
        pb = synth.config4(batch=16, horizon=60)
        u0 = synth.u0_time(pb)
    tk = pb.tasks[0]
    pr = np.array(tk.kp_precision, dtype=np.float64)
    pr[0, 0] = -0.05 * pr[0, 0]
    tk.kp_precision = pr
    nit = 4
    with Handle(pb) as h:
        out = h.solve_recursive(u0, nit, True, False)
    flagged = (out.status & cabi.ST_QUU_NOT_SPD) != 0
    assert flagged.sum() >= 4, "the crafted case must drive Quu indefinite (flagged %d of %d)" % (flagged.sum(), pb.batch)
    ref = oracle.solve_recursive(pb, u0, nit, True, False, threads=8)
    ok = np.isfinite(ref.cost_hist).all(axis=1) & pt.same_sequence(out.alpha_hist, ref.alpha_hist)
    assert (ok & flagged).sum() >= 2, "no flagged problem with a comparable step sequence"
    sel = ok & flagged
    assert rel(out.cost_hist[sel], ref.cost_hist[sel], floor=0.0) < 1e-6, rel(out.cost_hist[sel], ref.cost_hist[sel], floor=0.0)
    assert rel(out.X[sel], ref.X[sel]) < 1e-5 and rel(out.K[sel], ref.K[sel]) < 1e-4


def test_time_system_fast_path_equals_generic_path():
    """the same time-system solve under ILQRB_NO_FAST_TIME=1 (generic k_backward<.., TIME> + k_linesearch + k_accept): both GPU paths
    must agree with each other as closely as each agrees with the oracle"""
    code = (
        "import numpy as np\n"
        "from ilqr_planner_b200 import synth\n"
        "from ilqr_planner_b200.solver import Handle\n"
        "pb = synth.config4(batch=32, horizon=100)\n"
        "with Handle(pb, device=0) as h:\n"
        "    out = h.solve_recursive(synth.u0_time(pb), 8, True, False)\n"
        "    prof = h.launch_count()\n"
        "np.savez(%r, cost_hist=out.cost_hist, alpha_hist=out.alpha_hist, X=out.X, U=out.U, K=out.K, launches=prof)\n")
    import os
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        fa, fb = os.path.join(td, "fast.npz"), os.path.join(td, "generic.npz")
        _subprocess_case(code % fa, {})
        _subprocess_case(code % fb, {"ILQRB_NO_FAST_TIME": "1"})
        a, b = np.load(fa), np.load(fb)
        assert int(a["launches"]) < int(b["launches"]), "the fast path launches 3 kernels per iteration, the generic one 4"
        same = pt.same_sequence(a["alpha_hist"], b["alpha_hist"])
        assert same.sum() >= 16, "the two GPU paths choose different steps in %d of 32 problems" % (~same).sum()
        assert rel(a["cost_hist"][same], b["cost_hist"][same], floor=0.0) < 1e-5
        assert rel(a["X"][same], b["X"][same]) < 1e-4


@pytest.mark.parametrize("case", ["posorn", "time"])
def test_batch_ilqr_woodbury_option(oracle_lib, case):
    """the optional Woodbury solve (ILQRB_BATCH_WOODBURY: one (keypoint steps x n)^2 system instead of the N x N factorisation; not the
    reference's algorithm) gives the dense path's answer to the accuracy either has, and the oracle's within the batch tolerances"""
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle
    from oracle import oracle
    if case == "posorn":
        pb, u0, nit = synth.config2(batch=6, horizon=60), None, 4
    else:
        pb = synth.config4(batch=6, horizon=40)
        u0, nit = synth.u0_time(pb), 5
    uu = None if u0 is None else u0.reshape(pb.batch, -1)
    with Handle(pb) as h:
        dense = h.solve_batch(uu, nit)
        dU, dC, dA = dense.U.copy(), dense.cost_hist.copy(), dense.alpha_hist.copy()
        wb = h.solve_batch(uu, nit, woodbury=True)
    ref = oracle.solve_batch(pb, uu, nit, False, threads=8)
    same = pt.same_sequence(wb.alpha_hist, dA) & pt.same_sequence(wb.alpha_hist, ref.alpha_hist)
    assert same.sum() >= pb.batch // 2
    # the reduced system I + W Su R^-1 Su^T carries 1/R = 1e5 in every entry: its answer is good to ~1e-7 of the cost (measured 3.6e-7
    # on the PosOrn case), two digits fewer than the dense factorisation — the price of the 10^4-fold cheaper solve
    tol_c, tol_u = (5e-6, 1e-3) if case == "posorn" else (1e-5, 1e-3)
    assert rel(wb.cost_hist[same], dC[same], floor=0.0) < tol_c and rel(wb.U[same], dU[same]) < tol_u
    assert rel(wb.cost_hist[same], ref.cost_hist[same], floor=0.0) < tol_c and rel(wb.U[same], ref.U[same]) < tol_u


def test_batch_ilqr_time_system_vs_oracle(oracle_lib):
    """C4 shape at a size the oracle finishes in seconds: PosOrnTime, N = 8 * 39 = 312"""
    from ilqr_planner_b200 import synth
    pb = synth.config4(batch=6, horizon=40)
    _check_batch(pb, synth.u0_time(pb), 5, oracle_lib)


def test_batch_ilqr_cp_time_system_vs_oracle(oracle_lib):
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    pb = synth.config4(batch=6, horizon=40)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    _check_batch(pb, synth.u0_time(pb), 8, oracle_lib, psi=psi)


def test_batch_ilqr_posorn_vs_oracle(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config2(batch=6, horizon=50)
    _check_batch(pb, None, 5, oracle_lib)
    from ilqr_planner_b200.solver import build_psi
    psi = np.kron(build_psi("rbf", pb.horizon - 1, 4), np.identity(pb.m))
    _check_batch(pb, None, 6, oracle_lib, psi=psi, early_stop=True)


def test_batch_ilqr_double_integrator_vs_oracle(oracle_lib):
    """nb_deriv = 2: the column blocks of Su are A..A B products (closed form on the device, matrix recursion in the oracle)"""
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    pb = synth.config3(batch=6, horizon=40)
    _check_batch(pb, None, 5, oracle_lib)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 3), np.identity(pb.m))
    _check_batch(pb, None, 6, oracle_lib, psi=psi)


def test_batch_ilqr_dual_task_vs_oracle(oracle_lib):
    """two tasks on a 14-joint state (C5 shape): stacked residual rows, keypoints of the two sub-systems at the same steps"""
    from ilqr_planner_b200 import synth
    pb = synth.config5(batch=4, horizon=30)
    _check_batch(pb, None, 4, oracle_lib)


def test_double_integrator_runtime_dof_and_active_limits(oracle_lib):
    """5-joint chain, nb_deriv = 2: runtime-size kernels of the double-integrator path; with tight soft limits the line search
    takes its exact per-step branch and the backward pass carries the limit terms"""
    from ilqr_planner_b200 import synth
    _check_recursive(synth.config3_short(batch=16, tight_limits=False), None, 6, oracle_lib)
    _check_recursive(synth.config3_short(batch=16, tight_limits=True), None, 4, oracle_lib)


def test_async_solve_on_two_handles_matches_blocking_call():
    """ilqrb_solve_recursive_async: batches streamed through two handles (one batch's download overlapping the next
    batch's solve) return bit for bit what the blocking call returns, for every output buffer"""
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.problem import ResultBuffers
    pbs = [synth.config2(batch=512, offset=512 * i) for i in range(4)]
    want = []
    for pb in pbs:
        with _handle(pb) as h:
            want.append(h.solve_recursive(None, 5, True, False))
    hs = [_handle(pbs[0]), _handle(pbs[1])]
    try:
        outs = [ResultBuffers(pbs[0], 5, "recursive", pinned=True), ResultBuffers(pbs[1], 5, "recursive", pinned=True)]
        u0 = np.zeros((512, 99 * 7))
        # batches 0 and 1 in flight together; results must equal the blocking ones
        for k in range(2):
            hs[k].solve_recursive_async(u0, 5, True, False, out=outs[k])
        for k in range(2):
            hs[k].synchronize()
            for f in ("X", "fX", "U", "K", "d", "cost", "cost_hist", "alpha_hist", "iters", "status"):
                assert np.array_equal(getattr(outs[k], f), getattr(want[k], f), equal_nan=True), (k, f)
        # the same handle and buffers reused for a second round after synchronize()
        for k in range(2):
            hs[k].solve_recursive_async(u0, 5, True, False, out=outs[k])
        for k in range(2):
            hs[k].synchronize()
            assert np.array_equal(outs[k].K, want[k].K) and np.array_equal(outs[k].cost_hist, want[k].cost_hist)
        # ilqrb_chain_after_compute: each handle's solve ordered behind the other's kernels (its copies still in flight)
        for r in range(3):
            for k in range(2):
                hs[k].synchronize()
                outs[k].K[:] = np.nan
                hs[k].chain_after_compute(hs[1 - k])
                hs[k].solve_recursive_async(u0, 5, True, False, out=outs[k])
        for k in range(2):
            hs[k].synchronize()
            assert np.array_equal(outs[k].K, want[k].K) and np.array_equal(outs[k].X, want[k].X) and np.array_equal(outs[k].cost, want[k].cost)
    finally:
        for h in hs:
            h.close()


def test_dense_two_tile_backward_pass_still_matches():
    """ILQRB_NO_BLOCKS=1 (read once per process, hence the subprocess): the dense 16x16 backward kernel that the
    block-diagonal pass replaces for composite robots stays correct and gives the same answer as the block pass"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from ilqr_planner_b200 import synth\n"
        "from ilqr_planner_b200.solver import Handle\n"
        "from oracle import oracle\n"
        "pb = synth.config5(batch=8)\n"
        "with Handle(pb, device=0) as h: out = h.solve_recursive(None, 5, True, False)\n"
        "ref = oracle.solve_recursive(pb, None, 5, True, False, threads=8)\n"
        "sc = np.abs(ref.cost_hist).max()\n"
        "assert np.abs(out.cost_hist - ref.cost_hist).max() / sc < 1e-9, np.abs(out.cost_hist - ref.cost_hist).max() / sc\n"
        "assert np.array_equal(out.alpha_hist, ref.alpha_hist)\n"
        "assert np.abs(out.K - ref.K).max() / np.abs(ref.K).max() < 1e-8\n"
        "print('dense ok')\n" % root)
    env = dict(os.environ, ILQRB_NO_BLOCKS="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "dense ok" in r.stdout, r.stdout + r.stderr


# ---------------------------------------------------------------------------------------------------------------
# (4) C4 at its own size, the dense solve against a quad-precision truth, the full-size workload against the oracle
def test_c4_at_its_own_size(oracle_lib):
    """BASELINE configs[3] as it is: PosOrnTime, H = 200, N = 1592 (25 block columns of the left-looking Cholesky): BatchILQR
    2 iterations and BatchILQRCP 3 iterations of 2 problems against the oracle (BatchILQR.cpp:110-173); then the same BatchILQR
    run with the workspace budget cut so that the two problems go through two chunks: bit-identical to the one-chunk run."""
    import os
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    pb = synth.config4(batch=2, horizon=200)
    u0 = synth.u0_time(pb)
    assert (pb.horizon - 1) * pb.m == 1592
    out, ref = _check_batch_shifts(pb, u0, 2, oracle_lib)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    _check_batch_shifts(pb, u0, 3, oracle_lib, psi=psi)
    old = os.environ.get("ILQRB_BATCH_WS_GB")
    os.environ["ILQRB_BATCH_WS_GB"] = "0.03"  # one problem needs 21 MB of N x N matrix plus the rest: chunks of one problem
    try:
        with _handle(pb) as h:
            o2 = h.solve_batch(u0, 2, False)
    finally:
        if old is None:
            del os.environ["ILQRB_BATCH_WS_GB"]
        else:
            os.environ["ILQRB_BATCH_WS_GB"] = old
    assert np.array_equal(o2.U, out.U) and np.array_equal(o2.cost_hist, out.cost_hist) and np.array_equal(o2.alpha_hist, out.alpha_hist)


def _check_batch_shifts(pb, u0, nb_iter, oracle, psi=None):
    """_check_batch with fewer oracle neighbours (the N = 1592 inverse of the oracle takes tens of seconds per iteration)"""
    with _handle(pb) as h:
        out = h.solve_batch(u0, nb_iter, False, psi=psi)
    sens = pt.measure(lambda p, v=None: oracle.solve_batch(p, u0, nb_iter, False, psi=psi, threads=8, variant=v), pb, BATCH_FIELDS,
                      neighbours=((0, "fma"), (+1, None), (-1, None)), truth=False)
    ref = sens.ref
    pt.check_sequences(out, sens)
    keep = ~np.isnan(ref.cost_hist) & (np.arange(ref.cost_hist.shape[1])[None, :] <= sens.first_tie()[:, None])
    err_c = rel(np.where(keep, out.cost_hist, 0.0), np.where(keep, ref.cost_hist, 0.0), floor=0.0)
    assert err_c < sens.tol("cost_hist"), "cost trace %.2e (bound %.2e; %s)" % (err_c, sens.tol("cost_hist"), sens.summary())
    free = sens.tie_free()
    e = sens.err("U", out, free)
    assert e < sens.tol("U"), "u %.2e (bound %.2e; %s)" % (e, sens.tol("U"), sens.summary())
    return out, ref


@pytest.mark.parametrize("case", ["time", "time_cp", "posorn", "posorn_cp", "double_integrator", "dual_arm"])
def test_batch_step_against_quad_truth(oracle_lib, case):
    """A truth for the dense solve: the first BatchILQR / BatchILQRCP step du of the GPU (Cholesky / pivoted elimination) and of the
    oracle (the reference's explicit LU inverse) against lstq_A du = lstq_B formed and solved in __float128 from the oracle's
    double-precision Su, J, Q, L, R (oracle_batch_step_truth).  The GPU must be as close to the truth as the reference's own
    arithmetic is, within a factor 4 (plus the effect of its own J, Su being a few ulps away from the oracle's)."""
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    psi = None
    if case.startswith("time"):
        pb = synth.config4(batch=4, horizon=40)
        u0 = synth.u0_time(pb)
        if case == "time_cp":
            psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    elif case.startswith("posorn"):
        pb = synth.config2(batch=4, horizon=50)
        u0 = None
        if case == "posorn_cp":
            psi = np.kron(build_psi("rbf", pb.horizon - 1, 4), np.identity(pb.m))
    elif case == "double_integrator":
        pb, u0 = synth.config3(batch=4, horizon=40), None
    else:
        pb, u0 = synth.config5(batch=4, horizon=30), None
    N = (pb.horizon - 1) * pb.m
    u0f = np.zeros((pb.batch, N)) if u0 is None else np.asarray(u0).reshape(pb.batch, N)
    with _handle(pb) as h:
        out = h.solve_batch(u0, 1, False, psi=psi)
    du_ref, du_q = oracle_lib.batch_step_truth(pb, u0, psi=psi, threads=8)
    alpha = out.alpha_hist[:, 0]
    du_gpu = (out.U - u0f) / alpha[:, None]
    # the oracle's inputs moved by one ulp of q0: what the truth itself moves by (the GPU's J, Su differ from the oracle's by that much)
    _, du_q1 = oracle_lib.batch_step_truth(pt.ulp_shift(pb, +1), u0, psi=psi, threads=8)
    nq = np.linalg.norm(du_q, axis=1)
    e_gpu = np.linalg.norm(du_gpu - du_q, axis=1) / nq
    e_ref = np.linalg.norm(du_ref - du_q, axis=1) / nq
    e_in = np.linalg.norm(du_q1 - du_q, axis=1) / nq
    # u = u0 + alpha du is rounded to double: (u - u0) / alpha recovers du to eps * |u| / (alpha |du|)
    e_round = 2.3e-16 * np.linalg.norm(out.U, axis=1) / (alpha * nq) * np.sqrt(N)
    print("%s: |du_gpu - du_q|/|du_q| %s  reference %s  1-ulp input change %s" % (case, e_gpu, e_ref, e_in))
    assert np.all(e_gpu <= 4 * np.maximum(e_ref, e_in) + e_round + 1e-15), (e_gpu, e_ref, e_in, e_round)


def test_full_size_parity_against_measured_instability(oracle_lib):
    """The bench workload itself (configs[1]: 4096 problems, T = 100, 20 iterations) on the GPU against the oracle, with the
    tie rule and the tolerances MEASURED on the oracle instead of assumed: the oracle is run three times, on q0 and on q0 one
    ulp up and one ulp down.  Asserted:
      (i)   every problem without a tie (all line-search margins above the measured instability threshold) has the oracle's
            step sequence over all 20 iterations, and every iteration's cost within max(1e-9, 4 x oracle noise) of ITSELF;
      (ii)  the set of problems where the GPU's sequence differs from the oracle's is not larger than the set where the
            oracle's own sequence differs from the oracle's under a 1-ulp input change (by more than 10 % + 8);
      (iii) no GPU sequence parts from the oracle's before that problem's first tie;
      (iv)  the MEDIAN problem meets the literal 1e-9 bar in every iteration."""
    import os
    from ilqr_planner_b200 import synth
    B, nit = 4096, 20
    pb = synth.config2(batch=B)
    with _handle(pb) as h:
        out = h.solve_recursive(None, nit, True, False)
    threads = os.cpu_count() or 8
    fields = dict(REC_FIELDS)
    sens = pt.measure(lambda p, v=None: oracle_lib.solve_recursive(p, None, nit, True, False, threads=threads, variant=v), pb, fields,
                      neighbours=((0, "fma"), (+1, None), (-1, None)), tie_min=0.0)
    ref = sens.ref
    print("oracle vs itself:", sens.summary())
    gpu_diff = ~pt.same_sequence(out.alpha_hist, ref.alpha_hist)
    self_diff = [int((~pt.same_sequence(r.alpha_hist, ref.alpha_hist)).sum()) for r in sens.perturbed]
    free = sens.tie_free()
    with np.errstate(invalid="ignore", divide="ignore"):
        dev = np.abs(out.cost_hist - ref.cost_hist) / np.abs(ref.cost_hist)
    med_gpu, med_self = np.nanmedian(dev, axis=0), np.nanmedian(sens.cost_noise, axis=0)
    print("GPU sequences differing from the oracle's: %d ; oracle neighbours (FMA build, +1 ulp, -1 ulp) from the oracle: %s ; tie-free problems %d" % (gpu_diff.sum(), self_diff, free.sum()))
    print("per-iteration relative cost difference, median over problems: GPU %s\n   oracle noise %s" % (np.array2string(med_gpu, precision=1), np.array2string(med_self, precision=1)))
    print("worst over tie-free problems: GPU %.2e, oracle noise %.2e" % (np.nanmax(dev[free]), np.nanmax(sens.cost_noise[free & sens.stable])))
    # (iii) and (i): sequences
    pt.check_sequences(out, sens)
    assert not gpu_diff[free].any(), "%d tie-free problems take a different step sequence" % gpu_diff[free].sum()
    assert free.sum() >= B // 2
    # (i): every iteration's cost against itself, (and trajectories, gains) on the tie-free problems
    for f in ("cost_hist", "X", "U", "fX", "K"):
        e = sens.err(f, out, free)
        assert e < sens.tol(f), "%s %.2e (bound %.2e)" % (f, e, sens.tol(f))
    print("from the long-double truth (GPU):", {k: "%.1e" % v for k, v in sens.check_truth(out, free).items()})
    # (ii)
    assert gpu_diff.sum() <= 1.1 * max(self_diff) + 8, (gpu_diff.sum(), self_diff)
    # (iv)
    assert np.all(med_gpu < RTOL), med_gpu
    # the threshold the small-batch tests use as their floor must cover what was measured here
    assert sens.tie <= pt.TIE_MEASURED_FULL_SIZE * 2, "tests/_parity.py TIE_MEASURED_FULL_SIZE (%.1e) is stale: measured %.1e" % (pt.TIE_MEASURED_FULL_SIZE, sens.tie)

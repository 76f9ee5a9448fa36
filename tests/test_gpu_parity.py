"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI, against
  (1) the reference's stored notebook traces (tests/golden/notebook_goldens.json) and
  (2) the CPU oracle on seeded synthetic batches of every BASELINE.json config shape.

Tolerances (north_star: per iteration the same accepted alpha sequence; trajectories, gains and cost within 1e-9 relative):
  cost / X / U / K / fX : 1e-9 relative to the array's max magnitude
  d (feed-forward)      : 1e-9 relative to max|U| + max|d|  (d -> 0 at convergence, so it is scaled by the control it corrects)
  alpha                 : exact, for every iteration whose oracle decision is not a declared tie (see `decisive_prefix`)
"""
import numpy as np
import pytest

from tests import _cases as tc

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def _handle(pb):
    from ilqr_planner_b200.solver import Handle
    return Handle(pb, device=0)


def rel(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape
    nan_a, nan_b = np.isnan(a), np.isnan(b)
    assert np.array_equal(nan_a, nan_b), "NaN pattern differs"
    if nan_a.all():
        return 0.0
    s = scale if scale is not None else max(np.nanmax(np.abs(b)), 1e-300)
    return float(np.nanmax(np.abs(a - b)) / s)


TIE = 1e-9  # = the cost tolerance: a line-search comparison closer than this is not decided by the algorithm but by rounding


def decisive_prefix(ref, b):
    """Number of leading iterations of problem b whose line-search decisions are not ties.

    The oracle reports, per iteration, the smallest relative distance |candidate cost - incumbent cost| / |incumbent|
    over the candidates its sequential line search evaluated (`margin_hist`, oracle/ilqr_oracle.cpp trackMargin).
    `newCost >= cost0` with a margin below the 1e-9 cost tolerance is decided by FP64 rounding, and
    after such a tie the two runs may follow different (equally valid) branches: alpha sequences are compared up to the
    first tie, costs and trajectories over the whole solve."""
    n = int(ref.iters[b])
    for i in range(n):
        mg = ref.margin_hist[b, i]
        if np.isnan(ref.cost_hist[b, i]) or not (mg >= TIE):
            return i
    return n


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
def _check_recursive(pb, u0, nb_iter, oracle, early_stop=False):
    with _handle(pb) as h:
        out = h.solve_recursive(u0, nb_iter, True, early_stop)
    ref = oracle.solve_recursive(pb, u0, nb_iter, True, early_stop, threads=8)
    assert rel(out.cost_hist, ref.cost_hist) < RTOL
    # the final cost is held to the same absolute scale as the trace (a cost of 1e-7 is a sum of squared residuals of
    # 3e-4: a 1e-12 trajectory difference already moves it by 1e-8 of itself)
    assert rel(out.cost, ref.cost, scale=float(np.nanmax(np.abs(ref.cost_hist)))) < RTOL
    # problems whose every line-search decision is decisive must match in everything the reference returns;
    # after a declared tie the two runs may sit on different (equally valid) branches: those problems are held to the
    # cost tolerance above and to a loose trajectory bound only.
    dec = np.array([decisive_prefix(ref, b) == int(ref.iters[b]) for b in range(pb.batch)])
    assert dec.sum() >= pb.batch // 2, "too few decisive problems (%d of %d) for this test to mean anything" % (dec.sum(), pb.batch)
    for b in range(pb.batch):
        k = decisive_prefix(ref, b)
        assert np.array_equal(out.alpha_hist[b, :k], ref.alpha_hist[b, :k]), "problem %d alpha sequence" % b
    assert np.array_equal(out.iters[dec], ref.iters[dec])
    assert rel(out.X[dec], ref.X[dec]) < RTOL
    assert rel(out.U[dec], ref.U[dec]) < 10 * RTOL
    assert rel(out.fX[dec], ref.fX[dec]) < RTOL
    assert rel(out.K[dec], ref.K[dec]) < 10 * RTOL
    assert rel(out.d[dec], ref.d[dec], scale=np.abs(ref.U).max() + np.abs(ref.d).max()) < RTOL
    if (~dec).any():
        assert rel(out.X[~dec], ref.X[~dec]) < 1e-6
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


def test_config3_al_ilqr(oracle_lib):
    from ilqr_planner_b200 import synth
    pb = synth.config3(batch=12)
    al = synth.al_config3()
    nb = 12
    with _handle(pb) as h:
        out = h.solve_al(None, al["A"], al["b"], al["lambda0"], nb, al["lag_update_step"], al["penalty"], al["scaling_factor"], True, False)
    ref = oracle_lib.solve_al(pb, None, al["A"], al["b"], al["lambda0"], nb, al["lag_update_step"], al["penalty"], al["scaling_factor"], True, False, threads=8)
    assert rel(out.cost_hist, ref.cost_hist) < RTOL
    for b in range(pb.batch):
        k = decisive_prefix(ref, b)
        assert np.array_equal(out.alpha_hist[b, :k], ref.alpha_hist[b, :k])
    assert rel(out.X, ref.X) < RTOL and rel(out.U, ref.U) < 10 * RTOL
    assert rel(out.lam, ref.lam, scale=max(np.abs(ref.lam).max(), 1e-6)) < 1e-8


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


def _check_batch(pb, u0, nb_iter, oracle, psi=None, early_stop=False, ctol=1e-8, utol=1e-7):
    """Tolerances: the step comes out of a dense solve whose matrix R + Su^T W Su has condition number 1e5 (R = 1e-5
    against O(1) task curvature) to 1e7 (time systems: the time column of B is large).  The reference forms an explicit
    LU inverse there (relative error ~ cond * eps), this path a Cholesky / pivoted elimination: both answers are equally
    valid to ~cond * eps, and the difference compounds over iterations.  Hence 1e-8 / 1e-7 (cost / u) for plain systems
    and 1e-6 / 1e-5 for time systems instead of the Recursive solver's 1e-9."""
    with _handle(pb) as h:
        out = h.solve_batch(u0, nb_iter, early_stop, psi=psi)
    ref = oracle.solve_batch(pb, u0, nb_iter, early_stop, psi=psi, threads=8)
    assert rel(out.cost_hist, ref.cost_hist) < ctol
    for b in range(pb.batch):
        k = decisive_prefix(ref, b)
        assert np.array_equal(out.alpha_hist[b, :k], ref.alpha_hist[b, :k]), "problem %d alpha sequence" % b
    dec = np.array([decisive_prefix(ref, b) == int(ref.iters[b]) for b in range(pb.batch)])
    assert dec.sum() >= pb.batch // 2
    # u comes out of a dense N x N solve with condition number ~1e5 (R = 1e-5 against O(1) task curvature)
    assert rel(out.U[dec], ref.U[dec]) < utol
    return out, ref


def test_batch_ilqr_time_system_vs_oracle(oracle_lib):
    """C4 shape at a size the oracle finishes in seconds: PosOrnTime, N = 8 * 39 = 312"""
    from ilqr_planner_b200 import synth
    pb = synth.config4(batch=6, horizon=40)
    _check_batch(pb, synth.u0_time(pb), 5, oracle_lib, ctol=1e-6, utol=1e-5)


def test_batch_ilqr_cp_time_system_vs_oracle(oracle_lib):
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    pb = synth.config4(batch=6, horizon=40)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    _check_batch(pb, synth.u0_time(pb), 8, oracle_lib, psi=psi, ctol=1e-6, utol=1e-5)


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
    _check_batch(pb, None, 5, oracle_lib, ctol=1e-6, utol=1e-5)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 3), np.identity(pb.m))
    _check_batch(pb, None, 6, oracle_lib, psi=psi, ctol=1e-6, utol=1e-5)


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

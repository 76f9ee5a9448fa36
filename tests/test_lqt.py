"""solver::LQT (lqt.cpp:16-126).  The reference holds no test or tutorial for this class and cannot be built here, so the
oracle's restatement is pinned on an independent statement of the same mathematics: numpy's solve of the dense system solveLinAl
forms, and the DP recursion written with numpy matrices.  GPU: ilqrb_lqt_solve / ilqrb_lqt_command against the oracle through the C
ABI, the C++ class through PyLQR."""
import numpy as np
import pytest

from oracle import oracle


def integrator(nb_deriv, dim, dt=0.01):
    """chain of integrators in `dim` axes (the tutorials' linear system)"""
    n, m = nb_deriv * dim, dim
    A1 = np.zeros((nb_deriv, nb_deriv))
    B1 = np.zeros((nb_deriv, 1))
    for i in range(nb_deriv):
        for j in range(i, nb_deriv):
            A1[i, j] = dt ** (j - i) / float(np.prod(np.arange(1, j - i + 1))) if j > i else 1.0
        B1[nb_deriv - 1 - i, 0] = dt ** (i + 1) / float(np.prod(np.arange(1, i + 2)))
    return np.kron(A1, np.eye(dim)), np.kron(B1, np.eye(dim)), n, m


def problems(nq, T, nb_deriv, dim, seed, dense_q=False):
    rng = np.random.default_rng(seed)
    A, B, n, m = integrator(nb_deriv, dim)
    Qs = np.zeros((nq, T, n, n))
    for q in range(nq):
        for t in sorted(set([T - 1] + list(rng.integers(1, T, size=3)))):
            if dense_q:
                L = rng.standard_normal((n, n))
                Qs[q, t] = L @ L.T + np.eye(n)
            else:
                Qs[q, t] = np.diag(np.concatenate([np.full(dim, 10.0 ** rng.integers(0, 4))] + [np.full(dim, 1e-2)] * (nb_deriv - 1)))
    mu = rng.standard_normal((nq, T * n))
    r = float(np.float32(0.1)) ** nb_deriv  # lqt.cpp:17 with the constructor's float rfactor
    return A, B, Qs, mu, r, n, m


def numpy_lqt(A, B, Qs, mu, r):
    T, n = Qs.shape[0], A.shape[0]
    m = B.shape[1]
    Su = np.zeros((T * n, (T - 1) * m))
    Sx = np.zeros((T * n, n))
    Q = np.zeros((T * n, T * n))
    for t in range(T):
        Sx[t * n:(t + 1) * n] = np.linalg.matrix_power(A, t)
        Q[t * n:(t + 1) * n, t * n:(t + 1) * n] = Qs[t]
        for j in range(t):
            Su[t * n:(t + 1) * n, j * m:(j + 1) * m] = np.linalg.matrix_power(A, t - 1 - j) @ B
    x0 = mu[:n]
    u = np.linalg.solve(Su.T @ Q @ Su + r * np.eye((T - 1) * m), Su.T @ Q @ (mu - Sx @ x0))
    P = [None] * T
    d = [None] * T
    P[T - 1], d[T - 1] = Qs[T - 1], np.zeros(n)
    Rt = r * np.eye(m)
    for i in range(T - 2, -1, -1):
        Gi = np.linalg.inv(B.T @ P[i + 1] @ B + Rt)
        P[i] = Qs[i] - A.T @ (P[i + 1] @ B @ Gi @ B.T @ P[i + 1] - P[i + 1]) @ A
        d[i] = (A.T - A.T @ P[i + 1] @ B @ Gi @ B.T) @ (P[i + 1] @ (A @ mu[i * n:(i + 1) * n] - mu[(i + 1) * n:(i + 2) * n]) + d[i + 1])
    return u, Su @ u + Sx @ x0, np.array(P), np.array(d)


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.mark.parametrize("nb_deriv,dim,T", [(2, 2, 30), (1, 3, 20), (2, 7, 12)])
def test_oracle_lqt_against_numpy(nb_deriv, dim, T):
    A, B, Qs, mu, r, n, m = problems(2, T, nb_deriv, dim, seed=3)
    o = oracle.lqt(A, B, Qs, mu, r)
    for q in range(2):
        u, xh, P, d = numpy_lqt(A, B, Qs[q], mu[q], r)
        assert rel(o["u"][q], u) < 1e-9
        assert rel(o["xhat"][q], xh) < 1e-9
        assert rel(o["Ps"][q], P) < 1e-10
        assert rel(o["ds"][q], d) < 1e-10


def test_oracle_lqt_command_follows_the_reference_formula():
    """lqt.cpp:105-124: P, d and BOTH mu segments of timestep + 1"""
    A, B, Qs, mu, r, n, m = problems(1, 15, 2, 2, seed=5)
    o = oracle.lqt(A, B, Qs, mu, r)
    P, d = o["Ps"][0], o["ds"][0]
    for ts in (0, 6, 13):
        t = ts + 1
        Gi = np.linalg.inv(B.T @ P[t] @ B + r * np.eye(m))
        mt = mu[0, t * n:(t + 1) * n]
        assert rel(o["K"][0, ts], Gi @ B.T @ P[t] @ A) < 1e-10
        assert rel(o["f"][0, ts], -Gi @ B.T @ (P[t] @ (A @ mt - mt) + d[t])) < 1e-10


# ---- GPU ---------------------------------------------------------------------------------------------------------------------------
LQT_CASES = {"double_integrator_2d": (2, 2, 60, 64, False), "single_integrator_3d": (1, 3, 40, 16, False), "7_axes_second_order": (2, 7, 100, 8, False),
             "dense_weights": (2, 3, 50, 16, True), "third_order": (3, 2, 30, 8, False), "two_states": (2, 2, 2, 4, False)}


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(LQT_CASES))
def test_lqt_gpu_vs_oracle(case):
    from ilqr_planner_b200 import solver

    nb_deriv, dim, T, nq, dense_q = LQT_CASES[case]
    A, B, Qs, mu, r, n, m = problems(nq, T, nb_deriv, dim, seed=11, dense_q=dense_q)
    g = solver.lqt_solve(A, B, Qs, mu, r)
    nd = min(nq, 4)  # the oracle's dense (T-1)m inverse on a few problems, its DP on all of them
    o = oracle.lqt(A, B, Qs, mu, r, dense=False)
    od = oracle.lqt(A, B, Qs[:nd], mu[:nd], r, dense=True)
    assert (g["status"] == 0).all()
    # tolerance: the recursion multiplies T matrices of condition up to ~1e5 (weights 1e3 against r = 1e-2 .. 1e-3); measured 1e-12 .. 1e-10
    for k in ("Ps", "ds", "K", "f"):
        assert rel(g[k], o[k]) < 1e-8, (k, rel(g[k], o[k]))
    assert rel(g["u"][:nd], od["u"]) < 1e-7, rel(g["u"][:nd], od["u"])
    assert rel(g["xhat"][:nd], od["xhat"]) < 1e-8
    # getCommand(timestep, state)
    rng = np.random.default_rng(2)
    nc = 20
    pr = rng.integers(0, nq, size=nc).astype(np.int32)
    ts = rng.integers(0, T - 1, size=nc).astype(np.int32)
    st = rng.standard_normal((nc, n))
    uc = solver.lqt_command(g["K"], g["f"], mu, ts, st, problem=pr)
    for c in range(nc):
        want = o["K"][pr[c], ts[c]] @ (mu[pr[c], (ts[c] + 1) * n:(ts[c] + 2) * n] - st[c]) + o["f"][pr[c], ts[c]]
        assert np.max(np.abs(uc[c] - want)) <= 1e-8 * max(1.0, np.max(np.abs(want)))


@pytest.mark.gpu
def test_lqt_per_problem_dynamics_and_errors():
    from ilqr_planner_b200 import solver

    A, B, Qs, mu, r, n, m = problems(6, 25, 2, 2, seed=4)
    rng = np.random.default_rng(0)
    Aq = np.stack([A + 1e-2 * rng.standard_normal(A.shape) for _ in range(6)])
    Bq = np.stack([B * (1 + 0.1 * q) for q in range(6)])
    g = solver.lqt_solve(Aq, Bq, Qs, mu, r)
    for q in range(6):
        o = oracle.lqt(Aq[q], Bq[q], Qs[q:q + 1], mu[q:q + 1], r)
        assert rel(g["u"][q], o["u"][0]) < 1e-7
        assert rel(g["Ps"][q], o["Ps"][0]) < 1e-8
    with pytest.raises(RuntimeError):
        solver.lqt_command(g["K"], g["f"], mu, [24], np.zeros((1, n)))  # timestep nb_states - 1 has no command (lqt.cpp:106-114)
    with pytest.raises(RuntimeError):
        solver.lqt_solve(np.eye(17), np.ones((17, 1)), np.zeros((1, 3, 17, 17)), np.zeros((1, 51)), r)


@pytest.mark.gpu
def test_lqt_through_pylqr():
    import PyLQR

    A, B, Qs, mu, r, n, m = problems(3, 40, 2, 2, seed=8)
    o = oracle.lqt(A, B, Qs, mu, r)
    lqt = PyLQR.solver.LQT(A, B, [Qs[0, t] for t in range(40)], mu[0], 0.1, 2)
    assert lqt.get_nb_states() == 40
    with pytest.raises(RuntimeError, match="solveLinal"):
        lqt.get_command(0)
    with pytest.raises(RuntimeError, match="solveDP"):
        lqt.get_command(0, np.zeros(n))
    lqt.solve_lin_al()
    u = np.concatenate([lqt.get_command(t) for t in range(39)])
    assert rel(u, o["u"][0]) < 1e-7
    assert rel(lqt.get_predicted_states(), o["xhat"][0]) < 1e-8
    lqt.solve_DP()
    x = np.linspace(-1, 1, n)
    for ts in (0, 17, 38):
        want = o["K"][0, ts] @ (mu[0, (ts + 1) * n:(ts + 2) * n] - x) + o["f"][0, ts]
        assert np.max(np.abs(lqt.get_command(ts, x) - want)) <= 1e-8 * max(1.0, np.max(np.abs(want)))
    with pytest.raises(IndexError):
        lqt.get_command(39, x)
    assert rel(np.array(lqt.get_Ps()), o["Ps"][0]) < 1e-8
    b = PyLQR.solver.LQT.solve_batch(A, B, Qs, mu, 0.1, 2)
    assert rel(b["u"], o["u"]) < 1e-7 and rel(b["ds"], o["ds"]) < 1e-8

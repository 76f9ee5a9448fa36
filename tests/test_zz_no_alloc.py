"""ilqrb_kinematics_static — the kinematics hook that allocates nothing (its device staging area is a static array inside the
library; include/ilqrb.h) — against the existing ilqrb_kinematics, through the C ABI, on two calls in a row.

Collected last (pytest -x must not let a failure here hide the other tests); passes alone in a fresh
process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call_static(h, robot, q, dq):
    from ilqr_planner_b200 import cabi
    dof = h.pb.robots[robot].dof
    q = np.ascontiguousarray(np.asarray(q, dtype=np.float64).reshape(-1, dof))
    nq = q.shape[0]
    dq = None if dq is None else np.ascontiguousarray(np.asarray(dq, dtype=np.float64).reshape(-1, dof))
    x, quat, J, dx, w = np.full((nq, 3), np.nan), np.full((nq, 4), np.nan), np.full((nq, 6, dof), np.nan), np.full((nq, 3), np.nan), np.full((nq, 3), np.nan)
    rc = h.lib.ilqrb_kinematics_static(h.h, robot, nq, cabi.dptr(q), cabi.dptr(dq), cabi.dptr(x), cabi.dptr(quat), cabi.dptr(J), cabi.dptr(dx), cabi.dptr(w))
    return rc, (x, quat, J, dx, w)


def test_static_scratch_kinematics_matches_allocating_entry_point():
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import Handle
    from tests._parity import rel
    pb = synth.config3(batch=4)
    rng = np.random.default_rng(11)
    with Handle(pb, device=0) as h:
        # call 1: the largest size the static area takes, with joint velocities
        q1, dq1 = rng.uniform(-2, 2, (cabi.KIN_STATIC_MAX_NQ, 7)), rng.uniform(-1, 1, (cabi.KIN_STATIC_MAX_NQ, 7))
        rc, got1 = _call_static(h, 0, q1, dq1)
        assert rc == 0, h.lib.ilqrb_last_error()
        want1 = h.kinematics(0, q1, dq1)
        # call 2, right after: fewer configurations and dq = NULL — the array still holds call 1's velocities, which must not leak
        q2 = rng.uniform(-2, 2, (100, 7))
        rc, got2 = _call_static(h, 0, q2, None)
        assert rc == 0, h.lib.ilqrb_last_error()
        want2 = h.kinematics(0, q2, None)
        # call 3: the first inputs again give the first answer again
        rc, got3 = _call_static(h, 0, q1, dq1)
        assert rc == 0
        for got, want in ((got1, want1), (got2, want2), (got3, want1)):
            for a, b in zip(got, want):
                assert not np.isnan(a).any()
                assert rel(a, b) < 1e-12  # the tolerance of test_kinematics_and_rollout_hooks (same kernel: equal in fact)
        assert np.all(got2[3] == 0) and np.all(got2[4] == 0)  # dx, w with zero joint velocities
        # beyond the static area: a clear error, nothing written
        rc, _ = _call_static(h, 0, np.zeros((cabi.KIN_STATIC_MAX_NQ + 1, 7)), None)
        assert rc == -1 and b"ILQRB_KIN_STATIC_MAX_NQ" in h.lib.ilqrb_last_error()


def test_static_scratch_kinematics_matches_oracle(oracle_lib):
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle
    from tests._parity import rel
    pb = synth.config2(batch=2)
    rng = np.random.default_rng(12)
    q, dq = rng.uniform(-2, 2, (64, 7)), rng.uniform(-1, 1, (64, 7))
    with Handle(pb, device=0) as h:
        rc, got = _call_static(h, 0, q, dq)
    assert rc == 0
    for a, b in zip(got, oracle_lib.kinematics(pb, 0, q, dq)):
        assert rel(a, b) < 1e-12

"""SURVEY 8f N3 on the GPU: joint-space systems (JointSpacePlannerSys / JointSpaceTimePlannerSys with Angular(Time)Keypoints),
SequentialSystems mixing a joint-space and an end-effector sub-system (the HYBRID_* notebooks' structure) and dead-zone keypoints
(PosOrnKeypointDistFunct), against the oracle with the measures of tests/_parity.py.

Pinning: the four notebooks that use these classes draw their joint targets with an UNSEEDED np.random.uniform, so their stored
traces cannot be reproduced by anyone; the oracle's restatement of these classes is checked for the notebooks' trace SHAPE only
(tests/test_oracle_goldens.py) — parity for N3 is oracle-anchored, unpinned."""
import numpy as np
import pytest

from tests.test_gpu_parity import _check_batch, _check_recursive

pytestmark = pytest.mark.gpu


def _u0(pb):
    if not pb.time:
        return None
    u = np.zeros((pb.batch, pb.horizon - 1, pb.m))
    u[:, :, -1] = 0.1  # the notebooks' initial sqrt(dt)
    return u


CASES = {
    "joint_space": lambda s: s.config_joint_space(batch=16),
    "joint_space_time": lambda s: s.config_joint_space(batch=16, time=True),
    "hybrid": lambda s: s.config_hybrid(batch=16),
    "hybrid_time": lambda s: s.config_hybrid(batch=16, time=True),
    "deadzone": lambda s: s.config2_deadzone(batch=16),
    "deadzone_2nd_order": lambda s: s.config2_deadzone(batch=16, nb_deriv=2),
}


@pytest.mark.parametrize("case", list(CASES))
def test_n3_recursive_vs_oracle(oracle_lib, case):
    from ilqr_planner_b200 import synth
    pb = CASES[case](synth)
    _check_recursive(pb, _u0(pb), 5, oracle_lib, min_tie_free=0.25)


@pytest.mark.parametrize("case", ["joint_space", "joint_space_time", "hybrid", "deadzone"])
def test_n3_batch_solvers_vs_oracle(oracle_lib, case):
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import build_psi
    pb = CASES[case](synth).slice(0, 6)
    pb.horizon = pb.horizon  # N = (H - 1) m = 693 .. 952: seconds for the oracle
    # a joint-space system without a time state is linear-quadratic: the first step solves it, every later decision is a tie
    nit = 1 if case == "joint_space" else 3
    _check_batch(pb, _u0(pb), nit, oracle_lib)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    _check_batch(pb, _u0(pb), nit if case == "joint_space" else 4, oracle_lib, psi=psi)


def test_n3_refusals_are_loud():
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import Handle
    pb = synth.config_joint_space(batch=2)
    pb.nb_deriv = 2
    pb.tasks[0].kp_precision = np.array([[np.identity(14), np.identity(14)]])
    pb.tasks[0].state_max = pb.tasks[0].state_min = None
    with pytest.raises(RuntimeError, match="nb_deriv = 2 is not built"):
        Handle(pb)
    pb = synth.config4(batch=2, horizon=40)
    pb.tasks[0].kp_deadzone = np.zeros((2, 4))
    with pytest.raises(RuntimeError, match="no time variant"):
        Handle(pb)

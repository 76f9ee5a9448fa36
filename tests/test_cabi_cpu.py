"""CPU-side checks of the C ABI library: it loads, exports every symbol include/ilqrb.h declares, its host builders
(URDF -> chain, control primitives) agree with the independent test-side restatements, and it refuses to run
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from ilqr_planner_b200 import cabi, synth
from ilqr_planner_b200.solver import Handle, build_psi, chain_from_urdf
from tests.golden import psi_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    lib = cabi.load()
    hdr = open(os.path.join(ROOT, "include", "ilqrb.h")).read()
    declared = set(re.findall(r"\b(ilqrb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(cabi.SYMBOLS), declared ^ set(cabi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.ilqrb_abi_version() == cabi.ABI_VERSION


def test_struct_sizes_match_header_layout():
    assert C.sizeof(cabi.Segment) == 8 + 18 * 8
    assert C.sizeof(cabi.Robot) == 16 + 2 * 8
    assert C.sizeof(cabi.SolveOpts) == 16
    assert C.sizeof(cabi.Result) == 11 * 8


def test_chain_from_urdf_matches_test_side_parser():
    from oracle import urdf_chain
    text = open(synth.PANDA_URDF).read()
    for rpy, xyz in [((0, 0, 0), (0, 0, 0)), ((0.3, -0.2, 0.1), (0.01, 0.02, 0.03))]:
        segs, dof = chain_from_urdf(text, "panda_link0", "panda_tip", rpy, xyz)
        ref = urdf_chain.kdl_chain_from_urdf(text, "panda_link0", "panda_tip", rpy, xyz)
        assert dof == 7 and len(segs) == len(ref) == 11
        for a, b in zip(segs, ref):
            assert a["joint_type"] == b["joint_type"]
            for k in ("axis", "origin", "tip_R", "tip_p"):
                np.testing.assert_allclose(np.ravel(a[k]), np.ravel(b[k]), rtol=0, atol=1e-16)
    segs, dof = chain_from_urdf(synth.PANDA_URDF, "panda_link2", "panda_link6", is_path=True)
    assert dof == 4 and len(segs) == 5


def test_chain_matches_reference_tutorial_chain():
    """the kinematics-only URDF shipped here gives the same chain as the golden JSON's (reference model.urdf) chain"""
    from tests import _cases as tc
    g = tc.goldens()
    ref = g["chains"]["model.urdf:panda_link0->panda_tip:[0.0, 0.0, 0.0]:[0.0, 0.0, 0.0]"]
    segs = synth.panda_chain()
    assert len(segs) == len(ref)
    for a, b in zip(segs, ref):
        assert a["joint_type"] == b["joint_type"]
        for k in ("axis", "origin", "tip_R", "tip_p"):
            assert np.array_equal(np.ravel(a[k]), np.ravel(b[k])), k


def test_urdf_errors_carry_reference_messages():
    text = open(synth.PANDA_URDF).read()
    bad_axis = text.replace('<axis xyz="0 0 1"/>', '<axis xyz="0 1"/>', 1)
    with pytest.raises(RuntimeError, match="Expected a 3d array got less"):  # TUP tests/wrong_robot.urdf case
        chain_from_urdf(bad_axis, "panda_link0", "panda_tip")
    with pytest.raises(RuntimeError, match=r"\[KDLRobot\] Unable to build kinematic chain from panda_link0 to nowhere"):
        chain_from_urdf(synth.PANDA_URDF, "panda_link0", "nowhere", is_path=True)
    with pytest.raises(RuntimeError, match="Root element should be named <robot>"):
        chain_from_urdf("<model><joint/></model>", "a", "b")
    with pytest.raises(RuntimeError, match="Unknown joint type"):
        chain_from_urdf(text.replace('type="revolute"', 'type="floating"', 1), "panda_link0", "panda_tip")


@pytest.mark.parametrize("kind", ["rbf", "bernstein", "unitstep", "sawtooth", "linear"])
def test_build_psi(kind):
    for dim, K in [(99, 2), (399, 2), (49, 3), (199, 4)]:
        a, b = build_psi(kind, dim, K), psi_ref.build_psi(kind, dim, K)
        assert a.shape == b.shape
        if kind in ("unitstep", "sawtooth", "linear"):
            assert np.array_equal(a, b)
        else:
            np.testing.assert_allclose(a, b, rtol=1e-14, atol=0)


def test_get_dims_and_descriptor_validation():
    lib = cabi.load()
    for pb, exp in [(synth.config2(batch=2), (7, 7, 7, 6)), (synth.config3(batch=2), (14, 7, 14, 12)), (synth.config4(batch=2), (8, 8, 8, 7)),
                    (synth.config5(batch=2), (14, 14, 14, 12)), (synth.config1(batch=2), (3, 3, 2, 2))]:
        d = pb.to_desc()
        dims = cabi.Dims()
        cabi.check(lib.ilqrb_get_dims(C.byref(d), C.byref(dims)))
        assert (dims.n, dims.m, dims.nt, dims.r) == exp
        assert (pb.n, pb.m, pb.nt, pb.r) == exp
    pb = synth.config2(batch=2)
    d = pb.to_desc()
    d.nb_deriv = 3
    assert lib.ilqrb_get_dims(C.byref(d), C.byref(cabi.Dims())) == -1
    assert b"nb_deriv" in lib.ilqrb_last_error()


def test_no_cpu_fallback():
    """without a CUDA device the product must fail loudly instead of computing on the host"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Handle(synth.config2(batch=2))


def test_multi_plan_slices_tile_the_batch():
    """ilqrb_multi_plan (the slicing of ilqrb_create_multi): contiguous, complete, sizes within one of each other"""
    from ilqr_planner_b200.solver import multi_plan
    for B, G in ((65536, 8), (4096, 3), (7, 7), (10, 4), (1, 1), (4097, 8)):
        sl = multi_plan(B, G)
        assert sl[0][0] == 0 and sl[-1][1] == B and len(sl) == G
        assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))
        sizes = [hi - lo for lo, hi in sl]
        assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1
    with pytest.raises(RuntimeError, match="more slices than problems"):
        multi_plan(3, 4)


def test_create_multi_refuses_without_a_gpu_or_with_bad_devices():
    import torch
    from ilqr_planner_b200.solver import MultiHandle
    pb = synth.config2(batch=8)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CUDA device"):
            MultiHandle(pb, [0, 1])

"""ilqrb_create_multi: one batch sharded over several devices.  With one GPU the two slices share device 0 (the
slicing, the per-slice descriptors, the worker threads and the gather are the same code); with two or more GPUs the slices go
to different devices, the results are also gathered device-resident on GPU 0, and a second device runs the kernels that need
more than 48 KB of dynamic shared memory (the opt-in is per device)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FIELDS = ("X", "fX", "U", "K", "d", "cost", "cost_hist", "alpha_hist", "iters", "status")


def _ndev():
    import torch
    return torch.cuda.device_count()


def _devices():
    return [[0, 0], [0, 0, 0]] + ([[0, 1], list(range(min(_ndev(), 8)))] if _ndev() >= 2 else [])


def test_multi_recursive_equals_single_handle_bit_for_bit():
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle, MultiHandle
    pb = synth.config2(batch=301)  # uneven slices
    with Handle(pb, device=0) as h:
        want = h.solve_recursive(None, 5, True, False)
    for devs in _devices():
        with MultiHandle(pb, devs) as mh:
            sl = mh.slices()
            assert [s[0] for s in sl] == devs and sl[0][1] == 0 and sl[-1][2] == 301
            got = mh.solve_recursive(None, 5, True, False)
            for f in FIELDS:
                assert np.array_equal(getattr(got, f), getattr(want, f), equal_nan=True), (devs, f)
            # split form: upload / run on every device / synchronize / gather, twice in a row on the same handle
            from ilqr_planner_b200.problem import ResultBuffers
            for _ in range(2):
                out = ResultBuffers(pb, 5, "recursive")
                mh.upload_u0(None)
                mh.run_recursive(5, True, False)
                mh.synchronize()
                mh.download(out, 5)
                for f in FIELDS:
                    assert np.array_equal(getattr(out, f), getattr(want, f), equal_nan=True), (devs, f)


def test_multi_al_and_batch_solvers_equal_single_handle():
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle, MultiHandle, build_psi
    devs = [0, 1] if _ndev() >= 2 else [0, 0]
    pb = synth.config3_active(batch=33)
    al = synth.al_config3_active("mixed", lag_update_step=2)
    args = (None, al["A"], al["b"], al["lambda0"], 6, 2, al["penalty"], al["scaling_factor"], True, False)
    with Handle(pb, device=0) as h:
        want = h.solve_al(*args)
    with MultiHandle(pb, devs) as mh:
        got = mh.solve_al(*args)
    for f in ("X", "fX", "U", "cost_hist", "alpha_hist", "lam"):
        assert np.array_equal(getattr(got, f), getattr(want, f), equal_nan=True), f
    pb = synth.config4(batch=5, horizon=40)
    u0 = synth.u0_time(pb)
    psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
    for kw in ({}, {"psi": psi}):
        with Handle(pb, device=0) as h:
            want = h.solve_batch(u0, 3, False, **kw)
        with MultiHandle(pb, devs) as mh:
            got = mh.solve_batch(u0, 3, False, **kw)
        for f in ("U", "cost_hist", "alpha_hist"):
            assert np.array_equal(getattr(got, f), getattr(want, f), equal_nan=True), f


def test_device_resident_gather_to_gpu0():
    """the final gather without the host: every device's slice lands in whole-batch buffers on GPU 0 (peer copies)"""
    import torch
    from ilqr_planner_b200 import cabi, synth
    from ilqr_planner_b200.solver import Handle, MultiHandle
    devs = [0, 1] if _ndev() >= 2 else [0, 0]
    pb = synth.config2(batch=64)
    with Handle(pb, device=0) as h:
        want = h.solve_recursive(None, 4, True, False)
    B, H, n, m = 64, pb.horizon, pb.n, pb.m
    dev = torch.device("cuda", 0)
    t = {"X": torch.full((B, H, n), float("nan"), dtype=torch.float64, device=dev), "U": torch.full((B, H - 1, m), float("nan"), dtype=torch.float64, device=dev),
         "K": torch.full((B, H - 1, m, n), float("nan"), dtype=torch.float64, device=dev), "cost": torch.full((B,), float("nan"), dtype=torch.float64, device=dev),
         "cost_hist": torch.full((B, 4), float("nan"), dtype=torch.float64, device=dev)}
    res = cabi.Result()
    for k, v in t.items():
        setattr(res, k, C.cast(v.data_ptr(), cabi.c_double_p))
    with MultiHandle(pb, devs) as mh:
        mh.upload_u0(None)
        mh.run_recursive(4, True, False)
        cabi.check(mh.lib.ilqrb_multi_download(mh.h, C.byref(res), 4))
    torch.cuda.synchronize()
    for k, v in t.items():
        assert np.array_equal(v.cpu().numpy(), getattr(want, k)), k


def test_second_device_runs_the_large_shared_memory_kernels():
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per device: device 0 first, then the same kernels on device 1"""
    if _ndev() < 2:
        pytest.skip("needs two GPUs")
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.solver import Handle
    pb = synth.config4(batch=4, horizon=40)  # time system: generic backward kernel (> 48 KB), batch Cholesky (104 KB)
    u0 = synth.u0_time(pb)
    res = []
    for d in (0, 1):
        with Handle(pb, device=d) as h:
            res.append((h.solve_recursive(u0, 3, True, False), h.solve_batch(u0, 2, False)))
    assert np.array_equal(res[0][0].U, res[1][0].U) and np.array_equal(res[0][1].U, res[1][1].U)

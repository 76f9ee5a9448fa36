"""ctypes mirror of include/ilqrb.h and loader of the CUDA library (libilqrb.so).

This is plumbing: the product is the CUDA library behind the C ABI.  There is no CPU fallback — if the
library is missing or no CUDA device is present, loading / ilqrb_create fails loudly.
"""
import ctypes as C
import os

import numpy as np

ABI_VERSION = 2
SYS_POS_ORN, SYS_POS_ORN_TIME, SYS_POSITION = 0, 1, 2
ROBOT_KDL, ROBOT_2D = 0, 1
TASK_POSE, TASK_JOINT = 0, 1
JOINT_FIXED, JOINT_ROT, JOINT_TRANS = 0, 1, 2
KP_STRIDE = 15
N_ALPHA = 11
MAX_SEGMENTS = 24
BATCH_WOODBURY = 1
SD_LOGMAP, SD_EXPMAP, SD_DISTANCE, SD_TRANSPORT, SD_DQUAT_TO_W_JAC = range(5)
KIN_STATIC_MAX_NQ = 4096
ST_NAN_COST, ST_QUU_NOT_SPD, ST_EARLY_STOP = 1, 2, 4
PROF_SLOTS = 13

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


class Segment(C.Structure):
    _fields_ = [("joint_type", C.c_int32), ("pad_", C.c_int32), ("axis", C.c_double * 3), ("origin", C.c_double * 3),
                ("tip_R", C.c_double * 9), ("tip_p", C.c_double * 3)]


class Robot(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dof", C.c_int32), ("n_segments", C.c_int32), ("pad_", C.c_int32),
                ("segments", C.POINTER(Segment)), ("lengths", c_double_p)]


class Task(C.Structure):
    _fields_ = [("robot", C.c_int32), ("joint_offset", C.c_int32), ("has_transform", C.c_int32), ("n_keypoints", C.c_int32),
                ("T", C.c_double * 16), ("kp_timestep", c_int32_p), ("kp_target", c_double_p), ("kp_precision", c_double_p),
                ("precision_per_problem", C.c_int32), ("limits_set", C.c_int32), ("r_diag", c_double_p),
                ("state_max", c_double_p), ("state_min", c_double_p), ("limit_weight", c_int32_p),
                ("task_kind", C.c_int32), ("pad_", C.c_int32), ("kp_deadzone", c_double_p)]


class ProblemDesc(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("sys_kind", C.c_int32), ("nb_deriv", C.c_int32), ("horizon", C.c_int32),
                ("batch", C.c_int32), ("dof", C.c_int32), ("n_robots", C.c_int32), ("n_tasks", C.c_int32), ("dt", C.c_double),
                ("robots", C.POINTER(Robot)), ("tasks", C.POINTER(Task)), ("r_diag", c_double_p), ("q0", c_double_p),
                ("dq0", c_double_p)]


class Dims(C.Structure):
    _fields_ = [("n", C.c_int32), ("m", C.c_int32), ("nt", C.c_int32), ("r", C.c_int32), ("H", C.c_int32), ("B", C.c_int32),
                ("n_kp_total", C.c_int32)]


class ALDesc(C.Structure):
    _fields_ = [("nc", C.c_int32), ("per_step", C.c_int32), ("A", c_double_p), ("b", c_double_p), ("lambda0", c_double_p),
                ("lag_update_step", C.c_int32), ("pad_", C.c_int32), ("penalty", C.c_double), ("scaling_factor", C.c_double)]


class BatchDesc(C.Structure):
    _fields_ = [("Q", c_double_p), ("psi", c_double_p), ("nw", C.c_int32), ("flags", C.c_int32)]


class LqtDesc(C.Structure):
    _fields_ = [("nq", C.c_int32), ("n", C.c_int32), ("m", C.c_int32), ("T", C.c_int32), ("ab_per_problem", C.c_int32), ("pad_", C.c_int32),
                ("A", c_double_p), ("B", c_double_p), ("Qs", c_double_p), ("mu", c_double_p), ("r", C.c_double)]


class SolveOpts(C.Structure):
    _fields_ = [("nb_iter", C.c_int32), ("line_search", C.c_int32), ("early_stop", C.c_int32), ("pad_", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("X", c_double_p), ("fX", c_double_p), ("U", c_double_p), ("K", c_double_p), ("d", c_double_p), ("cost", c_double_p),
                ("cost_hist", c_double_p), ("alpha_hist", c_double_p), ("iters", c_int32_p), ("status", c_int32_p),
                ("lambda_", c_double_p)]


class Profile(C.Structure):
    _fields_ = [("ms", C.c_double * PROF_SLOTS), ("launches", C.c_int64 * PROF_SLOTS), ("name", C.c_char_p * PROF_SLOTS)]


def dptr(a):
    """pointer to a C-contiguous float64 array (None -> NULL)"""
    if a is None:
        return C.cast(None, c_double_p)
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], "need C-contiguous float64"
    return a.ctypes.data_as(c_double_p)


def iptr(a):
    if a is None:
        return C.cast(None, c_int32_p)
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"], "need C-contiguous int32"
    return a.ctypes.data_as(c_int32_p)


# every symbol include/ilqrb.h declares (tests/test_cabi_cpu.py checks the built library exports them all)
SYMBOLS = [
    "ilqrb_abi_version", "ilqrb_last_error", "ilqrb_get_dims", "ilqrb_create", "ilqrb_destroy", "ilqrb_set_stream",
    "ilqrb_synchronize", "ilqrb_update_problems", "ilqrb_solve_recursive", "ilqrb_solve_recursive_async", "ilqrb_solve_al", "ilqrb_solve_batch",
    "ilqrb_upload_u0", "ilqrb_run_recursive", "ilqrb_run_al", "ilqrb_run_batch", "ilqrb_download", "ilqrb_rollout",
    "ilqrb_kinematics", "ilqrb_set_profiling", "ilqrb_get_profile", "ilqrb_reset_profile", "ilqrb_launch_count",
    "ilqrb_chain_from_urdf", "ilqrb_build_psi", "ilqrb_fp64_peaks", "ilqrb_kinematics_static", "ilqrb_eval_point", "ilqrb_keypoint_diff", "ilqrb_sd", "ilqrb_host_alloc", "ilqrb_host_free", "ilqrb_lqt_solve", "ilqrb_lqt_command", "ilqrb_chain_after_compute",
    "ilqrb_multi_plan", "ilqrb_create_multi", "ilqrb_multi_destroy", "ilqrb_multi_count", "ilqrb_multi_slice", "ilqrb_multi_solve_recursive",
    "ilqrb_multi_solve_al", "ilqrb_multi_solve_batch", "ilqrb_multi_upload_u0", "ilqrb_multi_run_recursive", "ilqrb_multi_synchronize", "ilqrb_multi_download",
]

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libilqrb.so")
_lib = None


def load():
    """dlopen libilqrb.so (built by __graft_entry__.build()); raises if it is missing — no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libilqrb.so not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                           "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    lib.ilqrb_abi_version.restype = C.c_int
    lib.ilqrb_last_error.restype = C.c_char_p
    lib.ilqrb_get_dims.argtypes = [C.POINTER(ProblemDesc), C.POINTER(Dims)]
    lib.ilqrb_create.argtypes = [C.POINTER(ProblemDesc), C.c_int, C.POINTER(H)]
    lib.ilqrb_destroy.argtypes = [H]
    lib.ilqrb_set_stream.argtypes = [H, C.c_void_p]
    lib.ilqrb_synchronize.argtypes = [H]
    lib.ilqrb_chain_after_compute.argtypes = [H, H]
    lib.ilqrb_update_problems.argtypes = [H, c_double_p, c_double_p, C.c_int, c_double_p]
    lib.ilqrb_solve_recursive.argtypes = [H, c_double_p, C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_solve_recursive_async.argtypes = [H, c_double_p, C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_solve_al.argtypes = [H, c_double_p, C.POINTER(ALDesc), C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_solve_batch.argtypes = [H, c_double_p, C.POINTER(BatchDesc), C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_upload_u0.argtypes = [H, c_double_p]
    lib.ilqrb_run_recursive.argtypes = [H, C.POINTER(SolveOpts)]
    lib.ilqrb_run_al.argtypes = [H, C.POINTER(ALDesc), C.POINTER(SolveOpts)]
    lib.ilqrb_run_batch.argtypes = [H, C.POINTER(BatchDesc), C.POINTER(SolveOpts)]
    lib.ilqrb_download.argtypes = [H, C.POINTER(Result)]
    lib.ilqrb_rollout.argtypes = [H, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p]
    lib.ilqrb_kinematics.argtypes = [H, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p]
    lib.ilqrb_kinematics_static.argtypes = lib.ilqrb_kinematics.argtypes
    IP = C.POINTER(C.c_int)
    lib.ilqrb_eval_point.argtypes = [H, C.c_int, IP, IP] + [c_double_p] * 9
    lib.ilqrb_keypoint_diff.argtypes = [C.c_int] * 5 + [c_double_p, c_double_p, C.c_int, c_double_p, c_double_p]
    lib.ilqrb_sd.argtypes = [C.c_int, C.c_int, C.c_int] + [c_double_p] * 4
    lib.ilqrb_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    lib.ilqrb_lqt_solve.argtypes = [C.c_int, C.POINTER(LqtDesc)] + [c_double_p] * 6 + [c_int32_p]
    lib.ilqrb_lqt_command.argtypes = [C.c_int] * 5 + [c_double_p] * 3 + [C.c_int, c_int32_p, c_int32_p, c_double_p, c_double_p]
    lib.ilqrb_host_free.argtypes = [C.c_void_p]
    lib.ilqrb_multi_plan.argtypes = [C.c_int, C.c_int, IP, IP]
    lib.ilqrb_create_multi.argtypes = [C.POINTER(ProblemDesc), C.c_int, IP, C.POINTER(H)]
    lib.ilqrb_multi_destroy.argtypes = [H]
    lib.ilqrb_multi_count.argtypes = [H]
    lib.ilqrb_multi_slice.argtypes = [H, C.c_int, IP, IP, IP, C.POINTER(H)]
    lib.ilqrb_multi_solve_recursive.argtypes = [H, c_double_p, C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_multi_solve_al.argtypes = [H, c_double_p, C.POINTER(ALDesc), C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_multi_solve_batch.argtypes = [H, c_double_p, C.POINTER(BatchDesc), C.POINTER(SolveOpts), C.POINTER(Result)]
    lib.ilqrb_multi_upload_u0.argtypes = [H, c_double_p]
    lib.ilqrb_multi_run_recursive.argtypes = [H, C.POINTER(SolveOpts)]
    lib.ilqrb_multi_synchronize.argtypes = [H]
    lib.ilqrb_multi_download.argtypes = [H, C.POINTER(Result), C.c_int]
    lib.ilqrb_set_profiling.argtypes = [H, C.c_int]
    lib.ilqrb_get_profile.argtypes = [H, C.POINTER(Profile)]
    lib.ilqrb_reset_profile.argtypes = [H]
    lib.ilqrb_launch_count.argtypes = [H]
    lib.ilqrb_launch_count.restype = C.c_int64
    lib.ilqrb_chain_from_urdf.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_char_p, c_double_p, c_double_p,
                                          C.POINTER(Segment), C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.ilqrb_build_psi.argtypes = [C.c_char_p, C.c_int, C.c_int, c_double_p]
    lib.ilqrb_fp64_peaks.argtypes = [C.c_int, c_double_p, c_double_p]
    if lib.ilqrb_abi_version() != ABI_VERSION:
        raise RuntimeError("libilqrb.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().ilqrb_last_error()
        raise RuntimeError("ilqrb error %d: %s" % (rc, msg.decode() if msg else "?"))

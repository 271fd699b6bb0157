"""ctypes binding of libddm_b200.so (include/ddm_b200.h).  No compute happens in Python."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DDM_B200_LIB") or os.path.join(os.path.dirname(_HERE), "libddm_b200.so")  # override: kernel experiments
BUILD_SH = os.path.join(os.path.dirname(_HERE), "csrc", "build.sh")

SYMBOLS = [
    "ddm_last_error", "ddm_version", "ddm_ctx_create", "ddm_ctx_destroy", "ddm_comm_unique_id", "ddm_comm_init",
    "ddm_upload_operator", "ddm_upload_operator_bcsr3", "ddm_upload_halo", "ddm_upload_decomposition", "ddm_upload_factor", "ddm_upload_factor_lu", "ddm_factorize_subdomain", "ddm_upload_coarse",
    "ddm_finalize", "ddm_set_deterministic", "ddm_blockdiag_solve", "ddm_seg_dot", "ddm_seg_axpby", "ddm_update_operator_values", "ddm_update_coarse_matrix", "ddm_galerkin_assemble", "ddm_set_coarse_coloring", "ddm_vec_create", "ddm_vec_destroy", "ddm_vec_upload", "ddm_vec_download", "ddm_vec_fill",
    "ddm_vec_copy", "ddm_vec_axpy", "ddm_vec_scale", "ddm_vec_add", "ddm_vec_xpay", "ddm_vec_device_ptr",
    "ddm_op_apply", "ddm_op_applyscaleadd", "ddm_dot", "ddm_dot_noalloc", "ddm_norm", "ddm_schwarz_apply", "ddm_coarse_apply",
    "ddm_twolevel_apply", "ddm_pcg_solve", "ddm_gmres_solve", "ddm_gmres_solve_host", "ddm_op_apply_host", "ddm_op_applyscaleadd_host", "ddm_dot_host",
    "ddm_twolevel_apply_host", "ddm_pcg_solve_host", "ddm_host_register", "ddm_host_unregister", "ddm_get_stat", "ddm_timers_enable", "ddm_timers_reset",
    "ddm_get_timer", "ddm_stopwatch_start", "ddm_stopwatch_stop", "ddm_kernel_launches", "ddm_profiler_range", "ddm_sync",
]


class Result(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("converged", C.c_int32), ("reduction", C.c_double),
                ("conv_rate", C.c_double), ("elapsed", C.c_double), ("def0", C.c_double)]


class DDMError(RuntimeError):
    pass


_lib = None


def build(verbose=False):
    env = dict(os.environ)
    if verbose:
        env["DDM_PTXAS_V"] = "1"
    subprocess.run(["sh", BUILD_SH], check=True, env=env)


def load():
    """Load the CUDA library; there is no fallback if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DDMError(f"{LIB_PATH} is missing: build it with dune-ddm_b200/csrc/build.sh (the solve phase has no CPU path)")
    lib = C.CDLL(LIB_PATH)
    lib.ddm_last_error.restype = C.c_char_p
    lib.ddm_version.restype = C.c_char_p
    lib.ddm_vec_device_ptr.restype = C.c_void_p
    lib.ddm_kernel_launches.restype = C.c_int64
    _lib = lib
    return lib


def check(status):
    if status != 0:
        raise DDMError(load().ddm_last_error().decode())

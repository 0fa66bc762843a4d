"""ctypes binding of libcatenary_b200.so (include/catenary_b200.h).  No CPU fallback: a missing
library or a missing CUDA device raises."""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CAT_B200_LIB") or os.path.join(_HERE, "libcatenary_b200.so")   # env: a build elsewhere

CAT_OK, CAT_EINVAL, CAT_ECUDA, CAT_ENOMEM, CAT_ENODEV = 0, -1, -2, -3, -4
OUT_ORIGINAL, OUT_THETA_ROTATED, OUT_THETA_ALIGNED, OUT_FINAL = 1, 2, 4, 8
MPC_RNG_CONTROLS, MPC_CONTROLS_AOS, MPC_CONTROLS_F32 = 1, 2, 4
P2P_LOCKSTEP, P2P_PIPELINED = 0, 1
ABI_VERSION = 3
CTRL_ARGMIN, CTRL_MPPI = 0, 1
CTRL_LOG_W = 16
STATIC_MAX_BLOCKS, STATIC_MAX_PROBLEMS = 32768, 4096   # cat_mpc_step_static_dev_f64 (fixed device scratch)


class CatMpcParams(C.Structure):
    _fields_ = [("K", C.c_int64), ("k_offset", C.c_int64), ("H", C.c_int32), ("n_problems", C.c_int32),
                ("flags", C.c_int32), ("problem_offset", C.c_int32), ("eq_theta", C.c_int32),
                ("eq_gamma", C.c_int32), ("dt", C.c_double), ("L", C.c_double), ("w_trk", C.c_double),
                ("w_du", C.c_double), ("w_len", C.c_double), ("w_sag", C.c_double), ("w_ang", C.c_double),
                ("rho", C.c_double), ("z_floor", C.c_double), ("margin", C.c_double), ("c_lo", C.c_double),
                ("c_hi", C.c_double), ("nan_penalty", C.c_double), ("seed", C.c_uint64),
                ("sigma", C.c_double * 3), ("clip", C.c_double)]


class CatMpcProblem(C.Structure):
    _fields_ = [("P0", C.c_double * 3), ("P1", C.c_double * 3), ("V1", C.c_double * 3), ("A1", C.c_double * 3),
                ("theta", C.c_double), ("gamma", C.c_double), ("theta_prev", C.c_double),
                ("gamma_prev", C.c_double)]


class CatCtrlConfig(C.Structure):
    _fields_ = [("mode", C.c_int32), ("max_steps", C.c_int32), ("use_graph", C.c_int32), ("reserved", C.c_int32),
                ("lam", C.c_double), ("u_max", C.c_double * 3)]


PROBLEM_DTYPE = np.dtype([("P0", "f8", 3), ("P1", "f8", 3), ("V1", "f8", 3), ("A1", "f8", 3), ("theta", "f8"),
                          ("gamma", "f8"), ("theta_prev", "f8"), ("gamma_prev", "f8")])
assert PROBLEM_DTYPE.itemsize == C.sizeof(CatMpcProblem) == 128

_vp, _i64, _i32, _int = C.c_void_p, C.c_int64, C.c_int32, C.c_int
_PP = C.POINTER(CatMpcParams)
_PC = C.POINTER(CatCtrlConfig)

# name -> (restype, argtypes); every symbol include/catenary_b200.h declares
SIGNATURES = {
    "cat_abi_version": (_int, []),
    "cat_init": (_int, [_int, C.POINTER(_vp)]),
    "cat_destroy": (_int, [_vp]),
    "cat_sync": (_int, [_vp]),
    "cat_last_error": (C.c_char_p, [_vp]),
    "cat_launch_count": (_i64, [_vp]),
    "cat_stream": (_vp, [_vp]),
    "cat_set_stream": (_int, [_vp, _vp, _int]),
    "cat_host_alloc": (_int, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "cat_host_free": (_int, [_vp, _vp]),
    "cat_host_register": (_int, [_vp, _vp, C.c_size_t]),
    "cat_host_unregister": (_int, [_vp, _vp]),
    "cat_mpc_workspace_create": (_int, [_vp, _i64, _i32, C.POINTER(_vp)]),
    "cat_mpc_workspace_destroy": (_int, [_vp, _vp]),
    "cat_set_workspace": (_int, [_vp, _vp]),
    "cat_set_overlap": (_int, [_vp, _int]),
    "cat_mpc_stats": (_int, [_vp, _int, _int, _vp]),
    "cat_timer_start": (_int, [_vp]),
    "cat_timer_stop_ms": (_int, [_vp, C.POINTER(C.c_float)]),
    "cat_solve_param_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp]),
    "cat_solve_param_f32": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp]),
    "cat_solve_param_dev_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp]),
    "cat_solve_param_dev_f32": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp]),
    "cat_span_terms_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, C.c_double, _vp, _vp, _vp, _vp]),
    "cat_rodrigues_f64": (_int, [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp]),
    "cat_shape_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "cat_shape_f32": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "cat_shape_dev_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "cat_shape_dev_f32": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "cat_lowest_point_f64": (_int, [_vp, _i64, _i32, _vp, _vp, _vp]),
    "cat_equation_count": (_int, []),
    "cat_equation_shipped": (_int, [_int]),
    "cat_equation_info": (_int, [_int, C.POINTER(_int), C.POINTER(_int), C.POINTER(C.c_uint32)]),
    "cat_predict_f64": (_int, [_vp, _i64, _vp, _int, _int, _vp]),
    "cat_predict_f32": (_int, [_vp, _i64, _vp, _int, _int, _vp]),
    "cat_rk4_open_f64": (_int, [_vp, _i64, _i64, _vp, _vp, _vp, _int, _int, _vp]),
    "cat_rk4_open_f32": (_int, [_vp, _i64, _i64, _vp, _vp, _vp, _int, _int, _vp]),
    "cat_features_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _int, _vp]),
    "cat_mpc_step_f64": (_int, [_vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_f32": (_int, [_vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_dev_f64": (_int, [_vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_dev_f32": (_int, [_vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_static_dev_f64": (_int, [_vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_sample_controls_dev_f64": (_int, [_vp, _PP, _vp]),
    "cat_reduce_winners_f64": (_int, [_int, _vp, _vp, _vp, _vp]),
    "cat_reduce_winners_dev_f64": (_int, [_vp, _int, _vp, _vp]),
    "cat_catenary3d_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _i32, _vp]),
    "cat_catenary_equation_f64": (_int, [_vp, _i64, _vp, _vp, _i64, _vp]),
    "cat_resample_linear_f64": (_int, [_vp, _i64, _vp, _i32, _vp, _i64, _vp, _vp]),
    "cat_feature_bank_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, C.c_double, C.c_double, _vp]),
    "cat_fir_reflect_f64": (_int, [_vp, _i64, _i64, _vp, _i32, _vp, _vp]),
    "cat_savgol_f64": (_int, [_vp, _i64, _i64, _vp, _i32, _vp, _vp, _vp, _vp]),
    "cat_gradient_f64": (_int, [_vp, _i64, _i64, _vp, _vp, _i32, _vp]),
    "cat_features_valid_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _i32, _i32, _vp]),
    "cat_features_dd_f64": (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp]),
    "cat_expr_eval_f64": (_int, [_vp, _i64, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "cat_expr_eval_f32": (_int, [_vp, _i64, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "cat_expr_integrate_f64": (_int, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp]),
    "cat_mpc_step_expr_f64": (_int, [_vp, _PP, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_expr_f32": (_int, [_vp, _PP, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "cat_p2p_create": (_int, [_vp, _int, _int, C.POINTER(_vp), _vp]),
    "cat_p2p_connect": (_int, [_vp, _vp]),
    "cat_mpc_step_p2p_dev_f64": (_int, [_vp, _vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_p2p_dev_f32": (_int, [_vp, _vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_p2p_f64": (_int, [_vp, _vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_mpc_step_p2p_f32": (_int, [_vp, _vp, _PP, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cat_p2p_set_mode": (_int, [_vp, _int]),
    "cat_p2p_flush_dev": (_int, [_vp, _vp, _vp, _vp]),
    "cat_p2p_status": (_int, [_vp, C.POINTER(_int)]),
    "cat_p2p_destroy": (_int, [_vp]),
    "cat_ctrl_create_f64": (_int, [_vp, _PP, _PC, _vp, _vp, _vp, _i64, C.POINTER(_vp)]),
    "cat_ctrl_create_f32": (_int, [_vp, _PP, _PC, _vp, _vp, _vp, _i64, C.POINTER(_vp)]),
    "cat_ctrl_run": (_int, [_vp, _i32]),
    "cat_ctrl_set_state": (_int, [_vp, _vp]),
    "cat_ctrl_read": (_int, [_vp, C.POINTER(_i32), _vp, _vp, _vp, _i32, _i32]),
    "cat_ctrl_destroy": (_int, [_vp]),
    "cat_measure_peak": (_int, [_vp, _int, C.POINTER(C.c_double)]),
}

_lib = None
_lock = threading.Lock()


class CatenaryLibraryError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library and bind every declared symbol.  Raises if it is not built."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise CatenaryLibraryError(
                    f"{LIB_PATH} is not built -- run __graft_entry__.build() (there is no CPU fallback)")
            lib = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)      # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            if lib.cat_abi_version() != ABI_VERSION:
                raise CatenaryLibraryError(f"{LIB_PATH}: ABI {lib.cat_abi_version()}, this package needs {ABI_VERSION} "
                                           "-- rebuild with __graft_entry__.build()")
            _lib = lib
    return _lib


def ptr(a):
    """Address of a C-contiguous ndarray (None -> NULL)."""
    if a is None:
        return None
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("array must be C-contiguous")
    return a.ctypes.data

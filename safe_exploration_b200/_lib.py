"""ctypes binding of libsecem.so (include/secem.h).

There is no CPU fallback: if the shared library is missing, importing a symbol from it raises immediately.
"""
import ctypes as C
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsecem.so")

SE_F32, SE_F64 = 0, 1
SE_GEMM_AUTO, SE_GEMM_SIMT, SE_GEMM_MMA, SE_GEMM_TC = 0, 1, 2, 3
SE_OBJ_NEG_VARIANCE, SE_OBJ_ABS_STATE1, SE_OBJ_NEG_LAST_STATE = 0, 1, 2
SE_MAX_NS, SE_MAX_NU, SE_MAX_FACES, SE_MAX_ACT_FACES, SE_MAX_T = 6, 3, 16, 8, 64
SE_MAX_OBS_FACES = 8
MAX_ELITES = 4096

_dbl_p = C.POINTER(C.c_double)
_i32_p = C.POINTER(C.c_int32)


class SeProblem(C.Structure):
    """`se_problem` of include/secem.h."""
    _fields_ = [("n_s", C.c_int), ("n_u", C.c_int), ("k_fb", _dbl_p), ("a", _dbl_p), ("b", _dbl_p), ("l_mu", _dbl_p),
                ("l_sigma", _dbl_p), ("c_safety", C.c_double), ("m_safe", C.c_int), ("h_mat", _dbl_p),
                ("h_vec", _dbl_p), ("m_act", C.c_int), ("act_a", _dbl_p), ("act_b", _dbl_p),
                ("act_penalty", C.c_double), ("state_penalty", C.c_double), ("objective", C.c_int),
                ("objective_theta", C.c_double), ("all_states_constraint", C.c_int), ("m_obs", C.c_int),
                ("h_mat_obs", _dbl_p), ("h_vec_obs", _dbl_p), ("clamp_nonpositive", C.c_int)]


class SeCemConfig(C.Structure):
    _fields_ = [("num_instances", C.c_int), ("num_rollouts", C.c_int), ("time_horizon", C.c_int),
                ("num_elites", C.c_int), ("num_iterations", C.c_int), ("pick_best", C.c_int),
                ("first_rollout", C.c_int), ("local_rollouts", C.c_int), ("first_instance", C.c_int)]


class SeCemOutputs(C.Structure):
    _fields_ = [("actions", C.c_void_p), ("found", C.c_void_p), ("best_id", C.c_void_p), ("means", C.c_void_p),
                ("stds", C.c_void_p), ("elite_ids", C.c_void_p), ("objective", C.c_void_p),
                ("constraint", C.c_void_p), ("rollout_actions", C.c_void_p), ("trajectories", C.c_void_p),
                ("flags", C.c_void_p)]


# every symbol include/secem.h declares: name -> (restype, argtypes)
_vp, _i, _i64, _u64, _d = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double
SIGNATURES = {
    "se_last_error": (C.c_char_p, []),
    "se_version": (_i, []),
    "se_gp_create": (_i, [C.POINTER(_vp), _i, _i, _i, _i, _vp, _vp, _vp, _dbl_p, _dbl_p, _dbl_p, _vp]),
    "se_gp_destroy": (_i, [_vp]),
    "se_gp_set_workspace_limit": (_i, [_vp, _u64]),
    "se_gp_set_gemm_impl": (_i, [_vp, _i]),
    "se_gp_set_tc_digits": (_i, [_vp, _i]),
    "se_gp_get_gemm_info": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "se_debug_tc_probe": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "se_debug_eig_max": (_i, [_dbl_p, _i, _dbl_p]),
    "se_debug_exp": (_i, [_vp, _i64, _vp, _vp]),
    "se_debug_exp_table": (_i, [_vp, _i64, _vp, _vp]),
    "se_gp_predict": (_i, [_vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "se_gp_kernel_matrix": (_i, [_i, _i, _i64, _vp, _dbl_p, _dbl_p, _dbl_p, _vp, _vp]),
    "se_gp_mll_grad_workspace": (_i64, [_i, _i, _i64]),
    "se_gp_mll_grad": (_i, [_i, _i, _i64, _vp, _vp, _vp, _dbl_p, _dbl_p, _dbl_p, _vp, _vp, _vp]),
    "se_gp_factor_workspace": (_i64, [_i, _i64]),
    "se_gp_factor": (_i, [_i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "se_gp_kinv": (_i, [_i, _i64, _vp, _vp, _vp]),
    "se_gp_mll_grad_static_max_bytes": (_i64, []),
    "se_gp_mll_grad_static": (_i, [_i, _i, _i64, _vp, _vp, _vp, _dbl_p, _dbl_p, _dbl_p, _vp, _vp]),
    "se_reach_onestep": (_i, [_vp, C.POINTER(SeProblem), _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "se_reach_step": (_i, [_i, C.POINTER(SeProblem), _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "se_reach_multistep": (_i, [_vp, C.POINTER(SeProblem), _vp, _vp, _vp, _i64, _i, _dbl_p, _vp, _vp, _vp, _vp, _vp,
                                _vp, _vp]),
    "se_safety_distance": (_i, [_i, _i, _i, _dbl_p, _dbl_p, _d, _vp, _vp, _i64, _vp, _vp, _vp]),
    "se_cem_create": (_i, [C.POINTER(_vp), _vp, C.POINTER(SeProblem), C.POINTER(SeCemConfig)]),
    "se_cem_destroy": (_i, [_vp]),
    "se_cem_set_comm": (_i, [_vp, _vp, _i, _i]),
    "se_nccl_unique_id": (_i, [C.c_char_p]),
    "se_nccl_comm_init": (_i, [C.POINTER(_vp), _i, C.c_char_p, _i]),
    "se_nccl_comm_destroy": (_i, [_vp]),
    "se_cem_solve": (_i, [_vp, _vp, _vp, _u64, C.POINTER(SeCemOutputs), _vp]),
    "se_cem_noise": (_i, [_i, _u64, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "se_cem_select_refit": (_i, [_i, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "se_cem_pick": (_i, [_i, _vp, _vp, _i, _i, _vp, _vp]),
    "se_launch_count": (_u64, []),
    "se_launch_count_reset": (None, []),
    "se_measure_peak": (_i, [_i, _dbl_p]),
    "se_gp_timing_enable": (_i, [_vp, _i]),
    "se_gp_timing_read": (_i, [_vp, _dbl_p, C.POINTER(_i64), _dbl_p, C.POINTER(_i64), _vp]),
}

_lib = None


class SecemError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load libsecem.so and attach the signatures.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SecemError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                         f"or `make -C safe_exploration_b200/csrc` -- there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise SecemError(load().se_last_error().decode("utf-8", "replace"))


def dtype_code(dtype: torch.dtype) -> int:
    if dtype == torch.float64:
        return SE_F64
    if dtype == torch.float32:
        return SE_F32
    raise ValueError(f"unsupported dtype {dtype}: the CUDA path computes in float32 or float64")


def ptr(t):
    """Device pointer of a contiguous tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "tensor must be contiguous"
    return C.c_void_p(t.data_ptr())


def host_doubles(x):
    """Keep-alive numpy buffer + double* for host constants."""
    if x is None:
        return None, None
    arr = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
    return arr, arr.ctypes.data_as(_dbl_p)


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise SecemError(f"{name} must live on a CUDA device (got {t.device}); the B200 path has no CPU fallback")

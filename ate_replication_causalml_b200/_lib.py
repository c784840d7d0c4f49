"""ctypes binding of libate_b200.so (the C ABI declared in include/ate_b200.h).

There is no CPU fallback: if the shared library is missing, or no CUDA device is present when
a context is created, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# ATE_B200_LIB: load another build of the same library (same-box A/B timing of kernel variants)
SO_PATH = os.environ.get("ATE_B200_LIB") or os.path.join(_HERE, "libate_b200.so")

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)
c_ll = C.c_longlong


class AteError(RuntimeError):
    pass


class CrossfitOpts(C.Structure):
    _fields_ = [
        ("K", C.c_int), ("nfolds_inner", C.c_int), ("nlambda", C.c_int), ("thresh", C.c_double),
        ("maxit", C.c_int), ("use_1se", C.c_int), ("standardize_base", C.c_int),
        ("n_continuous", C.c_int), ("expand", C.c_int), ("B", C.c_int), ("boot_mode", C.c_int),
        ("boot_seed", C.c_uint64), ("boot_rep_offset", c_ll), ("boot_reps_here", C.c_int),
        ("r_seed", C.c_uint32), ("rng_kind", C.c_int), ("on_device", C.c_int),
        ("fold_begin", C.c_int), ("fold_end", C.c_int), ("skip_score", C.c_int),
        ("shard_rank", C.c_int), ("shard_count", C.c_int),
    ]


class CrossfitResult(C.Structure):
    _fields_ = [
        ("tau", C.c_double), ("se_sandwich", C.c_double), ("se_bootstrap", C.c_double),
        ("lower_ci", C.c_double), ("upper_ci", C.c_double), ("jerr", C.c_int),
        ("rounds_irls", C.c_int), ("gram_launches", c_ll), ("ms_design", C.c_double),
        ("ms_outcome", C.c_double), ("ms_propensity", C.c_double), ("ms_score", C.c_double),
        ("ms_bootstrap", C.c_double), ("irls_solves", C.c_double), ("strong_mean", C.c_double),
        ("strong_frac2", C.c_double), ("ms_folds", C.c_double), ("step_launches", C.c_double),
        ("row_fits", C.c_double), ("rows_staged", C.c_double), ("list_misses", C.c_double),
    ]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_double), C.c_longlong)

# every symbol include/ate_b200.h declares: name -> (restype, argtypes)
_vp = C.c_void_p
SIGNATURES = {
    "ate_version": (C.c_int, []),
    "ate_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "ate_ctx_destroy": (None, [_vp]),
    "ate_last_error": (C.c_char_p, [_vp]),
    "ate_ctx_stream": (_vp, [_vp]),
    "ate_ctx_set_allreduce": (C.c_int, [_vp, _vp, _vp]),
    "ate_nccl_unique_id": (C.c_int, [_vp]),
    "ate_ctx_nccl_init": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "ate_ctx_nccl_destroy": (C.c_int, [_vp]),
    "ate_launch_count": (c_ll, [_vp]),
    "ate_profile_enable": (C.c_int, [_vp, C.c_int]),
    "ate_profile_read": (C.c_int, [_vp, c_dp, C.POINTER(c_ll), c_dp, c_dp]),
    "ate_standardize": (C.c_int, [_vp, _vp, c_ll, C.c_int, _vp, C.c_int, _vp, _vp]),
    "ate_expand_deg2": (C.c_int, [_vp, _vp, c_ll, C.c_int, C.c_int, _vp]),
    "ate_glmnet": (C.c_int, [_vp, C.c_int, c_ll, C.c_int, _vp, _vp, _vp, C.c_int, C.c_double, _vp,
                             C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int), _vp, _vp, _vp, _vp,
                             C.POINTER(C.c_int), c_dp, C.POINTER(C.c_int)]),
    "ate_cv_glmnet": (C.c_int, [_vp, C.c_int, c_ll, C.c_int, _vp, _vp, _vp, C.c_int, _vp,
                                C.c_uint32, C.c_int, C.c_int, C.c_double, C.c_int,
                                C.POINTER(C.c_int), _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int),
                                C.POINTER(C.c_int), _vp, C.POINTER(C.c_int)]),
    "ate_cv_glmnet_device": (C.c_int, [_vp, C.c_int, c_ll, C.c_int, _vp, c_ll, _vp, C.c_int, _vp, C.c_int,
                                       C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), _vp, _vp, _vp,
                                       _vp, _vp, C.POINTER(C.c_int), C.POINTER(C.c_int), _vp, C.POINTER(C.c_int)]),
    "ate_glm_binomial": (C.c_int, [_vp, c_ll, C.c_int, _vp, _vp, C.c_int, C.c_double, _vp, c_dp,
                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ate_lm": (C.c_int, [_vp, c_ll, C.c_int, _vp, _vp, _vp, _vp, _vp, c_dp, c_dp, C.POINTER(C.c_int),
                         C.POINTER(c_ll), _vp]),
    "ate_predict": (C.c_int, [_vp, c_ll, C.c_int, _vp, C.c_double, _vp, C.c_int, _vp]),
    "ate_aipw_score": (C.c_int, [_vp, c_ll, _vp, _vp, _vp, _vp, _vp, c_dp, c_dp, _vp, _vp, C.c_int]),
    "ate_aipw_score_static": (C.c_int, [_vp, c_ll, _vp, _vp, _vp, _vp, _vp, c_dp, c_dp]),
    "ate_bootstrap": (C.c_int, [_vp, c_ll, _vp, _vp, _vp, _vp, _vp, C.c_int, C.c_int, _vp,
                                C.c_uint64, c_ll, _vp, c_dp, C.c_int]),
    "ate_tau_hat_dr_est": (C.c_int, [_vp, c_ll, _vp, _vp, _vp, _vp, _vp, _vp, c_dp]),
    "ate_r_set_seed": (C.c_int, [_vp, C.c_uint32]),
    "ate_r_sample": (C.c_int, [_vp, C.c_int, c_ll, C.c_int, _vp]),
    "ate_r_foldid": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp]),
    "ate_crossfit_default_opts": (None, [C.POINTER(CrossfitOpts)]),
    "ate_crossfit_aipw": (C.c_int, [_vp, C.POINTER(CrossfitOpts), c_ll, C.c_int, _vp, _vp, _vp, _vp,
                                    _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                    C.POINTER(CrossfitResult)]),
    "ate_cv_select": (C.c_int, [C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int),
                                C.POINTER(C.c_int)]),
    "ate_round_plan_describe": (C.c_int, [C.c_int, _vp, C.c_int, C.c_int, _vp]),
    "ate_round_kernel_bench": (C.c_int, [_vp, c_ll, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         c_dp, c_dp, c_dp, C.POINTER(C.c_int)]),
    "ate_weighted_gram": (C.c_int, [_vp, c_ll, C.c_int, _vp, c_ll, _vp, _vp, c_ll, _vp, C.c_int]),
}

_lib = None


def load():
    """dlopen the in-tree shared library and bind every declared symbol.  Raises if the
    library has not been built (python -m ate_replication_causalml_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise AteError(f"{SO_PATH} is missing: build it with __graft_entry__.build() "
                       "(there is no CPU fallback)")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def ptr(a):
    """void* of a numpy array (or None); device pointers are passed as plain ints."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return C.c_void_p(a.ctypes.data)


def f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel())


def fmat(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32).ravel())

"""ctypes binding of libsb200.so (include/sb200.h).  There is no CPU fallback: if the library is missing or a
call fails the error is raised to the caller."""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int32, c_int64, c_uint64, c_void_p

from .errors import SeldonB200Error

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsb200.so")

SB_F32, SB_I32 = 0, 1
SB_OPT_GRAM_TERMS, SB_OPT_FORCE_GENERIC, SB_OPT_PROFILE, SB_OPT_FUSED_PASS2, SB_OPT_LEGACY_MMA = 1, 2, 3, 4, 5
SB_OPT_NO_SUBSPACE = 6
SB_ROUTER_EPSILON_GREEDY, SB_ROUTER_THOMPSON = 0, 1
SB_REWARD_PER_CALL, SB_REWARD_TABLE, SB_REWARD_BERNOULLI = 0, 1, 2
SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD = 1

_uintptr = ctypes.c_size_t
_pp = POINTER(c_void_p)

# name -> (restype, argtypes); every symbol declared in include/sb200.h
SIGNATURES = {
    "sb_last_error": (c_char_p, []),
    "sb_version": (c_int32, []),
    "sb_launch_count": (c_int64, []),
    "sb_unpack_tensor_f64": (c_int32, [c_void_p, c_uint64, c_uint64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int64, _uintptr]),
    "sb_unpack_listvalue": (c_int32, [c_void_p, c_uint64, c_int64, c_int64, c_void_p, c_int64, c_void_p, _uintptr]),
    "sb_pack_tensor_f64": (c_int32, [c_void_p, c_int32, c_int64, c_void_p, _uintptr]),
    "sb_pack_listvalue": (c_int32, [c_void_p, c_int32, c_int64, c_int64, c_void_p, _uintptr]),
    "sb_listvalue_nbytes": (c_uint64, [c_int64, c_int64]),
    "sb_tensor_values_offset": (c_uint64, [c_int64, c_int64]),
    "sb_combiner_mean": (c_int32, [POINTER(c_void_p), c_int32, c_int64, c_void_p, _uintptr]),
    "sb_standardise": (c_int32, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int64, _uintptr]),
    "sb_json_workspace_bytes": (c_uint64, [c_uint64]),
    "sb_json_ndarray_count": (c_int32, [c_void_p, c_uint64, c_void_p, c_uint64, c_void_p, _uintptr]),
    "sb_unpack_json_ndarray": (c_int32, [c_void_p, c_uint64, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                         _uintptr]),
    "sb_json_text_capacity": (c_uint64, [c_int64, c_int64]),
    "sb_json_encode_workspace_bytes": (c_uint64, [c_int64, c_int64]),
    "sb_pack_json_ndarray": (c_int32, [c_void_p, c_int32, c_int64, c_int64, c_void_p, c_uint64, c_void_p, c_void_p, _uintptr]),
    "sb_feedback_state_bytes": (c_uint64, [c_int32]),
    "sb_feedback_reset": (c_int32, [c_void_p, c_int32, _uintptr]),
    "sb_feedback_update": (c_int32, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, _uintptr]),
    "sb_maha_create": (c_int32, [_pp, c_int32, c_int32, c_double, c_double, c_int64, c_int64]),
    "sb_maha_destroy": (c_int32, [c_void_p]),
    "sb_maha_set_input_transform": (c_int32, [c_void_p, c_void_p, c_void_p]),
    "sb_maha_workspace_bytes": (c_int32, [c_void_p, c_int64, POINTER(c_uint64)]),
    "sb_maha_stats_len": (c_int64, [c_void_p]),
    "sb_maha_pass1_stats": (c_int32, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_uint64, c_void_p, _uintptr]),
    "sb_maha_finalize": (c_int32, [c_void_p, c_void_p, c_int32, c_int32, c_int64, c_int64, c_void_p, c_uint64, _uintptr]),
    "sb_maha_pass2_score": (c_int32, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_uint64, _uintptr]),
    "sb_maha_commit": (c_int32, [c_void_p, _uintptr]),
    "sb_maha_abort_batch": (c_int32, [c_void_p]),
    "sb_maha_batch": (c_int32, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_uint64, _uintptr]),
    "sb_gram_stats_noalloc": (c_int32, [c_void_p, c_int64, c_int32, c_int64, c_void_p, c_void_p, _uintptr]),
    "sb_maha_get_state": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(c_int64), POINTER(c_int32), POINTER(c_int32), _uintptr]),
    "sb_maha_set_state": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, _uintptr]),
    "sb_maha_last_batch_sums": (c_int32, [c_void_p, POINTER(c_int64), POINTER(c_double), POINTER(c_int64), _uintptr]),
    "sb_maha_kernel_times": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_maha_debug_counters": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_maha_get_eigvecs": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_maha_set_option": (c_int32, [c_void_p, c_int32, c_int64]),
    "sb_router_create": (c_int32, [_pp, c_int32, c_int32, c_double, c_int64, c_int32]),
    "sb_router_destroy": (c_int32, [c_void_p]),
    "sb_router_run": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int32, c_void_p, _uintptr]),
    "sb_router_feedback": (c_int32, [c_void_p, c_int32, c_int64, c_double, _uintptr]),
    "sb_router_state_len": (c_int64, [c_void_p]),
    "sb_router_get_state": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_router_set_state": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_router_route": (c_int32, [c_void_p, POINTER(c_int32), _uintptr]),
    "sb_router_feedback_route": (c_int32, [c_void_p, c_int32, c_int64, c_double, POINTER(c_int32), _uintptr]),
    "sb_router_last_values": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_abtest_initial_state": (c_uint64, [c_int64]),
    "sb_abtest_route": (c_int32, [c_void_p, ctypes.c_float, c_int64, c_void_p, _uintptr]),
    "sb_router_counters": (c_int32, [c_void_p, c_void_p, _uintptr]),
    "sb_router_set_option": (c_int32, [c_void_p, c_int32, c_int64]),
}

_lib = None


def lib():
    """Load libsb200.so once.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise SeldonB200Error(
                "libsb200.so not found at %s - build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                "there is no CPU fallback" % LIB_PATH, status_code=500)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, status_code=400):
    if rc != 0:
        msg = lib().sb_last_error().decode("utf-8", "replace")
        raise SeldonB200Error(msg or "libsb200 error %d" % rc, status_code=status_code if rc == -1 else 500)


def current_stream(device=None):
    import torch

    return torch.cuda.current_stream(device).cuda_stream

"""ctypes binding of include/mcmcpp_b200/capi.h (libmcmcpp_b200.so).

No CPU fallback: if the shared library is missing this module raises at import of the
symbol table, and every entry point fails with MCMCPP_ERR_CUDA when no GPU is usable.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmcmcpp_b200.so")

OK, CHAIN_FULL = 0, 1
ERR_ARG, ERR_CUDA, ERR_NCCL, ERR_STATE, ERR_UNSUPPORTED = -1, -2, -3, -4, -5
RNG_SEQUENTIAL, RNG_PER_WALKER, RNG_PER_WALKER32 = 0, 1, 2
SHARD_NONE, SHARD_ALLGATHER, SHARD_PEER, SHARD_DATA, SHARD_DIM = 0, 1, 2, 3, 4
TARGET = {"skewed_gauss_2d": 0, "rosenbrock_2d": 1, "diag_gauss": 2, "dense_gauss": 3, "linreg": 4}


class Config(C.Structure):
    _fields_ = [("num_walkers", C.c_int32), ("num_params", C.c_int32), ("seed", C.c_int64),
                ("rng_mode", C.c_int32), ("alpha_num", C.c_int32), ("alpha_denom", C.c_int32),
                ("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("shard_mode", C.c_int32), ("trace_updates", C.c_int32), ("max_chain_bytes", C.c_uint64),
                ("mover_state_doubles", C.c_int32), ("chain_block_steps", C.c_int32),
                ("device_chain_bytes", C.c_uint64)]


class Trace(C.Structure):
    _fields_ = [("j", C.c_int32), ("accept", C.c_int32), ("z", C.c_double), ("e", C.c_double),
                ("newp", C.c_double), ("diff", C.c_double)]


# every symbol include/mcmcpp_b200/capi.h declares: name -> (restype, argtypes)
_dp = C.POINTER(C.c_double)
_h = C.c_void_p
SYMBOLS = {
    "mcmcpp_sampler_create": (C.c_int, [C.POINTER(Config), C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(_h)]),
    "mcmcpp_sampler_create_builtin": (C.c_int, [C.POINTER(Config), C.c_int, _dp, C.c_size_t, C.POINTER(_h)]),
    "mcmcpp_sampler_create_builtin_mover": (C.c_int, [C.POINTER(Config), C.c_int, _dp, C.c_size_t, C.c_int, _dp, C.c_size_t, C.POINTER(_h)]),
    "mcmcpp_sampler_destroy": (C.c_int, [_h]),
    "mcmcpp_sampler_set_initial": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_set_initial_device": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_set_initial_local": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_get_state_local": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_set_slicing": (C.c_int, [_h, C.c_int, C.c_int]),
    "mcmcpp_sampler_run": (C.c_int, [_h, C.c_int, C.POINTER(C.c_float)]),
    "mcmcpp_sampler_get_counts": (C.c_int, [_h, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
    "mcmcpp_sampler_get_stored_steps": (C.c_int, [_h]),
    "mcmcpp_sampler_reset": (C.c_int, [_h]),
    "mcmcpp_sampler_store_current": (C.c_int, [_h]),
    "mcmcpp_sampler_slice_and_burn": (C.c_int, [_h, C.c_int, C.c_int]),
    "mcmcpp_sampler_chain": (C.c_int, [_h, C.POINTER(_dp), C.POINTER(C.c_longlong)]),
    "mcmcpp_sampler_chain_consume": (C.c_int, [_h, C.c_longlong]),
    "mcmcpp_sampler_chain_device": (C.c_int, [_h, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "mcmcpp_sampler_chain_covariance": (C.c_int, [_h, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_longlong)]),
    "mcmcpp_sampler_chain_autocorr_times": (C.c_int, [_h, C.c_int, C.c_int, C.c_void_p]),
    "mcmcpp_sampler_get_state": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_state_device": (C.c_int, [_h, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "mcmcpp_sampler_get_trace": (C.c_int, [_h, C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "mcmcpp_sampler_rng_redraws": (C.c_int, [_h, C.POINTER(C.c_int)]),
    "mcmcpp_sampler_launch_count": (C.c_int, [_h, C.POINTER(C.c_ulonglong)]),
    "mcmcpp_sampler_stream": (C.c_void_p, [_h]),
    "mcmcpp_last_error": (C.c_char_p, []),
    "mcmcpp_device_count": (C.c_int, []),
    "mcmcpp_sampler_set_fused_launcher": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mcmcpp_fp64_tensor_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "mcmcpp_device_log_check": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "mcmcpp_covariance_device_noalloc": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                  C.c_void_p]),
    "mcmcpp_comm_unique_id": (C.c_int, [C.c_void_p]),
    "mcmcpp_sampler_comm_init": (C.c_int, [_h, C.c_void_p]),
    "mcmcpp_sampler_peer_export": (C.c_int, [_h, C.c_void_p]),
    "mcmcpp_sampler_peer_import": (C.c_int, [_h, C.c_void_p]),
    "mcmcpp_sampler_set_group_launcher": (C.c_int, [_h, C.c_void_p]),
    "mcmcpp_sampler_peer_import_local": (C.c_int, [C.POINTER(_h), C.c_int]),
    "mcmcpp_sampler_run_group": (C.c_int, [C.POINTER(_h), C.c_int, C.c_int]),
    "mcmcpp_sampler_cov_begin": (C.c_int, [_h, C.c_int, C.c_int]),
    "mcmcpp_sampler_cov_result": (C.c_int, [_h, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_longlong)]),
    "mcmcpp_sampler_cov_end": (C.c_int, [_h]),
    "mcmcpp_covariance": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_longlong)]),
    "mcmcpp_sampler_set_step_action": (C.c_int, [_h, C.c_void_p, C.c_void_p]),
    "mcmcpp_sampler_checkpoint_bytes": (C.c_int, [_h, C.POINTER(C.c_size_t)]),
    "mcmcpp_sampler_checkpoint_save": (C.c_int, [_h, C.c_void_p, C.c_size_t]),
    "mcmcpp_sampler_checkpoint_restore": (C.c_int, [_h, C.c_void_p, C.c_size_t]),
    "mcmcpp_autocorr_times": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "mcmcpp_dim_slice": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
}

MOVER_STRETCH, MOVER_DIFF_EVOLUTION, MOVER_WALK = 0, 1, 2
STEP_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p)

_lib = None


class McmcppError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("mcmcpp_b200 error %d: %s" % (code, msg))
        self.code = code


def lib():
    """Load libmcmcpp_b200.so (built in-tree by __graft_entry__.build()); loud failure if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code):
    """Raise on negative status; return the (non-negative) status otherwise."""
    if code < 0:
        raise McmcppError(code, lib().mcmcpp_last_error().decode("utf-8", "replace"))
    return code

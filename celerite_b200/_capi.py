"""ctypes binding of the C ABI (include/celerite_b200.h).  No compute happens here.

The shared library must have been built in-tree (``python -m celerite_b200.build``); a missing
library or a missing GPU raises — there is deliberately no CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcelerite_b200.so")

OK, DIMENSION_MISMATCH, NOT_COMPUTED, NOT_POSITIVE_DEFINITE, CUDA_ERROR, INVALID_ARGUMENT, UNSUPPORTED = range(7)
EXCEEDS_STATIC_SCRATCH = 7
STATIC_SCRATCH_BYTES = 32 << 20
HOST_POINTERS = 1

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p
_i64 = C.c_int64

_COEFS = [C.c_double, C.c_int, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp]

# name -> (restype, argtypes); must list every symbol declared in include/celerite_b200.h
SIGNATURES = {
    "cel_create": (C.c_int, [C.POINTER(_vp), C.c_int]),
    "cel_destroy": (C.c_int, [_vp]),
    "cel_last_error": (C.c_char_p, [_vp]),
    "cel_version": (C.c_char_p, []),
    "cel_set_stream": (C.c_int, [_vp, _vp]),
    "cel_compute": (C.c_int, [_vp] + _COEFS + [C.c_int, _vp, _vp]),
    "cel_computed": (C.c_int, [_vp]),
    "cel_log_determinant": (C.c_int, [_vp, _dp]),
    "cel_dims": (C.c_int, [_vp, _ip, _ip]),
    "cel_dot_solve": (C.c_int, [_vp, C.c_int, _vp, _dp]),
    "cel_solve": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp]),
    "cel_dot_L": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp]),
    "cel_dot": (C.c_int, [_vp] + _COEFS + [C.c_int, _vp, C.c_int, _vp, _vp]),
    "cel_predict": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _vp, _vp]),
    "cel_grad_log_likelihood": (C.c_int, [_vp] + _COEFS + [C.c_int, _vp, _vp, _vp, _dp, _vp]),
    "cel_get_state": (C.c_int, [_vp, _ip, _ip, _ip, _dp, _vp, _vp, _vp, _vp]),
    "cel_set_state": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_double, _vp, _vp, _vp, _vp]),
    "cel_batched_log_likelihood": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                             _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint]),
    "cel_batched_log_likelihood_ragged": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                                    _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint]),
    "cel_batched_grad_log_likelihood": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                                  _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint]),
    "cel_batched_grad_log_likelihood_ragged": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                                         _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cel_batched_compute": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                      _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cel_batched_compute_ragged": (C.c_int, [_vp, _i64, C.c_int, _i64, C.c_int, C.c_int, _vp, _i64,
                                             _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cel_grad_log_likelihood_static_scratch": (C.c_int, [C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp,
                                                         _vp, _vp, _vp, _vp]),
    "cel_solve_device": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "cel_dot_L_device": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "cel_span_trim": (C.c_int, [_vp, C.c_int]),
    "cel_span_summarize": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp]),
    "cel_span_apply": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, C.c_int]),
    "cel_term_layout": (C.c_int, [C.c_int, _vp, _ip, _ip, _ip]),
    "cel_term_coefficients": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int64, _vp, C.c_int64, _vp, C.c_int64, _vp]),
    "cel_term_backprop": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int64, _vp, C.c_int64, _vp, C.c_int64, _vp, C.c_int64]),
    "cel_span_fold": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp]),
    "cel_span_log_determinant": (C.c_int, [_vp, _vp]),
    "cel_span_quad": (C.c_int, [_vp, _vp, _vp]),
    "cel_span_dims": (C.c_int, [_vp, _ip, _ip, _ip, _ip]),
    "cel_comm_unique_id": (C.c_int, [_vp]),
    "cel_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "cel_comm_destroy": (C.c_int, [_vp]),
    "cel_mgpu_bounds": (C.c_int, [_i64, C.c_int, C.c_int, C.POINTER(_i64), C.POINTER(_i64)]),
    "cel_compute_mgpu": (C.c_int, [_vp, C.c_double, C.c_int, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _i64]),
    "cel_solve_mgpu": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "cel_dot_L_mgpu": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "cel_log_likelihood_mgpu": (C.c_int, [_vp, _vp, _dp]),
    "cel_debug_math": (C.c_int, [C.c_int, _i64, _vp, _vp, _vp, _vp]),
    "cel_debug_guards_intact": (C.c_int, [_vp]),
    "cel_launch_count": (_i64, [_vp]),
}
COMM_ID_BYTES = 128

_lib = None


class CeleriteB200Error(RuntimeError):
    """Any non-OK status that has no counterpart among the reference's exceptions."""


class LinAlgError(Exception):
    """celerite::linalg_exception — registered as ``solver.LinAlgError`` (solver.cpp:87)."""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CeleriteB200Error(
                "libcelerite_b200.so is not built (run `python -m celerite_b200.build`); "
                "celerite_b200 has no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def last_error(handle):
    msg = lib().cel_last_error(handle)
    return msg.decode() if msg else ""


def check(rc, handle=None):
    """Map a C-ABI status onto the exception the reference raises at the Python boundary:
    linalg_exception -> LinAlgError, everything else -> RuntimeError (solver.cpp:87)."""
    if rc == OK:
        return
    msg = last_error(handle)
    if rc == NOT_POSITIVE_DEFINITE:
        raise LinAlgError(msg or "failed to factorize or solve matrix")
    if rc == DIMENSION_MISMATCH:
        raise RuntimeError(msg or "dimension mismatch")
    if rc == NOT_COMPUTED:
        raise RuntimeError(msg or "you must call 'compute' first")
    raise CeleriteB200Error("celerite_b200 status %d: %s" % (rc, msg))


def create(device=-1):
    h = _vp()
    rc = lib().cel_create(C.byref(h), device)
    if rc != OK:
        raise CeleriteB200Error("cel_create failed: " + last_error(None))
    return h

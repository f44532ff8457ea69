"""ctypes binding of libcno_b200.so (include/cno_b200.h).  Loaded lazily; a missing library is a
hard error on the product path (no fallback)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('CNO_B200_LIB') or os.path.join(_HERE, 'libcno_b200.so')

F32, BF16, F16, F64 = 0, 1, 2, 3
SIGN_NONE, SIGN_WRITE, SIGN_READ, SIGN_READ_SIGNONLY = 0, 1, 2, 3

_P, _I, _L, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
SIGNATURES = {
    'cno_version': (C.c_int, []),
    'cno_last_error': (C.c_char_p, []),
    'cno_bias_act': (_I, [_P, _P, _P, _P, _P, _P, _I, _L, _L, _L, _I, _I, _F, _F, _F, _P]),
    'cno_upfirdn2d': (_I, [_P, _P, _P, _I, _I, _L] + [_I] * 13 + [_F, _P]),
    'cno_filtered_lrelu': (_I, [_P] * 7 + [_I] * 19 + [_F, _F, _F, _I, _P]),
    'cno_filtered_lrelu_query': (_I, [_I] * 11),
    'cno_filtered_lrelu_plane_kernel': (_I, [_I]),
    'cno_filtered_lrelu_act': (_I, [_P, _P, _I, _L] + [_I] * 8 + [_F, _F, _F, _P]),
    'cno_cin_fwd': (_I, [_P] * 8 + [_I] * 4 + [_F, _P]),
    'cno_cin_bwd': (_I, [_P] * 9 + [_I] * 4 + [_P]),
    'cno_film_fwd': (_I, [_P] * 16 + [_I] * 5 + [_F, _P]),
    'cno_cin_fwd_res': (_I, [_P] * 9 + [_I] * 4 + [_F, _P]),
    'cno_film_fwd_res': (_I, [_P] * 17 + [_I] * 5 + [_F, _P]),
    'cno_film_stats': (_I, [_P] * 17 + [_I] * 5 + [_F, _P]),
    'cno_filtered_lrelu_affine': (_I, [_P] * 7 + [_I] * 19 + [_F, _F, _F, _I, _P, _P, _P]),
    'cno_film_bwd_finish': (_I, [_P] * 22 + [_I] * 3 + [_P]),
    'cno_cin_fwd_noalloc': (_I, [_P] * 6 + [_I] * 4 + [_F, _P]),
    'cno_adamw_chunk': (_I, []),
    'cno_adamw_multi': (_I, [_P, _P, _I, _P, C.c_double, C.c_double, C.c_double, _P]),
    'cno_nchw_to_nhwc': (_I, [_P, _P, _I, _I, _I, _I, _P]),
    'cno_nhwc_to_nchw': (_I, [_P, _P, _I, _I, _I, _I, _P]),
    'cno_probe_fma': (_I, [_P, _I, _I, _I, _I, _P]),
}

_lib = None


class CnoError(RuntimeError):
    pass


def load():
    """Returns the loaded library; raises if it has not been built (python pregenerating-pde_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CnoError(f'{LIB_PATH} is missing: build it with `python pregenerating-pde_b200/build.py` '
                           '(nvcc, sm_100a).  There is no fallback implementation.')
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc, what):
    """0 ok; -1 is returned to the caller (no specialised kernel); >0 raises like TORCH_CHECK / AT_CUDA_CHECK."""
    if rc > 0:
        msg = load().cno_last_error().decode()
        raise RuntimeError(f'{what}: {msg or "CUDA error"} (code {rc})')
    return rc

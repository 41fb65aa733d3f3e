"""ctypes binding of libob200.so (include/ob200.h).  No CPU fallback: a missing library raises.

The handle lives in a module global, never inside an algorithm instance: Orion deep-copies the
algorithm object on every producer update (`src/orion/core/worker/producer.py:161`).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OB200_LIB") or os.path.join(HERE, "libob200.so")  # OB200_LIB: tuning variants

OK, EINVAL, ECUDA, ENOTSPD, ENOMEM, ESTATE, EBOUNDS = 0, -1, -2, -3, -4, -5, -6
KIND_REAL, KIND_INT, KIND_CAT = 0, 1, 2
PRIOR_UNIFORM, PRIOR_LOGUNIFORM, PRIOR_NORMAL, PRIOR_CHOICE = 0, 1, 2, 3
KERNEL_MATERN52, KERNEL_RBF = 0, 1
ACQ_EI, ACQ_PI, ACQ_LCB = 0, 1, 2


class Slot(C.Structure):
    _fields_ = [("kind", C.c_int32), ("prior", C.c_int32), ("out_col", C.c_int32), ("width", C.c_int32),
                ("ncat", C.c_int32), ("cdf_off", C.c_int32), ("p0", C.c_double), ("p1", C.c_double),
                ("low", C.c_double), ("high", C.c_double), ("nlo", C.c_double), ("nhi", C.c_double)]


class Kernel(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_length_scale", C.c_int32), ("amplitude", C.c_double),
                ("length_scale_host", C.POINTER(C.c_double))]


class Best(C.Structure):
    _fields_ = [("value", C.c_double), ("index", C.c_int64), ("mu", C.c_double), ("sigma", C.c_double)]


class Ob200Error(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libob200 error %d: %s" % (code, message))
        self.code = code


class NotPositiveDefinite(Ob200Error):
    """Counterpart of numpy.linalg.LinAlgError raised by sklearn's fit (sklearn:_gpr.py:351-361)."""


_vp, _i, _i64, _u64, _d = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double
_pd = C.POINTER(C.c_double)

# every symbol declared in include/ob200.h: (restype, argtypes)
SIGNATURES = {
    "ob200_version": (_i, []),
    "ob200_ctx_create": (_i, [_i, C.POINTER(_vp)]),
    "ob200_ctx_destroy": (None, [_vp]),
    "ob200_last_error": (C.c_char_p, [_vp]),
    "ob200_set_stream": (_i, [_vp, _vp]),
    "ob200_sync": (_i, [_vp]),
    "ob200_launch_count": (_i64, [_vp]),
    "ob200_space_set": (_i, [_vp, C.POINTER(Slot), _i, _pd, _i]),
    "ob200_sample_transform": (_i, [_vp, _u64, _u64, _i64, _i, _vp, _i64, _vp]),
    "ob200_sample_status": (_i, [_vp]),
    "ob200_sample_transform_host": (_i, [_vp, _u64, _u64, _i64, _i, _vp, _i64, _vp]),
    "ob200_gp_fit": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(Kernel), _d, _i]),
    "ob200_gp_fit_search": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(Kernel), _d, _i]),
    "ob200_gp_fit_host": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(Kernel), _d, _i]),
    "ob200_gp_append": (_i, [_vp, _vp, _vp, _i, _i]),
    "ob200_gp_lml": (_i, [_vp, _pd]),
    "ob200_gp_lml_grad": (_i, [_vp, _pd, _pd, _i]),
    "ob200_gp_predict_grad_host": (_i, [_vp, _vp, _i64, _i, _vp]),
    "ob200_gp_get_factor": (_i, [_vp, _vp, _vp]),
    "ob200_gp_get_ystats": (_i, [_vp, _pd, _pd, _pd]),
    "ob200_gp_state_bytes": (_i, [_vp, _i, _i, C.POINTER(_i64)]),
    "ob200_gp_export": (_i, [_vp, _vp, _i64]),
    "ob200_gp_import": (_i, [_vp, _vp, _i64]),
    "ob200_score": (_i, [_vp, _vp, _i64, _i64, _u64, _u64, _i, _d, _d, _d, C.POINTER(Best), _vp, _vp, _vp]),
    "ob200_score_host": (_i, [_vp, _vp, _i64, _i64, _i, _d, _d, _d, C.POINTER(Best), _vp, _vp, _vp]),
    "ob200_score_async": (_i, [_vp, _vp, _i64, _i64, _u64, _u64, _i, _d, _d, _d, _vp, _vp, _vp, _vp]),
    "ob200_score_pruned": (_i, [_vp, _vp, _i64, _i64, _u64, _u64, _i, _d, _d, _d, C.POINTER(Best), C.POINTER(_i64)]),
    "ob200_score_pruned_multi": (_i, [_vp, _vp, _i64, _i64, _u64, _u64, _i, _d, _d, _d, C.POINTER(Best), C.POINTER(_i64)]),
    "ob200_score_bounds": (_i, [_vp, _vp, _i64, _i64, _vp, _vp, C.POINTER(_i)]),
    "ob200_prune_stats": (_i, [_vp, C.POINTER(_i64)]),
    "ob200_acq_argmax": (_i, [_vp, _vp, _vp, _i64, _u64, _i, _d, _d, _d, C.POINTER(Best), _vp]),
    "ob200_acq_argmax_static": (_i, [_vp, _vp, _i64, _u64, _i, _d, _d, _d, _vp, _vp, _vp, C.c_char_p, _i]),
}

_lib = None


def load():
    """Load libob200.so (built in-tree by `python -m metaopt_b200.build`).  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libob200.so is missing at %s: build it with `python -m metaopt_b200.build` "
                          "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(ctx, rc):
    if rc == OK:
        return
    msg = load().ob200_last_error(ctx)
    msg = msg.decode() if msg else ""
    if rc == ENOTSPD:
        raise NotPositiveDefinite(rc, msg)
    if rc == EBOUNDS:
        raise ValueError(msg)
    raise Ob200Error(rc, msg)

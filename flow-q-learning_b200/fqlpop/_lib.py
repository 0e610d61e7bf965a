"""ctypes loader for libfqlpop.so (the C-ABI declared in include/fqlpop.h).

There is no CPU fallback: if the shared library is missing the import fails loudly.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FQLPOP_LIB: another build of the same library (profiling tools load lib/libfqlpop_trace.so, `make TRACE=1`)
LIB_PATH = os.environ.get("FQLPOP_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libfqlpop.so")


class FqlPopError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise FqlPopError(
            f"{LIB_PATH} not found: build it with `make -C flow-q-learning_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback."
        )
    return ctypes.CDLL(LIB_PATH)


lib = _load()
lib.fqlpop_last_error.restype = ctypes.c_char_p
lib.fqlpop_version.restype = ctypes.c_int


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib.fqlpop_last_error().decode("utf-8", "replace")
        raise FqlPopError(f"{what} failed (status {rc}): {msg}")


class GemmTestDesc(ctypes.Structure):
    _fields_ = [
        ("M", ctypes.c_int32), ("N", ctypes.c_int32), ("K", ctypes.c_int32), ("P", ctypes.c_int32),
        ("a_mn", ctypes.c_int32), ("b_mn", ctypes.c_int32), ("n_sub", ctypes.c_int32),
        ("epi_kind", ctypes.c_int32), ("act", ctypes.c_int32), ("ln_eps", ctypes.c_float),
        ("A", ctypes.c_void_p), ("a_ld", ctypes.c_int64), ("a_ms", ctypes.c_int64),
        ("B", ctypes.c_void_p), ("b_ld", ctypes.c_int64), ("b_ms", ctypes.c_int64),
        ("bias", ctypes.c_void_p), ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
        ("out_h", ctypes.c_void_p), ("out_gp", ctypes.c_void_p), ("out_xhat", ctypes.c_void_p),
        ("out_rstd", ctypes.c_void_p),
        ("gp", ctypes.c_void_p), ("xhat", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
        ("out_f32", ctypes.c_void_p),
        ("dbias_part", ctypes.c_void_p), ("dgamma_part", ctypes.c_void_p), ("dbeta_part", ctypes.c_void_p),
        ("n_active", ctypes.c_void_p), ("member_list", ctypes.c_void_p),
    ]


lib.fqlpop_gemm_test.argtypes = [ctypes.POINTER(GemmTestDesc), ctypes.c_void_p]
lib.fqlpop_gemm_test.restype = ctypes.c_int

EPI_STORE_F32, EPI_STORE_BF16, EPI_FWD, EPI_FWD_LN, EPI_DGRAD, EPI_DGRAD_LN = range(6)
ACT_NONE, ACT_GELU, ACT_RELU = range(3)


def gemm_test(**kw):
    """Launch one grouped GEMM. Tensor arguments are torch CUDA tensors (or None)."""
    d = GemmTestDesc()
    for k, v in kw.items():
        if hasattr(v, "data_ptr"):
            v = v.data_ptr()
        setattr(d, k, v)
    check(lib.fqlpop_gemm_test(ctypes.byref(d), None), "fqlpop_gemm_test")


def default_device():
    """The device a handle lands on when the caller names none: one process per GPU, so LOCAL_RANK (torchrun) if set,
    else torch's current device -- never a hard-coded cuda:0."""
    import os
    import torch
    lr = os.environ.get("LOCAL_RANK")
    if lr is not None and torch.cuda.is_available() and int(lr) < torch.cuda.device_count():
        return torch.device("cuda", int(lr))
    return torch.device("cuda", torch.cuda.current_device() if torch.cuda.is_available() else 0)

"""ctypes binding of libvitb200.so (the C ABI declared in include/vitb200.h).

There is no CPU or eager fallback: if the shared object is missing, or the device is not sm_100, every
entry point raises.  Tensors are passed as raw device pointers on torch's current CUDA stream.
"""
import ctypes
import os

import torch

F32, BF16 = 0, 1
MAJOR_K, MAJOR_MN = 0, 1

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libvitb200.so")

c_void_p, c_int, c_int64, c_float, c_uint64 = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float,
                                               ctypes.c_uint64)


class Dropout(ctypes.Structure):
    _fields_ = [("p", c_float), ("seed", c_uint64), ("stream", c_uint64)]


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ("M", c_int), ("N", c_int), ("K", c_int),
        ("a_major", c_int), ("b_major", c_int),
        ("in_dtype", c_int), ("out_dtype", c_int),
        ("A", c_void_p), ("lda", c_int64),
        ("B", c_void_p), ("ldb", c_int64),
        ("C", c_void_p), ("ldc", c_int64),
        ("bias", c_void_p),
        ("relu", c_int),
        ("mask_aux", c_void_p),
        ("dropout", Dropout),
        ("row_add", c_void_p),
        ("row_period", c_int),
        ("row_group_stride", c_int),
        ("residual", c_void_p),
        ("splits", c_int),
        ("accumulate", c_int),
    ]


_lib = None


def lib():
    """Load (once) and return the shared library; raise if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python vision-transformer_b200/build.py` "
                "(libvitb200 has no CPU / PyTorch fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.vitb200_last_error.restype = ctypes.c_char_p
        _declare(_lib)
    return _lib


def _declare(L):
    L.vitb200_version.restype = c_int
    L.vitb200_device_ok.restype = c_int
    L.vitb200_gemm.argtypes = [ctypes.POINTER(GemmArgs), c_void_p]
    L.vitb200_gemm.restype = c_int


def check(rc):
    if rc != 0:
        raise RuntimeError("libvitb200: " + lib().vitb200_last_error().decode())


def stream_ptr():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def dtype_code(dt):
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"libvitb200 supports float32 and bfloat16, got {dt}")


def ptr(t):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libvitb200: CPU tensors are not supported (no CPU fallback); move the model and "
                           "inputs to a CUDA device")
    return c_void_p(t.data_ptr())


def gemm(A, B, C, *, a_major=MAJOR_K, b_major=MAJOR_K, M, N, K, bias=None, relu=False, mask_aux=None,
         dropout=(0.0, 0, 0), row_add=None, row_period=0, row_group_stride=0, residual=None, splits=0,
         accumulate=False, lda=None, ldb=None, ldc=None):
    """C[m,n] = epilogue(sum_k A(m,k) B(n,k)); see vitb200_gemm_args in include/vitb200.h."""
    a = GemmArgs()
    a.M, a.N, a.K = M, N, K
    a.a_major, a.b_major = a_major, b_major
    a.in_dtype, a.out_dtype = dtype_code(A.dtype), dtype_code(C.dtype)
    a.A, a.lda = ptr(A), lda if lda is not None else A.stride(0)
    a.B, a.ldb = ptr(B), ldb if ldb is not None else B.stride(0)
    a.C, a.ldc = ptr(C), ldc if ldc is not None else C.stride(0)
    a.bias = ptr(bias)
    a.relu = int(relu)
    a.mask_aux = ptr(mask_aux)
    a.dropout = Dropout(float(dropout[0]), int(dropout[1]), int(dropout[2]))
    a.row_add = ptr(row_add)
    a.row_period, a.row_group_stride = row_period, row_group_stride
    a.residual = ptr(residual)
    a.splits, a.accumulate = splits, int(accumulate)
    check(lib().vitb200_gemm(ctypes.byref(a), stream_ptr()))
    return C

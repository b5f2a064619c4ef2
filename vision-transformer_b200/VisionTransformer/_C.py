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
        ("col_sum", c_void_p),
        ("bits_out", c_void_p),
        ("mask_bits", c_void_p),
        ("bits_rows", c_int),
        ("dot_aux", c_void_p),
        ("dot_out", c_void_p),
        ("dot_T", c_int),
        ("dot_H", c_int),
    ]


class AdamWHyper(ctypes.Structure):
    _fields_ = [("lr", c_float), ("beta1", c_float), ("beta2", c_float), ("eps", c_float), ("weight_decay", c_float),
                ("step", c_int), ("grad_scale", c_float)]


class Dims(ctypes.Structure):
    _fields_ = [("B", c_int), ("T", c_int), ("E", c_int), ("H", c_int),
                ("C", c_int), ("P", c_int), ("Hi", c_int), ("Wi", c_int),
                ("classes", c_int), ("dtype", c_int), ("dropout_p", c_float), ("seed", c_uint64),
                ("ln_eps", c_float), ("seed_dev", c_void_p)]


class ProfEntry(ctypes.Structure):
    _fields_ = [("ms", ctypes.c_double), ("work", ctypes.c_double), ("launches", c_int64)]


PROF_CLASSES = 8


def _ptr_struct(name, fields):
    return type(name, (ctypes.Structure,), {"_fields_": [(f, c_void_p) for f in fields]})


EMBED_FIELDS_P = ("w_f32", "w_lp", "bias", "cls", "pos")
EMBED_FIELDS_G = ("w", "bias", "cls", "pos")
BLOCK_FIELDS_P = ("ln1_w", "ln1_b", "ln2_w", "ln2_b", "w_qkv", "w_proj", "b_proj", "w_fc1", "b_fc1", "w_fc2", "b_fc2")
BLOCK_FIELDS_G = ("ln1_w", "ln1_b", "ln2_w", "ln2_b", "w_qkv", "w_proj", "b_proj", "w_fc1", "b_fc1", "w_fc2", "b_fc2")
BLOCK_GRADS_STRUCT = BLOCK_FIELDS_G + ("b_fc2_prev",)  # vitb200_block_grads: own gradients + the chained bias pointer
HEAD_FIELDS = ("w1", "b1", "ln_w", "ln_b", "w3", "b3")
HEAD_PARAMS_STRUCT = HEAD_FIELDS + ("w1_lp",)  # vitb200_head_params: fp32 parameters + the optional bf16 copy of w1
EmbedParams = _ptr_struct("EmbedParams", EMBED_FIELDS_P)
EmbedGrads = _ptr_struct("EmbedGrads", EMBED_FIELDS_G)
BlockParams = _ptr_struct("BlockParams", BLOCK_FIELDS_P)
BlockGrads = _ptr_struct("BlockGrads", BLOCK_GRADS_STRUCT)
HeadParams = _ptr_struct("HeadParams", HEAD_PARAMS_STRUCT)
HeadGrads = _ptr_struct("HeadGrads", HEAD_FIELDS)

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
    L.vitb200_launch_count.restype = c_int64
    L.vitb200_gemm_auto_splits.argtypes = [c_int, c_int, c_int, c_int]
    L.vitb200_gemm_auto_splits.restype = c_int
    L.vitb200_abi_sizeof.argtypes = [c_int]
    L.vitb200_abi_sizeof.restype = c_int64
    L.vitb200_prof_class_name.argtypes = [c_int]
    L.vitb200_prof_class_name.restype = ctypes.c_char_p
    L.vitb200_prof_begin.restype = c_int
    L.vitb200_prof_end.argtypes = [ctypes.POINTER(ProfEntry)]
    L.vitb200_prof_end.restype = c_int
    P = ctypes.POINTER
    sig = {
        "vitb200_ln_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p],
        "vitb200_ln_bwd": [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                           c_void_p, c_void_p, Dropout, c_int, c_int, c_void_p],
        "vitb200_attn_fwd": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p],
        "vitb200_attn_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                             c_float, c_void_p],
        "vitb200_attn_bwd_noalloc": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                     c_float, c_void_p],
        "vitb200_attn_probs": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p],
        "vitb200_im2col": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
        "vitb200_colsum": [c_void_p, c_int, c_int64, c_void_p, c_int, c_int, c_void_p],
        "vitb200_cross_entropy": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p],
        "vitb200_cast_f32_bf16": [c_void_p, c_void_p, c_int64, c_void_p],
        "vitb200_dropout_mask": [c_void_p, c_int64, Dropout, c_void_p],
        "vitb200_mask_cast": [c_void_p, c_void_p, c_int, Dropout, c_int64, c_void_p],
        "vitb200_adamw_flat": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, AdamWHyper, c_void_p],
        "vitb200_resize_bilinear_u8": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
                                       c_void_p, c_void_p, c_int, c_void_p],
        "vitb200_grad_pack": [c_void_p, c_void_p, c_int64, c_void_p],
        "vitb200_adamw_range": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, AdamWHyper,
                                c_void_p],
        "vitb200_adamw_range_dev": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                    AdamWHyper, c_void_p, c_void_p],
        "vitb200_step_state_advance": [c_void_p, c_float, c_float, c_uint64, c_void_p],
        "vitb200_adamw_multi": [c_void_p] * 8 + [c_int, AdamWHyper, c_void_p],
        "vitb200_embed_fwd": [P(Dims), P(EmbedParams), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_embed_fwd_u8": [P(Dims), P(EmbedParams), c_void_p, c_int, c_void_p, c_void_p, c_void_p],
        "vitb200_im2col_u8": [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
        "vitb200_embed_bwd": [P(Dims), P(EmbedGrads), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_block_fwd": [P(Dims), c_int, P(BlockParams), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_block_bwd": [P(Dims), c_int, P(BlockParams), P(BlockGrads), c_void_p, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_void_p],
        "vitb200_block_fwd_row0": [P(Dims), c_int, P(BlockParams), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_block_bwd_row0": [P(Dims), c_int, P(BlockParams), P(BlockGrads), c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_head_fwd_row0": [P(Dims), P(HeadParams), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_head_bwd_row0": [P(Dims), c_int, P(HeadParams), P(HeadGrads), c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p],
        "vitb200_head_fwd": [P(Dims), P(HeadParams), c_void_p, c_void_p, c_void_p, c_void_p],
        "vitb200_head_bwd": [P(Dims), c_int, P(HeadParams), P(HeadGrads), c_void_p, c_void_p, c_void_p, c_void_p,
                             c_void_p, c_void_p],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = c_int
    for name in ("vitb200_embed_act_bytes", "vitb200_block_act_bytes", "vitb200_block_scratch_bytes",
                 "vitb200_head_act_bytes"):
        fn = getattr(L, name)
        fn.argtypes = [P(Dims)]
        fn.restype = c_int64
    L.vitb200_block_act_offset.argtypes = [P(Dims), c_int]
    L.vitb200_block_act_offset.restype = c_int64


EXPORTS = (
    "vitb200_last_error", "vitb200_version", "vitb200_device_ok", "vitb200_launch_count", "vitb200_prof_class_name",
    "vitb200_prof_begin", "vitb200_prof_end", "vitb200_abi_sizeof", "vitb200_gemm", "vitb200_gemm_auto_splits", "vitb200_ln_fwd", "vitb200_ln_bwd",
    "vitb200_attn_fwd", "vitb200_attn_bwd", "vitb200_attn_bwd_noalloc", "vitb200_attn_probs", "vitb200_im2col", "vitb200_colsum",
    "vitb200_cross_entropy", "vitb200_cast_f32_bf16", "vitb200_dropout_mask", "vitb200_mask_cast", "vitb200_adamw_flat",
    "vitb200_resize_bilinear_u8", "vitb200_grad_pack", "vitb200_adamw_range", "vitb200_adamw_range_dev",
    "vitb200_step_state_advance", "vitb200_adamw_multi", "vitb200_embed_act_bytes", "vitb200_embed_fwd", "vitb200_embed_fwd_u8", "vitb200_im2col_u8", "vitb200_embed_bwd",
    "vitb200_block_act_bytes", "vitb200_block_scratch_bytes", "vitb200_block_act_offset", "vitb200_block_fwd", "vitb200_block_bwd",
    "vitb200_head_act_bytes", "vitb200_head_fwd", "vitb200_head_bwd", "vitb200_block_fwd_row0",
    "vitb200_block_bwd_row0", "vitb200_head_fwd_row0", "vitb200_head_bwd_row0",
)


def launch_count():
    return int(lib().vitb200_launch_count())


def prof_begin():
    check(lib().vitb200_prof_begin())


def prof_end():
    """-> {class name: {"ms", "work", "launches"}} for the kernels launched since prof_begin()."""
    arr = (ProfEntry * PROF_CLASSES)()
    check(lib().vitb200_prof_end(arr))
    return {lib().vitb200_prof_class_name(i).decode(): {"ms": arr[i].ms, "work": arr[i].work,
                                                         "launches": int(arr[i].launches)}
            for i in range(PROF_CLASSES)}


def check(rc):
    if rc != 0:
        raise RuntimeError("libvitb200: " + lib().vitb200_last_error().decode())


ACT_QKV, ACT_ATTN, ACT_LSE, ACT_XMID = 0, 1, 2, 3
NOALLOC_MAX_DELTA = 1 << 20   # VITB200_NOALLOC_MAX_DELTA
STEP_STATE_BYTES = 32         # VITB200_STEP_STATE_BYTES


def stream_ptr(device=None):
    """torch's current stream of `device` (default: the current device).  Callers that hold tensors on a device other
    than the current one must run inside `torch.cuda.device(dev)` (the Functions and DPTrainer do)."""
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


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
         accumulate=False, col_sum=None, bits_out=None, mask_bits=None, bits_rows=0, lda=None, ldb=None, ldc=None,
         dot_aux=None, dot_out=None, dot_T=0, dot_H=0):
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
    a.col_sum = ptr(col_sum)
    a.bits_out, a.mask_bits, a.bits_rows = ptr(bits_out), ptr(mask_bits), int(bits_rows)
    a.dot_aux, a.dot_out, a.dot_T, a.dot_H = ptr(dot_aux), ptr(dot_out), int(dot_T), int(dot_H)
    check(lib().vitb200_gemm(ctypes.byref(a), stream_ptr()))
    return C

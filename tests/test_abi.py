"""CPU: the C-ABI shared library loads and exports every symbol include/vitb200.h declares (no compute calls)."""
import ctypes
import os
import re

from conftest import ROOT
from VisionTransformer import _C


def _declared():
    src = open(os.path.join(ROOT, "include", "vitb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vitb200_[a-z0-9_]+)\s*\(", src)))


def test_library_exists_and_loads():
    assert os.path.exists(_C.LIB_PATH), "run `python __graft_entry__.py` (build) first"
    L = ctypes.CDLL(_C.LIB_PATH)
    assert L.vitb200_version() >= 100


def test_every_declared_symbol_is_exported():
    L = ctypes.CDLL(_C.LIB_PATH)
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/vitb200.h but not exported"


def test_python_binding_table_covers_header():
    assert set(_C.EXPORTS) == set(_declared())


def test_struct_layouts_match_header_sizes():
    # sizes computed from the C declarations (LP64): guards against silent ctypes / header drift
    assert ctypes.sizeof(_C.Dropout) == 24
    assert ctypes.sizeof(_C.Dims) == 72
    assert ctypes.sizeof(_C.AdamWHyper) == 28
    assert ctypes.sizeof(_C.BlockParams) == 11 * 8
    assert ctypes.sizeof(_C.GemmArgs) == 216
    # and against the sizes the library itself was compiled with
    lib = _C.lib()
    lib.vitb200_abi_sizeof.restype = ctypes.c_int64
    for which, cls in enumerate((_C.Dims, _C.Dropout, _C.GemmArgs, _C.AdamWHyper, _C.BlockParams, _C.BlockGrads,
                                 _C.HeadParams, _C.ProfEntry)):
        assert lib.vitb200_abi_sizeof(which) == ctypes.sizeof(cls), cls.__name__
    assert lib.vitb200_abi_sizeof(99) == -1


def test_error_channel_without_gpu():
    L = _C.lib()
    # no device in the CPU container -> refuses, never crashes
    assert L.vitb200_device_ok() in (0, 1)
    assert isinstance(L.vitb200_last_error(), bytes)


def test_split_k_rule_of_the_weight_gradient_gemms():
    """Host arithmetic of vitb200_gemm_auto_splits (no GPU): on the static persistent walk over 74 CTA pairs the split count
    must leave no partly filled extra round — the 'about two waves' rule of round 1 ran ViT-B/16's dW sites as three rounds
    at 69-81 % occupancy (profiles/r02_gemm_sites_split_rule_{old,new}.log)."""
    L = ctypes.CDLL(_C.LIB_PATH)
    L.vitb200_gemm_auto_splits.argtypes = [ctypes.c_int] * 4
    L.vitb200_gemm_auto_splits.restype = ctypes.c_int
    cd = lambda a, b: -(-a // b)  # noqa: E731
    tok = 256 * 197
    # ViT-B/16 at batch 256: dW1 / dW2 (36 tiles) -> 2 splits = 72 units, dWqkv (27) -> 8 = 216, dWproj (9) -> 8 = 72
    assert L.vitb200_gemm_auto_splits(3072, 768, tok, 148) == 2
    assert L.vitb200_gemm_auto_splits(768, 3072, tok, 148) == 2
    assert L.vitb200_gemm_auto_splits(2304, 768, tok, 148) == 8
    assert L.vitb200_gemm_auto_splits(768, 768, tok, 148) == 8
    assert L.vitb200_gemm_auto_splits(0, 8, 8, 148) == -1
    # every weight-gradient site of every BASELINE config: occupancy of the walk >= 0.85 and >= 4 k-blocks per split
    for E, tokens in ((384, 64 * 197), (768, tok), (1024, tok), (1280, 128 * 577), (1280, 64 * 730)):
        for M, N in ((4 * E, E), (E, 4 * E), (3 * E, E), (E, E)):
            s = L.vitb200_gemm_auto_splits(M, N, tokens, 148)
            kb = cd(tokens, 64)
            per = cd(kb, s)
            units = cd(M, 256) * cd(N, 256) * cd(kb, per)
            assert 1 <= s <= 64 and per >= 4
            assert units / (cd(units, 74) * 74) >= 0.85, (E, M, N, s, units)

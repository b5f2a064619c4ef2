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

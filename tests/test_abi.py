"""CPU: the C-ABI library loads and exports every symbol include/objcavit_b200.h declares."""
import ctypes
import os

import pytest

from objcavit_b200 import _lib


def test_header_prototypes_cover_each_other():
    declared = _lib.declared_symbols()
    assert declared, "no symbols parsed from the header"
    assert sorted(_lib.PROTOTYPES) == declared


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        from objcavit_b200 import build
        build.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in _lib.declared_symbols():
        assert hasattr(L, name), f"{name} is declared in the header but not exported"
    lib = _lib.lib()
    assert lib.ocv_abi_version() == 1
    assert b"sm_100" in lib.ocv_status_string(-4)


def test_argument_validation_without_gpu():
    """Error convention: bad arguments are rejected with a negative status before anything touches the device."""
    lib = _lib.lib()
    assert lib.ocv_linear_f32(None, None, None, None, None, 1, 1, 1, 1, 1, 1, 1, 0, 1, 0, 0, 0, None) == -1
    assert lib.ocv_bins_f32(1, 1, None, None, 1, 5000, 0, 0.0, 1.0, None) == -2
    assert lib.ocv_mha_core_f32(1, 1, 1, 1, None, 1, 1, 1, 4, 64, 0, 0, 0, 0, 0, 0, 0, 0, None) == -2

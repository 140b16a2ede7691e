"""CPU: the C-ABI library loads, exports every symbol include/acai_b200.h declares, and the
product path refuses to run without a CUDA device (no fallback)."""
import ctypes

import pytest
import torch

from acai_omr_b200 import _lib, ops


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _lib.header_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/acai_b200.h but not exported"
    assert set(declared) == set(_lib.SIGNATURES), "ctypes signatures out of sync with the header"
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        getattr(raw, name)
    assert lib.acai_version() >= 100


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("this check is for the GPU-less container")
    with pytest.raises(_lib.AcaiError):
        ops.linear(torch.zeros(2, 4), torch.zeros(3, 4))
    with pytest.raises(_lib.AcaiError):
        ops.layernorm(torch.zeros(2, 4), torch.ones(4), torch.zeros(4), 1e-5)
    assert _lib.load().acai_device_ok() == 0

"""CPU: the C-ABI library loads and exports exactly the symbols include/mft_b200.h declares (no compute)."""
import os
import re

from conftest import ROOT


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "mft_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(mft_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    import mobilefinetuner_b200.capi as capi
    decl = declared_symbols()
    assert len(decl) >= 35
    for name in decl:
        assert hasattr(capi.lib, name), f"{name} declared in mft_b200.h but not exported"
        assert name in capi.SIGNATURES, f"{name} has no ctypes signature"
    assert sorted(capi.SIGNATURES) == decl
    assert capi.lib.mft_version() >= 100


def test_struct_layouts_match_header():
    import ctypes as C
    import mobilefinetuner_b200.capi as capi
    assert C.sizeof(capi.TensorSlot) == 56
    assert C.sizeof(capi.GemmArgs) == 240      # 224 + c_bf16, ldc_bf16


def test_no_cpu_fallback_on_missing_device():
    """Without a GPU the device check must fail loudly rather than compute on the host."""
    import torch
    import mobilefinetuner_b200.capi as capi
    if torch.cuda.is_available():
        return
    assert capi.lib.mft_device_check() != 0
    assert capi.lib.mft_last_error()

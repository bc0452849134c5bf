"""CPU-side checks of the C-ABI library: it loads, exports every symbol the header
declares, agrees with the oracle on the flat layout, and fails loudly (no CPU
fallback) when asked to compute without a GPU."""
import ctypes as C
import os
import re

import pytest
import torch

from multidisttorch_b200 import _lib
from oracle import vae_numpy as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "mdt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mdt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/mdt_b200.h but not exported"
    assert sorted(_lib.exported_symbols()) == names, "ctypes prototypes and header disagree"
    assert lib.mdt_version() == 200


@pytest.mark.parametrize("H", [400, 4096])
def test_layout_matches_oracle(H):
    lib = _lib.load()
    cfg = _lib.mdt_cfg(784, H, 20, 128, 0, 0)
    off, num, tot = (C.c_int64 * 8)(), (C.c_int64 * 8)(), C.c_int64()
    assert lib.mdt_param_layout(C.byref(cfg), off, num, C.byref(tot)) == 0
    lay, total = O.layout(O.Dims(H=H))
    assert tot.value == total
    for i, k in enumerate(O.SEGMENTS):
        assert (off[i], num[i]) == lay[k]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    lib = _lib.load()
    cfg = _lib.mdt_cfg(784, 400, 20, 128, 0, 0)
    h = C.c_void_p()
    rc = lib.mdt_engine_create(C.byref(cfg), None, 1, 0, None, C.byref(h))
    assert rc == -2 and b"CUDA" in lib.mdt_last_error()          # MDT_ERR_CUDA, never a silent CPU path
    from multidisttorch_b200.engine import VAEEngine
    with pytest.raises(RuntimeError, match="no CPU path"):
        VAEEngine()


def test_bad_arguments_are_rejected_with_messages():
    lib = _lib.load()
    cfg = _lib.mdt_cfg(0, 400, 20, 128, 0, 0)
    assert lib.mdt_param_layout(C.byref(cfg), None, None, None) == -1
    assert b"bad cfg" in lib.mdt_last_error()
    cfg = _lib.mdt_cfg(784, 400, 20, 128, 7, 0)
    h = C.c_void_p()
    assert lib.mdt_engine_create(C.byref(cfg), None, 1, 0, None, C.byref(h)) == -1
    assert b"gemm_mode" in lib.mdt_last_error()

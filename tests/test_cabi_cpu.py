"""CPU: the C-ABI shared library builds for sm_100a, loads, and exports every symbol include/sailrecon_b200.h
declares; the ctypes structs mirror the C structs; compute entry points fail loudly without a B200."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "sailrecon_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sr_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from sailrecon_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in _lib.SIGNATURES, f"{n} declared in the header but not bound in _lib.SIGNATURES"
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.sr_version() >= 100


def test_struct_layouts_match_header():
    from sailrecon_b200 import _lib
    # sizes / offsets printed by a C program including include/sailrecon_b200.h (gcc, LP64)
    assert ctypes.sizeof(_lib.SrEpilogue) == 152
    assert ctypes.sizeof(_lib.SrAttn) == 176
    assert _lib.SrEpilogue.pos.offset == 144 and _lib.SrAttn.carry_in.offset == 168


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from sailrecon_b200 import _lib, ops
    lib = _lib.load()
    assert lib.sr_arch_check() == _lib.SR_ERR_ARCH
    assert b"no CPU fallback" in lib.sr_last_error()
    with pytest.raises(RuntimeError):
        ops.layernorm(torch.zeros(4, 64), None, None, 1e-5, torch.zeros(4, 64))

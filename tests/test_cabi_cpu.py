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


def test_struct_layouts_match_header(tmp_path):
    """ctypes mirrors of the C-ABI structs against what gcc computes from include/sailrecon_b200.h (LP64)."""
    import subprocess
    from sailrecon_b200 import _lib
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "sailrecon_b200.h")
    src = tmp_path / "layout.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "%s"\n'
        'int main(void) { printf("%%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu\\n", sizeof(sr_epilogue_t), offsetof(sr_epilogue_t, pos), '
        'sizeof(sr_attn_t), offsetof(sr_attn_t, carry_in), offsetof(sr_attn_t, lse_pad), sizeof(sr_attn_bwd_t), '
        'offsetof(sr_attn_bwd_t, num_heads), offsetof(sr_attn_bwd_t, scale), sizeof(sr_reproj_loss_t), '
        'offsetof(sr_reproj_loss_t, min_val), offsetof(sr_reproj_loss_t, tables)); return 0; }\n' % hdr)
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-o", str(exe), str(src)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    assert got == [ctypes.sizeof(_lib.SrEpilogue), _lib.SrEpilogue.pos.offset, ctypes.sizeof(_lib.SrAttn), _lib.SrAttn.carry_in.offset,
                   _lib.SrAttn.lse_pad.offset, ctypes.sizeof(_lib.SrAttnBwd), _lib.SrAttnBwd.num_heads.offset,
                   _lib.SrAttnBwd.scale.offset, ctypes.sizeof(_lib.SrReprojLoss), _lib.SrReprojLoss.min_val.offset,
                   _lib.SrReprojLoss.tables.offset]
    assert ctypes.sizeof(_lib.SrEpilogue) == 160 and _lib.SrAttn.carry_in.offset == 168


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from sailrecon_b200 import _lib, ops
    lib = _lib.load()
    assert lib.sr_arch_check() == _lib.SR_ERR_ARCH
    assert b"no CPU fallback" in lib.sr_last_error()
    with pytest.raises(RuntimeError):
        ops.layernorm(torch.zeros(4, 64), None, None, 1e-5, torch.zeros(4, 64))


def test_attention_work_decomposition_plan():
    """sr_attn_plan (host only): four-tile CTAs first, three-tile CTAs in the tail wave.  16-anchor global shape on 148 SMs:
    172 query tiles per head = 37 x 4 + 8 x 3 (592 heavy CTAs = exactly 4 waves) — and every plan covers all the tiles."""
    import ctypes as C
    from sailrecon_b200 import _lib
    lib = _lib.load()

    def plan(q_len, heads, batch, sms=148):
        n4, n3 = C.c_int32(), C.c_int32()
        assert lib.sr_attn_plan(q_len, heads, batch, sms, C.byref(n4), C.byref(n3)) == 0
        return n4.value, n3.value

    assert plan(16 * 1374, 16, 1) == (37, 8)
    for q_len, heads, batch in ((1374, 16, 32), (1374, 16, 16), (1, 16, 1), (128, 2, 3), (512 * 1374, 16, 1), (70, 16, 3), (600, 16, 1)):
        n4, n3 = plan(q_len, heads, batch)
        t = (q_len + 127) // 128
        assert n4 >= 0 and n3 >= 0 and n4 + n3 >= 1
        assert 4 * n4 + 3 * n3 >= t and (4 * n4 + 3 * (n3 - 1) < t if n3 else 4 * (n4 - 1) < t), (q_len, n4, n3)

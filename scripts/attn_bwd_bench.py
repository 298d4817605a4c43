#!/usr/bin/env python
"""Isolated timing of sr_attn_fwd_bf16 (with lse) and sr_attn_bwd_bf16 on the shapes of the 24-frame training step
(BASELINE.json configs[4]: 12 anchors + 12 queries, 518 px): frame blocks (24 x 1374), global block (12 x 1374 rows),
reloc block (12 query frames against a 12 x 305 scene prefix).  TFLOP/s on algorithmic flops: forward 4 L_q L_k C,
backward 10 L_q L_k C (five of the seven GEMMs the two-pass kernel runs are counted)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sailrecon_b200 import ops  # noqa: E402

H, d, P = 16, 64, 1374
C = H * d
dev = torch.device("cuda")
cases = [("frame24", 24, P, 0), ("global12", 1, 12 * P, 0), ("reloc12", 12, P, 12 * 305)]
for name, batch, L, pre in cases:
    g = torch.Generator(device="cuda").manual_seed(1)
    qkv = torch.randn(batch * L, 3 * C, device=dev, generator=g).bfloat16()
    dout = torch.randn(batch * L, C, device=dev, generator=g).bfloat16()
    pkv = torch.randn(max(pre, 1), 2 * C, device=dev, generator=g).bfloat16()
    q, k, v = (qkv[:, i * C:(i + 1) * C] for i in range(3))
    out = torch.empty(batch * L, C, dtype=torch.bfloat16, device=dev)
    lse = ops.new_lse(H, batch, L, dev)
    dqkv = torch.empty_like(qkv)
    dpkv = torch.empty_like(pkv)
    kw = dict(prefix_k=pkv[:, :C], prefix_v=pkv[:, C:], prefix_len=pre) if pre else {}
    kwb = dict(dprefix_k=dpkv[:, :C], dprefix_v=dpkv[:, C:], **kw) if pre else {}
    fwd = lambda: ops.attention(q, k, v, out, H, d, batch, L, L, L, L, lse=lse, **kw)
    bwd = lambda: ops.attention_bwd(q, k, v, out, dout, lse, dqkv[:, :C], dqkv[:, C:2 * C], dqkv[:, 2 * C:], H, d, batch, L, L,
                                    L, L, **kwb)
    res = {}
    for label, fn, mult in (("fwd", fwd, 4.0), ("bwd", bwd, 10.0)):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        fl = mult * batch * L * (L + pre) * C
        if label == "bwd" and pre:      # scene rows do not attend in this call; their K/V gradient pass is counted above
            pass
        res[label] = (ms, fl / ms / 1e9)
    print(f"{name:9s} fwd {res['fwd'][0]:8.3f} ms {res['fwd'][1]:7.1f} TF/s | bwd {res['bwd'][0]:8.3f} ms {res['bwd'][1]:7.1f} TF/s", flush=True)

#!/usr/bin/env python
"""Dump the pipeline trace of sr_attn_fwd_bf16 (SR_ATTN_TRACE=1) on the 16-anchor global shape: per K/V tile, the SM-clock
deltas between the hand-offs of softmax warp 4, the MMA issuer of tile 0 and the TMA producer of CTA (0,0,0)."""
import ctypes as C
import os
import sys

os.environ["SR_ATTN_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sailrecon_b200 import _lib, ops

dev = "cuda"
H, D, Cc = 16, 64, 1024
L = 16 * 1374
g = torch.Generator(device=dev).manual_seed(0)
qkv = torch.randn(L, 3 * Cc, device=dev, generator=g).to(torch.bfloat16)
out = torch.empty(L, Cc, device=dev, dtype=torch.bfloat16)
QL = int(os.environ.get("TRACE_Q_LEN", L))   # 128 -> one query tile per CTA: a lone softmax warp per sub-partition
for _ in range(3):
    ops.attention(qkv[:, :Cc], qkv[:, Cc:2 * Cc], qkv[:, 2 * Cc:], out, H, D, 1, QL, L)
torch.cuda.synchronize()
buf = (C.c_longlong * (3 * 64 * 8))()
_lib.check(_lib.load().sr_debug_trace_read(buf, 3 * 64 * 8))
t = torch.tensor(list(buf)).view(3, 64, 8)
t0 = int(t[0, 0, 0])
names = {0: ["s_full", "(unused)", "max", "exp_done", "pv_done", "p_full->"],
         1: ["k_full", "s_free", "S_issued", "v_full", "p_full", "PV_issued"], 2: ["k_empty", "v_empty"]}
for j in range(20, 30):
    for role in (0, 1, 2):
        ev = [(names[role][e], int(t[role, j, e]) - t0) for e in range(len(names[role]))]
        print(f"tile {j} role {role}: " + "  ".join(f"{n}={v}" for n, v in ev))
sm = t[0, 1:60, 0] - t[0, 0:59, 0]
print("softmax tile period (clk): mean %.0f min %d max %d" % (sm[10:].float().mean(), sm[10:].min(), sm[10:].max()))
for a, b_ in ((0, 2), (2, 3), (3, 4), (4, 5)):
    d = (t[0, 10:60, b_] - t[0, 10:60, a]).float().mean()
    print(f"  softmax {names[0][a]} -> {names[0][b_]}: {d:.0f}")
for e in range(5):
    d = (t[1, 10:60, e + 1] - t[1, 10:60, e]).float().mean()
    print(f"  mma {names[1][e]} -> {names[1][e+1]}: {d:.0f}")

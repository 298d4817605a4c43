#!/usr/bin/env python
"""Pipeline trace of the four-tile attention kernel (SR_ATTN_TRACE=1, CTA (0,0,0)) on the 16-anchor global shape."""
import ctypes as C
import os
import sys

os.environ["SR_ATTN_TRACE"] = "1"
os.environ["SR_ATTN_KERNEL"] = "4"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sailrecon_b200 import _lib, ops

dev = "cuda"
H, D, Cc = 16, 64, 1024
L = 16 * 1374
g = torch.Generator(device=dev).manual_seed(0)
qkv = torch.randn(L, 3 * Cc, device=dev, generator=g).to(torch.bfloat16)
out = torch.empty(L, Cc, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    ops.attention(qkv[:, :Cc], qkv[:, Cc:2 * Cc], qkv[:, 2 * Cc:], out, H, D, 1, L, L)
torch.cuda.synchronize()
buf = (C.c_longlong * (3 * 64 * 8))()
_lib.check(_lib.load().sr_debug_trace_read(buf, 3 * 64 * 8))
t = torch.tensor(list(buf)).view(3, 64, 8)
t0 = int(t[0, 0, 0])
for j in range(30, 34):
    sm = "  ".join(f"{n}={int(t[0, j, e]) - t0}" for e, n in ((0, "s_full"), (2, "max"), (3, "exp_done"), (5, "p_full->")))
    s3 = "  ".join(f"{n}={int(t[2, j, e]) - t0}" for e, n in ((0, "s_full"), (5, "p_full->")))
    mm = "  ".join(f"{n}={int(t[1, j, e]) - t0}" for e, n in ((0, "top"), (1, "p0"), (2, "p1"), (5, "end")))
    print(f"tile {j}: q0 {sm} | q3 {s3}\n          issuer {mm}")
per = (t[0, 11:60, 0] - t[0, 10:59, 0]).float()
print("kv64 step period (clk): mean %.0f min %d max %d" % (per.mean(), per.min(), per.max()))
for a, b_, n in ((0, 2, "s_full -> max"), (2, 3, "max -> exp_done"), (3, 5, "exp_done -> p_full arrive")):
    print("  softmax q0 %s: %.0f" % (n, (t[0, 10:60, b_] - t[0, 10:60, a]).float().mean()))
print("  softmax q0 p_full arrive -> next s_full: %.0f" % (t[0, 11:60, 0] - t[0, 10:59, 5]).float().mean())
for e, e0, n in ((1, 0, "kv_full seen -> p0 seen"), (2, 1, "p0 -> p1 seen"), (5, 2, "p1 seen -> end")):
    print("  issuer(0-1) %s: %.0f" % (n, (t[1, 10:60, e] - t[1, 10:60, e0]).float().mean()))
print("  issuer(0-1) end -> next kv_full seen: %.0f" % (t[1, 11:60, 0] - t[1, 10:59, 5]).float().mean())

#!/usr/bin/env python
"""Per-shape timing of sr_gemm_bf16 on the GEMMs of one aggregator block (M = 32 x 1374 and 16 x 1374 token rows).
CUDA events, 3 warm-up + 10 timed launches; A rotates over 4 buffers (> L2 together).  One JSON line per shape."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sailrecon_b200 import ops
from sailrecon_b200._lib import SR_EPI_BIAS, SR_EPI_BIAS_GELU, SR_EPI_LS_RESIDUAL, SR_EPI_QKV

dev = "cuda"
C = 1024


def bench(name, M, N, K, kind, iters=10):
    g = torch.Generator(device=dev).manual_seed(0)
    As = [torch.randn(M, K, device=dev, generator=g).to(torch.bfloat16) for _ in range(4)]
    W = (torch.randn(N, K, device=dev, generator=g) * 0.03).to(torch.bfloat16)
    bias = torch.randn(N, device=dev, generator=g)
    kw = {}
    if kind == SR_EPI_LS_RESIDUAL:
        out = torch.randn(M, N, device=dev, generator=g)
        kw = dict(gamma=torch.rand(N, device=dev, generator=g), resid=out)
    else:
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    if kind == SR_EPI_QKV:
        cos, sin = ops.rope_tables(40, 16, 100.0, dev)
        pos = torch.randint(0, 38, (M, 2), device=dev, generator=g, dtype=torch.int32)
        ln = [torch.rand(64, device=dev, generator=g) for _ in range(4)]
        kw = dict(qkv=ops.QkvEpilogue(C, 0, 64, (ln[0], ln[1]), (ln[2], ln[3]), 1e-5, cos, sin, pos))
    for i in range(3):
        ops.gemm(As[i % 4], W, bias, out, kind=kind, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        ops.gemm(As[i % 4], W, bias, out, kind=kind, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps({"gemm": name, "M": M, "N": N, "K": K, "ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 1)}), flush=True)


if __name__ == "__main__":
    for M in (32 * 1374, 16 * 1374):
        bench("qkv+qknorm+rope", M, 3 * C, C, SR_EPI_QKV)
        bench("qkv bias (dino)", M, 3 * C, C, SR_EPI_BIAS)
        bench("proj ls_residual", M, C, C, SR_EPI_LS_RESIDUAL)
        bench("fc1 gelu", M, 4 * C, C, SR_EPI_BIAS_GELU)
        bench("fc2 ls_residual", M, C, 4 * C, SR_EPI_LS_RESIDUAL)

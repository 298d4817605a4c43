#!/usr/bin/env python
"""Per-op answer to "is the hand-written kernel faster than the library kernel on this B200?" (VERDICT r02, item 12).

The three attention shapes and the four GEMM shapes of one agg32 aggregator layer, each timed two ways on the same box:
  ours     sr_attn_fwd_bf16 (attn_tc4, structured prefix instead of a mask) / sr_gemm_bf16 (gemm_tc, SR_EPI_BIAS)
  library  F.scaled_dot_product_attention (torch picks flash / cuDNN; NO mask — the best case for the library: the
           reference really passes a dense (S*P)^2 boolean mask, which rules the flash backend out) /
           F.linear in bf16 (cuBLASLt with its bias epilogue)
The two alternate inside every repetition (fair under clock / power drift); the median of the repetitions is reported.
The reloc shape gives the library each segment's [scene prefix ; own frame] keys already concatenated (the copy is not timed).
Library calls are measurement only: nothing under sailrecon_b200/ calls SDPA or cuBLAS.  One JSON line per shape."""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from sailrecon_b200 import ops
from sailrecon_b200._lib import SR_EPI_BIAS

H, D, C, P = 16, 64, 1024, 1374
dev = "cuda"
REPS, ITERS = 7, 6


def _time(fn, iters=ITERS):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def _ab(ours, lib):
    for i in range(3):
        ours(i)
        lib(i)
    torch.cuda.synchronize()
    to, tl = [], []
    for _ in range(REPS):
        to.append(_time(ours))
        tl.append(_time(lib))
    return statistics.median(to), statistics.median(tl)


def attention_shape(name, batch, L, prefix_len):
    T = batch * L
    g = torch.Generator(device=dev).manual_seed(0)
    nbuf = 4 if T * 3 * C * 2 < 600e6 else 2
    qkvs = [torch.randn(T, 3 * C, device=dev, generator=g).to(torch.bfloat16) for _ in range(nbuf)]
    pre = torch.randn(prefix_len, 2 * C, device=dev, generator=g).to(torch.bfloat16) if prefix_len else None
    out = torch.empty(T, C, device=dev, dtype=torch.bfloat16)

    def ours(i):
        qkv = qkvs[i % nbuf]
        ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, H, D, batch, L, L, L, L,
                      prefix_k=None if pre is None else pre[:, :C], prefix_v=None if pre is None else pre[:, C:],
                      prefix_len=prefix_len)

    # library operands: [batch, H, L, D] views of the same rows (strided, as attention.py:103-109 hands them over after its
    # reshape / permute); reloc: keys / values of every segment = [prefix ; own], concatenated beforehand
    lib_in = []
    for qkv in qkvs:
        v5 = qkv.view(batch, L, 3, H, D)
        q, k, v = (v5[:, :, j].permute(0, 2, 1, 3) for j in range(3))
        if pre is not None:
            pk = pre[:, :C].view(1, prefix_len, H, D).permute(0, 2, 1, 3).expand(batch, -1, -1, -1)
            pv = pre[:, C:].view(1, prefix_len, H, D).permute(0, 2, 1, 3).expand(batch, -1, -1, -1)
            k, v = torch.cat([pk, k], 2), torch.cat([pv, v], 2)
        lib_in.append((q, k, v))
    lib_out = [None]

    def lib(i):
        q, k, v = lib_in[i % nbuf]
        lib_out[0] = F.scaled_dot_product_attention(q, k, v)

    t_o, t_l = _ab(ours, lib)
    ours(0)
    lib(0)
    torch.cuda.synchronize()
    ref = lib_out[0].permute(0, 2, 1, 3).reshape(T, C).float()
    err = float((out.float() - ref).norm() / ref.norm())
    flops = 4.0 * batch * L * (prefix_len + L) * C
    print(json.dumps({"op": "attention", "shape": name, "batch": batch, "L": L, "prefix": prefix_len,
                      "ours_ms": round(t_o, 4), "library_ms": round(t_l, 4), "ours_tflops": round(flops / t_o / 1e9, 1),
                      "library_tflops": round(flops / t_l / 1e9, 1), "speedup_vs_library": round(t_l / t_o, 3),
                      "rel_l2_ours_vs_library": err, "library": "F.scaled_dot_product_attention, no mask"}), flush=True)


def gemm_shape(name, M, N, K):
    g = torch.Generator(device=dev).manual_seed(0)
    As = [torch.randn(M, K, device=dev, generator=g).to(torch.bfloat16) for _ in range(4)]
    W = (torch.randn(N, K, device=dev, generator=g) * 0.03).to(torch.bfloat16)
    bias = torch.randn(N, device=dev, generator=g)
    bias_bf = bias.to(torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    lib_out = [None]

    def ours(i):
        ops.gemm(As[i % 4], W, bias, out, kind=SR_EPI_BIAS)

    def lib(i):
        lib_out[0] = F.linear(As[i % 4], W, bias_bf)

    t_o, t_l = _ab(ours, lib)
    ours(0)
    lib(0)
    torch.cuda.synchronize()
    err = float((out.float() - lib_out[0].float()).norm() / lib_out[0].float().norm())
    flops = 2.0 * M * N * K
    print(json.dumps({"op": "gemm", "shape": name, "M": M, "N": N, "K": K, "ours_ms": round(t_o, 4), "library_ms": round(t_l, 4),
                      "ours_tflops": round(flops / t_o / 1e9, 1), "library_tflops": round(flops / t_l / 1e9, 1),
                      "speedup_vs_library": round(t_l / t_o, 3), "rel_l2_ours_vs_library": err,
                      "library": "F.linear bf16 (cuBLASLt, bias epilogue)"}), flush=True)


if __name__ == "__main__":
    S, Na = 32, 16
    attention_shape("frame  (32 segments x 1374)", S, P, 0)
    attention_shape("global (1 segment x 21984)", 1, Na * P, 0)
    attention_shape("reloc  (16 segments x 1374 + 4880 scene prefix)", S - Na, P, Na * 305)
    T = S * P
    gemm_shape("qkv  [T,1024]x[1024,3072]", T, 3 * C, C)
    gemm_shape("proj [T,1024]x[1024,1024]", T, C, C)
    gemm_shape("fc1  [T,1024]x[1024,4096]", T, 4 * C, C)
    gemm_shape("fc2  [T,4096]x[4096,1024]", T, C, 4 * C)

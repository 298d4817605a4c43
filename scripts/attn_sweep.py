#!/usr/bin/env python
"""Per-shape timing of sr_attn_fwd_bf16 on the three aggregator attention shapes of agg32 (+ optional extra shapes).
CUDA events on the launching stream, 3 warm-up + 10 timed launches per shape; inputs (>= 135 MB for global) rotate over
4 buffers so that consecutive launches do not hit in L2.  Prints one JSON line per shape."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sailrecon_b200 import ops

H, D, C = 16, 64, 1024
dev = "cuda"


def bench(name, batch, L, prefix_len=0, iters=10, check=False, variant=0):
    T = batch * L
    g = torch.Generator(device=dev).manual_seed(0)
    nbuf = 4 if T * 3 * C * 2 < 600e6 else 1
    qkvs = [(torch.randn(T, 3 * C, device=dev, generator=g) * 1.0).to(torch.bfloat16) for _ in range(nbuf)]
    pre = None
    if prefix_len:
        pre = (torch.randn(prefix_len, 2 * C, device=dev, generator=g)).to(torch.bfloat16)
    out = torch.empty(T, C, device=dev, dtype=torch.bfloat16)

    def run(i):
        qkv = qkvs[i % nbuf]
        ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, H, D, batch, L, L, L, L,
                      prefix_k=None if pre is None else pre[:, :C], prefix_v=None if pre is None else pre[:, C:],
                      prefix_len=prefix_len, variant=variant)
    for i in range(3):
        run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flops = 4.0 * batch * L * (prefix_len + L) * C
    rec = {"shape": name, "variant": variant, "batch": batch, "L": L, "prefix": prefix_len, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}
    if check:
        # fp32 reference on a few heads / segments (bf16 inputs, fp32 math)
        qkv = qkvs[(iters - 1) % nbuf].float()
        errs = []
        for b in (0, batch - 1):
            for h in (0, H - 1):
                q = qkv[b * L:(b + 1) * L, h * D:(h + 1) * D]
                k = qkv[b * L:(b + 1) * L, C + h * D:C + (h + 1) * D]
                v = qkv[b * L:(b + 1) * L, 2 * C + h * D:2 * C + (h + 1) * D]
                if pre is not None:
                    k = torch.cat([pre[:, h * D:(h + 1) * D].float(), k])
                    v = torch.cat([pre[:, C + h * D:C + (h + 1) * D].float(), v])
                rows = slice(0, min(L, 2048))
                ref = torch.softmax(q[rows] @ k.t() * D ** -0.5, -1) @ v
                got = out[b * L:(b + 1) * L, h * D:(h + 1) * D].float()[rows]
                errs.append(float((got - ref).norm() / ref.norm()))
        rec["rel_l2_max"] = max(errs)
    print(json.dumps(rec), flush=True)


def interleaved(shapes, variants, reps=6, iters=8):
    """A/B that is fair under clock / power drift: the variants alternate inside every repetition (a sweep that times one
    variant after the other favours whichever runs first, at boost clocks); reports the median of the per-repetition times."""
    import statistics
    for name, batch, L, prefix_len in shapes:
        T = batch * L
        g = torch.Generator(device=dev).manual_seed(0)
        qkvs = [(torch.randn(T, 3 * C, device=dev, generator=g)).to(torch.bfloat16) for _ in range(4 if T * 3 * C * 2 < 600e6 else 1)]
        pre = torch.randn(prefix_len, 2 * C, device=dev, generator=g).to(torch.bfloat16) if prefix_len else None
        out = torch.empty(T, C, device=dev, dtype=torch.bfloat16)

        def run(i, variant):
            qkv = qkvs[i % len(qkvs)]
            ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, H, D, batch, L, L, L, L,
                          prefix_k=None if pre is None else pre[:, :C], prefix_v=None if pre is None else pre[:, C:],
                          prefix_len=prefix_len, variant=variant)
        for i in range(30):                                  # reach the steady power / clock state first
            run(i, variants[i % len(variants)])
        torch.cuda.synchronize()
        times = {v: [] for v in variants}
        for rep in range(reps):
            for v in (variants if rep % 2 == 0 else variants[::-1]):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(iters):
                    run(i, v)
                e1.record()
                torch.cuda.synchronize()
                times[v].append(e0.elapsed_time(e1) / iters)
        flops = 4.0 * batch * L * (prefix_len + L) * C
        print(json.dumps({"shape": name, "interleaved": True,
                          "tflops_median": {str(v): round(flops / statistics.median(t) / 1e9, 1) for v, t in times.items()}}), flush=True)


if __name__ == "__main__":
    check = "--check" in sys.argv
    if "--interleaved" in sys.argv:
        interleaved([("global16", 1, 16 * 1374, 0), ("frame32", 32, 1374, 0), ("reloc16", 16, 1374, 16 * 305), ("global12", 1, 12 * 1374, 0)],
                    [4, 5, 6, 3])
        sys.exit(0)
    for variant in (0, 5, 6, 4, 3):                      # 0 = mixed four- / three-tile decomposition (default), 4 / 3 = uniform
        bench("global16", 1, 16 * 1374, check=check, variant=variant)
        bench("frame32", 32, 1374, check=check, variant=variant)
        bench("reloc16", 16, 1374, prefix_len=16 * 305, check=check, variant=variant)
        bench("global12", 1, 12 * 1374, variant=variant)
        bench("reloc1x100", 1, 1374, prefix_len=100 * 305, variant=variant)
    if "--big" in sys.argv:
        bench("global64", 1, 64 * 1374, iters=3)

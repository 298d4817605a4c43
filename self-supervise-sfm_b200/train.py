"""Training step of one transformer Block on the C-ABI kernels: forward that keeps what the backward needs, and the
backward itself (autograd of sailrecon/layers/block.py:86-112 = attention.py:70-122 + mlp.py:33-40 + layer_scale.py:22,
as train/train_imc.py:358-380 differentiates it under bf16 autocast).

The reference wraps every aggregator block in torch.utils.checkpoint (models/aggregator.py:654-657,726-729,758-761):
only the block input survives the forward pass, the block is re-run inside backward.  The same schedule here:

    forward pass      engine.run_block(...)                       (inference kernels, nothing kept but x)
    backward pass     saved = block_forward_saving(blk, x, ...)   (re-run with un-fused epilogues that keep u, z, raw qkv, lse)
                      dx = block_backward(blk, saved, dy, grads)

Every contraction runs on the tcgen05 kernels: dX = dY W and dW = dY^T X read dY, X and W as they lie in memory
(sr_gemm_bf16_t: MN-major operand descriptors, no transposed copies), attention on sr_attn_bwd_bf16; the rest are the row
kernels of csrc/train_rows.cu.
Parameter gradients are accumulated in fp32 into `grads[param]` (the role of .grad under DDP, train_imc.py:472).
There is no torch fallback.  Blocks whose attention call reads an external K/V prefix (the reloc blocks: query frames
against the scene tokens, aggregator.py:672-741) return the prefix gradient in `saved.dprefix`; scene_forward_saving /
scene_backward carry it through the scene-token path (row gather + norm1 + K|V projection of the reloc block,
aggregator.py:593-626) back into the anchor stream.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from . import engine, ops
from ._lib import SR_EPI_BIAS

BF = torch.bfloat16


@dataclass
class BlockSaved:
    x: torch.Tensor                    # block input, fp32 [T, C]
    h1: torch.Tensor                   # LN1(x) bf16
    raw: torch.Tensor                  # qkv Linear output before q/k norm and RoPE, bf16 [T, 3C]
    qkv: torch.Tensor                  # after q/k norm and RoPE (what attention consumed)
    o: torch.Tensor                    # attention output bf16 [T, C]
    lse: List[torch.Tensor]            # one per attention call
    z1: torch.Tensor                   # proj output (rounded) bf16
    x1: torch.Tensor                   # x + g1 z1, fp32
    h2: torch.Tensor                   # LN2(x1) bf16
    u: torch.Tensor                    # fc1 output before GELU, bf16 [T, 4C]
    g: torch.Tensor                    # gelu(u) bf16
    z2: torch.Tensor                   # fc2 output (rounded) bf16
    calls: Sequence[engine.AttnCall] = field(default_factory=list)
    epi: Optional[ops.QkvEpilogue] = None
    dprefix: Optional[torch.Tensor] = None    # set by block_backward: [prefix rows, 3C] bf16, k / v thirds hold d prefix K / V


def _qkv_epilogue(attn: nn.Module, rows_pos: Optional[engine.RowPositions], device, T: int) -> Optional[ops.QkvEpilogue]:
    C = attn.qkv.weight.shape[1]
    d = C // attn.num_heads
    qn = kn = None
    if isinstance(attn.q_norm, nn.LayerNorm):
        qn = (attn.q_norm.weight.detach(), attn.q_norm.bias.detach())
        kn = (attn.k_norm.weight.detach(), attn.k_norm.bias.detach())
    cos = sin = pos32 = None
    if attn.rope is not None and rows_pos is not None:
        cos, sin = attn.rope.tables(device, rows_pos.max_pos, d)
        pos32 = rows_pos.pos32
        if pos32.shape[0] != T:
            raise ValueError("sailrecon_b200: one (y, x) position per token row is required")
    if qn is None and cos is None:
        return None
    if d != 64:
        raise NotImplementedError("sailrecon_b200: q/k LayerNorm and RoPE are implemented for head_dim 64")
    return ops.QkvEpilogue(C, 0, d, qn, kn, attn.q_norm.eps if qn is not None else 1e-5, cos, sin, pos32)


def _check_calls(calls: Sequence[engine.AttnCall]) -> None:
    for c in calls:
        if c.prefix is not None and c.prefix.length > 0:
            if c.prefix.k is None:
                raise NotImplementedError("sailrecon_b200: training attention with an in-buffer prefix is not supported")
            if len(calls) != 1 or c.prefix.row0 != 0:
                raise NotImplementedError("sailrecon_b200: a training block takes at most one call with an external prefix")


def block_forward_saving(blk: nn.Module, x: torch.Tensor, calls: Sequence[engine.AttnCall],
                         rows_pos: Optional[engine.RowPositions]) -> BlockSaved:
    """Block.forward on rows x [T, C] (fp32, not modified) keeping every intermediate the backward reads.  Same arithmetic
    as engine.run_block in bf16 mode: each fused epilogue is split at the point where its input has to be kept."""
    engine.require_cuda_f32(x, "x")
    _check_calls(calls)
    T, C = x.shape
    dev = x.device
    attn, mlp = blk.attn, blk.mlp
    H = attn.num_heads
    d = C // H
    w1, b1, eps1 = engine.ln_params(blk.norm1)
    h1 = ops.layernorm(x, w1, b1, eps1, torch.empty(T, C, dtype=BF, device=dev))
    raw = torch.empty(T, 3 * C, dtype=BF, device=dev)
    ops.gemm(h1, engine.weight_bf16(attn.qkv), engine.bias_for(attn.qkv, "bf16"), raw, kind=SR_EPI_BIAS)
    epi = _qkv_epilogue(attn, rows_pos, dev, T)
    qkv = raw if epi is None else ops.qknorm_rope_fwd(raw, torch.empty_like(raw), epi)
    o = torch.empty(T, C, dtype=BF, device=dev)
    lses = []
    q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
    for c in calls:
        stride = c.stride or c.seg_len
        r1 = c.q_row0 + (c.batch - 1) * stride + c.seg_len
        lse = ops.new_lse(H, c.batch, c.seg_len, dev)
        pk = pv = None
        plen = 0
        if c.prefix is not None and c.prefix.length > 0:
            pk, pv, plen = c.prefix.k[:c.prefix.length], c.prefix.v[:c.prefix.length], c.prefix.length
        ops.attention(q[c.q_row0:r1], k[c.q_row0:r1], v[c.q_row0:r1], o[c.q_row0:r1], H, d, c.batch, c.seg_len, c.seg_len, stride,
                      stride, prefix_k=pk, prefix_v=pv, prefix_len=plen, scale=attn.scale, lse=lse)
        lses.append(lse)
    z1 = torch.empty(T, C, dtype=BF, device=dev)
    ops.gemm(o, engine.weight_bf16(attn.proj), engine.bias_for(attn.proj, "bf16"), z1, kind=SR_EPI_BIAS)
    x1 = ops.ls_residual(x, z1, engine.gamma_of(blk.ls1), torch.empty_like(x))
    w2, b2, eps2 = engine.ln_params(blk.norm2)
    h2 = ops.layernorm(x1, w2, b2, eps2, torch.empty(T, C, dtype=BF, device=dev))
    hid = mlp.fc1.weight.shape[0]
    u = torch.empty(T, hid, dtype=BF, device=dev)
    ops.gemm(h2, engine.weight_bf16(mlp.fc1), engine.bias_for(mlp.fc1, "bf16"), u, kind=SR_EPI_BIAS)
    g = ops.gelu_fwd(u, torch.empty_like(u))
    z2 = torch.empty(T, C, dtype=BF, device=dev)
    ops.gemm(g, engine.weight_bf16(mlp.fc2), engine.bias_for(mlp.fc2, "bf16"), z2, kind=SR_EPI_BIAS)
    return BlockSaved(x, h1, raw, qkv, o, lses, z1, x1, h2, u, g, z2, calls, epi)


def block_output(blk: nn.Module, s: BlockSaved, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x1 + g2 z2: the block output of a saving forward, into `out` (contiguous rows) or a fresh tensor."""
    return ops.ls_residual(s.x1, s.z2, engine.gamma_of(blk.ls2), torch.empty_like(s.x1) if out is None else out)


def _acc(grads: Dict[nn.Parameter, torch.Tensor], p: Optional[nn.Parameter]) -> Optional[torch.Tensor]:
    if p is None:
        return None
    if p not in grads:
        grads[p] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
    return grads[p]


# One-pass LayerScale / GELU backward kernels that also produce the bias gradients (sr_ls_bwd_bf16, sr_gelu_bwd_colsum_bf16).
# False = the separate colsum / scale_cast / gelu_bwd launches they replace (kept as the A/B reference of the tests).
FUSED_ROW_BACKWARD = True


def linear_backward(lin: nn.Module, a_in: torch.Tensor, dz: torch.Tensor, grads: Dict[nn.Parameter, torch.Tensor],
                    need_input_grad: bool = True, use_transposes: bool = False, bias_done: bool = False) -> Optional[torch.Tensor]:
    """Gradients of z = a_in W^T + b (nn.Linear): dW += dz^T a_in, db += colsum(dz) (fp32 accumulators in `grads`), returns
    d a_in = dz W (bf16).  a_in [T, K] bf16, dz [T, N] bf16 — both GEMMs on the tcgen05 kernel, reading dz, a_in and W AS THEY
    LIE in memory (MN-major operand descriptors, sr_gemm_bf16_t): no transposed copies, and dW accumulates in fp32 straight
    into the gradient bucket.  use_transposes=True is the round-1 path (explicit bf16 transposes + the K-major kernel), kept
    as the A/B reference of tests/test_train_gpu.py.  bias_done: the kernel that produced dz already added its column sums to
    the bias gradient (ops.ls_bwd / ops.gelu_bwd_colsum)."""
    T, K = a_in.shape
    N = dz.shape[1]
    dev = a_in.device
    gb = _acc(grads, lin.bias) if (lin.bias is not None and not bias_done) else None
    gw = _acc(grads, lin.weight)
    if N % 128 or K % 128:
        raise NotImplementedError("sailrecon_b200: Linear gradients need in / out features that are multiples of 128")
    fresh = getattr(grads, "fresh", None)
    first = fresh is not None and lin.weight in fresh              # first contribution since zero_grad: write, do not add
    if first:
        fresh.discard(lin.weight)
    if use_transposes:
        dzT = ops.transpose_bf16(dz, colsum=gb)                    # [N, T_pad]; db += colsum(dz) in the same pass
        aT = ops.transpose_bf16(a_in)                              # [K, T_pad]
        if first:
            ops.gemm(dzT, aT, None, gw.view(N, K), kind=SR_EPI_BIAS)   # [N, T_pad] x [K, T_pad]^T
        else:
            tmp = torch.empty(N, K, dtype=torch.float32, device=dev)
            ops.gemm(dzT, aT, None, tmp, kind=SR_EPI_BIAS)
            gw.view(N, K).add_(tmp)
    else:
        if gb is not None:
            ops.colsum(dz, None, gb, accumulate=True)
        ops.gemm_t(dz, a_in, gw.view(N, K), a_mn_major=True, accumulate=not first)    # dW[N, K] (+)= dz[T, N]^T a_in[T, K]
    if not need_input_grad:
        return None
    da = torch.empty(T, K, dtype=BF, device=dev)
    if use_transposes:
        wT = engine._cached(lin, "wT_bf16", lin.weight, lambda w: ops.transpose_bf16(engine.weight_bf16(lin)))   # [K, N]
        ops.gemm(dz, wT, None, da, kind=SR_EPI_BIAS)
    else:
        ops.gemm_t(dz, engine.weight_bf16(lin), da, a_mn_major=False)                # da[T, K] = dz[T, N] W[N, K]
    return da


def block_backward(blk: nn.Module, s: BlockSaved, dy: torch.Tensor, grads: Dict[nn.Parameter, torch.Tensor],
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dy = gradient w.r.t. the block output (fp32 [T, C], not modified) -> gradient w.r.t. the block input (fp32, written to
    `out` or a fresh tensor); parameter gradients accumulate into `grads`."""
    engine.require_cuda_f32(dy, "dy")
    T, C = dy.shape
    dev = dy.device
    attn, mlp = blk.attn, blk.mlp
    H = attn.num_heads
    d = C // H
    dx1 = torch.empty_like(dy)                                     # = dy (residual path of x2 = x1 + g2 z2) + MLP branch
    # ---- MLP branch
    g2 = engine.gamma_of(blk.ls2)
    dz2 = torch.empty(T, C, dtype=BF, device=dev)
    fused2 = FUSED_ROW_BACKWARD and g2 is not None and mlp.fc2.bias is not None
    if fused2:                                                     # d gamma2, dz2 and fc2's bias gradient in one pass
        ops.ls_bwd(dy, s.z2, g2, dz2, _acc(grads, blk.ls2.gamma), _acc(grads, mlp.fc2.bias))
    else:
        if g2 is not None:
            ops.colsum(dy, s.z2, _acc(grads, blk.ls2.gamma), accumulate=True)
        ops.scale_cast_bf16(dy, g2, dz2)
    dg = linear_backward(mlp.fc2, s.g, dz2, grads, bias_done=fused2)
    fused_u = FUSED_ROW_BACKWARD and mlp.fc1.bias is not None
    if fused_u:                                                    # in place over dg, fc1's bias gradient in the same pass
        du = ops.gelu_bwd_colsum(s.u, dg, dg, _acc(grads, mlp.fc1.bias))
    else:
        du = ops.gelu_bwd(s.u, dg, dg)
    dh2 = linear_backward(mlp.fc1, s.h2, du, grads, bias_done=fused_u)
    w2, _, eps2 = engine.ln_params(blk.norm2)
    ops.layernorm_bwd(s.x1, w2, dh2, dx1, _acc(grads, blk.norm2.weight), _acc(grads, blk.norm2.bias), eps2, dres=dy)
    # ---- attention branch
    dx = out if out is not None else torch.empty_like(dy)          # = dx1 (residual path of x1 = x + g1 z1) + attention branch
    g1 = engine.gamma_of(blk.ls1)
    fused1 = FUSED_ROW_BACKWARD and g1 is not None and attn.proj.bias is not None
    if fused1:
        dz1 = ops.ls_bwd(dx1, s.z1, g1, dz2, _acc(grads, blk.ls1.gamma), _acc(grads, attn.proj.bias))   # dz2 is free again
    else:
        if g1 is not None:
            ops.colsum(dx1, s.z1, _acc(grads, blk.ls1.gamma), accumulate=True)
        dz1 = ops.scale_cast_bf16(dx1, g1, dz2)
    do = linear_backward(attn.proj, s.o, dz1, grads, bias_done=fused1)
    dqkv = torch.empty(T, 3 * C, dtype=BF, device=dev)
    q, k, v = s.qkv[:, :C], s.qkv[:, C:2 * C], s.qkv[:, 2 * C:]
    dq, dk, dv = dqkv[:, :C], dqkv[:, C:2 * C], dqkv[:, 2 * C:]
    covered = 0
    for c, lse in zip(s.calls, s.lse):
        stride = c.stride or c.seg_len
        r1 = c.q_row0 + (c.batch - 1) * stride + c.seg_len
        if stride != c.seg_len:
            raise NotImplementedError("sailrecon_b200: training attention calls must cover their rows densely")
        sl = slice(c.q_row0, r1)
        kw = {}
        if c.prefix is not None and c.prefix.length > 0:
            n = c.prefix.length
            s.dprefix = torch.zeros(n, 3 * C, dtype=BF, device=dev)                # q third stays zero (scene rows are K/V only)
            kw = dict(prefix_k=c.prefix.k[:n], prefix_v=c.prefix.v[:n], prefix_len=n, dprefix_k=s.dprefix[:, C:2 * C],
                      dprefix_v=s.dprefix[:, 2 * C:])
        ops.attention_bwd(q[sl], k[sl], v[sl], s.o[sl], do[sl], lse, dq[sl], dk[sl], dv[sl], H, d, c.batch, c.seg_len, c.seg_len,
                          stride, stride, scale=attn.scale, **kw)
        covered += r1 - c.q_row0
    if covered != T:
        raise ValueError("sailrecon_b200: the attention calls of a training block must cover every row exactly once")
    if s.epi is not None:
        dqn = dkn = None
        if s.epi.q_norm is not None:
            dqn = (_acc(grads, attn.q_norm.weight), _acc(grads, attn.q_norm.bias))
            dkn = (_acc(grads, attn.k_norm.weight), _acc(grads, attn.k_norm.bias))
        ops.qknorm_rope_bwd(s.raw, dqkv, s.epi, dqn, dkn)
    dh1 = linear_backward(attn.qkv, s.h1, dqkv, grads)
    w1, _, eps1 = engine.ln_params(blk.norm1)
    ops.layernorm_bwd(s.x, w1, dh1, dx, _acc(grads, blk.norm1.weight), _acc(grads, blk.norm1.bias), eps1, dres=dx1)
    return dx


# ------------------------------------------------------------------------------------------------ scene-token path
@dataclass
class SceneSaved:
    idx: torch.Tensor                  # int32 [Ns] rows of the anchor stream
    a_s: torch.Tensor                  # norm1(X[idx]) bf16 [Ns, C]
    raw: torch.Tensor                  # qkv Linear on a_s, bf16 [Ns, 3C] (the q third is computed but never attended)
    post: torch.Tensor                 # after k norm + RoPE; k = post[:, C:2C], v = post[:, 2C:]
    epi: Optional[ops.QkvEpilogue]


def scene_forward_saving(rblk: nn.Module, X: torch.Tensor, idx: torch.Tensor, spos: engine.RowPositions) -> SceneSaved:
    """Scene tokens of one layer (aggregator.py:351-360,593-626 + block.py:87 + attention.py:73-82 of the reloc block):
    gather rows idx of the anchor stream X, norm1, qkv projection, k norm + RoPE.  Returns K/V in `post`."""
    Ns = idx.numel()
    C = X.shape[1]
    dev = X.device
    w1, b1, eps1 = engine.ln_params(rblk.norm1)
    a_s = ops.layernorm(X, w1, b1, eps1, torch.empty(Ns, C, dtype=BF, device=dev), idx=idx)
    raw = torch.empty(Ns, 3 * C, dtype=BF, device=dev)
    ops.gemm(a_s, engine.weight_bf16(rblk.attn.qkv), engine.bias_for(rblk.attn.qkv, "bf16"), raw, kind=SR_EPI_BIAS)
    epi = _qkv_epilogue(rblk.attn, spos, dev, Ns)
    post = raw if epi is None else ops.qknorm_rope_fwd(raw, torch.empty_like(raw), epi)
    return SceneSaved(idx, a_s, raw, post, epi)


def scene_backward(rblk: nn.Module, s: SceneSaved, dprefix: torch.Tensor, X: torch.Tensor, dX: torch.Tensor,
                   grads: Dict[nn.Parameter, torch.Tensor]) -> None:
    """dprefix [Ns, 3C] bf16 (k / v thirds = gradient of the scene K / V, q third zero) -> dX[idx] += ... (fp32 anchor-stream
    gradient) and the reloc block's norm1 / qkv / k_norm parameter gradients."""
    attn = rblk.attn
    Ns, C = s.a_s.shape
    if s.epi is not None:
        dqn = dkn = None
        if s.epi.q_norm is not None:
            dqn = (_acc(grads, attn.q_norm.weight), _acc(grads, attn.q_norm.bias))
            dkn = (_acc(grads, attn.k_norm.weight), _acc(grads, attn.k_norm.bias))
        ops.qknorm_rope_bwd(s.raw, dprefix, s.epi, dqn, dkn)
    dh = linear_backward(attn.qkv, s.a_s, dprefix, grads)
    xg = ops.gather_rows(X, s.idx, torch.empty(Ns, C, dtype=torch.float32, device=X.device))
    tmp = torch.zeros(Ns, C, dtype=torch.float32, device=X.device)
    w1, _, eps1 = engine.ln_params(rblk.norm1)
    ops.layernorm_bwd(xg, w1, dh, tmp, _acc(grads, rblk.norm1.weight), _acc(grads, rblk.norm1.bias), eps1)
    ops.scatter_add_rows(dX, s.idx, tmp)


# ------------------------------------------------------------------------------------------------ one aggregator layer
@dataclass
class LayerCtx:
    """Row structure of one scene: Na anchor frames then Nq query frames, P tokens each (aggregator.py:242-433)."""
    Na: int
    Nq: int
    P: int
    rows_pos: engine.RowPositions      # all (Na + Nq) * P rows

    def __post_init__(self):
        n = self.Na * self.P
        self.pos_a = engine.RowPositions(self.rows_pos.pos32[:n], self.rows_pos.max_pos)
        self.pos_q = engine.RowPositions(self.rows_pos.pos32[n:], self.rows_pos.max_pos)
        self.frame_calls = engine.full_segments(self.Na + self.Nq, self.P)
        self.global_calls = engine.full_segments(1, n)


def layer_forward(fblk: nn.Module, gblk: nn.Module, rblk: nn.Module, X: torch.Tensor, ctx: LayerCtx, idx: torch.Tensor,
                  spos: engine.RowPositions) -> torch.Tensor:
    """One aggregator layer in place on X [(Na + Nq) P, C] with the inference kernels (aggregator.py:343-392): frame block on
    every frame, scene K|V from rows idx of the anchors, reloc block on the queries, global block on the anchors.
    Returns the copy of X after the frame block (the checkpoint the backward re-runs the reloc / global blocks from)."""
    n = ctx.Na * ctx.P
    C = X.shape[1]
    engine.run_block(fblk, X, ctx.frame_calls, ctx.rows_pos, "bf16", tag="train.")
    X_mid = X.clone()
    Xa, Xq = X[:n], X[n:]
    if ctx.Nq > 0:
        Ns = idx.numel()
        w1, b1, eps1 = engine.ln_params(rblk.norm1)
        a_s = ops.layernorm(X, w1, b1, eps1, torch.empty(Ns, C, dtype=BF, device=X.device), idx=idx)
        skv = torch.empty(Ns, 2 * C, dtype=BF, device=X.device)
        engine.qkv_rows(rblk.attn, a_s, skv, "bf16", spos, col_offset=C)
        pre = engine.Prefix(skv[:, :C], skv[:, C:], 0, Ns)
        engine.run_block(rblk, Xq, [engine.AttnCall(0, ctx.Nq, ctx.P, ctx.P, pre)], ctx.pos_q, "bf16", tag="train.")
    engine.run_block(gblk, Xa, ctx.global_calls, ctx.pos_a, "bf16", tag="train.")
    return X_mid


def layer_backward(fblk: nn.Module, gblk: nn.Module, rblk: nn.Module, X_in: torch.Tensor, X_mid: torch.Tensor, dX: torch.Tensor,
                   ctx: LayerCtx, idx: torch.Tensor, spos: engine.RowPositions, grads: Dict[nn.Parameter, torch.Tensor],
                   d_mid: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dX = gradient w.r.t. the layer output -> gradient w.r.t. the layer input.  X_in / X_mid are the checkpoints (layer
    input, state after the frame block); d_mid is an extra gradient on the state after the frame block (the first half of
    the [frame | global] concat the heads read, aggregator.py:403-413)."""
    n = ctx.Na * ctx.P
    Xa_mid, Xq_mid = X_mid[:n], X_mid[n:]
    d_all = torch.empty_like(dX)
    sg = block_forward_saving(gblk, Xa_mid, ctx.global_calls, ctx.pos_a)
    dXa = block_backward(gblk, sg, dX[:n], grads, out=d_all[:n])
    del sg
    if ctx.Nq > 0:
        scene = scene_forward_saving(rblk, Xa_mid, idx, spos)
        C = X_in.shape[1]
        pre = engine.Prefix(scene.post[:, C:2 * C], scene.post[:, 2 * C:], 0, idx.numel())
        sr = block_forward_saving(rblk, Xq_mid, [engine.AttnCall(0, ctx.Nq, ctx.P, ctx.P, pre)], ctx.pos_q)
        block_backward(rblk, sr, dX[n:], grads, out=d_all[n:])
        scene_backward(rblk, scene, sr.dprefix, Xa_mid, dXa, grads)
        del sr, scene
    if d_mid is not None:
        d_all += d_mid
    sf = block_forward_saving(fblk, X_in, ctx.frame_calls, ctx.rows_pos)
    return block_backward(fblk, sf, d_all, grads)


# ------------------------------------------------------------------------------------------------ the 3 x depth block stack
class GradViews(dict):
    """parameter -> fp32 gradient view.  `fresh` = the 2-D (Linear weight) parameters that have not received a gradient since
    zero_grad: linear_backward lets the wgrad GEMM write those views directly instead of GEMM -> temporary -> add
    (the reloc block's qkv weight gets two contributions per step, scene tokens and query frames: the second one adds)."""
    fresh: set = frozenset()


class TrainState:
    """Optimizer-side state of the aggregator's block stacks (frame / global / global_reloc, aggregator.py:99-149), one flat
    fp32 bucket per layer index i holding the parameters of frame_blocks[i], global_blocks[i] and global_reloc_blocks[i]:

        param  the modules' parameters are re-pointed into it (p.data = view), so the Adam kernel updates them in place
        grad   what DDP's bucket is in the reference (train_imc.py:472): block_backward accumulates into views of it and
               ONE all-reduce per layer runs on it while the backward of the next (lower) layer computes
        m, v   Adam moments (train_imc.py:475: Adam(lr, betas=(0.9, 0.999), eps=1e-8))

    ~37.8 M parameters per layer, 4 x 151 MB; 14.5 GB for the 24 layers."""

    def __init__(self, agg: nn.Module, lr: float = 1e-5, betas=(0.9, 0.999), eps: float = 1e-8):
        self.lr, self.betas, self.eps, self.step = lr, betas, eps, 0
        self.param: List[torch.Tensor] = []
        self.grad: List[torch.Tensor] = []
        self.m: List[torch.Tensor] = []
        self.v: List[torch.Tensor] = []
        self.grads: "GradViews" = GradViews()
        self.params_of_layer: List[List[nn.Parameter]] = []
        for i in range(agg.depth):
            ps = [p for blk in (agg.frame_blocks[i], agg.global_blocks[i], agg.global_reloc_blocks[i]) for p in blk.parameters()]
            offs, n = [], 0
            for p in ps:
                offs.append(n)
                n += (p.numel() + 63) // 64 * 64                                       # 256-byte aligned views
            dev = ps[0].device
            flat = torch.zeros(n, dtype=torch.float32, device=dev)
            g = torch.zeros(n, dtype=torch.float32, device=dev)
            with torch.no_grad():
                for p, o in zip(ps, offs):
                    view = flat[o:o + p.numel()].view(p.shape)
                    view.copy_(p.detach())
                    p.data = view
                    self.grads[p] = g[o:o + p.numel()].view(p.shape)
            self.param.append(flat)
            self.grad.append(g)
            self.m.append(torch.zeros_like(flat))
            self.v.append(torch.zeros_like(flat))
            self.params_of_layer.append(ps)

    def zero_grad(self) -> None:
        for g in self.grad:
            g.zero_()
        self.grads.fresh = {p for ps in self.params_of_layer for p in ps if p.dim() == 2}

    def adam_step(self, grad_scale: float = 1.0) -> None:
        self.step += 1
        for i in range(len(self.param)):
            ops.adam_step(self.param[i], self.grad[i], self.m[i], self.v[i], self.lr, self.step, self.betas, self.eps, grad_scale)
            for p in self.params_of_layer[i]:
                torch.autograd.graph.increment_version(p)                             # engine's bf16 weight copies are stale now


def layer_forward_saving(fblk: nn.Module, gblk: nn.Module, rblk: nn.Module, X: torch.Tensor, ctx: LayerCtx, idx: torch.Tensor,
                         spos: engine.RowPositions):
    """One aggregator layer keeping every activation its backward reads (no re-run in the backward): X is not modified,
    returns (layer output, saved).  ~2.8 GB per layer at 24 frames of 518 px, 68 GB for the 24 layers — the schedule the
    180 GB of a B200 allow; layer_forward / layer_backward are the reference's checkpointed schedule."""
    n = ctx.Na * ctx.P
    C = X.shape[1]
    sf = block_forward_saving(fblk, X, ctx.frame_calls, ctx.rows_pos)
    X_mid = block_output(fblk, sf)
    Y = torch.empty_like(X_mid)
    sg = block_forward_saving(gblk, X_mid[:n], ctx.global_calls, ctx.pos_a)
    block_output(gblk, sg, out=Y[:n])
    scene = sr = None
    if ctx.Nq > 0:
        scene = scene_forward_saving(rblk, X_mid[:n], idx, spos)
        pre = engine.Prefix(scene.post[:, C:2 * C], scene.post[:, 2 * C:], 0, idx.numel())
        sr = block_forward_saving(rblk, X_mid[n:], [engine.AttnCall(0, ctx.Nq, ctx.P, ctx.P, pre)], ctx.pos_q)
        block_output(rblk, sr, out=Y[n:])
    return Y, (sf, sg, scene, sr, X_mid)


def layer_backward_saved(fblk: nn.Module, gblk: nn.Module, rblk: nn.Module, saved, dX: torch.Tensor, ctx: LayerCtx,
                         grads: Dict[nn.Parameter, torch.Tensor], d_mid: Optional[torch.Tensor] = None) -> torch.Tensor:
    """layer_backward on the activations layer_forward_saving kept."""
    sf, sg, scene, sr, X_mid = saved
    n = ctx.Na * ctx.P
    d_all = torch.empty_like(dX)
    dXa = block_backward(gblk, sg, dX[:n], grads, out=d_all[:n])
    if sr is not None:
        block_backward(rblk, sr, dX[n:], grads, out=d_all[n:])
        scene_backward(rblk, scene, sr.dprefix, X_mid[:n], dXa, grads)
    if d_mid is not None:
        d_all += d_mid
    return block_backward(fblk, sf, d_all, grads)


def stack_forward(agg: nn.Module, X: torch.Tensor, ctx: LayerCtx, draw_scene, keep_activations: bool = False,
                  on_layer=None) -> list:
    """The 24 aggregator layers (aggregator.py:339-423) on X.  draw_scene() -> (idx int32 [Na P'], RowPositions) is called
    once per layer, in order (the per-layer scene-token draw, :351-360).
    keep_activations False: the reference's schedule — inference kernels in place on X, per layer only what
    torch.utils.checkpoint keeps (the block inputs: layer input and the state after the frame block, :654-657,726-729,
    758-761); the backward re-runs each block.  True: every block runs its saving forward once and nothing is re-run.
    on_layer(i, X_mid, X_out) sees the two states the reference collects per layer (read them before the next layer runs)."""
    ck = []
    for i in range(agg.depth):
        idx, spos = draw_scene()
        blocks = (agg.frame_blocks[i], agg.global_blocks[i], agg.global_reloc_blocks[i])
        if keep_activations:
            X, saved = layer_forward_saving(*blocks, X, ctx, idx, spos)
            ck.append((saved,))
            X_mid = saved[4]
        else:
            X_in = X.clone()
            X_mid = layer_forward(*blocks, X, ctx, idx, spos)
            ck.append((X_in, X_mid, idx, spos))
        if on_layer is not None:
            on_layer(i, X_mid, X)            # state after the frame block / after the whole layer (aggregator.py:403-423)
    return ck


def stack_backward(agg: nn.Module, ck: list, d_out: Dict[int, tuple], ctx: LayerCtx, grads: Dict[nn.Parameter, torch.Tensor],
                   on_layer_done=None) -> torch.Tensor:
    """Backward through the 24 layers.  d_out[i] = (d_frame_half, d_global_half), each fp32 [T, C] or None: the loss gradient
    on the state after layer i's frame block and after the whole layer (the two halves of output_dict[i],
    aggregator.py:403-413; the last layer's anchor rows carry the camera-token gradient, :414-423).  Checkpoints /
    activations are released layer by layer; on_layer_done(i) runs when layer i's parameter gradients are complete."""
    dX = None
    for i in reversed(range(agg.depth)):
        d_mid, d_post = d_out.get(i, (None, None))
        ref = ck[i][0][4] if len(ck[i]) == 1 else ck[i][0]
        if dX is None:
            dX = torch.zeros_like(ref) if d_post is None else d_post.clone()
        elif d_post is not None:
            dX += d_post
        blocks = (agg.frame_blocks[i], agg.global_blocks[i], agg.global_reloc_blocks[i])
        if len(ck[i]) == 1:
            dX = layer_backward_saved(*blocks, ck[i][0], dX, ctx, grads, d_mid=d_mid)
        else:
            X_in, X_mid, idx, spos = ck[i]
            dX = layer_backward(*blocks, X_in, X_mid, dX, ctx, idx, spos, grads, d_mid=d_mid)
        ck[i] = None
        if on_layer_done is not None:
            on_layer_done(i)
    return dX


def activation_bytes(ctx: LayerCtx, depth: int, C: int, scene_rows: int = 0) -> int:
    """HBM needed to keep every activation the backward reads (layer_forward_saving) for `depth` layers.  Per token row of a
    block: bf16 h1, raw qkv, qkv, o, z1, h2, u, g, z2 = 19 C elements and fp32 x1 + block output = 2 C elements; every row
    goes through the frame block and through one of the global / reloc blocks; scene rows keep a_s, raw and post (7 C bf16).
    The fp32 log-sum-exp rows (H floats per row and block) are included with H = C / 64."""
    T = (ctx.Na + ctx.Nq) * ctx.P
    per_row = 19 * C * 2 + 2 * C * 4 + (C // 64) * 4
    return depth * (2 * T * per_row + scene_rows * 7 * C * 2)


def choose_keep_activations(ctx: LayerCtx, depth: int, C: int, scene_rows: int, free_bytes: int, headroom: float = 0.85) -> bool:
    """True when the kept activations (plus one layer of backward temporaries, ~ the size of one layer's activations) fit in
    `headroom` of the free HBM; otherwise the reference's checkpointed schedule is used."""
    need = activation_bytes(ctx, depth + 1, C, scene_rows)
    return need <= headroom * free_bytes


def training_step(agg: nn.Module, state: TrainState, X0: torch.Tensor, ctx: LayerCtx, draw_scene, d_out: Dict[int, tuple],
                  world: int = 1, fwd_bwd=None, optimizer_step=None, keep_activations: Optional[bool] = None,
                  skip_nonfinite: bool = False) -> torch.Tensor:
    """zero_grad + forward + backward + gradient all-reduce + Adam for the block stacks on one scene per rank
    (train_imc.py:361-381 with DDP, :472).  The all-reduce of layer i's bucket (sum; the 1 / world of DDP's mean is
    folded into the Adam kernel) is issued as soon as layer i's backward is enqueued and overlaps the layers below it.
    Returns the gradient w.r.t. X0 (what would flow on into the patch embed).
    keep_activations: True keeps every block's activations from the forward (68 GB at 24 frames — fits a B200, no re-run);
    False is the reference's torch.utils.checkpoint schedule (block inputs only, each block re-run inside the backward);
    None (default) keeps them when activation_bytes() fits the free HBM of X0's device (choose_keep_activations).
    fwd_bwd(on_layer_done) -> dX0 and optimizer_step(grad_scale) replace the CUDA kernels in the CPU (gloo) tests of this
    schedule; the product path leaves them None.
    skip_nonfinite: GradScaler.step semantics (train_imc.py:380-381) — after the all-reduce the gradient buckets are checked
    and the optimizer step is SKIPPED (returning with `state.skipped_steps` incremented) when any gradient is inf / nan, so
    one bad scene cannot corrupt the fp32 masters and the Adam moments; costs one read of the gradient buckets (host sync).
    Learning-rate schedules (the reference's CosineWarmupScheduler, train_imc.py:62-90,384) drive `state.lr` between steps.
    Note: Adam runs over whole layer buckets; with no query frames (Nq = 0) the reloc blocks' gradients are zero, not None,
    so their moments decay where torch.optim would skip them (no reference script trains without queries)."""
    import torch.distributed as dist
    state.zero_grad()
    works = []

    def reduce_layer(i):
        if world > 1:
            works.append(dist.all_reduce(state.grad[i], op=dist.ReduceOp.SUM, async_op=True))

    if fwd_bwd is None:
        if keep_activations is None:
            free, _ = torch.cuda.mem_get_info(X0.device)
            free += torch.cuda.memory_reserved(X0.device) - torch.cuda.memory_allocated(X0.device)   # cached blocks are reusable
            scene_rows = ctx.Na * min(ctx.P, 305) if ctx.Nq > 0 else 0        # upper bound at the reference's fix_rank 300
            keep_activations = choose_keep_activations(ctx, agg.depth, X0.shape[1], scene_rows, free)
        X = X0 if keep_activations else X0.clone()                 # the checkpointed forward works in place
        ck = stack_forward(agg, X, ctx, draw_scene, keep_activations)
        dX0 = stack_backward(agg, ck, d_out, ctx, state.grads, reduce_layer)
    else:
        dX0 = fwd_bwd(reduce_layer)
    for w in works:
        w.wait()
    if skip_nonfinite and not all(bool(torch.isfinite(g).all()) for g in state.grad):
        state.skipped_steps = getattr(state, "skipped_steps", 0) + 1
        return dX0
    (optimizer_step or state.adam_step)(grad_scale=1.0 / world)
    return dX0


# ------------------------------------------------------------------------------------------------ patch embed (DINOv2 ViT / conv)
@dataclass
class EmbedSaved:
    a: torch.Tensor                    # normalised im2col rows of the images, bf16 [S*hw, Kp] (operand of the projection's wgrad)
    blocks: list                       # BlockSaved per ViT block ([] for the "conv" patch embed)
    x_final: Optional[torch.Tensor]    # ViT residual rows before the final LayerNorm, fp32 [S*Pd, D]
    S: int
    Pd: int
    hw: int
    HW: tuple


# The im2col rows are the K operand of the projection GEMM (K % 64) AND the N operand of its weight gradient dW = dz^T a
# (sr_gemm_bf16_t: N % 128), so the training path pads them to 128 columns: 588 -> 640 for patch 14 (what ceil-to-64 gives
# too, i.e. unchanged), 192 -> 256 for patch 8, where ceil-to-64 used to fail with a shape error in the middle of backward.
_IM2COL_PAD = 128


def _im2col_width(K: int) -> int:
    return (K + _IM2COL_PAD - 1) // _IM2COL_PAD * _IM2COL_PAD


def embed_forward_saving(agg: nn.Module, imgs: torch.Tensor, X0: torch.Tensor, S: int, P: int, C: int):
    """aggregator.py:267-274 + vision_transformer.py:242-307 in the bf16 autocast arithmetic, keeping what the backward reads.
    Writes the patch-token rows of X0 [S*P, C] (rows psi.. of every frame) and returns EmbedSaved.  Same arithmetic as
    Aggregator._embed (inference kernels); the ViT blocks run block_forward_saving instead of engine.run_block."""
    from .layers.patch_embed import PatchEmbed
    from .models.aggregator import _RESNET_MEAN, _RESNET_STD
    pe_mod = agg.patch_embed
    dev = imgs.device
    _, _, H, W = imgs.shape
    psi = agg.patch_start_idx
    if isinstance(pe_mod, PatchEmbed):                                   # patch_embed="conv": tokens = proj(patches)
        ph = pe_mod.patch_size[0]
        hw = (H // ph) * (W // ph)
        K = 3 * ph * ph
        Kp = _im2col_width(K)
        a = torch.empty(S * hw, Kp, dtype=BF, device=dev)
        ops.patchify(imgs.contiguous(), a, ph, _RESNET_MEAN, _RESNET_STD)
        y = torch.empty(S * hw, C, dtype=BF, device=dev)
        ops.gemm(a, engine.weight_bf16(pe_mod.proj, pad_k=_IM2COL_PAD), engine.bias_for(pe_mod.proj, "bf16"), y, kind=SR_EPI_BIAS)
        X0.view(S, P, C)[:, psi:] = y.view(S, hw, C)
        return EmbedSaved(a, [], None, S, P, hw, (H, W))
    vit = pe_mod
    D, R = vit.embed_dim, vit.num_register_tokens
    ph = vit.patch_size
    hw = (H // ph) * (W // ph)
    Pd = 1 + R + hw
    if Pd != P or D != C:
        raise NotImplementedError("sailrecon_b200: the ViT patch embed must have the aggregator's width and register count")
    pe = vit.interpolated_pos_embed(H, W)
    K = 3 * ph * ph
    Kp = _im2col_width(K)
    a = torch.empty(S * hw, Kp, dtype=BF, device=dev)
    ops.patchify(imgs.contiguous(), a, ph, _RESNET_MEAN, _RESNET_STD)
    x = torch.empty(S * Pd, D, dtype=torch.float32, device=dev)
    proj = vit.patch_embed.proj
    from ._lib import SR_EPI_LS_RESIDUAL
    for s_ in range(S):                                                  # x[s, 1+R:] = pos_embed[1:] + proj(patches)
        ops.gemm(a[s_ * hw:(s_ + 1) * hw], engine.weight_bf16(proj, pad_k=_IM2COL_PAD), engine.bias_for(proj, "bf16"),
                 x[s_ * Pd + 1 + R:(s_ + 1) * Pd], kind=SR_EPI_LS_RESIDUAL, gamma=None, resid=pe[1:])
    xv = x.view(S, Pd, D)
    xv[:, 0] = vit.cls_token.detach()[0, 0] + pe[0]
    if R:
        xv[:, 1:1 + R] = vit.register_tokens.detach()[0]
    calls = engine.full_segments(S, Pd)
    saved = []
    for blk in vit.blocks:
        sb = block_forward_saving(blk, x, calls, None)
        x = block_output(blk, sb)
        saved.append(sb)
    w, b, eps = engine.ln_params(vit.norm)
    y = torch.empty(S * Pd, D, dtype=torch.float32, device=dev)
    ops.layernorm(x, w, b, eps, y)
    X0.view(S, P, C)[:, psi:] = y.view(S, Pd, D)[:, 1 + R:]
    return EmbedSaved(a, saved, x, S, Pd, hw, (H, W))


def _proj_wgrad(proj: nn.Module, a: torch.Tensor, dz: torch.Tensor, grads: Dict[nn.Parameter, torch.Tensor]) -> None:
    """Conv2d(3, D, 14, stride 14) as a Linear on im2col rows: dW += dz^T a (a zero-padded to Kp columns), db += colsum(dz)."""
    D = dz.shape[1]
    K = proj.weight[0].numel()
    if proj.bias is not None:
        ops.colsum(dz, None, _acc(grads, proj.bias), accumulate=True)
    if a.shape[1] % _IM2COL_PAD or a.shape[1] < K:
        raise ValueError(f"sailrecon_b200: im2col rows must be padded to a multiple of {_IM2COL_PAD} columns >= {K}, got {a.shape[1]}")
    tmp = torch.empty(D, a.shape[1], dtype=torch.float32, device=dz.device)
    ops.gemm_t(dz, a, tmp, a_mn_major=True)                              # [D, Kp] = dz[T, D]^T a[T, Kp], operands in place
    _acc(grads, proj.weight).view(D, K).add_(tmp[:, :K])


def embed_backward(agg: nn.Module, sv: EmbedSaved, dX0: torch.Tensor, grads: Dict[nn.Parameter, torch.Tensor]) -> None:
    """dX0 [S*P, C] = gradient of the token rows the stack starts from; its patch rows flow into the patch embed."""
    from .layers.patch_embed import PatchEmbed
    pe_mod = agg.patch_embed
    S, P, hw = sv.S, sv.Pd, sv.hw
    C = dX0.shape[1]
    psi = agg.patch_start_idx
    dev = dX0.device
    if isinstance(pe_mod, PatchEmbed):
        dz = ops.cast_bf16(dX0.view(S, P, C)[:, psi:].reshape(S * hw, C))
        _proj_wgrad(pe_mod.proj, sv.a, dz, grads)
        return
    vit = pe_mod
    D, R = vit.embed_dim, vit.num_register_tokens
    dY = dX0.clone().view(S, P, D)
    dY[:, :1 + R] = 0                                                    # only x_norm_patchtokens is consumed (aggregator.py:273-274)
    dy16 = ops.cast_bf16(dY.view(S * P, D))
    w, _, eps = engine.ln_params(vit.norm)
    dX = torch.zeros(S * P, D, dtype=torch.float32, device=dev)
    ops.layernorm_bwd(sv.x_final, w, dy16, dX, _acc(grads, vit.norm.weight), _acc(grads, vit.norm.bias), eps)
    for blk, sb in zip(reversed(list(vit.blocks)), reversed(sv.blocks)):
        dX = block_backward(blk, sb, dX, grads)
    # rows: patch tokens -> projection + pos_embed[1:], row 0 -> cls_token + pos_embed[0], rows 1..R -> register tokens;
    # the sums over the S frames are one column-sum over the [S, Pd * D] view
    gsum = torch.zeros(P * D, dtype=torch.float32, device=dev)
    ops.colsum(dX.view(S, P * D), None, gsum)
    gsum = gsum.view(P, D)
    _acc(grads, vit.cls_token).view(D).add_(gsum[0])
    if R:
        _acc(grads, vit.register_tokens).view(R, D).add_(gsum[1:1 + R])
    H, W = sv.HW
    N = vit.pos_embed.shape[1] - 1
    gpe = _acc(grads, vit.pos_embed)
    if hw == N and H == W:
        gpe[0, 0].add_(gsum[0])
        gpe[0, 1:].add_(gsum[1 + R:])
    else:
        # bicubic / antialias resampling of the parameter grid (vision_transformer.py:206-240) is a function of the parameter
        # alone, evaluated once per image size: its (linear) adjoint is taken from the same torch op
        import math
        M = int(math.sqrt(N))
        h0, w0 = H // vit.patch_size, W // vit.patch_size
        with torch.enable_grad():
            leaf = vit.pos_embed.detach().float()[:, 1:].reshape(1, M, M, D).permute(0, 3, 1, 2).clone().requires_grad_(True)
            kw = {"scale_factor": (float(h0 + vit.interpolate_offset) / M, float(w0 + vit.interpolate_offset) / M)} \
                if vit.interpolate_offset else {"size": (h0, w0)}
            grid = nn.functional.interpolate(leaf, mode="bicubic", antialias=vit.interpolate_antialias, **kw)
            grid.backward(gsum[1 + R:].view(1, h0, w0, D).permute(0, 3, 1, 2))
        gpe[0, 0].add_(gsum[0])
        gpe[0, 1:].add_(leaf.grad.permute(0, 2, 3, 1).reshape(N, D))
    dz = ops.cast_bf16(dX.view(S, P, D)[:, 1 + R:].reshape(S * hw, D))
    _proj_wgrad(vit.patch_embed.proj, sv.a, dz, grads)


# ------------------------------------------------------------------------------------------------ autograd bridge
class AggregatorFunction(torch.autograd.Function):
    """Aggregator.forward as ONE autograd node, so the reference's training loop runs unchanged through the overlay
    (train_imc.py:361-381: `predictions = model(...)` under autocast(bf16), `scaler.scale(loss).backward()`, DDP hooks,
    `scaler.step(optimizer)`): forward = patch embed + stack_forward on the C-ABI kernels, backward = stack_backward.
    Differentiated: the patch embed (DINOv2 ViT blocks + patch projection + position embedding / cls / register tokens, or
    the "conv" projection; embed_forward_saving / embed_backward), the four special-token parameters and the 3 x depth
    aggregator blocks — every parameter of the Aggregator except the ViT's unused mask_token.  Not differentiated: the
    images.  Inputs after the six non-tensor arguments are exactly the tensors that receive gradients, in
    `AggregatorFunction.parameters_of(agg)` order."""

    @staticmethod
    def parameters_of(agg: nn.Module) -> List[nn.Parameter]:
        ps = [agg.camera_token, agg.register_token, agg.camera_token_reloc, agg.register_token_reloc]
        ps.extend(p for p in agg.patch_embed.parameters() if p.requires_grad)
        for i in range(agg.depth):
            for blk in (agg.frame_blocks[i], agg.global_blocks[i], agg.global_reloc_blocks[i]):
                ps.extend(blk.parameters())
        return ps

    @staticmethod
    def forward(ctx, agg, images, no_reloc_list, reloc_list, fix_rank, keep_activations, *params):
        B, S, C_in, H, W = images.shape
        Na, Nq = len(no_reloc_list), len(reloc_list)
        if C_in != 3:
            raise ValueError(f"Expected 3 input channels, got {C_in}")
        if B != 1 or Na == 0 or sorted(list(no_reloc_list) + list(reloc_list)) != list(range(S)):
            raise NotImplementedError("sailrecon_b200: training forward needs B = 1, at least one anchor and disjoint lists "
                                      "covering range(S)")
        if engine.compute_mode() != "bf16":
            raise NotImplementedError("sailrecon_b200: the differentiated path runs under torch.autocast(bfloat16), as "
                                      "train/train_imc.py:363 does")
        engine.require_cuda_f32(images, "images")
        dev = images.device
        C = agg.camera_token.shape[-1]
        psi = agg.patch_start_idx
        h, w = H // agg.patch_size, W // agg.patch_size
        hw, P = h * w, psi + h * w
        agg.num_recon = Na
        agg.rank = agg._draw_rank(fix_rank, hw, clamp=True)
        l_prime = min(agg.rank, hw)
        order = list(no_reloc_list) + list(reloc_list)
        imgs = images[0] if order == list(range(S)) else images[0][torch.tensor(order, device=dev)]
        T = S * P
        X0 = torch.empty(T, C, dtype=torch.float32, device=dev)
        ctx.embed = embed_forward_saving(agg, imgs.contiguous(), X0, S, P, C)
        X0.view(S, P, C)[:, :psi] = agg._special_rows(order, set(reloc_list), False)
        rows_pos, frame_pos_cpu = agg._positions(S, h, w, dev)
        lctx = LayerCtx(Na, Nq, P, rows_pos)

        def draw_scene():
            idx, spos, drawn = agg._scene_rows(Na, P, hw, l_prime, frame_pos_cpu, dev)
            agg.last_scene_indices.append(drawn)
            return idx, engine.RowPositions(spos, rows_pos.max_pos)

        if keep_activations is None:
            free, _ = torch.cuda.mem_get_info(dev)
            free += torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev)
            keep_activations = choose_keep_activations(lctx, agg.depth, C, Na * (psi + l_prime) if Nq else 0, free)
        outs, cams = {}, {}

        def on_layer(i, X_mid, X_out):
            if i in agg.intermediate_layer_idx and Nq > 0:                               # aggregator.py:403-413
                o = torch.empty(1, Nq, P, 2 * C, dtype=torch.float32, device=dev)
                ops.copy_cols(X_mid[Na * P:], o.view(Nq * P, 2 * C), 0)
                ops.copy_cols(X_out[Na * P:], o.view(Nq * P, 2 * C), C)
                outs[i] = o
            if i == agg.depth - 1:                                                       # :414-423
                cams[0] = torch.cat([X_mid.view(S, P, C)[:Na, 0], X_out.view(S, P, C)[:Na, 0]], dim=-1)[None].clone()

        agg.last_scene_indices = []
        ck = stack_forward(agg, X0 if keep_activations else X0.clone(), lctx, draw_scene, keep_activations, on_layer)
        ctx.agg, ctx.ck, ctx.lctx, ctx.order = agg, ck, lctx, order
        ctx.geom = (S, P, C, Na, Nq, psi)
        ctx.layers = sorted(outs)
        ctx.params = params
        ctx.reloc = set(reloc_list)
        return tuple(outs[i] for i in ctx.layers) + (cams[0],)

    @staticmethod
    def backward(ctx, *gouts):
        agg, lctx = ctx.agg, ctx.lctx
        S, P, C, Na, Nq, psi = ctx.geom
        T = S * P
        dev = ctx.params[0].device
        d_out = {}
        for i, g in zip(ctx.layers, gouts[:-1]):
            if g is None:
                continue
            g2 = g.reshape(Nq * P, 2 * C).float()
            d_mid = torch.zeros(T, C, dtype=torch.float32, device=dev)
            d_post = torch.zeros(T, C, dtype=torch.float32, device=dev)
            d_mid[Na * P:] = g2[:, :C]
            d_post[Na * P:] = g2[:, C:]
            d_out[i] = (d_mid, d_post)
        g_cam = gouts[-1]
        if g_cam is not None:
            last = agg.depth - 1
            if last not in d_out:
                d_out[last] = (torch.zeros(T, C, dtype=torch.float32, device=dev), torch.zeros(T, C, dtype=torch.float32, device=dev))
            d_out[last][0].view(S, P, C)[:Na, 0] += g_cam[0, :, :C].float()
            d_out[last][1].view(S, P, C)[:Na, 0] += g_cam[0, :, C:].float()
        grads: Dict[nn.Parameter, torch.Tensor] = {}
        dX0 = stack_backward(agg, ctx.ck, d_out, lctx, grads)
        ctx.ck = None
        embed_backward(agg, ctx.embed, dX0, grads)                  # DINOv2 ViT blocks / conv projection, pos-embed, cls, registers
        ctx.embed = None
        # gradient of the special-token rows (aggregator.py:287-299): frame 0 -> slot 0, other anchors -> slot 1, queries -> *_reloc
        d_sp = dX0.view(S, P, C)[:, :psi]
        g_cam_tok = torch.zeros_like(agg.camera_token)
        g_reg_tok = torch.zeros_like(agg.register_token)
        g_cam_rel = torch.zeros_like(agg.camera_token_reloc)
        g_reg_rel = torch.zeros_like(agg.register_token_reloc)
        for pos_in_order, f in enumerate(ctx.order):
            row = d_sp[pos_in_order]
            if f in ctx.reloc:
                g_cam_rel[0, 0] += row[:1]
                g_reg_rel[0, 0] += row[1:]
            else:
                slot = 0 if f == 0 else 1
                g_cam_tok[0, slot] += row[:1]
                g_reg_tok[0, slot] += row[1:]
        grads[agg.camera_token], grads[agg.register_token] = g_cam_tok, g_reg_tok
        grads[agg.camera_token_reloc], grads[agg.register_token_reloc] = g_cam_rel, g_reg_rel
        by_id = {id(p): g for p, g in grads.items()}
        plist = AggregatorFunction.parameters_of(agg)
        return (None,) * 6 + tuple(by_id.get(id(p)) for p in plist)


def aggregator_forward_autograd(agg: nn.Module, images: torch.Tensor, no_reloc_list, reloc_list, fix_rank=None,
                                keep_activations: Optional[bool] = None):
    """Aggregator.forward (aggregator.py:242-433) with a grad_fn: what Aggregator.forward dispatches to in train() mode
    with autograd enabled.  Returns (output_dict, patch_start_idx, cam_token_last_layer) like the inference forward."""
    params = AggregatorFunction.parameters_of(agg)
    res = AggregatorFunction.apply(agg, images, list(no_reloc_list), list(reloc_list), fix_rank, keep_activations, *params)
    layers = sorted(i for i in agg.intermediate_layer_idx) if len(reloc_list) else []
    output_dict = dict(zip(layers, res[:-1]))
    assert (agg.depth - 1 in output_dict) or len(reloc_list) == 0, \
        f"Please make sure the last layer ({agg.depth - 1}) is in the output_dict: {output_dict.keys()}"
    if len(reloc_list):
        output_dict[-1] = output_dict[agg.depth - 1]
    return output_dict, agg.patch_start_idx, res[-1]

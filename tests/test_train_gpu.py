"""GPU parity of the training-step kernels (csrc/train_rows.cu, csrc/attn_bwd.cu via sailrecon_b200.train) against
plain fp32 PyTorch autograd of the same Block (sailrecon/layers/block.py:86-112 with attention.py:70-122, rope.py:136-207,
mlp.py:33-40, layer_scale.py:22).  The kernels run the bf16 autocast arithmetic of train/train_imc.py:358-380, the
reference here is fp32: tolerances are the bf16 ones of north_star (2e-2 relative L2), 3e-2 for parameter gradients.
"""
import math
import os

import pytest
import torch
import torch.nn.functional as F

from golden_common import rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF = torch.bfloat16


def _g(seed):
    return torch.Generator(device=DEV).manual_seed(seed)


def _randn(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, device=DEV, generator=_g(seed)) * scale


# ------------------------------------------------------------------------------------------------ row kernels
def test_transpose_colsum_scale_cast_ls_residual():
    from sailrecon_b200 import ops
    a = _randn(333, 1024, seed=1).to(BF)
    t = ops.transpose_bf16(a)
    assert t.shape == (1024, 384)
    assert torch.equal(t[:, :333], a.t()) and bool((t[:, 333:] == 0).all())
    cs = torch.ones(1024, device=DEV)
    t2 = ops.transpose_bf16(a, colsum=cs)                                    # fused column sum (bias gradient), accumulating
    assert torch.equal(t2, t) and rel_l2(cs.cpu(), (1.0 + a.double().sum(0)).float().cpu()) < 1e-5
    wide = _randn(200, 3072, seed=4).to(BF)
    for view in (wide[:, 1024:2048], wide[:, 8:72], wide[3:131, :1000]):     # strided, narrow and row-offset sources
        cs = torch.zeros(view.shape[1], device=DEV)
        tv = ops.transpose_bf16(view, colsum=cs)
        rp = (view.shape[0] + 63) // 64 * 64
        assert tv.shape == (view.shape[1], rp) and torch.equal(tv[:, :view.shape[0]], view.t()) and bool((tv[:, view.shape[0]:] == 0).all())
        assert rel_l2(cs.cpu(), view.double().sum(0).float().cpu()) < 1e-5
    odd = wide[:, 3:67]                                                      # not 16-byte aligned: scalar kernel, no fused sum
    assert torch.equal(ops.transpose_bf16(odd)[:, :200], odd.t())
    with pytest.raises(ValueError):
        ops.transpose_bf16(odd, colsum=torch.zeros(64, device=DEV))
    x = _randn(333, 1024, seed=2)
    out = torch.full((1024,), float("nan"), device=DEV)
    ops.colsum(x, a, out)
    ref = (x.double() * a.double()).sum(0)
    assert rel_l2(out.cpu(), ref.float().cpu()) < 1e-5
    ops.colsum(a, None, out, accumulate=True)
    assert rel_l2(out.cpu(), (ref + a.double().sum(0)).float().cpu()) < 1e-5
    gam = _randn(1024, seed=3)
    o = ops.scale_cast_bf16(x, gam, torch.empty(333, 1024, dtype=BF, device=DEV))
    assert torch.equal(o, (x * gam).to(BF))
    y = ops.ls_residual(x, a, gam, torch.empty_like(x))
    assert torch.allclose(y, x + gam * a.float(), rtol=1e-6, atol=1e-6)


def test_gelu_fwd_bwd():
    from sailrecon_b200 import ops
    u = _randn(77, 4096, seed=4, scale=2.0).to(BF)
    dg = _randn(77, 4096, seed=5).to(BF)
    g = ops.gelu_fwd(u, torch.empty_like(u))
    uf = u.float().requires_grad_(True)
    ref = F.gelu(uf)
    assert rel_l2(g.float().cpu(), ref.detach().cpu()) < 4e-3
    ref.backward(dg.float())
    du = ops.gelu_bwd(u, dg, torch.empty_like(u))
    assert rel_l2(du.float().cpu(), uf.grad.cpu()) < 4e-3


def test_fused_ls_bwd_and_gelu_bwd_colsum_equal_the_separate_launches():
    """sr_ls_bwd_bf16 == colsum(dy, z) + scale_cast(dy, gamma) + colsum(dz); sr_gelu_bwd_colsum_bf16 == gelu_bwd + colsum(du):
    the bf16 outputs bit-identical, the column sums to fp32 summation order.  Ragged row counts, strided inputs, in place."""
    from sailrecon_b200 import ops
    for R in (1, 255, 333, 1374 * 3):
        dy, gam = _randn(R, 1024, seed=R), _randn(1024, seed=7)
        zwide = _randn(R, 2048, seed=R + 1).to(BF)
        z = zwide[:, 512:1536]                                               # strided view
        dg0, db0 = torch.full((1024,), 0.5, device=DEV), torch.full((1024,), -0.25, device=DEV)
        dz = ops.ls_bwd(dy, z, gam, torch.empty(R, 1024, dtype=BF, device=DEV), dg0, db0)
        ref_dz = ops.scale_cast_bf16(dy, gam, torch.empty(R, 1024, dtype=BF, device=DEV))
        assert torch.equal(dz, ref_dz)
        rg, rb = torch.full((1024,), 0.5, device=DEV), torch.full((1024,), -0.25, device=DEV)
        ops.colsum(dy, z.contiguous(), rg, accumulate=True)
        ops.colsum(ref_dz, None, rb, accumulate=True)
        assert rel_l2(dg0.cpu(), rg.cpu()) < 1e-5 and rel_l2(db0.cpu(), rb.cpu()) < 1e-5
        assert rel_l2(dg0.cpu(), (0.5 + (dy.double() * z.double()).sum(0)).float().cpu()) < 1e-5
        dg1 = torch.zeros(1024, device=DEV)
        ops.ls_bwd(dy, z, gam, dz, dg1, None)                                # no bias: skipped
        u, dgr = _randn(R, 4096, seed=R + 2, scale=2.0).to(BF), _randn(R, 4096, seed=R + 3).to(BF)
        ref_du = ops.gelu_bwd(u, dgr, torch.empty_like(u))
        rb4 = torch.zeros(4096, device=DEV)
        ops.colsum(ref_du, None, rb4, accumulate=True)
        b4 = torch.zeros(4096, device=DEV)
        work = dgr.clone()
        du = ops.gelu_bwd_colsum(u, work, work, b4)                          # in place over the incoming gradient
        assert du.data_ptr() == work.data_ptr() and torch.equal(du, ref_du)
        assert rel_l2(b4.cpu(), rb4.cpu()) < 1e-5


def test_block_backward_fused_row_kernels_equal_the_separate_ones():
    from sailrecon_b200 import engine, train
    blk = _rand_block(9)
    T = 450
    x, dy = _randn(T, 1024, seed=30), _randn(T, 1024, seed=31)
    pos = _positions(3, 150)
    rp = engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1)
    res = []
    for fused in (True, False):
        train.FUSED_ROW_BACKWARD = fused
        try:
            saved = train.block_forward_saving(blk, x, engine.full_segments(3, 150), rp)
            grads = {}
            dx = train.block_backward(blk, saved, dy, grads)
        finally:
            train.FUSED_ROW_BACKWARD = True
        res.append((dx, grads))
    assert torch.equal(res[0][0], res[1][0])
    for p_, g_ in res[0][1].items():
        assert rel_l2(g_.cpu(), res[1][1][p_].cpu()) < 1e-5


def test_layernorm_bwd():
    from sailrecon_b200 import ops
    T, C = 203, 1024
    x = _randn(T, C, seed=6, scale=1.7) + 0.3
    w = _randn(C, seed=7) * 0.2 + 1.0
    b = _randn(C, seed=8) * 0.1
    dy = _randn(T, C, seed=9).to(BF)
    base = _randn(T, C, seed=10)
    xr = x.clone().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    F.layer_norm(xr, (C,), wr, br, 1e-5).backward(dy.float())
    dx = base.clone()
    dw = torch.zeros(C, device=DEV)
    db = torch.zeros(C, device=DEV)
    ops.layernorm_bwd(x, w, dy, dx, dw, db, 1e-5)
    assert rel_l2((dx - base).cpu(), xr.grad.cpu()) < 1e-4
    assert rel_l2(dw.cpu(), wr.grad.cpu()) < 1e-4
    assert rel_l2(db.cpu(), br.grad.cpu()) < 1e-4


def _rope_ref(x, pos, freq=100.0):
    """x [T, H, 64], pos int [T, 2] (y, x): rope.py:136-207 — halves rotate with y / x, pairs (i, i + 16), 16 frequencies."""
    inv = 1.0 / freq ** (torch.arange(16, device=x.device, dtype=torch.float32) / 16.0)
    parts = []
    for half in range(2):
        ang = pos[:, half].float()[:, None] * inv[None]
        c, s = ang.cos()[:, None, :], ang.sin()[:, None, :]
        a, b = x[..., half * 32:half * 32 + 16], x[..., half * 32 + 16:half * 32 + 32]
        parts += [a * c - b * s, b * c + a * s]
    return torch.cat(parts, dim=-1)


def _positions(frames, P):
    """5 special tokens at (0, 0), then a patch grid offset by 1 (aggregator.py:319-328)."""
    side = int(math.ceil(math.sqrt(P - 5)))
    ys = torch.arange(P - 5) // side + 1
    xs = torch.arange(P - 5) % side + 1
    one = torch.cat([torch.zeros(5, 2, dtype=torch.long), torch.stack([ys, xs], -1)])
    return one.repeat(frames, 1).to(DEV)


def test_qknorm_rope_fwd_equals_fused_epilogue_and_bwd_matches_autograd():
    from sailrecon_b200 import engine, ops
    from sailrecon_b200.layers import RotaryPositionEmbedding2D
    T, H, d = 300, 16, 64
    C = H * d
    rope = RotaryPositionEmbedding2D(frequency=100.0)
    pos = _positions(2, 150)
    rp = engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1)
    cos, sin = rope.tables(DEV, rp.max_pos, d)
    qn = (_randn(d, seed=11) * 0.2 + 1.0, _randn(d, seed=12) * 0.1)
    kn = (_randn(d, seed=13) * 0.2 + 1.0, _randn(d, seed=14) * 0.1)
    epi = ops.QkvEpilogue(C, 0, d, qn, kn, 1e-5, cos, sin, rp.pos32)
    raw = _randn(T, 3 * C, seed=15, scale=1.5).to(BF)
    out = ops.qknorm_rope_fwd(raw, torch.empty_like(raw), epi)

    def ref_fwd(r):
        q, k, v = (r[:, i * C:(i + 1) * C].reshape(T, H, d) for i in range(3))
        q = _rope_ref(F.layer_norm(q, (d,), qw, qb, 1e-5), pos)
        k = _rope_ref(F.layer_norm(k, (d,), kw, kb, 1e-5), pos)
        return torch.cat([q.reshape(T, C), k.reshape(T, C), v.reshape(T, C)], dim=1)

    qw, qb, kw, kb = (t.clone().requires_grad_(True) for t in (*qn, *kn))
    rr = raw.float().requires_grad_(True)
    ref = ref_fwd(rr)
    assert rel_l2(out.float().cpu(), ref.detach().cpu()) < 4e-3
    assert torch.equal(out[:, 2 * C:], raw[:, 2 * C:])
    dpost = _randn(T, 3 * C, seed=16).to(BF)
    ref.backward(dpost.float())
    dq = [torch.zeros(d, device=DEV) for _ in range(4)]
    draw = ops.qknorm_rope_bwd(raw, dpost.clone(), epi, (dq[0], dq[1]), (dq[2], dq[3]))
    assert rel_l2(draw.float().cpu(), rr.grad.cpu()) < 6e-3
    for got, want in zip(dq, (qw, qb, kw, kb)):
        assert rel_l2(got.cpu(), want.grad.cpu()) < 2e-3



@pytest.mark.parametrize("ph,Kp", [(14, 640), (8, 256)])
def test_patch_projection_weight_gradient_at_im2col_widths_that_are_not_multiples_of_128(ph, Kp):
    """train._proj_wgrad: dW of Conv2d(3, D, ph, stride ph) as dz^T im2col on the transposed-operand GEMM, which needs the
    im2col width % 128 == 0 — patch 14 gets 640 from ceil-to-64 by luck, patch 8 (192 -> 256) needs the explicit padding."""
    from sailrecon_b200 import ops, train
    from sailrecon_b200.models.aggregator import _RESNET_MEAN, _RESNET_STD
    torch.manual_seed(ph)
    D, S, H, W = 128, 2, 8 * ph, 12 * ph
    proj = torch.nn.Conv2d(3, D, ph, ph).to(DEV)
    imgs = torch.rand(S, 3, H, W, device=DEV)
    T, K = S * 8 * 12, 3 * ph * ph
    assert train._im2col_width(K) == Kp
    a = torch.empty(T, Kp, dtype=BF, device=DEV)
    ops.patchify(imgs, a, ph, _RESNET_MEAN, _RESNET_STD)
    dz = _randn(T, D, seed=3).to(BF)
    grads = {}
    train._proj_wgrad(proj, a, dz, grads)
    mean = torch.tensor(_RESNET_MEAN, device=DEV).view(1, 3, 1, 1)
    std = torch.tensor(_RESNET_STD, device=DEV).view(1, 3, 1, 1)
    x = ((imgs - mean) / std).to(BF).float()                              # the kernel's operand rounding
    z = F.conv2d(x, proj.weight, proj.bias, stride=ph).permute(0, 2, 3, 1).reshape(T, D)
    (z * dz.float()).sum().backward()
    assert rel_l2(grads[proj.weight], proj.weight.grad) < 1e-3 and rel_l2(grads[proj.bias], proj.bias.grad) < 1e-3
    with pytest.raises(ValueError, match="padded"):
        train._proj_wgrad(proj, a[:, :Kp - 64].contiguous(), dz, grads)

@pytest.mark.parametrize("T,K,N", [(333, 1024, 3072), (2748, 1024, 4096), (1374, 4096, 1024), (70, 1024, 1024)])
def test_linear_backward_on_tensor_cores(T, K, N):
    """dW = dz^T a and da = dz W on the tensor-core GEMM reading its operands in place (MN-major descriptors, sr_gemm_bf16_t)
    against fp32 torch, and against the round-1 path (explicit transposes + the K-major kernel): same products, same fp32
    accumulation order along the token axis -> identical up to the last bit of the accumulate."""
    import torch.nn as nn
    from sailrecon_b200 import train
    lin = nn.Linear(K, N).to(DEV)
    a = _randn(T, K, seed=17).to(BF)
    dz = _randn(T, N, seed=18).to(BF)
    grads = {}
    da = train.linear_backward(lin, a, dz, grads)
    w = lin.weight.detach().to(BF).float()
    assert rel_l2(da.float().cpu(), (dz.float() @ w).cpu()) < 6e-3
    assert rel_l2(grads[lin.weight].cpu(), (dz.float().t() @ a.float()).cpu()) < 1e-3
    assert rel_l2(grads[lin.bias].cpu(), dz.float().sum(0).cpu()) < 1e-4
    old = {}
    da_old = train.linear_backward(lin, a, dz, old, use_transposes=True)
    assert rel_l2(da.float(), da_old.float()) < 1e-6 and rel_l2(grads[lin.weight], old[lin.weight]) < 1e-6
    train.linear_backward(lin, a, dz, grads)                                   # accumulates (exact fp32 add in the epilogue)
    assert rel_l2(grads[lin.weight].cpu(), (2 * dz.float().t() @ a.float()).cpu()) < 1e-3
    assert rel_l2(grads[lin.bias].cpu(), 2 * dz.float().sum(0).cpu()) < 1e-4


# ------------------------------------------------------------------------------------------------ whole block
def _qkv_ref(P, h1, pos, H):
    T, C = h1.shape
    d = C // H
    qkv = h1 @ P["attn.qkv.weight"].t() + P["attn.qkv.bias"]
    q, k, v = (qkv[:, i * C:(i + 1) * C].reshape(T, H, d) for i in range(3))
    q = _rope_ref(F.layer_norm(q, (d,), P["attn.q_norm.weight"], P["attn.q_norm.bias"], 1e-5), pos)
    k = _rope_ref(F.layer_norm(k, (d,), P["attn.k_norm.weight"], P["attn.k_norm.bias"], 1e-5), pos)
    return q, k, v


def _block_ref(P, x, pos, batch, L, H, prefix=None):
    """prefix = (k, v) [Ns, H, d] every segment also attends to (the scene tokens of the reloc block)."""
    T, C = x.shape
    d = C // H
    h1 = F.layer_norm(x, (C,), P["norm1.weight"], P["norm1.bias"], 1e-5)
    q, k, v = _qkv_ref(P, h1, pos, H)
    outs = []
    for b in range(batch):
        sl = slice(b * L, (b + 1) * L)
        kk, vv = (k[sl], v[sl]) if prefix is None else (torch.cat([prefix[0], k[sl]]), torch.cat([prefix[1], v[sl]]))
        s = torch.einsum("qhd,khd->hqk", q[sl], kk) * d ** -0.5
        outs.append(torch.einsum("hqk,khd->qhd", torch.softmax(s, -1), vv).reshape(L, C))
    o = torch.cat(outs)
    x1 = x + P["ls1.gamma"] * (o @ P["attn.proj.weight"].t() + P["attn.proj.bias"])
    h2 = F.layer_norm(x1, (C,), P["norm2.weight"], P["norm2.bias"], 1e-5)
    g = F.gelu(h2 @ P["mlp.fc1.weight"].t() + P["mlp.fc1.bias"])
    return x1 + P["ls2.gamma"] * (g @ P["mlp.fc2.weight"].t() + P["mlp.fc2.bias"])


@pytest.mark.parametrize("batch,L", [(3, 150), (1, 450)])
def test_block_forward_backward_matches_fp32_autograd(batch, L):
    """One aggregator Block (dim 1024, 16 heads, q/k norm, 2D RoPE, LayerScale) as a frame block (3 segments) and as a global
    block (one segment): saving forward == inference forward, dx and every parameter gradient against fp32 autograd."""
    from sailrecon_b200 import engine, train
    from sailrecon_b200.layers import Block, RotaryPositionEmbedding2D
    C, H = 1024, 16
    T = batch * L
    torch.manual_seed(5)
    blk = Block(C, H, qk_norm=True, rope=RotaryPositionEmbedding2D(frequency=100.0), init_values=0.01).to(DEV)
    with torch.no_grad():                                   # O(1) LayerScale and non-trivial norms so every path shows up
        for n, p in blk.named_parameters():
            if n.endswith("gamma"):
                p.uniform_(0.3, 0.9)
            elif "norm" in n and n.endswith("weight"):
                p.uniform_(0.7, 1.3)
            elif n.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    x = _randn(T, C, seed=20)
    dy = _randn(T, C, seed=21)
    pos = _positions(T // 150, 150)
    rp = engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1)
    calls = engine.full_segments(batch, L)

    saved = train.block_forward_saving(blk, x, calls, rp)
    y = train.block_output(blk, saved)
    y_inf = engine.run_block(blk, x.clone(), calls, rp, "bf16")
    assert rel_l2(y.cpu(), y_inf.cpu()) < 2e-3              # fused and un-fused epilogues round at the same places
    grads = {}
    dx = train.block_backward(blk, saved, dy, grads)
    torch.cuda.synchronize()

    P = {n: p.detach().clone().requires_grad_(True) for n, p in blk.named_parameters()}
    xr = x.clone().requires_grad_(True)
    y_ref = _block_ref(P, xr, pos, batch, L, H)
    y_ref.backward(dy)
    assert rel_l2(y.cpu(), y_ref.detach().cpu()) < 2e-2
    assert rel_l2(dx.cpu(), xr.grad.cpu()) < 2e-2
    named = dict(blk.named_parameters())
    assert set(grads) == set(named.values())
    for n, p in named.items():
        e = rel_l2(grads[p].cpu(), P[n].grad.cpu())
        assert e < 3e-2, (n, e)


def _rand_block(seed):
    from sailrecon_b200.layers import Block, RotaryPositionEmbedding2D
    torch.manual_seed(seed)
    blk = Block(1024, 16, qk_norm=True, rope=RotaryPositionEmbedding2D(frequency=100.0), init_values=0.01).to(DEV)
    with torch.no_grad():
        for n, p in blk.named_parameters():
            if n.endswith("gamma"):
                p.uniform_(0.3, 0.9)
            elif "norm" in n and n.endswith("weight"):
                p.uniform_(0.7, 1.3)
            elif n.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    return blk


def test_aggregator_layer_forward_backward_matches_fp32_autograd():
    """One aggregator layer (aggregator.py:343-392): frame block on 2 anchors + 2 queries, scene tokens gathered from the
    anchors (35 rows each) -> K|V prefix of the reloc block on the queries, global block on the anchors; gradients w.r.t.
    the layer input and all three blocks' parameters, incl. the path scene K/V -> k_norm / RoPE -> qkv -> norm1 -> gather."""
    from sailrecon_b200 import engine, train
    Na, Nq, Pn, H, C = 2, 2, 150, 16, 1024
    S, n = Na + Nq, Na * Pn
    fblk, gblk, rblk = _rand_block(31), _rand_block(32), _rand_block(33)
    pos = _positions(S, Pn)
    ctx = train.LayerCtx(Na, Nq, Pn, engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1))
    gidx = torch.Generator().manual_seed(3)
    idx = torch.cat([a * Pn + torch.cat([torch.arange(5), 5 + torch.randperm(Pn - 5, generator=gidx)[:30]]) for a in range(Na)])
    idx_d = idx.to(torch.int32).to(DEV)
    spos = engine.RowPositions(ctx.rows_pos.pos32[idx_d.long()].contiguous(), ctx.rows_pos.max_pos)
    X0 = _randn(S * Pn, C, seed=34)
    dY = _randn(S * Pn, C, seed=35)
    d_mid = _randn(S * Pn, C, seed=36) * 0.5

    X = X0.clone()
    X_mid = train.layer_forward(fblk, gblk, rblk, X, ctx, idx_d, spos)
    grads = {}
    dX0 = train.layer_backward(fblk, gblk, rblk, X0, X_mid, dY, ctx, idx_d, spos, grads, d_mid=d_mid)
    torch.cuda.synchronize()

    Pf, Pg, Pr = ({k: p.detach().clone().requires_grad_(True) for k, p in b.named_parameters()} for b in (fblk, gblk, rblk))
    xr = X0.clone().requires_grad_(True)
    mid = _block_ref(Pf, xr, pos, S, Pn, H)
    ya = _block_ref(Pg, mid[:n], pos[:n], 1, n, H)
    hs = F.layer_norm(mid[idx.to(DEV)], (C,), Pr["norm1.weight"], Pr["norm1.bias"], 1e-5)
    _, sk, sv = _qkv_ref(Pr, hs, pos[idx.to(DEV)], H)
    yq = _block_ref(Pr, mid[n:], pos[n:], Nq, Pn, H, prefix=(sk, sv))
    y_ref = torch.cat([ya, yq])
    ((y_ref * dY).sum() + (mid * d_mid).sum()).backward()
    assert rel_l2(X.cpu(), y_ref.detach().cpu()) < 2e-2
    assert rel_l2(X_mid.cpu(), mid.detach().cpu()) < 2e-2
    assert rel_l2(dX0.cpu(), xr.grad.cpu()) < 2e-2
    for blk, Pref, tag in ((fblk, Pf, "frame"), (gblk, Pg, "global"), (rblk, Pr, "reloc")):
        for name, p in blk.named_parameters():
            e = rel_l2(grads[p].cpu(), Pref[name].grad.cpu())
            assert e < 3e-2, (tag, name, e)


def test_adam_step_matches_torch_optim():
    """sr_adam_step_f32 against torch.optim.Adam(lr, betas=(0.9, 0.999), eps=1e-8) (train_imc.py:475) over three steps, with the
    1 / world gradient scale folded in."""
    from sailrecon_b200 import ops
    n = 4 * 100003
    p0 = _randn(n, seed=51)
    p, m, v = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref], lr=3e-3, betas=(0.9, 0.999), eps=1e-8)
    for step in range(1, 4):
        g = _randn(n, seed=60 + step) * (10.0 ** (step - 2))
        ops.adam_step(p, g, m, v, 3e-3, step, grad_scale=0.5)
        ref.grad = g * 0.5
        opt.step()
    torch.cuda.synchronize()
    assert (p - ref.detach()).abs().max().item() < 2e-6
    st = opt.state[ref]
    assert rel_l2(m.cpu(), st["exp_avg"].cpu()) < 1e-6 and rel_l2(v.cpu(), st["exp_avg_sq"].cpu()) < 1e-6
    with pytest.raises(ValueError):
        ops.adam_step(p[:6], g[:6], m[:6], v[:6], 1e-3, 1)


class _Stacks(torch.nn.Module):
    """The three block stacks of an Aggregator (aggregator.py:99-149) without the patch embed."""

    def __init__(self, depth, seed):
        super().__init__()
        self.depth = depth
        self.frame_blocks = torch.nn.ModuleList([_rand_block(seed + 3 * i) for i in range(depth)])
        self.global_blocks = torch.nn.ModuleList([_rand_block(seed + 3 * i + 1) for i in range(depth)])
        self.global_reloc_blocks = torch.nn.ModuleList([_rand_block(seed + 3 * i + 2) for i in range(depth)])


@pytest.mark.parametrize("keep_activations", [True, False, None])
def test_training_step_two_layers_matches_fp32_autograd_and_adam(keep_activations):
    """train.training_step over a 2-layer stack (checkpointed forward, backward through both layers with gradients on both
    halves of the layer-0 output and on the final state, flat per-layer buckets, Adam): input gradient and every parameter
    gradient against fp32 autograd; updated parameters against torch.optim.Adam fed with the reference gradients."""
    from sailrecon_b200 import engine, train
    Na, Nq, Pn, H, C, depth = 2, 2, 150, 16, 1024, 2
    S, n = Na + Nq, Na * Pn
    agg = _Stacks(depth, 71)
    before = {k: p.detach().clone() for k, p in agg.named_parameters()}
    state = train.TrainState(agg, lr=1e-3)
    for k, p in agg.named_parameters():
        assert torch.equal(p.detach(), before[k])                                      # re-pointing keeps the values
    pos = _positions(S, Pn)
    ctx = train.LayerCtx(Na, Nq, Pn, engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1))
    gidx = torch.Generator().manual_seed(5)
    draws = []
    for _ in range(depth):
        idx = torch.cat([a * Pn + torch.cat([torch.arange(5), 5 + torch.randperm(Pn - 5, generator=gidx)[:30]]) for a in range(Na)])
        idx_d = idx.to(torch.int32).to(DEV)
        draws.append((idx, idx_d, engine.RowPositions(ctx.rows_pos.pos32[idx_d.long()].contiguous(), ctx.rows_pos.max_pos)))
    it = iter(draws)

    def draw_scene():
        _, idx_d, spos = next(it)
        return idx_d, spos

    X0 = _randn(S * Pn, C, seed=72)
    d_out = {0: (_randn(S * Pn, C, seed=73) * 0.5, _randn(S * Pn, C, seed=74) * 0.5),
             1: (None, _randn(S * Pn, C, seed=75))}
    X0_in = X0.clone()
    dX0 = train.training_step(agg, state, X0, ctx, draw_scene, d_out, keep_activations=keep_activations)
    torch.cuda.synchronize()
    assert torch.equal(X0, X0_in)                                                  # the step does not touch its input

    P = {k: v.clone().requires_grad_(True) for k, v in before.items()}
    sub = lambda stack, i: {k[len(f"{stack}.{i}."):]: v for k, v in P.items() if k.startswith(f"{stack}.{i}.")}
    xr = X0.clone().requires_grad_(True)
    x, loss = xr, 0.0
    for i in range(depth):
        Pf, Pg, Pr = sub("frame_blocks", i), sub("global_blocks", i), sub("global_reloc_blocks", i)
        idx = draws[i][0].to(DEV)
        mid = _block_ref(Pf, x, pos, S, Pn, H)
        ya = _block_ref(Pg, mid[:n], pos[:n], 1, n, H)
        hs = F.layer_norm(mid[idx], (C,), Pr["norm1.weight"], Pr["norm1.bias"], 1e-5)
        _, sk, sv = _qkv_ref(Pr, hs, pos[idx], H)
        yq = _block_ref(Pr, mid[n:], pos[n:], Nq, Pn, H, prefix=(sk, sv))
        x = torch.cat([ya, yq])
        dm, dp = d_out[i]
        if dm is not None:
            loss = loss + (mid * dm).sum()
        loss = loss + (x * dp).sum()
    loss.backward()
    assert rel_l2(dX0.cpu(), xr.grad.cpu()) < 3e-2
    named = dict(agg.named_parameters())
    for k, p in named.items():
        e = rel_l2(state.grads[p].cpu(), P[k].grad.cpu())
        assert e < 4e-2, (k, e)
    # Adam on our own gradients reproduces torch.optim.Adam exactly (first step: p -= lr * sign-like update)
    for k, p in named.items():
        ref = torch.nn.Parameter(before[k].clone())
        ref.grad = state.grads[p].clone()
        torch.optim.Adam([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8).step()
        assert (p.detach() - ref.detach()).abs().max().item() < 1e-6, k
    # the cached bf16 weights follow the update
    w = agg.frame_blocks[0].mlp.fc1
    assert torch.equal(engine.weight_bf16(w), w.weight.detach().to(torch.bfloat16))
    # GradScaler.step semantics (train_imc.py:380-381): a non-finite gradient skips the optimizer step entirely
    snap = [p_.clone() for p_ in state.param]
    step_before = state.step
    bad = {0: d_out[0], 1: (None, d_out[1][1].clone())}
    bad[1][1][7, 3] = float("inf")
    it = iter(draws)
    train.training_step(agg, state, X0, ctx, draw_scene, bad, keep_activations=keep_activations, skip_nonfinite=True)
    assert state.step == step_before and state.skipped_steps == 1
    assert all(torch.equal(a, b) for a, b in zip(snap, state.param))


# ------------------------------------------------------------------------------------------------ data parallel over NCCL
def _dp_setup(seed_scene):
    """A 1-layer stack + one small scene (2 anchors + 2 queries of 150 tokens); the scene differs per seed."""
    from sailrecon_b200 import engine, train
    Na, Nq, Pn, C = 2, 2, 150, 1024
    S = Na + Nq
    agg = _Stacks(1, 91)
    pos = _positions(S, Pn)
    ctx = train.LayerCtx(Na, Nq, Pn, engine.RowPositions(pos.to(torch.int32).contiguous(), int(pos.max()) + 1))
    gidx = torch.Generator().manual_seed(7)
    idx = torch.cat([a * Pn + torch.cat([torch.arange(5), 5 + torch.randperm(Pn - 5, generator=gidx)[:30]]) for a in range(Na)])
    idx_d = idx.to(torch.int32).to(DEV)
    spos = engine.RowPositions(ctx.rows_pos.pos32[idx_d.long()].contiguous(), ctx.rows_pos.max_pos)
    X0 = _randn(S * Pn, C, seed=100 + seed_scene)
    d_out = {0: (_randn(S * Pn, C, seed=200 + seed_scene) * 0.5, _randn(S * Pn, C, seed=300 + seed_scene))}
    return agg, ctx, (lambda: (idx_d, spos)), X0, d_out


def _dp_worker(rank, world, port, out_path):
    import torch.distributed as dist
    from sailrecon_b200 import train
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        global DEV
        DEV = f"cuda:{rank}"
        agg, ctx, draw, X0, d_out = _dp_setup(rank)
        state = train.TrainState(agg, lr=1e-3)
        train.training_step(agg, state, X0, ctx, draw, d_out, world=world)
        torch.cuda.synchronize()
        torch.save({k: p.detach().cpu() for k, p in agg.named_parameters()}, f"{out_path}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")
def test_training_step_two_processes_nccl(tmp_path):
    """DDP semantics of train_imc.py:472 on real kernels: two ranks, one (different) scene each, per-layer NCCL all-reduce.
    Both ranks must end with identical parameters, equal to ONE Adam step on the mean of the two scenes' gradients
    (computed here in a single process with the same kernels)."""
    import socket

    import torch.multiprocessing as mp
    from sailrecon_b200 import train
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "dp")
    mp.spawn(_dp_worker, args=(2, port, out), nprocs=2, join=True)
    got = [torch.load(f"{out}.{r}") for r in range(2)]
    for k in got[0]:
        assert torch.equal(got[0][k], got[1][k]), k                                # replicas stay bit-identical
    # single-process reference: gradients of both scenes, mean, torch.optim.Adam
    mean = None
    for scene in range(2):
        agg, ctx, draw, X0, d_out = _dp_setup(scene)
        before = {k: p.detach().clone() for k, p in agg.named_parameters()}
        state = train.TrainState(agg, lr=1e-3)
        train.training_step(agg, state, X0, ctx, draw, d_out, optimizer_step=lambda grad_scale: None)
        g = {k: state.grads[p].clone() for k, p in agg.named_parameters()}
        mean = g if mean is None else {k: (mean[k] + g[k]) * 0.5 for k in g}
    worst = 0.0
    for k, w0 in before.items():
        ref = torch.nn.Parameter(w0.clone())
        ref.grad = mean[k]
        torch.optim.Adam([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8).step()
        # first Adam step = lr * g / (|g| + eps): a gradient that differs in the last bits (atomic summation order) moves the
        # update by far less than lr
        worst = max(worst, (got[0][k].to(DEV) - ref.detach()).abs().max().item())
    assert worst < 2e-4, worst

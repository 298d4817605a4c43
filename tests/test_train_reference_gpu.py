"""Backward parity PINNED TO THE REFERENCE: gradients of sailrecon_b200.train (stack_forward /
stack_backward: tcgen05 attention backward, dgrad / wgrad GEMMs, row kernels) against REFERENCE AUTOGRAD through the
unmodified `sailrecon.models.aggregator.Aggregator` (oracle/_ref) under the same torch.autocast(bfloat16) that
train/train_imc.py:361-381 uses — production width (C = 1024, 16 heads, head_dim 64), P = 1374 tokens per frame (518 px),
2 anchors + 2 queries, scene-token prefix of 2 x 305 rows, depth 2 — and, loosely, against the reference in fp32.

How the two are tied together: the reference runs from images with patch_embed="conv"; a forward hook replaces the patch
tokens by an fp32 LEAF holding the same values, so the reference's autograd yields d(loss)/d(patch tokens) and the
gradients of every block parameter; our stack starts from exactly those token rows (X0) and the same special tokens, with
the same scene-token draw (same generator seed).  Loss = sum_i <output_dict[i], G_i> + <cam_token_last_layer, G_cam>.

Tolerances (stated per tensor): bf16 arithmetic on both sides, different accumulation orders and rounding points inside
the fused kernels: input gradient 2e-2, parameter gradients 3e-2 (q/k-norm parameters 4e-2: 64-element reductions of
bf16-rounded gradients over 5 500 rows x 16 heads).
"""
import pytest
import torch

import ref_live
from golden_common import rel_l2

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_live.available(), reason="oracle/_ref absent (run oracle/make_ref.py in the build container)")]
DEV = "cuda"
C, H, DEPTH, NA, NQ, SIZE = 1024, 16, 2, 2, 2, 518


def _pair(patch_embed="conv", vit_blocks=None):
    from sailrecon_b200.models import Aggregator
    R = ref_live.ref()
    kw = dict(depth=DEPTH, patch_embed=patch_embed, intermediate_layer_idx=list(range(DEPTH)))
    torch.manual_seed(3)
    with torch.device(DEV):
        r_agg = R.Aggregator(**kw).eval().to(DEV)
        o_agg = Aggregator(**kw).eval()
    if vit_blocks is not None:                                   # full-width DINOv2 ViT-L, first `vit_blocks` blocks only
        r_agg.patch_embed.blocks = r_agg.patch_embed.blocks[:vit_blocks]
        o_agg.patch_embed.blocks = o_agg.patch_embed.blocks[:vit_blocks]
    g = torch.Generator().manual_seed(4)
    with torch.no_grad():                                        # O(1) LayerScale, non-trivial norms / biases / special tokens
        for n, p in sorted(r_agg.named_parameters()):
            leaf = n.split(".")[-1]
            if leaf == "gamma":
                p.copy_((0.3 + 0.6 * torch.rand(p.shape, generator=g)).to(DEV))
            elif "norm" in n and leaf == "weight":
                p.copy_((0.7 + 0.6 * torch.rand(p.shape, generator=g)).to(DEV))
            elif leaf == "bias":
                p.copy_((0.2 * torch.rand(p.shape, generator=g) - 0.1).to(DEV))
            elif "token" in leaf:
                p.copy_(torch.randn(p.shape, generator=g).to(DEV))
    o_agg.load_state_dict(r_agg.state_dict())
    return r_agg, o_agg


def _reference_grads(r_agg, images, G, Gc, bf16, detach_tokens=True, hw=(SIZE, SIZE)):
    """Reference forward (eval mode: same arithmetic as train mode minus torch.utils.checkpoint) + autograd.
    detach_tokens: replace the patch tokens by an fp32 leaf (gradient w.r.t. the tokens, patch embed not differentiated);
    False: differentiate the whole Aggregator including its patch embed."""
    leaf = {}

    def hook(mod, inp, out):
        if not detach_tokens:
            return None
        t = out.detach().float().requires_grad_(True)
        leaf["pt"] = t
        return t

    h = r_agg.patch_embed.register_forward_hook(hook)
    for p in r_agg.parameters():
        p.grad = None
    r_agg.generator.manual_seed(1234)
    try:
        with torch.enable_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out, psi, cam = r_agg(images, list(range(NA)), list(range(NA, NA + NQ)), fix_rank=300)
    finally:
        h.remove()
    loss = (cam.float() * Gc).sum()
    for i in range(DEPTH):
        loss = loss + (out[i].float() * G[i]).sum()
    loss.backward()
    grads = {n: p.grad.detach().clone() for n, p in r_agg.named_parameters() if p.grad is not None}
    if not detach_tokens:
        return None, None, {k: v.detach() for k, v in out.items()}, cam.detach(), grads
    return leaf["pt"].detach(), leaf["pt"].grad.detach(), {k: v.detach() for k, v in out.items()}, cam.detach(), grads


@pytest.mark.parametrize("keep", [True, False])
def test_stack_backward_matches_reference_autograd(keep):
    from sailrecon_b200 import train
    from sailrecon_b200.engine import RowPositions
    r_agg, o_agg = _pair()
    images = torch.rand(1, NA + NQ, 3, SIZE, SIZE, generator=torch.Generator().manual_seed(8)).to(DEV)
    S, hw = NA + NQ, (SIZE // 14) ** 2
    P, psi = hw + 5, 5
    T = S * P
    gg = torch.Generator(device=DEV).manual_seed(9)
    G = {i: torch.randn(1, NQ, P, 2 * C, device=DEV, generator=gg) for i in range(DEPTH)}
    Gc = torch.randn(1, NA, 2 * C, device=DEV, generator=gg)
    pt, d_pt, r_out, r_cam, r_grads = _reference_grads(r_agg, images, G, Gc, bf16=True)
    _, d_pt32, _, _, r_grads32 = _reference_grads(r_agg, images, G, Gc, bf16=False)

    # ---- our stack on the same token rows
    X0 = torch.empty(S, P, C, device=DEV)
    X0[:, psi:] = pt
    with torch.no_grad():
        X0[:, :psi] = o_agg._special_rows(list(range(S)), set(range(NA, S)), False)
    X0 = X0.view(T, C)
    rows_pos, frame_pos_cpu = o_agg._positions(S, SIZE // 14, SIZE // 14, DEV)
    ctx = train.LayerCtx(NA, NQ, P, rows_pos)

    def draw_scene():
        idx, spos, _ = o_agg._scene_rows(NA, P, hw, 300, frame_pos_cpu, DEV)
        return idx, RowPositions(spos, rows_pos.max_pos)

    d_out = {}
    for i in range(DEPTH):
        d_mid = torch.zeros(T, C, device=DEV)
        d_post = torch.zeros(T, C, device=DEV)
        d_mid[NA * P:] = G[i][0, :, :, :C].reshape(NQ * P, C)        # first half of the concat = frame-block output (:403-413)
        d_post[NA * P:] = G[i][0, :, :, C:].reshape(NQ * P, C)       # second half = reloc-block output
        if i == DEPTH - 1:                                          # cam_token_last_layer = [frame | global] anchors' token 0
            d_mid.view(S, P, C)[:NA, 0] = Gc[0, :, :C]
            d_post.view(S, P, C)[:NA, 0] = Gc[0, :, C:]
        d_out[i] = (d_mid, d_post)
    o_agg.generator.manual_seed(1234)
    grads = {}
    X = X0.clone()
    ck = train.stack_forward(o_agg, X, ctx, draw_scene, keep)
    dX0 = train.stack_backward(o_agg, ck, d_out, ctx, grads)
    torch.cuda.synchronize()

    # ---- input gradient: patch-token rows against the leaf's gradient, special rows against the token parameters' gradients
    e_in = rel_l2(dX0.view(S, P, C)[:, psi:], d_pt)
    e_in32 = rel_l2(dX0.view(S, P, C)[:, psi:], d_pt32)
    d_special = dX0.view(S, P, C)[:, :psi]
    r_special = torch.stack([
        torch.cat([r_grads["camera_token"][0, 0], r_grads["register_token"][0, 0]], 0),                       # frame 0 slot
        torch.cat([r_grads["camera_token"][0, 1], r_grads["register_token"][0, 1]], 0),                       # other anchors
        torch.cat([r_grads["camera_token_reloc"][0, 0], r_grads["register_token_reloc"][0, 0]], 0)])           # queries
    o_special = torch.stack([d_special[0], d_special[1:NA].sum(0), d_special[NA:].sum(0)])
    e_tok = rel_l2(o_special, r_special)
    named = {n: p for n, p in o_agg.named_parameters() if n.split(".")[0] in ("frame_blocks", "global_blocks", "global_reloc_blocks")}
    assert set(grads) == set(named.values()), "every block parameter receives a gradient"
    errs, errs32 = {}, {}
    for n, p in named.items():
        errs[n] = rel_l2(grads[p], r_grads[n])
        errs32[n] = rel_l2(grads[p], r_grads32[n])
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:5]
    print("train parity vs reference autograd (bf16 autocast): dX0", f"{e_in:.2e}", "tokens", f"{e_tok:.2e}", "worst params",
          [(n, f"{e:.2e}") for n, e in worst], "| vs fp32 reference: dX0", f"{e_in32:.2e}", "max param", f"{max(errs32.values()):.2e}")
    assert e_in < 2e-2 and e_tok < 2e-2, (e_in, e_tok)
    for n, e in errs.items():
        assert e < (4e-2 if ("q_norm" in n or "k_norm" in n) else 3e-2), (n, e)
    assert e_in32 < 3e-2 and max(errs32.values()) < 6e-2


@pytest.mark.parametrize("patch_embed,vit_blocks,H,W", [("conv", None, SIZE, SIZE), ("dinov2_vitl14_reg", 2, 392, 518)])
def test_autograd_bridge_runs_the_reference_training_calls(patch_embed, vit_blocks, H, W):
    """train_imc.py:361-381 through the module API: model.train(); under autocast(bf16) `out = aggregator(images, ...)` carries a
    grad_fn (train.AggregatorFunction), `loss.backward()` fills .grad of EVERY Aggregator parameter — block stacks, special
    tokens and the patch embed (conv projection, or DINOv2 ViT-L blocks + projection + cls / register tokens + the
    position embedding through its bicubic interpolation at a non-square size) — compared with the reference's own .grad
    after the same calls; torch.optim.Adam then steps on them.  GradScaler-style loss scaling (x 1024) must scale the
    gradients linearly."""
    r_agg, o_agg = _pair(patch_embed, vit_blocks)
    images = torch.rand(1, NA + NQ, 3, H, W, generator=torch.Generator().manual_seed(8)).to(DEV)
    P = (H // 14) * (W // 14) + 5
    gg = torch.Generator(device=DEV).manual_seed(9)
    G = {i: torch.randn(1, NQ, P, 2 * C, device=DEV, generator=gg) for i in range(DEPTH)}
    Gc = torch.randn(1, NA, 2 * C, device=DEV, generator=gg)
    _, _, r_out, r_cam, r_grads = _reference_grads(r_agg, images, G, Gc, bf16=True, detach_tokens=False)
    o_agg.train()
    o_agg.generator.manual_seed(1234)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out, psi, cam = o_agg(images, list(range(NA)), list(range(NA, NA + NQ)), fix_rank=300)
    assert psi == 5 and set(out) == {0, 1, -1} and out[-1] is out[DEPTH - 1]
    assert out[0].requires_grad and cam.requires_grad and out[0].grad_fn is not None
    for i in range(DEPTH):
        assert rel_l2(out[i].detach(), r_out[i].float()) < 2e-2
    assert rel_l2(cam.detach(), r_cam.float()) < 2e-2
    loss = (cam * Gc).sum()
    for i in range(DEPTH):
        loss = loss + (out[i] * G[i]).sum()
    (1024.0 * loss).backward()
    named = dict(o_agg.named_parameters())
    errs = {}
    for n, p in named.items():
        if n == "patch_embed.mask_token":
            assert p.grad is None                      # unused by the forward (requires_grad False, aggregator.py:238-240)
            continue
        assert p.grad is not None, n
        assert n in r_grads, n
        errs[n] = rel_l2(p.grad / 1024.0, r_grads[n])
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:6]
    print("bridge parity", patch_embed, [(n, f"{e:.2e}") for n, e in worst])
    for n, e in errs.items():
        assert e < (4e-2 if ("q_norm" in n or "k_norm" in n) else 3e-2), (n, e)
    assert len(errs) >= 4 + 3 * DEPTH * 18 + 2
    before = named["frame_blocks.0.attn.qkv.weight"].detach().clone()
    torch.optim.Adam([p for p in o_agg.parameters() if p.grad is not None], lr=1e-4).step()
    assert not torch.equal(named["frame_blocks.0.attn.qkv.weight"].detach(), before)
    # the cached bf16 weight copies follow the update (engine._cached is keyed on the parameter version)
    o_agg.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        o_agg.generator.manual_seed(1234)
        out2, _, _ = o_agg(images, list(range(NA)), list(range(NA, NA + NQ)), fix_rank=300)
    assert rel_l2(out2[-1], out[-1].detach()) > 1e-6

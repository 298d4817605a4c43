"""CameraHead under autograd — what `scaler.scale(loss).backward()` (train/train_imc.py:380) does to
sailrecon/heads/camera_head.py:85-186, as ONE autograd node on fp32 kernels (csrc/head_train.cu + sr_gemm_f32).

The reference runs the head with autocast disabled (sail_recon.py:119-121): forward and backward are fp32.  The node takes the
T = anchors + queries camera-token rows and returns the `num_iterations` activated pose encodings; backward walks every
iteration whose output received a gradient (SailRecon only reads the last one, sail_recon.py:122; the prediction fed to the
next iteration is detached in the reference, :142-144, so the iterations are independent in the backward).

Per iteration: pose_branch (2 Linear + GELU) <- trunk_norm <- 4 trunk blocks (LayerScale, MLP, proj, masked attention over the
T rows, qkv, LayerNorms) <- adaLN-modulate-gate-residual <- poseLN_modulation (SiLU + Linear) <- embed_pose, and finally
token_norm.  Every Linear gradient streams its fp32 weight once (dX = dY W through a transposed copy) and writes its gradient
once (dW = dY^T X, a sum of T outer products): the step is bound by the 216 M parameters, not by the T rows.
"""
from __future__ import annotations

from typing import Dict, List

import torch

from .. import engine, ops


def _lin(a: torch.Tensor, lin: torch.nn.Module) -> torch.Tensor:
    out = torch.empty(a.shape[0], lin.weight.shape[0], dtype=torch.float32, device=a.device)
    return ops.gemm(a, lin.weight.detach(), None if lin.bias is None else lin.bias.detach(), out)


def _ln(x: torch.Tensor, norm: torch.nn.Module) -> torch.Tensor:
    w, b, eps = engine.ln_params(norm)
    return ops.layernorm(x, w, b, eps, torch.empty_like(x))


class _Grads:
    """name -> gradient; a parameter that is reached from several iterations is summed."""

    def __init__(self, head: torch.nn.Module):
        self.names = {id(p): n for n, p in head.named_parameters()}
        self.g: Dict[str, torch.Tensor] = {}

    def add(self, param: torch.Tensor, grad: torch.Tensor) -> None:
        n = self.names[id(param)]
        grad = grad.view(param.shape)
        if n in self.g:
            self.g[n] += grad
        else:
            self.g[n] = grad

    def buffer(self, param: torch.Tensor) -> torch.Tensor:
        """A zero-initialised (first use) accumulation buffer for kernels that ADD into their parameter gradients."""
        n = self.names[id(param)]
        if n not in self.g:
            self.g[n] = torch.zeros_like(param)
        return self.g[n]


def _linear_bwd(dY: torch.Tensor, X: torch.Tensor, lin: torch.nn.Module, grads: _Grads, need_dx: bool = True):
    """Y = X W^T + b  ->  dX = dY W,  dW = dY^T X,  db = column sums of dY."""
    W = lin.weight.detach()
    T = dY.shape[0]
    dX = None
    if need_dx:
        dX = ops.gemm(dY, ops.transpose_f32(W), None, torch.empty(T, W.shape[1], dtype=torch.float32, device=dY.device))
    dW = ops.gemm(ops.transpose_f32(dY), ops.transpose_f32(X), None, torch.empty_like(W))
    grads.add(lin.weight, dW)
    if lin.bias is not None:
        grads.add(lin.bias, ops.colsum_f32(dY))
    return dX


def _block_fwd(blk: torch.nn.Module, x: torch.Tensor, calls) -> dict:
    """Block.forward (block.py:86-112) un-fused, keeping what the backward reads."""
    attn, mlp = blk.attn, blk.mlp
    if getattr(attn, "qk_norm", False) or getattr(attn, "rope", None) is not None:
        raise NotImplementedError("sailrecon_b200: the camera-head trunk blocks have no q/k norm and no RoPE (camera_head.py:51-60)")
    s = {"x": x}
    s["a1"] = _ln(x, blk.norm1)
    s["qkv"] = _lin(s["a1"], attn.qkv)
    s["o"] = engine.attention_rows(attn, s["qkv"], torch.empty_like(x), calls)
    s["z1"] = _lin(s["o"], attn.proj)
    s["x_mid"] = ops.ls_residual_f32(x, s["z1"], engine.gamma_of(blk.ls1))
    s["a2"] = _ln(s["x_mid"], blk.norm2)
    s["u"] = _lin(s["a2"], mlp.fc1)
    s["g"] = ops.act_f32(s["u"], "gelu")
    s["z2"] = _lin(s["g"], mlp.fc2)
    s["out"] = ops.ls_residual_f32(s["x_mid"], s["z2"], engine.gamma_of(blk.ls2))
    return s


def _ls_bwd(ls: torch.nn.Module, dy: torch.Tensor, z: torch.Tensor, grads: _Grads) -> torch.Tensor:
    gamma = engine.gamma_of(ls)
    if gamma is None:
        return dy
    grads.add(ls.gamma, ops.colsum_f32(dy, z))
    return ops.scale_cols(dy, gamma)


def _ln_bwd(norm: torch.nn.Module, x: torch.Tensor, dy: torch.Tensor, grads: _Grads, dres=None) -> torch.Tensor:
    w, _, eps = engine.ln_params(norm)
    return ops.layernorm_bwd_f32(x, w, dy, eps, dres, grads.buffer(norm.weight), grads.buffer(norm.bias))


def _block_bwd(blk: torch.nn.Module, s: dict, d_out: torch.Tensor, n_anchor: int, grads: _Grads) -> torch.Tensor:
    attn, mlp = blk.attn, blk.mlp
    d_g = _linear_bwd(_ls_bwd(blk.ls2, d_out, s["z2"], grads), s["g"], mlp.fc2, grads)
    d_a2 = _linear_bwd(ops.act_bwd_f32(s["u"], d_g, "gelu"), s["a2"], mlp.fc1, grads)
    d_mid = _ln_bwd(blk.norm2, s["x_mid"], d_a2, grads, dres=d_out)
    d_o = _linear_bwd(_ls_bwd(blk.ls1, d_mid, s["z1"], grads), s["o"], attn.proj, grads)
    d_qkv = ops.attn_small_bwd_f32(s["qkv"], d_o, attn.num_heads, n_anchor, float(attn.scale))
    d_a1 = _linear_bwd(d_qkv, s["a1"], attn.qkv, grads)
    return _ln_bwd(blk.norm1, s["x"], d_a1, grads, dres=d_mid)


class CameraHeadFunction(torch.autograd.Function):
    """forward(head, num_iterations, n_anchor, x_in [T, C], *parameters) -> num_iterations tensors [T, 9]."""

    @staticmethod
    def forward(ctx, head, num_iterations: int, n_anchor: int, x_in: torch.Tensor, *params):
        T, C = x_in.shape
        x_in = x_in.contiguous()
        x = _ln(x_in, head.token_norm)                                                     # camera_head.py:110
        calls = engine.MaskStructure(n_anchor, 1, T - n_anchor).calls() if T > n_anchor else engine.full_segments(1, T)
        empty = head.empty_pose_tokens.detach()[0].expand(T, -1).contiguous()
        pred = None
        saved: List[dict] = []
        acts = []
        for it in range(num_iterations):                                                   # trunk_fn :143-184
            s = {"inp": empty if it == 0 else pred}
            s["e_pre"] = _lin(s["inp"], head.embed_pose)
            s["emb"] = ops.act_f32(s["e_pre"], "silu")
            s["mod"] = _lin(s["emb"], head.poseLN_modulation[1])
            y = ops.adaln_modulate(x, s["mod"], float(head.adaln_norm.eps), torch.empty_like(x))
            s["blocks"] = []
            for blk in head.trunk:
                b = _block_fwd(blk, y, calls)
                s["blocks"].append(b)
                y = b["out"]
            s["y"] = y
            s["yn"] = _ln(y, head.trunk_norm)
            s["hu"] = _lin(s["yn"], head.pose_branch.fc1)
            s["hg"] = ops.act_f32(s["hu"], "gelu")
            delta = _lin(s["hg"], head.pose_branch.fc2)
            new_pred = torch.empty_like(delta)
            act = torch.empty_like(delta)
            if it == 0:
                new_pred.copy_(delta)                       # pose_update accumulates in place: keep every iteration's value
            else:
                new_pred.copy_(pred)
            ops.pose_update(new_pred, delta, act, it == 0)                                  # :172-184
            pred = new_pred
            s["pred"] = pred
            saved.append(s)
            acts.append(act)
        ctx.head, ctx.n_anchor, ctx.saved, ctx.x_in, ctx.x = head, n_anchor, saved, x_in, x
        ctx.set_materialize_grads(False)
        return tuple(acts)

    @staticmethod
    def backward(ctx, *d_acts):
        head, saved, x = ctx.head, ctx.saved, ctx.x
        grads = _Grads(head)
        dx = torch.zeros_like(x)
        touched = False
        for it, d_act in enumerate(d_acts):
            if d_act is None:
                continue
            s = saved[it]
            d_pred = ops.pose_act_bwd(s["pred"], d_act.contiguous().float())
            d_hg = _linear_bwd(d_pred, s["hg"], head.pose_branch.fc2, grads)
            d_yn = _linear_bwd(ops.act_bwd_f32(s["hu"], d_hg, "gelu"), s["yn"], head.pose_branch.fc1, grads)
            d_y = _ln_bwd(head.trunk_norm, s["y"], d_yn, grads)
            for blk, b in zip(reversed(list(head.trunk)), reversed(s["blocks"])):
                d_y = _block_bwd(blk, b, d_y, ctx.n_anchor, grads)
            d_mod = ops.adaln_modulate_bwd(x, s["mod"], d_y, float(head.adaln_norm.eps), dx, accumulate_dx=touched)
            touched = True
            d_emb = _linear_bwd(d_mod, s["emb"], head.poseLN_modulation[1], grads)
            d_inp = _linear_bwd(ops.act_bwd_f32(s["e_pre"], d_emb, "silu"), s["inp"], head.embed_pose, grads, need_dx=(it == 0))
            if it == 0:                                     # the learned empty pose, broadcast over the rows (:145-146)
                grads.add(head.empty_pose_tokens, ops.colsum_f32(d_inp))
        d_x_in = None
        if touched:
            d_x_in = _ln_bwd(head.token_norm, ctx.x_in, dx, grads)
        out = [grads.g.get(n) for n, _ in head.named_parameters()]
        ctx.saved = None
        return (None, None, None, d_x_in, *out)


def forward_autograd(head, aggregated_tokens_list, cam_token_last_layer: torch.Tensor, num_iterations: int = 4) -> list:
    """CameraHead.forward (camera_head.py:85-121) in train() mode with autograd enabled."""
    tokens = aggregated_tokens_list[-1]
    pose_q = tokens[:, :, 0]
    if pose_q.shape[0] != 1:
        raise NotImplementedError("sailrecon_b200: batch size 1")
    if not pose_q.is_cuda:
        raise RuntimeError("sailrecon_b200: CameraHead needs CUDA tensors (no CPU fallback)")
    n_anchor = cam_token_last_layer.shape[1]
    x_in = torch.cat([cam_token_last_layer[0], pose_q[0]], dim=0).float()                  # :105-109, [T, C]
    if x_in.shape[0] > 128:
        raise NotImplementedError("sailrecon_b200: the differentiated camera head handles at most 128 camera tokens")
    params = [p for _, p in head.named_parameters()]
    acts = CameraHeadFunction.apply(head, num_iterations, n_anchor, x_in, *params)
    return [a[None, n_anchor:] for a in acts]                                              # :118-121

"""CameraHead — API mirror of sailrecon/heads/camera_head.py (CameraHead :19-186, modulate :189-194,
build_lr_mask :197-228) on the B200 kernel library.

Runs in fp32, as the reference does (SailRecon disables autocast around the heads, sail_recon.py:119-121).  The
boolean mask of build_lr_mask is consumed structurally: anchors attend to anchors; every query camera token attends
to the anchors and to itself — two launches of the structured attention kernel instead of a dense masked SDPA.
Per refinement iteration: embed_pose+SiLU (one GEMM epilogue), the 2048->6144 modulation GEMM, one fused
adaLN-modulate-gate-residual kernel, 4 trunk blocks (7 launches each), trunk_norm, pose_branch (2 GEMMs) and one
accumulate+activate kernel.
In train() mode with autograd enabled `forward` is one autograd node with an fp32 kernel backward (camera_head_train.py).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import engine, ops
from .._lib import SR_EPI_BIAS, SR_EPI_BIAS_GELU, SR_EPI_BIAS_SILU
from ..layers import Mlp
from ..layers.block import Block


class CameraHead(nn.Module):
    def __init__(self, dim_in: int = 2048, trunk_depth: int = 4, pose_encoding_type: str = "absT_quaR_FoV",
                 num_heads: int = 16, mlp_ratio: int = 4, init_values: float = 0.01, trans_act: str = "linear",
                 quat_act: str = "linear", fl_act: str = "relu"):
        super().__init__()
        if pose_encoding_type == "absT_quaR_FoV":
            self.target_dim = 9
        else:
            raise ValueError(f"Unsupported camera encoding type: {pose_encoding_type}")    # camera_head.py:43
        if (trans_act, quat_act, fl_act) != ("linear", "linear", "relu"):
            raise NotImplementedError("sailrecon_b200: CameraHead implements the activations SAIL-Recon configures "
                                      "(trans/quat 'linear', FoV 'relu')")
        self.trans_act, self.quat_act, self.fl_act = trans_act, quat_act, fl_act
        self.trunk_depth = trunk_depth
        self.trunk = nn.Sequential(*[Block(dim=dim_in, num_heads=num_heads, mlp_ratio=mlp_ratio, init_values=init_values)
                                     for _ in range(trunk_depth)])
        self.token_norm = nn.LayerNorm(dim_in)
        self.trunk_norm = nn.LayerNorm(dim_in)
        self.empty_pose_tokens = nn.Parameter(torch.zeros(1, 1, self.target_dim))
        self.embed_pose = nn.Linear(self.target_dim, dim_in)
        self.poseLN_modulation = nn.Sequential(nn.SiLU(), nn.Linear(dim_in, 3 * dim_in, bias=True))
        self.adaln_norm = nn.LayerNorm(dim_in, elementwise_affine=False, eps=1e-6)
        self.pose_branch = Mlp(in_features=dim_in, hidden_features=dim_in // 2, out_features=self.target_dim, drop=0)

    def forward(self, aggregated_tokens_list, cam_token_last_layer: torch.Tensor, num_iterations: int = 4) -> list:
        if self.training and torch.is_grad_enabled():          # train_imc.py:361-381: one autograd node, fp32 backward kernels
            from .camera_head_train import forward_autograd
            return forward_autograd(self, aggregated_tokens_list, cam_token_last_layer, num_iterations)
        tokens = aggregated_tokens_list[-1]                                                # camera_head.py:103-109
        pose_q = tokens[:, :, 0]
        if pose_q.shape[0] != 1:
            raise NotImplementedError("sailrecon_b200: batch size 1")
        engine.require_cuda_f32(pose_q, "aggregated tokens")
        engine.require_cuda_f32(cam_token_last_layer, "cam_token_last_layer")
        Na = cam_token_last_layer.shape[1]
        x_in = torch.cat([cam_token_last_layer[0], pose_q[0]], dim=0).contiguous()          # [T, C]
        w, b, eps = engine.ln_params(self.token_norm)
        x = torch.empty_like(x_in)
        ops.layernorm(x_in, w, b, eps, x)                                                  # :110
        # ~build_lr_mask (:112-116, :197-228) as structure: anchors | each query behind the anchors
        calls = engine.MaskStructure(Na, 1, pose_q.shape[1]).calls()
        return [p[None, Na:] for p in self._trunk_rows(x, calls, num_iterations)]          # query rows only, :118-121

    # ---------------------------------------------------------------------------------------------- cached anchors
    def prime_anchors(self, cam_token_last_layer: torch.Tensor, num_iterations: int = 4) -> None:
        """Localisation fast path (localize.Localizer).  build_lr_mask (camera_head.py:197-228) lets anchor tokens attend to
        anchors only, so the anchors' rows of the trunk do not depend on the query frames: they are run ONCE here and the
        keys / values they present to the queries are kept per (iteration, trunk block).  `forward_queries` then runs the
        query rows alone (M = Nq rows instead of Na + Nq: at one query image per call the head becomes a stream of GEMVs
        over its fp32 weights) and returns what `forward` returns."""
        engine.forbid_autograd(self, "CameraHead.prime_anchors")
        engine.require_cuda_f32(cam_token_last_layer, "cam_token_last_layer")
        x_in = cam_token_last_layer[0].contiguous()
        w, b, eps = engine.ln_params(self.token_norm)
        x = torch.empty_like(x_in)
        ops.layernorm(x_in, w, b, eps, x)
        kv: list = []
        self._trunk_rows(x, engine.full_segments(1, x.shape[0]), num_iterations, capture_kv=kv)
        n = len(self.trunk)
        self._anchor_kv = [kv[i * n:(i + 1) * n] for i in range(num_iterations)]
        self._anchor_rows = x.shape[0]

    def forward_queries(self, aggregated_tokens_list, num_iterations: int = 4) -> list:
        """== forward(aggregated_tokens_list, cam_token_last_layer) for the anchors given to prime_anchors."""
        engine.forbid_autograd(self, "CameraHead.forward_queries")
        kv = getattr(self, "_anchor_kv", None)
        if kv is None or len(kv) != num_iterations:
            raise RuntimeError("sailrecon_b200: CameraHead.forward_queries needs prime_anchors(cam_token_last_layer, num_iterations) first")
        pose_q = aggregated_tokens_list[-1][:, :, 0]
        if pose_q.shape[0] != 1:
            raise NotImplementedError("sailrecon_b200: batch size 1")
        engine.require_cuda_f32(pose_q, "aggregated tokens")
        x_in = pose_q[0].contiguous()
        Nq, Na = x_in.shape[0], self._anchor_rows
        w, b, eps = engine.ln_params(self.token_norm)
        x = torch.empty_like(x_in)
        ops.layernorm(x_in, w, b, eps, x)

        def calls_for(it, bi):
            k, v = kv[it][bi]
            return [engine.AttnCall(0, Nq, 1, 1, engine.Prefix(k, v, 0, Na))]

        return [p[None] for p in self._trunk_rows(x, calls_for, num_iterations)]

    def _trunk_rows(self, x: torch.Tensor, calls, num_iterations: int, capture_kv=None) -> list:
        """trunk_fn (camera_head.py:123-186) on the normalised camera-token rows x [T, C] (fp32, autocast off): per
        refinement iteration embed_pose + SiLU (one GEMM epilogue), the modulation GEMM, one fused adaLN-modulate-gate-
        residual kernel, the trunk blocks, trunk_norm, pose_branch and one accumulate + activate kernel.  -> [T, 9] each."""
        T, C = x.shape
        mode = "f32"
        f32 = dict(dtype=torch.float32, device=x.device)
        emb = torch.empty(T, C, **f32)
        mod = torch.empty(T, 3 * C, **f32)
        y = torch.empty(T, C, **f32)
        yn = torch.empty(T, C, **f32)
        hid = torch.empty(T, self.pose_branch.fc1.weight.shape[0], **f32)
        delta = torch.empty(T, self.target_dim, **f32)
        pred = torch.empty(T, self.target_dim, **f32)
        empty = self.empty_pose_tokens.detach()[0].expand(T, -1).contiguous()
        tw, tb, teps = engine.ln_params(self.trunk_norm)
        out = []
        for it in range(num_iterations):                                                   # trunk_fn :143-184
            inp = empty if it == 0 else pred
            engine.linear_rows(self.embed_pose, inp, emb, mode, SR_EPI_BIAS_SILU)          # embed_pose + the SiLU of :72-74
            engine.linear_rows(self.poseLN_modulation[1], emb, mod, mode, SR_EPI_BIAS)     # shift | scale | gate
            ops.adaln_modulate(x, mod, float(self.adaln_norm.eps), y)                      # :157-161
            for bi, blk in enumerate(self.trunk):
                engine.run_block(blk, y, calls(it, bi) if callable(calls) else calls, None, mode, tag="cam.",
                                 capture_kv=capture_kv)                                    # :163-166
            ops.layernorm(y, tw, tb, teps, yn)
            engine.linear_rows(self.pose_branch.fc1, yn, hid, mode, SR_EPI_BIAS_GELU)      # :168-170
            engine.linear_rows(self.pose_branch.fc2, hid, delta, mode, SR_EPI_BIAS)
            act = torch.empty(T, self.target_dim, **f32)
            ops.pose_update(pred, delta, act, it == 0)                                     # :172-184
            out.append(act)
        return out

    def trunk_fn(self, pose_tokens: torch.Tensor, attention_mask: torch.Tensor, num_iterations: int) -> list:
        """camera_head.py:123-186: pose_tokens [1, S, C] already token_norm'ed; attention_mask as build_lr_mask returns it
        (True = masked; the blocks receive ~mask).  Returns `num_iterations` activated encodings [1, S, 9] (all rows)."""
        engine.forbid_autograd(self, "CameraHead.trunk_fn")
        if pose_tokens.dim() != 3 or pose_tokens.shape[0] != 1:
            raise NotImplementedError("sailrecon_b200: batch size 1")
        engine.require_cuda_f32(pose_tokens, "pose_tokens")
        S = pose_tokens.shape[1]
        calls = engine.full_segments(1, S) if attention_mask is None else engine.structure_from_mask(~attention_mask, S).calls()
        return [p[None] for p in self._trunk_rows(pose_tokens[0].contiguous(), calls, num_iterations)]


def modulate(x: torch.Tensor, shift: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """camera_head.py:189-194: x * (1 + scale) + shift (stand-alone form; inside CameraHead.forward it is fused with the
    LayerNorm, the gate and the residual in sr_adaln_modulate_f32)."""
    if not x.is_cuda:
        raise RuntimeError("sailrecon_b200: modulate runs on a CUDA device only (no CPU fallback)")
    return ops.modulate(x.float(), shift.float(), scale.float())


def build_lr_mask(S: int, no_reloc_list, device="cpu"):
    """camera_head.py:197-228 (True = masked).  Kept for API parity; the head itself uses the structure directly."""
    r = [i for i in range(S) if i not in no_reloc_list]
    mask = torch.zeros(S, S, dtype=torch.bool, device=device)
    l_idx = torch.as_tensor(sorted(set(no_reloc_list)), dtype=torch.long, device=device)
    r_idx = torch.tensor(r, dtype=torch.long, device=device)
    if l_idx.numel() and r_idx.numel():
        mask[l_idx[:, None], r_idx[None, :]] = True
    if r_idx.numel() > 1:
        mask[r_idx[:, None], r_idx[None, :]] = True
        mask[r_idx, r_idx] = False
    return mask.unsqueeze(0).unsqueeze(0)

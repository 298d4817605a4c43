"""Block — API mirror of sailrecon/layers/block.py (Block :26-112, NestedTensorBlock :237-290, tensor path).

x + ls1(attn(norm1(x), pos, mask)); x + ls2(mlp(norm2(x))) as 7 launches of the C-ABI library (engine.run_block).
Stochastic depth (block.py:93-108) is inactive on this path (drop_path = 0 everywhere) and is not implemented.
"""
from __future__ import annotations

from typing import Callable

import torch
import torch.nn as nn
from torch import Tensor

from .. import engine, ops
from .attention import Attention, MemEffAttention
from .layer_scale import LayerScale
from .mlp import Mlp


class Block(nn.Module):
    def __init__(self, dim: int, num_heads: int, mlp_ratio: float = 4.0, qkv_bias: bool = True, proj_bias: bool = True,
                 ffn_bias: bool = True, drop: float = 0.0, attn_drop: float = 0.0, init_values=None, drop_path: float = 0.0,
                 act_layer: Callable[..., nn.Module] = nn.GELU, norm_layer: Callable[..., nn.Module] = nn.LayerNorm,
                 attn_class: Callable[..., nn.Module] = Attention, ffn_layer: Callable[..., nn.Module] = Mlp,
                 qk_norm: bool = False, fused_attn: bool = True, rope=None, kv_cache: bool = False) -> None:
        super().__init__()
        if drop_path != 0.0 or drop != 0.0 or attn_drop != 0.0:
            raise NotImplementedError("sailrecon_b200: dropout / stochastic depth are 0 on the SAIL-Recon path")
        self.norm1 = norm_layer(dim)
        self.attn = attn_class(dim, num_heads=num_heads, qkv_bias=qkv_bias, proj_bias=proj_bias, attn_drop=attn_drop,
                               proj_drop=drop, qk_norm=qk_norm, fused_attn=fused_attn, rope=rope, kv_cache=kv_cache)
        self.ls1 = LayerScale(dim, init_values=init_values) if init_values else nn.Identity()
        self.drop_path1 = nn.Identity()
        self.norm2 = norm_layer(dim)
        self.mlp = ffn_layer(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop=drop,
                             bias=ffn_bias)
        self.ls2 = LayerScale(dim, init_values=init_values) if init_values else nn.Identity()
        self.drop_path2 = nn.Identity()
        self.sample_drop_ratio = drop_path

    def forward(self, x: Tensor, pos=None, attn_mask=None) -> Tensor:
        """x [Bx, N, C] fp32 -> fp32 [Bx, N, C] (a fresh tensor, like the reference)."""
        engine.forbid_autograd(self, "Block.forward")
        engine.require_cuda_f32(x, "x")
        mode = engine.compute_mode()
        Bx, N, C = x.shape
        rows = x.reshape(Bx * N, C).clone()
        attn = self.attn
        priming = attn.kv_cache and attn.k_cache is None
        rows_pos = attn._rows_pos(pos, Bx * N)
        if priming:
            # attention.py:94-100: store K/V of these tokens, the attention branch contributes zeros
            act = torch.bfloat16 if mode == "bf16" else torch.float32
            a = engine.WS.get("a", (Bx * N, C), act, x.device)
            w1, b1, eps1 = engine.ln_params(self.norm1)
            ops.layernorm(rows, w1, b1, eps1, a)
            qkv = engine.WS.get("qkv", (Bx * N, 3 * C), act, x.device)
            engine.qkv_rows(attn, a, qkv, mode, rows_pos)
            attn.k_cache = qkv[:, C:2 * C].clone()
            attn.v_cache = qkv[:, 2 * C:].clone()
            engine.run_block(self, rows, [], rows_pos, mode, attn_zero=True)
        else:
            cache_len = attn.k_cache.shape[0] if (attn.kv_cache and attn.k_cache is not None) else 0
            engine.run_block(self, rows, attn._calls(Bx, N, attn_mask, cache_len), rows_pos, mode)
        return rows.view(Bx, N, C)


class NestedTensorBlock(Block):
    """DINOv2's block type (block.py:237-290).  Only the tensor path exists on this code base: the nested-list path
    needs xformers, which the reference disables (block.py:23)."""

    def forward(self, x_or_x_list):
        if isinstance(x_or_x_list, Tensor):
            return super().forward(x_or_x_list)
        if isinstance(x_or_x_list, list):
            raise AssertionError("xFormers is required for using nested tensors")
        raise AssertionError

"""Attention — API mirror of sailrecon/layers/attention.py (Attention :21-122, MemEffAttention :125-143).

Same constructor, parameter names (qkv, proj, q_norm, k_norm) and ``forward(x, pos=None, attn_mask=None)``.
Differences that follow from the B200 design (documented in INTEGRATION.md):
  * the boolean mask is consumed structurally (engine.structure_from_mask) — SAIL-Recon only ever builds
    [shared prefix | block diagonal] masks;
  * the kv cache (attention.py:43-46,84-100) stays RESIDENT IN HBM as bf16/fp32 row matrices [N, H*d] instead of being
    moved to host memory and back on every call.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
from torch import Tensor

from .. import engine, ops
from .._lib import SR_EPI_BIAS


class Attention(nn.Module):
    def __init__(self, dim: int, num_heads: int = 8, qkv_bias: bool = True, proj_bias: bool = True, attn_drop: float = 0.0,
                 proj_drop: float = 0.0, norm_layer: nn.Module = nn.LayerNorm, qk_norm: bool = False,
                 fused_attn: bool = True, rope=None, kv_cache: bool = False) -> None:
        super().__init__()
        assert dim % num_heads == 0, "dim should be divisible by num_heads"        # attention.py:37
        if attn_drop != 0.0 or proj_drop != 0.0:
            raise NotImplementedError("sailrecon_b200: dropout is 0 everywhere on the SAIL-Recon path")
        self.num_heads = num_heads
        self.head_dim = dim // num_heads
        self.scale = self.head_dim ** -0.5
        self.fused_attn = fused_attn
        self.kv_cache = kv_cache
        self.k_cache: Optional[Tensor] = None
        self.v_cache: Optional[Tensor] = None
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.q_norm = norm_layer(self.head_dim) if qk_norm else nn.Identity()
        self.k_norm = norm_layer(self.head_dim) if qk_norm else nn.Identity()
        self.attn_drop = nn.Dropout(attn_drop)
        self.proj = nn.Linear(dim, dim, bias=proj_bias)
        self.proj_drop = nn.Dropout(proj_drop)
        self.rope = rope

    def clear_kv_cache(self) -> None:
        """attention.py:56-68 (the reference also calls torch.cuda.empty_cache(); the allocator is left alone here)."""
        self.k_cache = None
        self.v_cache = None

    # ---------------------------------------------------------------------------------------------- helpers
    def _rows_pos(self, pos: Optional[Tensor], n_rows: int) -> Optional[engine.RowPositions]:
        if pos is None or self.rope is None:
            return None
        if isinstance(pos, engine.RowPositions):
            return pos
        p32 = pos.reshape(-1, 2).to(torch.int32).contiguous()
        return engine.RowPositions(p32, int(pos.max()) + 1)                      # rope.py:190 does this sync on every call

    def _calls(self, Bx: int, N: int, attn_mask, cache_len: int):
        """Attention structure of a [Bx, N] call: independent sequences, optionally under a structured mask and
        optionally behind cached prefix keys."""
        if attn_mask is None:
            if cache_len:
                return [engine.AttnCall(0, Bx, N, N, engine.Prefix(self.k_cache, self.v_cache, 0, cache_len))]
            return engine.full_segments(Bx, N)
        if isinstance(attn_mask, engine.MaskStructure):
            st = attn_mask
            if cache_len:                            # the structure describes the NEW tokens; all of them see the cache
                if st.prefix_len:
                    raise NotImplementedError("sailrecon_b200: a MaskStructure with its own prefix cannot be combined with a kv cache")
                return [engine.AttnCall(0, st.num_segs, st.seg_len, st.seg_len,
                                        engine.Prefix(self.k_cache, self.v_cache, 0, cache_len))]
        else:
            if Bx != 1:
                raise NotImplementedError("sailrecon_b200: masks are supported for batch 1 (as the reference uses them)")
            if cache_len:
                # forward_with_cache mask (aggregator.py:496-508): rows = the new tokens, cols = [cache ; new tokens]
                m = attn_mask
                while m.dim() > 2:
                    m = m[0]
                if m.shape[-1] != cache_len + N or not bool(m[:, :cache_len].all()):
                    raise NotImplementedError("sailrecon_b200: unsupported mask for cached attention")
                sub = engine.structure_from_mask(m[:, cache_len:], N)
                if sub.prefix_len == N and sub.num_segs == 0:
                    sub = engine.MaskStructure(0, N, 1)          # one new sequence (S = 1): the all-True sub-mask is ONE segment
                if sub.prefix_len:
                    raise NotImplementedError("sailrecon_b200: unsupported mask for cached attention")
                st = engine.MaskStructure(0, sub.seg_len, sub.num_segs)
                return [engine.AttnCall(0, st.num_segs, st.seg_len, st.seg_len,
                                        engine.Prefix(self.k_cache, self.v_cache, 0, cache_len))]
            st = engine.structure_from_mask(attn_mask, N)
        return st.calls()

    # ---------------------------------------------------------------------------------------------- forward
    def forward(self, x: Tensor, pos=None, attn_mask=None) -> Tensor:
        """x [Bx, N, C] (already normalised by the caller, block.py:87) -> proj(attention(x)) [Bx, N, C].
        Under autocast(bf16) the result is bf16, as the reference's proj Linear returns it."""
        engine.forbid_autograd(self, "Attention.forward")
        mode = engine.compute_mode()
        if not x.is_cuda:
            raise RuntimeError("sailrecon_b200: Attention runs on a CUDA device only (no CPU fallback)")
        Bx, N, C = x.shape
        T = Bx * N
        act = torch.bfloat16 if mode == "bf16" else torch.float32
        rows = x.reshape(T, C)
        if rows.dtype == torch.float32 and mode == "bf16":
            a = ops.cast_bf16(rows.contiguous())
        elif rows.dtype == act:
            a = rows.contiguous()
        else:
            raise TypeError(f"sailrecon_b200: Attention input must be fp32 (or bf16 under autocast), got {x.dtype}")
        qkv = engine.WS.get("attn.qkv", (T, 3 * C), act, x.device)
        engine.qkv_rows(self, a, qkv, mode, self._rows_pos(pos, T))
        if self.kv_cache and self.k_cache is None:
            # priming pass (attention.py:94-100): keep post-norm / post-RoPE K, V on the device, return zeros
            self.k_cache = qkv[:, C:2 * C].clone()
            self.v_cache = qkv[:, 2 * C:].clone()
            return torch.zeros_like(x)
        cache_len = self.k_cache.shape[0] if (self.kv_cache and self.k_cache is not None) else 0
        if cache_len and self.k_cache.dtype != act:
            raise RuntimeError("sailrecon_b200: kv cache was filled in a different precision mode")
        o = torch.empty(T, C, dtype=act, device=x.device)
        engine.attention_rows(self, qkv, o, self._calls(Bx, N, attn_mask, cache_len))
        y = torch.empty(T, C, dtype=act, device=x.device)
        engine.linear_rows(self.proj, o, y, mode, SR_EPI_BIAS)
        return y.view(Bx, N, C)


class MemEffAttention(Attention):
    """attention.py:125-143: with xformers disabled (`XFORMERS_AVAILABLE = False`, :18) this is Attention.forward(x)."""

    def forward(self, x: Tensor, attn_bias=None, pos=None) -> Tensor:
        assert pos is None
        if attn_bias is not None:
            raise AssertionError("xFormers is required for using nested tensors")
        return super().forward(x)

"""DINOv2 ViT — API mirror of sailrecon/layers/vision_transformer.py (DinoVisionTransformer :53-376, vit_* :395-448)
for the way the aggregator uses it as its patch embed (models/aggregator.py:196-240,271-274): tensor input,
`forward(x)["x_norm_patchtokens"]`, block_chunks = 0.  Parameter names match the reference state_dict
(`patch_embed.proj.*`, `cls_token`, `pos_embed`, `register_tokens`, `mask_token`, `blocks.{i}.*`, `norm.*`).

All S frames run as ONE row buffer [S*(1 + R + h*w), D]: the 24 blocks are engine.run_block calls with batch = S
segments (no RoPE, no QK-norm, LayerNorm eps 1e-6, LayerScale gamma init 1).
"""
from __future__ import annotations

import math
from functools import partial
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn
from torch import Tensor

from .. import engine, ops
from .block import Block, NestedTensorBlock
from .attention import MemEffAttention
from .mlp import Mlp
from .patch_embed import IDENTITY_MEAN, IDENTITY_STD, PatchEmbed


class DinoVisionTransformer(nn.Module):
    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768, depth=12, num_heads=12, mlp_ratio=4.0,
                 qkv_bias=True, ffn_bias=True, proj_bias=True, drop_path_rate=0.0, drop_path_uniform=False,
                 init_values=None, embed_layer=PatchEmbed, act_layer=nn.GELU, block_fn=Block, ffn_layer="mlp",
                 block_chunks=1, num_register_tokens=0, interpolate_antialias=False, interpolate_offset=0.1,
                 qk_norm=False):
        super().__init__()
        if drop_path_rate != 0.0:
            raise NotImplementedError("sailrecon_b200: stochastic depth is 0 on the SAIL-Recon path")
        if ffn_layer != "mlp":
            raise NotImplementedError("sailrecon_b200: only the 'mlp' FFN is used by SAIL-Recon (vit_large)")
        if block_chunks != 0:
            raise NotImplementedError("sailrecon_b200: block_chunks must be 0 (aggregator.py:204)")
        norm_layer = partial(nn.LayerNorm, eps=1e-6)
        self.num_features = self.embed_dim = embed_dim
        self.num_tokens = 1
        self.n_blocks = depth
        self.num_heads = num_heads
        self.patch_size = patch_size
        self.num_register_tokens = num_register_tokens
        self.interpolate_antialias = interpolate_antialias
        self.interpolate_offset = interpolate_offset
        self.patch_embed = embed_layer(img_size=img_size, patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim)
        num_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, num_patches + self.num_tokens, embed_dim))
        assert num_register_tokens >= 0
        self.register_tokens = nn.Parameter(torch.zeros(1, num_register_tokens, embed_dim)) if num_register_tokens else None
        self.chunked_blocks = False
        self.blocks = nn.ModuleList([
            block_fn(dim=embed_dim, num_heads=num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, proj_bias=proj_bias,
                     ffn_bias=ffn_bias, drop_path=0.0, norm_layer=norm_layer, act_layer=act_layer, ffn_layer=Mlp,
                     init_values=init_values, qk_norm=qk_norm)
            for _ in range(depth)])
        self.norm = norm_layer(embed_dim)
        self.head = nn.Identity()
        self.mask_token = nn.Parameter(torch.zeros(1, embed_dim))
        self._pos_cache: Dict[Tuple, Tensor] = {}
        self.init_weights()

    def init_weights(self):
        nn.init.trunc_normal_(self.pos_embed, std=0.02)
        nn.init.normal_(self.cls_token, std=1e-6)
        if self.register_tokens is not None:
            nn.init.normal_(self.register_tokens, std=1e-6)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    # ---------------------------------------------------------------------------------------------- position embedding
    def interpolated_pos_embed(self, H: int, W: int) -> Tensor:
        """vision_transformer.py:206-240 — a function of the parameter and (H, W) only, so it is evaluated once per
        image size and cached (bicubic resampling of a [1, D, 37, 37] parameter grid; not part of the per-frame path)."""
        pe = self.pos_embed
        key = (pe.data_ptr(), pe._version, pe.device, H, W)
        hit = self._pos_cache.get(key)
        if hit is not None:
            return hit
        with torch.no_grad():
            N = pe.shape[1] - 1
            h0, w0 = H // self.patch_size, W // self.patch_size
            if h0 * w0 == N and H == W:
                out = pe.detach().float()[0].contiguous()
            else:
                M = int(math.sqrt(N))
                assert N == M * M
                kwargs = {}
                if self.interpolate_offset:
                    kwargs["scale_factor"] = (float(h0 + self.interpolate_offset) / M, float(w0 + self.interpolate_offset) / M)
                else:
                    kwargs["size"] = (h0, w0)
                grid = nn.functional.interpolate(pe.detach().float()[:, 1:].reshape(1, M, M, -1).permute(0, 3, 1, 2),
                                                 mode="bicubic", antialias=self.interpolate_antialias, **kwargs)
                assert (h0, w0) == tuple(grid.shape[-2:])
                out = torch.cat([pe.detach().float()[0, :1], grid.permute(0, 2, 3, 1).reshape(h0 * w0, -1)], dim=0).contiguous()
        self._pos_cache = {key: out}
        return out

    # ---------------------------------------------------------------------------------------------- forward
    def forward_rows(self, images: Tensor, mode: str, mean=IDENTITY_MEAN, std=IDENTITY_STD,
                     out: Optional[Tensor] = None) -> Tuple[Tensor, int]:
        """images fp32 [S, 3, H, W] -> (LayerNorm-ed token rows fp32 [S*Pd, D], Pd) with Pd = 1 + R + h*w.
        Row s*Pd is the cls token, the next R rows the registers, then the patch tokens (vision_transformer.py:242-307)."""
        S, _, H, W = images.shape
        D, R = self.embed_dim, self.num_register_tokens
        hw = (H // self.patch_size) * (W // self.patch_size)
        Pd = 1 + R + hw
        dev = images.device
        pe = self.interpolated_pos_embed(H, W)                                            # [1 + hw, D]
        x = engine.WS.get("dino.x", (S * Pd, D), torch.float32, dev)
        # patch rows: x[s, 1+R:] = proj(patches) + pos_embed[1:]   (conv output is bf16 under autocast, the sum fp32)
        self.patch_embed.project_rows(images, mode, mean, std, out=x, resid=pe[1:], frame_rows=Pd)
        xv = x.view(S, Pd, D)
        with torch.no_grad():
            xv[:, 0] = self.cls_token.detach()[0, 0] + pe[0]                              # vision_transformer.py:250-251
            if R:
                xv[:, 1:1 + R] = self.register_tokens.detach()[0]                         # :253-257 (no pos embed)
        calls = engine.full_segments(S, Pd)
        for blk in self.blocks:
            engine.run_block(blk, x, calls, None, mode, tag="dino.")
        w, b, eps = engine.ln_params(self.norm)
        y = out if out is not None else torch.empty(S * Pd, D, dtype=torch.float32, device=dev)
        ops.layernorm(x, w, b, eps, y)
        return y, Pd

    def forward_features(self, x: Tensor, masks=None) -> dict:
        if masks is not None or isinstance(x, list):
            raise NotImplementedError("sailrecon_b200: masked / list inputs are not used by SAIL-Recon")
        engine.require_cuda_f32(x, "x")
        mode = engine.compute_mode()
        S = x.shape[0]
        y, Pd = self.forward_rows(x, mode)
        y = y.view(S, Pd, self.embed_dim)
        R = self.num_register_tokens
        return {"x_norm_clstoken": y[:, 0], "x_norm_regtokens": y[:, 1:R + 1], "x_norm_patchtokens": y[:, R + 1:],
                "x_prenorm": None, "masks": masks}

    def forward(self, *args, is_training=True, **kwargs):
        ret = self.forward_features(*args, **kwargs)
        if is_training:
            return ret
        return self.head(ret["x_norm_clstoken"])


def _vit(embed_dim, depth, num_heads, patch_size=16, num_register_tokens=0, **kwargs):
    return DinoVisionTransformer(patch_size=patch_size, embed_dim=embed_dim, depth=depth, num_heads=num_heads, mlp_ratio=4,
                                 block_fn=partial(NestedTensorBlock, attn_class=MemEffAttention),
                                 num_register_tokens=num_register_tokens, **kwargs)


def vit_small(patch_size=16, num_register_tokens=0, **kwargs):
    return _vit(384, 12, 6, patch_size, num_register_tokens, **kwargs)


def vit_base(patch_size=16, num_register_tokens=0, **kwargs):
    return _vit(768, 12, 12, patch_size, num_register_tokens, **kwargs)


def vit_large(patch_size=16, num_register_tokens=0, **kwargs):
    return _vit(1024, 24, 16, patch_size, num_register_tokens, **kwargs)


def vit_giant2(patch_size=16, num_register_tokens=0, **kwargs):
    raise NotImplementedError("sailrecon_b200: vit_giant2 uses a SwiGLU FFN, which SAIL-Recon never instantiates")

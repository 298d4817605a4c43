"""PatchEmbed — API mirror of sailrecon/layers/patch_embed.py:25-84 (2D image -> patch embedding, conv k = s = patch).

The stride-patch convolution is an im2col + GEMM here: csrc/misc.cu (sr_patchify) writes normalised patch rows and the
projection runs on the same tcgen05 GEMM kernel as every other Linear.  ``proj`` stays an nn.Conv2d so the
state_dict keys (`proj.weight [D, 3, p, p]`, `proj.bias`) match the reference checkpoint.
"""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch
import torch.nn as nn
from torch import Tensor

from .. import engine, ops
from .._lib import SR_EPI_BIAS

IDENTITY_MEAN = (0.0, 0.0, 0.0)
IDENTITY_STD = (1.0, 1.0, 1.0)


def make_2tuple(x):
    if isinstance(x, tuple):
        assert len(x) == 2
        return x
    assert isinstance(x, int)
    return (x, x)


class PatchEmbed(nn.Module):
    def __init__(self, img_size: Union[int, Tuple[int, int]] = 224, patch_size: Union[int, Tuple[int, int]] = 16,
                 in_chans: int = 3, embed_dim: int = 768, norm_layer=None, flatten_embedding: bool = True) -> None:
        super().__init__()
        image_HW = make_2tuple(img_size)
        patch_HW = make_2tuple(patch_size)
        if in_chans != 3 or patch_HW[0] != patch_HW[1]:
            raise NotImplementedError("sailrecon_b200: PatchEmbed implements 3-channel images and square patches")
        if norm_layer is not None:
            raise NotImplementedError("sailrecon_b200: PatchEmbed norm_layer is unused on the SAIL-Recon path")
        self.img_size = image_HW
        self.patch_size = patch_HW
        self.patches_resolution = (image_HW[0] // patch_HW[0], image_HW[1] // patch_HW[1])
        self.num_patches = self.patches_resolution[0] * self.patches_resolution[1]
        self.in_chans = in_chans
        self.embed_dim = embed_dim
        self.flatten_embedding = flatten_embedding
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_HW, stride=patch_HW)
        self.norm = nn.Identity()

    def project_rows(self, images: Tensor, mode: str, mean=IDENTITY_MEAN, std=IDENTITY_STD, out: Optional[Tensor] = None,
                     resid: Optional[Tensor] = None, frame_rows: int = 0) -> Tensor:
        """images fp32 [S, 3, H, W] -> patch rows [S*h*w, D].  (mean, std) fuse the ImageNet normalisation of
        aggregator.py:267 into the im2col.  With `resid` ([h*w, D] fp32, e.g. the position embedding) the result
        is written as fp32 `resid + proj(x)` per frame into `out` rows [s*frame_rows + (frame_rows - h*w), ...)."""
        S, _, H, W = images.shape
        ph = self.patch_size[0]
        assert H % ph == 0, f"Input image height {H} is not a multiple of patch height {ph}"       # patch_embed.py:71-76
        assert W % ph == 0, f"Input image width {W} is not a multiple of patch width: {ph}"
        hw = (H // ph) * (W // ph)
        K = 3 * ph * ph
        D = self.embed_dim
        dev = images.device
        if mode == "bf16":
            Kp = (K + 63) // 64 * 64
            a = engine.WS.get("patch.a", (S * hw, Kp), torch.bfloat16, dev)
            w = engine.weight_bf16(self.proj, pad_k=64)
        else:
            Kp = K
            a = engine.WS.get("patch.a", (S * hw, Kp), torch.float32, dev)
            w = engine.weight_f32(self.proj)
        ops.patchify(images.contiguous(), a, ph, mean, std)
        b = engine.bias_for(self.proj, mode)
        if resid is None:
            y = torch.empty(S * hw, D, dtype=a.dtype, device=dev) if out is None else out
            ops.gemm(a, w, b, y, kind=SR_EPI_BIAS)
            return y
        from .._lib import SR_EPI_LS_RESIDUAL
        for s in range(S):
            dst = out[s * frame_rows + (frame_rows - hw):(s + 1) * frame_rows]
            ops.gemm(a[s * hw:(s + 1) * hw], w, b, dst, kind=SR_EPI_LS_RESIDUAL, gamma=None, resid=resid)
        return out

    def forward(self, x: Tensor) -> Tensor:
        engine.require_cuda_f32(x, "x")
        mode = engine.compute_mode()
        if mode == "bf16" and self.embed_dim % 128:
            mode = "f32"
        S, _, H, W = x.shape
        y = self.project_rows(x, mode)
        y = y.view(S, -1, self.embed_dim)
        if not self.flatten_embedding:
            y = y.reshape(-1, H // self.patch_size[0], W // self.patch_size[1], self.embed_dim)
        return y

"""Shared by oracle/make_golden.py (runs in the build container, where /root/reference exists) and the tests
(run anywhere): deterministic parameter / input generation so golden fixtures only need to store OUTPUTS.

TEST INFRASTRUCTURE ONLY — never imported by the product package.
"""
from __future__ import annotations

import math
from typing import Dict, Sequence, Tuple

import torch


def make_params(shapes: Dict[str, Sequence[int]], seed: int) -> Dict[str, torch.Tensor]:
    """One CPU generator, names in sorted order -> reproducible on any machine with the same torch.  Scales are
    chosen so every sub-path matters at random init (LayerScale gamma O(1), special tokens O(1)); the reference's own
    init (gamma 0.01, tokens 1e-6, models/aggregator.py:80,179-182) would hide kernel bugs behind the residual."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name in sorted(shapes):
        shape = tuple(shapes[name])
        leaf = name.split(".")[-1]
        if leaf == "gamma":
            t = 0.5 + torch.rand(shape, generator=g)
        elif "token" in leaf and leaf != "empty_pose_tokens":
            t = torch.randn(shape, generator=g)
        elif leaf == "empty_pose_tokens":
            t = 0.1 * torch.randn(shape, generator=g)
        elif leaf == "pos_embed":
            t = 0.02 * torch.randn(shape, generator=g)
        elif len(shape) == 1 and leaf == "weight":
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif len(shape) == 1:
            t = 0.1 * torch.randn(shape, generator=g)
        else:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            t = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        out[name] = t
    return out


def make_images(S: int, H: int, W: int, seed: int) -> torch.Tensor:
    """torch.rand(1, S, 3, H, W, generator=Generator().manual_seed(seed))."""
    return torch.rand(1, S, 3, H, W, generator=torch.Generator().manual_seed(seed))


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


# name -> (aggregator kwargs, camera-head kwargs, S, Na, H, W, fix_rank)
GOLDEN_CASES = {
    # conv patch embed, 2 anchors + 2 queries, 4x5 patch grid (P = 25), rank 6
    "tiny_conv": dict(agg=dict(img_size=56, patch_size=14, embed_dim=128, depth=2, num_heads=2, patch_embed="conv",
                               intermediate_layer_idx=[0, 1]),
                      head=dict(dim_in=256, trunk_depth=2, num_heads=2), S=4, Na=2, H=56, W=70, fix_rank=6, seed=11),
    # rank larger than the patch count (clamped, aggregator.py:277-278), 3 anchors + 1 query, anchors not first
    "tiny_clamp": dict(agg=dict(img_size=56, patch_size=14, embed_dim=128, depth=3, num_heads=2, patch_embed="conv",
                                intermediate_layer_idx=[1, 2]),
                       head=dict(dim_in=256, trunk_depth=1, num_heads=2), S=4, Na=3, H=42, W=56, fix_rank=300, seed=12),
    # DINOv2 ViT-S/14+reg patch embed (12 blocks, 6 heads x 64) with position-embedding interpolation (non-square)
    "small_dino": dict(agg=dict(img_size=70, patch_size=14, embed_dim=384, depth=2, num_heads=6,
                                patch_embed="dinov2_vits14_reg", intermediate_layer_idx=[0, 1]),
                       head=dict(dim_in=768, trunk_depth=1, num_heads=6), S=3, Na=2, H=56, W=84, fix_rank=5, seed=13),
}

"""DPTHead — API mirror of sailrecon/heads/dpt_head.py (DPTHead :22-344, FeatureFusionBlock / ResidualConvUnit :430-563,
custom_interpolate :565-594) on the B200 kernel library.

Same constructor, module / parameter names (state_dict compatible with `sailrecon.pt`: norm, projects.{0..3},
resize_layers.{0,1,3}, scratch.layer{1..4}_rn, scratch.refinenet{1..4}.{out_conv,resConfUnit1,resConfUnit2}.{conv1,conv2},
scratch.output_conv1, scratch.output_conv2.{0,2}) and `forward(aggregated_tokens_list, images, patch_start_idx,
frames_chunk_size=8) -> (preds [B, S, H, W, out_dim - 1], conf [B, S, H, W])`.

B200-first layout: every feature map is a CHANNELS-LAST fp32 row matrix [S*H*W, C], so every convolution is a GEMM on the
tcgen05 kernel (1x1 / ConvTranspose(k = stride) directly on the rows, 3x3 through im2col rows) with split-precision
operands — fp32 semantics as the reference runs the heads with autocast disabled (sail_recon.py:119), see csrc/dpt.cu.
The nn.Conv2d / nn.ConvTranspose2d members only HOLD the parameters; their forward is never called.  No torch fallback.
"""
from __future__ import annotations

from typing import List, Tuple, Union

import torch
import torch.nn as nn

from .. import engine, ops


def _conv_weights(conv: nn.Module):
    """(w_hh [Np, 2K] bf16, w_lo [Np, K] bf16, bias fp32 [Np] or None, N) for a Conv2d (rows = output channels, columns
    (ky, kx, cin)) or a ConvTranspose2d with kernel = stride (rows (ky, kx, cout), columns cin; its bias is added by the
    pixel shuffle).  Np = N rounded up to 128 (zero rows); cached until the parameter changes."""
    def make(w):
        if isinstance(conv, nn.ConvTranspose2d):
            w2 = w.permute(2, 3, 1, 0).reshape(-1, w.shape[0])            # [(ky, kx, cout), cin]
        else:
            w2 = w.permute(0, 2, 3, 1).reshape(w.shape[0], -1)            # [cout, (ky, kx, cin)]
        n, k = w2.shape
        npad = (n + 127) // 128 * 128
        full = torch.zeros(npad, k, dtype=torch.float32, device=w.device)
        full[:n] = w2
        sp = ops.im2col_split(full.view(npad, 1, 1, k), 1)               # [npad, 2k] = [hi | lo]
        return torch.cat([sp[:, :k], sp[:, :k]], dim=1).contiguous(), sp[:, k:].contiguous(), n
    w_hh, w_lo, n = engine._cached(conv, "w_split", conv.weight, make)
    bias = None
    if conv.bias is not None and not isinstance(conv, nn.ConvTranspose2d):
        def make_b(b):
            full = torch.zeros(w_lo.shape[0], dtype=torch.float32, device=b.device)
            full[:n] = b
            return full
        bias = engine._cached(conv, "b_pad", conv.bias, make_b)
    return w_hh, w_lo, bias, n


def conv_rows(conv: nn.Module, x: torch.Tensor, relu_in: bool = False, resid: torch.Tensor = None) -> torch.Tensor:
    """nn.Conv2d (kernel 1 or 3, padding k // 2, stride 1 or 2) on a channels-last fp32 map x [N, H, W, Cin] -> [N, Ho, Wo,
    Cout] (+ resid).  relu_in applies the ResidualConvUnit's pre-activation inside the im2col pass."""
    k, s = conv.kernel_size[0], conv.stride[0]
    N, H, W, _ = x.shape
    w_hh, w_lo, bias, n = _conv_weights(conv)
    a2 = ops.im2col_split(x, k, s, conv.padding[0], relu_in)
    Ho, Wo = (H + 2 * conv.padding[0] - k) // s + 1, (W + 2 * conv.padding[0] - k) // s + 1
    npad = w_lo.shape[0]
    out = torch.empty(N * Ho * Wo, npad, dtype=torch.float32, device=x.device)
    if resid is not None and npad != n:
        raise ValueError("sailrecon_b200: a fused residual needs an output width that is a multiple of 128")
    ops.gemm_split(a2, w_hh, w_lo, bias, out, None if resid is None else resid.reshape(N * Ho * Wo, n))
    y = out.view(N, Ho, Wo, npad)
    return y if npad == n else y[..., :n]


def conv3_pad(conv: nn.Module, x: torch.Tensor, in_pad: bool, relu_in: bool = False, resid: torch.Tensor = None) -> torch.Tensor:
    """nn.Conv2d(kernel 3, stride 1, padding 1) WITHOUT an im2col matrix: x fp32 [S, H (+2), W (+2), Cin] (in_pad: it carries the
    one-pixel border) -> fp32 [S, H+2, W+2, Np] in the padded row numbering (Np = Cout rounded up to 128; border rows are
    garbage).  One split pass (8 B per element) + one shifted-GEMM launch with the three split-precision parts inside."""
    w_hh, w_lo, bias, n = _conv_weights(conv)
    w27 = engine._cached(conv, "w27", conv.weight, lambda w: torch.cat([w_hh, w_lo], dim=1).contiguous())   # Whi | Whi | Wlo
    xs = ops.split_pad(x, in_pad, relu_in)
    return ops.conv3x3(xs, w27, bias, resid)


class ResidualConvUnit(nn.Module):
    """dpt_head.py:430-482.  The reference's activation is nn.ReLU(inplace=True) (:357), so `out = self.activation(x)`
    rectifies x IN PLACE and the skip connection adds the RECTIFIED input: RCU(x) = relu(x) + conv2(relu(conv1(relu(x)))).
    `rows` therefore takes xr = relu(x) (the callers fuse that ReLU into the addition that produces x)."""

    def __init__(self, features, activation=None, bn=False, groups=1):
        super().__init__()
        if bn or groups != 1:
            raise NotImplementedError("sailrecon_b200: DPT runs bn=False, groups=1")
        self.bn, self.groups = bn, groups
        self.conv1 = nn.Conv2d(features, features, kernel_size=3, stride=1, padding=1, bias=True)
        self.conv2 = nn.Conv2d(features, features, kernel_size=3, stride=1, padding=1, bias=True)
        self.norm1 = self.norm2 = None

    def rows(self, xr: torch.Tensor) -> torch.Tensor:
        """xr = relu(x) as a zero-bordered map [S, H+2, W+2, C] -> RCU(x) in the same padded layout."""
        h = conv3_pad(self.conv1, xr, in_pad=True)
        return conv3_pad(self.conv2, h, in_pad=True, relu_in=True, resid=xr)     # skip_add fused into the GEMM epilogue


class FeatureFusionBlock(nn.Module):
    """dpt_head.py:485-563."""

    def __init__(self, features, activation=None, deconv=False, bn=False, expand=False, align_corners=True, size=None,
                 has_residual=True, groups=1):
        super().__init__()
        if deconv or expand or not align_corners:
            raise NotImplementedError("sailrecon_b200: DPT runs deconv=False, expand=False, align_corners=True")
        self.deconv, self.align_corners, self.groups, self.expand = deconv, align_corners, groups, expand
        self.out_conv = nn.Conv2d(features, features, kernel_size=1, stride=1, padding=0, bias=True)
        if has_residual:
            self.resConfUnit1 = ResidualConvUnit(features, activation, bn, groups)
        self.has_residual = has_residual
        self.resConfUnit2 = ResidualConvUnit(features, activation, bn, groups)
        self.size = size

    def rows(self, x0: torch.Tensor, x0_padded: bool, x1: torch.Tensor = None, size=None) -> torch.Tensor:
        """forward(*xs, size) (dpt_head.py:536-563): x0 = xs[0] (dense, or padded when x0_padded), x1 = xs[1] = a layer_rn output
        in the padded layout.  Returns the DENSE map after interpolation + out_conv."""
        if self.has_residual:
            res = self.resConfUnit1.rows(ops.map_add(x1, True, None, False, True, relu=True))
            out_r = ops.map_add(x0, x0_padded, res, True, True, relu=True)                # relu(xs[0] + res): all RCU2 reads
        else:
            out_r = ops.map_add(x0, x0_padded, None, False, True, relu=True)
        output = self.resConfUnit2.rows(out_r)                                           # padded
        Hh, Ww = output.shape[1] - 2, output.shape[2] - 2
        if size is None and self.size is None:
            Ho, Wo = int(Hh * 2), int(Ww * 2)
        else:
            Ho, Wo = (self.size if size is None else size)
        output = ops.bilinear_nhwc(output, int(Ho), int(Wo), in_pad=True)
        return conv_rows(self.out_conv, output)


def _make_scratch(in_shape: List[int], out_shape: int) -> nn.Module:
    scratch = nn.Module()
    for i, c in enumerate(in_shape):
        setattr(scratch, f"layer{i + 1}_rn", nn.Conv2d(c, out_shape, kernel_size=3, stride=1, padding=1, bias=False))
    return scratch


class DPTHead(nn.Module):
    def __init__(self, dim_in: int, patch_size: int = 14, output_dim: int = 4, activation: str = "inv_log",
                 conf_activation: str = "expp1", features: int = 256, out_channels: List[int] = [256, 512, 1024, 1024],
                 intermediate_layer_idx: List[int] = [4, 11, 17, 23], pos_embed: bool = True, feature_only: bool = False,
                 down_ratio: int = 1) -> None:
        super().__init__()
        self.patch_size, self.activation, self.conf_activation = patch_size, activation, conf_activation
        self.pos_embed, self.feature_only, self.down_ratio = pos_embed, feature_only, down_ratio
        self.intermediate_layer_idx = intermediate_layer_idx
        self.norm = nn.LayerNorm(dim_in)
        self.projects = nn.ModuleList([nn.Conv2d(dim_in, oc, kernel_size=1, stride=1, padding=0) for oc in out_channels])
        self.resize_layers = nn.ModuleList([
            nn.ConvTranspose2d(out_channels[0], out_channels[0], kernel_size=4, stride=4, padding=0),
            nn.ConvTranspose2d(out_channels[1], out_channels[1], kernel_size=2, stride=2, padding=0),
            nn.Identity(),
            nn.Conv2d(out_channels[3], out_channels[3], kernel_size=3, stride=2, padding=1)])
        self.scratch = _make_scratch(out_channels, features)
        self.scratch.stem_transpose = None
        self.scratch.refinenet1 = FeatureFusionBlock(features)
        self.scratch.refinenet2 = FeatureFusionBlock(features)
        self.scratch.refinenet3 = FeatureFusionBlock(features)
        self.scratch.refinenet4 = FeatureFusionBlock(features, has_residual=False)
        if feature_only:
            self.scratch.output_conv1 = nn.Conv2d(features, features, kernel_size=3, stride=1, padding=1)
        else:
            self.scratch.output_conv1 = nn.Conv2d(features, features // 2, kernel_size=3, stride=1, padding=1)
            self.scratch.output_conv2 = nn.Sequential(nn.Conv2d(features // 2, 32, kernel_size=3, stride=1, padding=1),
                                                      nn.ReLU(inplace=True),
                                                      nn.Conv2d(32, output_dim, kernel_size=1, stride=1, padding=0))
        self._pe_cache = {}

    # ---------------------------------------------------------------------------------------------- forward
    def forward(self, aggregated_tokens_list, images: torch.Tensor, patch_start_idx: int,
                frames_chunk_size: int = 8) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        """dpt_head.py:162-229: frames are processed in chunks of `frames_chunk_size` (bounds the im2col scratch)."""
        engine.forbid_autograd(self, "DPTHead.forward")
        B, S, _, H, W = images.shape
        if B != 1:
            raise NotImplementedError("sailrecon_b200: batch size 1 (every SAIL-Recon script runs B = 1)")
        if frames_chunk_size is None or frames_chunk_size >= S:
            return self._forward_impl(aggregated_tokens_list, images, patch_start_idx, 0, S)
        assert frames_chunk_size > 0
        parts = [self._forward_impl(aggregated_tokens_list, images, patch_start_idx, f0, min(f0 + frames_chunk_size, S))
                 for f0 in range(0, S, frames_chunk_size)]
        if self.feature_only:
            return torch.cat(parts, dim=1)
        return torch.cat([p[0] for p in parts], dim=1), torch.cat([p[1] for p in parts], dim=1)

    def _pos_table(self, h: int, w: int, C: int, W_img: int, H_img: int, device) -> torch.Tensor:
        """_apply_pos_embed (dpt_head.py:296-311): uv grid with aspect W / H -> sin / cos table [h, w, C] (cached)."""
        key = (h, w, C, W_img, H_img, torch.device(device))
        t = self._pe_cache.get(key)
        if t is None:
            t = ops.uv_posembed_table(h, w, C, W_img / H_img, device)
            if len(self._pe_cache) > 16:
                self._pe_cache.clear()
            self._pe_cache[key] = t
        return t

    def _add_pos(self, x: torch.Tensor, W_img: int, H_img: int, ratio: float = 0.1) -> torch.Tensor:
        N, h, w, C = x.shape
        x = x.contiguous()
        return ops.add_bcast(x, self._pos_table(h, w, C, W_img, H_img, x.device), ratio, out=x)

    def _forward_impl(self, aggregated_tokens_list, images, patch_start_idx, f0: int, f1: int):
        """dpt_head.py:231-294 for frames [f0, f1)."""
        _, _, _, H, W = images.shape
        S = f1 - f0
        ph, pw = H // self.patch_size, W // self.patch_size
        dev = images.device
        out = []
        for dpt_idx, layer_idx in enumerate(self.intermediate_layer_idx):
            tok = aggregated_tokens_list[layer_idx]                                   # [1, S_all, P, C]
            engine.require_cuda_f32(tok, "aggregated tokens")
            P, Cin = tok.shape[2], tok.shape[3]
            rows = tok[0, f0:f1].reshape(S * P, Cin)
            idx = (torch.arange(S, device=dev, dtype=torch.int32)[:, None] * P
                   + torch.arange(patch_start_idx, P, device=dev, dtype=torch.int32)[None]).reshape(-1).contiguous()
            x = torch.empty(S * ph * pw, Cin, dtype=torch.float32, device=dev)
            w_, b_, eps = engine.ln_params(self.norm)
            ops.layernorm(rows, w_, b_, eps, x, idx=idx)                              # drop the special tokens + self.norm
            x = conv_rows(self.projects[dpt_idx], x.view(S, ph, pw, Cin))             # 1x1 projection
            if self.pos_embed:
                x = self._add_pos(x, W, H)
            rl = self.resize_layers[dpt_idx]
            if isinstance(rl, nn.ConvTranspose2d):
                k = rl.kernel_size[0]
                w_hh, w_lo, _, n = _conv_weights(rl)
                C = x.shape[-1]
                a2 = ops.im2col_split(x.contiguous(), 1)
                g = torch.empty(S * ph * pw, w_lo.shape[0], dtype=torch.float32, device=dev)
                ops.gemm_split(a2, w_hh, w_lo, None, g)
                x = ops.pixel_shuffle(g if w_lo.shape[0] == n else g[:, :n].contiguous(), rl.bias.detach(), S, ph, pw,
                                      n // (k * k), k)
            elif isinstance(rl, nn.Conv2d):
                x = conv_rows(rl, x.contiguous())
            out.append(x)
        out = self.scratch_forward(out)                                               # padded [S, h1*2 + 2, w1*2 + 2, 128 | 256]
        Ho, Wo = int(ph * self.patch_size / self.down_ratio), int(pw * self.patch_size / self.down_ratio)
        out = ops.bilinear_nhwc(out, Ho, Wo, in_pad=True)
        if self.pos_embed:
            out = self._add_pos(out, W, H)
        if self.feature_only:
            return out.permute(0, 3, 1, 2)[None]                                     # [B, S, C, H, W] view, as the reference
        conv_a, conv_b = self.scratch.output_conv2[0], self.scratch.output_conv2[2]
        hmap = conv3_pad(conv_a, out, in_pad=False)                                   # [S, Ho+2, Wo+2, 128]: 32 valid channels
        pred, conf = ops.dpt_out_head(hmap.view(-1, hmap.shape[-1]), conv_a.out_channels, conv_b.weight.detach(),
                                      conv_b.bias.detach(), self.activation, self.conf_activation, padded_map=(S, Ho, Wo))
        return pred.view(1, S, Ho, Wo, -1), conf.view(1, S, Ho, Wo)

    def scratch_forward(self, features: List[torch.Tensor]) -> torch.Tensor:
        """dpt_head.py:313-344 on channels-last maps; the 3x3 convolutions keep their maps in the zero-bordered layout."""
        l1, l2, l3, l4 = features
        sc = self.scratch
        l1rn, l2rn = conv3_pad(sc.layer1_rn, l1.contiguous(), False), conv3_pad(sc.layer2_rn, l2.contiguous(), False)
        l3rn, l4rn = conv3_pad(sc.layer3_rn, l3.contiguous(), False), conv3_pad(sc.layer4_rn, l4.contiguous(), False)
        dense = lambda t: (t.shape[1] - 2, t.shape[2] - 2)
        out = sc.refinenet4.rows(l4rn, True, None, size=dense(l3rn))
        out = sc.refinenet3.rows(out, False, l3rn, size=dense(l2rn))
        out = sc.refinenet2.rows(out, False, l2rn, size=dense(l1rn))
        out = sc.refinenet1.rows(out, False, l1rn)
        return conv3_pad(sc.output_conv1, out.contiguous(), False)

"""Mlp — API mirror of sailrecon/layers/mlp.py:16-40: fc2(GELU_erf(fc1(x))); dropout p = 0 on this path."""
from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.nn as nn
from torch import Tensor

from .. import engine, ops
from .._lib import SR_EPI_BIAS, SR_EPI_BIAS_GELU


class Mlp(nn.Module):
    def __init__(self, in_features: int, hidden_features: Optional[int] = None, out_features: Optional[int] = None,
                 act_layer: Callable[..., nn.Module] = nn.GELU, drop: float = 0.0, bias: bool = True) -> None:
        super().__init__()
        if act_layer is not nn.GELU:
            raise NotImplementedError("sailrecon_b200: Mlp implements the exact-erf GELU the reference uses (mlp.py:22)")
        if drop != 0.0:
            raise NotImplementedError("sailrecon_b200: dropout is 0 everywhere on the SAIL-Recon path")
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features, bias=bias)
        self.fc2 = nn.Linear(hidden_features, out_features, bias=bias)

    def forward(self, x: Tensor) -> Tensor:
        """Standalone call (camera head pose_branch, camera_head.py:78-83,168).  The tensor-core path needs
        128-multiples of the output widths; narrower layers (pose_branch: 9 outputs) run on the fp32 kernels."""
        engine.require_cuda_f32(x, "x")
        mode = engine.compute_mode()
        n1, n2 = self.fc1.weight.shape[0], self.fc2.weight.shape[0]
        if mode == "bf16" and (n1 % 128 or n2 % 128 or self.fc1.weight.shape[1] % 64 or n1 % 64):
            mode = "f32"
        rows = x.reshape(-1, x.shape[-1]).contiguous()
        act = torch.bfloat16 if mode == "bf16" else torch.float32
        a = ops.cast_bf16(rows) if mode == "bf16" else rows
        h = torch.empty(rows.shape[0], n1, dtype=act, device=x.device)
        engine.linear_rows(self.fc1, a, h, mode, SR_EPI_BIAS_GELU)
        y = torch.empty(rows.shape[0], n2, dtype=act, device=x.device)
        engine.linear_rows(self.fc2, h, y, mode, SR_EPI_BIAS)
        return y.view(*x.shape[:-1], n2)

"""LayerScale — API mirror of sailrecon/layers/layer_scale.py:14-23.  Inside a Block gamma is folded into the
proj / fc2 GEMM epilogue (SR_EPI_LS_RESIDUAL); the standalone forward is one elementwise kernel."""
from __future__ import annotations

from typing import Union

import torch
import torch.nn as nn
from torch import Tensor

from .. import ops


class LayerScale(nn.Module):
    def __init__(self, dim: int, init_values: Union[float, Tensor] = 1e-5, inplace: bool = False) -> None:
        super().__init__()
        self.inplace = inplace
        self.gamma = nn.Parameter(init_values * torch.ones(dim))

    def forward(self, x: Tensor) -> Tensor:
        if not x.is_cuda:
            raise RuntimeError("sailrecon_b200: LayerScale runs on a CUDA device only (no CPU fallback)")
        y = ops.scale_cols(x.float().contiguous(), self.gamma.detach())
        if self.inplace:
            x.copy_(y)
            return x
        return y

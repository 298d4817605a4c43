"""2D rotary position embedding — API mirror of sailrecon/layers/rope.py (PositionGetter :25-66,
RotaryPositionEmbedding2D :69-207).

Inside a Block the rotation is applied by the qkv GEMM epilogue (csrc/gemm_tc.cu, SR_EPI_QKV); this module owns the
cos/sin tables those kernels read and offers the reference's standalone ``forward(tokens, positions)``.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn as nn

from .. import ops


class PositionGetter:
    """(y, x) grid coordinates of every patch; cached per (h, w) like the reference (rope.py:36-66)."""

    def __init__(self):
        self.position_cache: Dict[Tuple[int, int], torch.Tensor] = {}

    def __call__(self, batch_size: int, height: int, width: int, device) -> torch.Tensor:
        key = (height, width)
        if key not in self.position_cache:
            ys = torch.arange(height, device=device).repeat_interleave(width)
            xs = torch.arange(width, device=device).repeat(height)
            self.position_cache[key] = torch.stack([ys, xs], dim=-1)            # int64 [h*w, 2], same as cartesian_prod
        cached = self.position_cache[key]
        return cached.view(1, height * width, 2).expand(batch_size, -1, -1).clone()


class RotaryPositionEmbedding2D(nn.Module):
    """Same constructor and call signature as the reference (rope.py:86,165).  No parameters, no state_dict entries."""

    def __init__(self, frequency: float = 100.0, scaling_factor: float = 1.0):
        super().__init__()
        self.base_frequency = frequency
        self.scaling_factor = scaling_factor
        self._tables: Dict[Tuple[torch.device, int], Tuple[torch.Tensor, torch.Tensor, int]] = {}

    def tables(self, device, max_pos: int, head_dim: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """cos / sin fp32 [>= max_pos, head_dim/4] on `device` (rope.py:107-123 without the :119 duplication)."""
        quarter = head_dim // 4
        key = (torch.device(device), quarter)
        hit = self._tables.get(key)
        if hit is None or hit[2] < max_pos:
            n = max(64, 1 << (int(max_pos) - 1).bit_length())
            cos, sin = ops.rope_tables(n, quarter, self.base_frequency, device)
            hit = (cos, sin, n)
            self._tables[key] = hit
        return hit[0], hit[1]

    def forward(self, tokens: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        # argument checks follow rope.py:181-184
        assert tokens.size(-1) % 2 == 0, "Feature dimension must be even"
        assert positions.ndim == 3 and positions.shape[-1] == 2, "Positions must have shape (batch_size, n_tokens, 2)"
        if not tokens.is_cuda:
            raise RuntimeError("sailrecon_b200: RotaryPositionEmbedding2D runs on a CUDA device only (no CPU fallback)")
        max_pos = int(positions.max()) + 1                                        # rope.py:190 (standalone call only)
        cos, sin = self.tables(tokens.device, max_pos, tokens.shape[-1])
        out = ops.rope2d(tokens.float().contiguous(), positions.to(torch.int32).contiguous(), cos, sin)
        return out.to(tokens.dtype)

"""Mirror of sailrecon/layers/__init__.py (same exported names; SwiGLU is never instantiated on this path)."""
from .attention import Attention, MemEffAttention
from .block import Block, NestedTensorBlock
from .layer_scale import LayerScale
from .mlp import Mlp
from .patch_embed import PatchEmbed
from .rope import PositionGetter, RotaryPositionEmbedding2D

__all__ = ["Attention", "MemEffAttention", "Block", "NestedTensorBlock", "LayerScale", "Mlp", "PatchEmbed",
           "PositionGetter", "RotaryPositionEmbedding2D"]

"""sailrecon_b200 — B200-native (sm_100a) hot path of SAIL-Recon behind the reference's nn.Module API.

The directory is named ``self-supervise-sfm_b200`` (not importable as such); ``import sailrecon_b200`` works through
the shim ``sailrecon_b200.py`` at the repository root, which loads this package under that name.

Layout:
  csrc/      hand-written CUDA kernels + the C-ABI (include/sailrecon_b200.h) -> lib/libsailrecon_b200.so
  _lib.py    ctypes binding;  ops.py  torch-tensor wrappers (pointers + strides + stream, nothing else)
  engine.py  the fused per-block launch schedule and the structured-attention description
  layers/    Block / Attention / Mlp / LayerScale / RotaryPositionEmbedding2D / PositionGetter / DINOv2 ViT  (sailrecon/layers)
  models/    Aggregator                                                                           (sailrecon/models)
  heads/     CameraHead                                                                           (sailrecon/heads)
  parallel.py  frame-sharded (sequence-parallel) global attention with ring K/V exchange
"""
__version__ = "0.1.0"

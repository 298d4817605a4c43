"""Training loss of the reference's self-supervised step (train/train_imc.py:141-256) on the kernels of
csrc/loss.cu: the per-frame CDF loss module, the reprojection geometry and the fused `compute_loss`."""
from .cdf_loss import CDFLossIndexPytorch, CDFLossTorchWrapper
from .geometry import (backproject_and_reproject, backproject_and_reproject_with_approximation, compute_projective_residual,
                       compute_relative_pose)
from .reproj_loss import compute_loss, reprojection_cdf_loss

__all__ = ["CDFLossIndexPytorch", "CDFLossTorchWrapper", "backproject_and_reproject",
           "backproject_and_reproject_with_approximation", "compute_projective_residual", "compute_relative_pose",
           "compute_loss", "reprojection_cdf_loss"]

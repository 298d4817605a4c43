"""pose_encoding_to_extri_intri — API mirror of sailrecon/utils/pose_enc.py:68-135 on one fused kernel
(sr_pose_to_extri_intri: quaternion -> rotation, [R | T], FoV -> pinhole intrinsics)."""
from __future__ import annotations

import torch

from .. import ops


def pose_encoding_to_extri_intri(pose_encoding: torch.Tensor, image_size_hw=None, pose_encoding_type: str = "absT_quaR_FoV",
                                 build_intrinsics: bool = True):
    """pose_encoding [B, S, 9] = (T, quat xyzw, fov_h, fov_w) -> (extrinsics [B, S, 3, 4], intrinsics [B, S, 3, 3] | None).
    Called right after the camera head in every forward (sail_recon.py:122-124, :235-237)."""
    if pose_encoding_type != "absT_quaR_FoV":
        raise NotImplementedError                                                          # pose_enc.py:133-134
    if not pose_encoding.is_cuda:
        raise RuntimeError("sailrecon_b200: pose_encoding_to_extri_intri runs on a CUDA device only (no CPU fallback)")
    return ops.pose_to_extri_intri(pose_encoding.float(), image_size_hw, build_intrinsics)

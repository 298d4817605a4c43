"""quat_to_mat — API mirror of sailrecon/utils/rotation.py:14-44 (scalar-last XYZW quaternion -> 3x3 rotation)."""
from __future__ import annotations

import torch

from .. import ops


def quat_to_mat(quaternions: torch.Tensor) -> torch.Tensor:
    """quaternions [..., 4] (real part last) -> [..., 3, 3].  Runs the fused pose kernel with a zero translation and no
    intrinsics (sr_pose_to_extri_intri) and returns the rotation block."""
    if quaternions.shape[-1] != 4:
        raise ValueError(f"Invalid quaternion shape {tuple(quaternions.shape)}.")
    pose = torch.zeros(*quaternions.shape[:-1], 9, dtype=torch.float32, device=quaternions.device)
    pose[..., 3:7] = quaternions.float()
    ext, _ = ops.pose_to_extri_intri(pose, None, build_intrinsics=False)
    return ext[..., :3].contiguous()

"""pose_encoding_to_extri_intri — API mirror of sailrecon/utils/pose_enc.py:68-135 on one fused kernel
(sr_pose_to_extri_intri: quaternion -> rotation, [R | T], FoV -> pinhole intrinsics)."""
from __future__ import annotations

import torch

from .. import ops


class _PoseToCameras(torch.autograd.Function):
    """Differentiable form for training (the loss reads the predicted cameras, train_imc.py:156,175)."""

    @staticmethod
    def forward(ctx, pose, image_size_hw, build_intrinsics):
        ext, intr = ops.pose_to_extri_intri(pose, image_size_hw, build_intrinsics)
        ctx.save_for_backward(pose)
        ctx.hw, ctx.has_intr = image_size_hw, intr is not None
        return (ext, intr) if intr is not None else (ext, None)

    @staticmethod
    def backward(ctx, d_ext, d_intr):
        (pose,) = ctx.saved_tensors
        if d_ext is None:
            d_ext = torch.zeros(*pose.shape[:-1], 3, 4, dtype=torch.float32, device=pose.device)
        return ops.pose_to_extri_intri_bwd(pose, d_ext, d_intr if ctx.has_intr else None, ctx.hw), None, None


def pose_encoding_to_extri_intri(pose_encoding: torch.Tensor, image_size_hw=None, pose_encoding_type: str = "absT_quaR_FoV",
                                 build_intrinsics: bool = True):
    """pose_encoding [B, S, 9] = (T, quat xyzw, fov_h, fov_w) -> (extrinsics [B, S, 3, 4], intrinsics [B, S, 3, 3] | None).
    Called right after the camera head in every forward (sail_recon.py:122-124, :235-237)."""
    if pose_encoding_type != "absT_quaR_FoV":
        raise NotImplementedError                                                          # pose_enc.py:133-134
    if not pose_encoding.is_cuda:
        raise RuntimeError("sailrecon_b200: pose_encoding_to_extri_intri runs on a CUDA device only (no CPU fallback)")
    if pose_encoding.requires_grad and torch.is_grad_enabled():
        ctx_free = _PoseToCameras.apply(pose_encoding.float(), image_size_hw, build_intrinsics)
        return ctx_free[0], ctx_free[1]
    return ops.pose_to_extri_intri(pose_encoding.float(), image_size_hw, build_intrinsics)

"""Correspondence geometry of the training loss — train/utils/geometry.py:89-303 — on the kernels of csrc/loss.cu.

Same four entry points and argument meaning as the reference; each is one autograd node whose backward is a kernel too
(gradients flow to the intrinsics and the poses, which is all the trainer differentiates: coordinates, depths and depth
scales are data, train_imc.py:178-197 — asking for their gradient raises).  The reference chains inverse / bmm /
transpose / cat launches over [pairs, K, 3..4] intermediates; here a point is projected in registers.
"""
from __future__ import annotations

from typing import Tuple

import torch

from .. import ops


def to_homogeneous(points: torch.Tensor) -> torch.Tensor:
    """(..., N) -> (..., N + 1) with a trailing 1 (geometry.py:6-22)."""
    return torch.cat([points, points.new_ones(points.shape[:-1] + (1,))], dim=-1)


def from_homogeneous(points: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """(..., N + 1) -> (..., N): divide by the last coordinate + eps (geometry.py:24-34)."""
    return points[..., :-1] / (points[..., -1:] + eps)


def pad_poses(pose: torch.Tensor) -> torch.Tensor:
    """(..., 3, 4) -> (..., 4, 4) with the row (0, 0, 0, 1) (geometry.py:36-57)."""
    last = pose.new_tensor([0.0, 0.0, 0.0, 1.0]).expand(pose.shape[:-2] + (1, 4))
    return torch.cat([pose[..., :3, :4], last], dim=-2)


def unpad_poses(pose: torch.Tensor) -> torch.Tensor:
    return pose[..., :3, :4]


def pad_intrinsic44(intrinsic: torch.Tensor) -> torch.Tensor:
    """[B, 3, 3] -> [B, 4, 4] with a unit fourth diagonal entry (geometry.py:59-72)."""
    out = torch.eye(4, device=intrinsic.device, dtype=intrinsic.dtype).repeat(intrinsic.shape[0], 1, 1)
    out[:, :3, :3] = intrinsic
    return out


def _no_grad_inputs(**tensors) -> None:
    for name, t in tensors.items():
        if t is not None and t.requires_grad:
            raise NotImplementedError(f"sailrecon_b200.losses.geometry: no gradient w.r.t. {name} (it is data in the reference's "
                                      "trainer, train_imc.py:178-197)")


class _RelativePose(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, dst):
        ctx.save_for_backward(src, dst)
        return ops.relative_pose(src, dst)

    @staticmethod
    def backward(ctx, g):
        src, dst = ctx.saved_tensors
        return ops.relative_pose_bwd(src, dst, g)


def compute_relative_pose(src_extrinsic: torch.Tensor, dst_extrinsic: torch.Tensor) -> torch.Tensor:
    """World-to-camera [BZ, 3, 4] or [BZ, 4, 4] of two cameras -> src-to-dst transform dst @ inverse(src), [BZ, 4, 4]
    (geometry.py:252-287)."""
    return _RelativePose.apply(src_extrinsic.float(), dst_extrinsic.float())


class _Reproject(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale):
        ctx.save_for_backward(src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale)
        return ops.reproject(src_coords, src_depth, src_k, dst_k, rel, src_scale, dst_depth, dst_scale)

    @staticmethod
    def backward(ctx, g):
        src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale = ctx.saved_tensors
        dks, dkd, dr = ops.reproject_bwd(src_coords, src_depth, src_k, dst_k, rel, g, src_scale, dst_depth, dst_scale)
        return None, None, None, dks, dkd, dr, None, None


def backproject_and_reproject(src_coords, src_depth, src_intrinsic, dst_intrinsic, relative_pose,
                              src_depth_scale) -> Tuple[torch.Tensor, torch.Tensor]:
    """Lift source pixels [BZ, N, 2] with their depths [BZ, N] through K_src^-1, move them with the src-to-dst pose, project
    with K_dst and divide by the projected depth + 1e-6 -> (dst pixels [BZ, N, 2], all-true mask [BZ, N]); geometry.py:89-158."""
    _no_grad_inputs(src_coords=src_coords, src_depth=src_depth, src_depth_scale=src_depth_scale)
    out = _Reproject.apply(src_coords, src_depth, None, src_intrinsic, dst_intrinsic, relative_pose, src_depth_scale, None)
    return out, torch.ones(out.shape[:2], dtype=torch.bool, device=out.device)


def backproject_and_reproject_with_approximation(src_coords, src_depth, dst_depth, src_intrinsic, dst_intrinsic, relative_pose,
                                                 src_depth_scale, dst_depth_scale) -> Tuple[torch.Tensor, torch.Tensor]:
    """As above, but the perspective division uses the MATCHED destination depth (dst_depth * dst_depth_scale + 1e-6) instead
    of the projected one: linear in the projected point (geometry.py:160-250)."""
    _no_grad_inputs(src_coords=src_coords, src_depth=src_depth, dst_depth=dst_depth, src_depth_scale=src_depth_scale,
                    dst_depth_scale=dst_depth_scale)
    out = _Reproject.apply(src_coords, src_depth, dst_depth, src_intrinsic, dst_intrinsic, relative_pose, src_depth_scale,
                           dst_depth_scale)
    return out, torch.ones(out.shape[:2], dtype=torch.bool, device=out.device)


class _ProjectiveResidual(torch.autograd.Function):
    @staticmethod
    def forward(ctx, predicted, actual):
        ctx.save_for_backward(predicted, actual)
        return ops.projective_residual(predicted, actual)

    @staticmethod
    def backward(ctx, g):
        predicted, actual = ctx.saved_tensors
        d = ops.projective_residual_bwd(predicted, actual, g)
        return d, (-d if ctx.needs_input_grad[1] else None)


def compute_projective_residual(predicted_dst_coords: torch.Tensor, actual_dst_coords: torch.Tensor) -> torch.Tensor:
    """Pixel distance between predicted and matched destination coordinates, [BZ, N] (geometry.py:289-303)."""
    return _ProjectiveResidual.apply(predicted_dst_coords, actual_dst_coords)

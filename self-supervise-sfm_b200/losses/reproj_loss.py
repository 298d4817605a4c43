"""`compute_loss` of train/train_imc.py:141-256 as ONE autograd node over csrc/loss.cu's fused kernels.

The reference chains: recover the intrinsics (K'_to_K @ K, optional shared-focal average) -> gather per pair -> relative
poses -> two reprojections (exact / matched-depth division) -> pixel residuals -> log1p -> CDFLossIndexPytorch twice ->
means.  `sr_reproj_cdf_loss_f32` evaluates that chain AND its gradient w.r.t. the predicted intrinsics / extrinsics in a
memset + 7 launches during the forward; backward() only scales the stored gradients by the incoming one (the GradScaler
factor at train_imc.py:380).
"""
from __future__ import annotations

import torch

from .. import ops
from .cdf_loss import CDFLossIndexPytorch


class _ReprojCDFLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, intrinsics, extrinsics, kprime_to_k, shared_focal, src_idx, dst_idx, src_coords, dst_coords, src_depth,
                dst_depth, module, want_tables):
        loss, d_k, d_e, tables = ops.reproj_cdf_loss(
            intrinsics, extrinsics, kprime_to_k, shared_focal, src_idx, dst_idx, src_coords, dst_coords, src_depth, dst_depth,
            float(module.min_val), module._inv_bin_width, module.num_bins, module._sobel_w,
            module.smooth_conv.weight.detach().reshape(-1), want_tables)
        ctx.save_for_backward(d_k, d_e)
        ctx.mark_non_differentiable(*([tables] if tables is not None else []))
        return (loss.reshape(()), tables) if tables is not None else (loss.reshape(()), None)

    @staticmethod
    def backward(ctx, g, _g_tables=None):
        d_k, d_e = ctx.saved_tensors
        return (d_k * g, d_e * g) + (None,) * 10


def reprojection_cdf_loss(pred_intrinsics, pred_extrinsics, kprime_to_k, shared_focal, src_idx, dst_idx, src_coords, dst_coords,
                          src_depth, dst_depth, cdf_loss_module: CDFLossIndexPytorch, want_tables: bool = False):
    """pred_intrinsics [N, 3, 3], pred_extrinsics [N, 3, 4] (or [N, 4, 4]) -> scalar loss (and, optionally, the per-frame tables
    [3 (pmf, cdf, pdf), 2 (exact, approximated), N, bins])."""
    for name, t in (("src_coords", src_coords), ("dst_coords", dst_coords), ("src_depth", src_depth), ("dst_depth", dst_depth)):
        if t.requires_grad:
            raise NotImplementedError(f"sailrecon_b200.losses: no gradient w.r.t. {name} (data in train_imc.py:178-183)")
    loss, tables = _ReprojCDFLoss.apply(pred_intrinsics.float(), pred_extrinsics.float(), kprime_to_k, bool(shared_focal),
                                        src_idx, dst_idx, src_coords, dst_coords, src_depth, dst_depth, cdf_loss_module,
                                        want_tables)
    return (loss, tables) if want_tables else loss


def compute_loss(predictions, batch, device, cdf_loss_module: CDFLossIndexPytorch, do_record: bool = False) -> dict:
    """Drop-in for train_imc.py:141-270: `predictions` is the list SailRecon.forward returns, `batch` the data loader's
    dictionary (K_prime_to_K, shared_focal, src/dst coords, depths and frame indices of every pair)."""
    intrinsics = torch.cat([p["intrinsic"] for p in predictions], dim=0)
    extrinsics = torch.cat([p["extrinsic"] for p in predictions], dim=0)
    b = {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in batch.items()}
    out = reprojection_cdf_loss(intrinsics, extrinsics, b["K_prime_to_K"].float(), b["shared_focal"], b["src_idx"].long(),
                                b["dst_idx"].long(), b["src_coords"], b["dst_coords"], b["src_depth"], b["dst_depth"],
                                cdf_loss_module, want_tables=do_record)
    if not do_record:
        return {"loss": out}
    loss, tables = out
    return {"loss": loss,
            "plot_data": {"frame_cdfs": tables[1, 0].cpu().numpy(), "frame_pdfs": tables[2, 0].cpu().numpy(),
                          "max_val": cdf_loss_module.max_val, "min_val": cdf_loss_module.min_val,
                          "num_bins": cdf_loss_module.num_bins}}

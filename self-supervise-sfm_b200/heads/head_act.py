"""activate_pose — API mirror of sailrecon/heads/head_act.py:12-60 for the activations SAIL-Recon configures
(translation / quaternion "linear", field of view "relu", camera_head.py:34-36).  Inside CameraHead the activation is
fused with the pose accumulation (csrc/misc.cu, sr_pose_update); this function serves direct callers."""
import torch

from .. import ops


def activate_pose(pred_pose_enc, trans_act="linear", quat_act="linear", fl_act="linear"):
    for name, act in (("trans_act", trans_act), ("quat_act", quat_act)):
        if act != "linear":
            raise NotImplementedError(f"sailrecon_b200: {name}={act!r} is not used by SAIL-Recon (only 'linear')")
    if fl_act == "linear":
        return pred_pose_enc
    if fl_act != "relu":
        raise ValueError(f"Unknown act_type: {fl_act}")
    if not pred_pose_enc.is_cuda:
        raise RuntimeError("sailrecon_b200: activate_pose runs on a CUDA device only (no CPU fallback)")
    p = pred_pose_enc.float().contiguous()
    out = torch.empty_like(p)
    ops.pose_update(p.clone(), p, out, True)
    return out

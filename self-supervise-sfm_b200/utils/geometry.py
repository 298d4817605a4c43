"""unproject_depth_map_to_point_map — API mirror of sailrecon/utils/geometry.py:19-128.

The reference converts its inputs to numpy and unprojects on the host (a device->host sync of every depth map inside
SailRecon.forward, sail_recon.py:139-141).  Here one kernel does it in HBM (sr_unproject_depth).  With the default
`as_numpy=True` the result is returned as the reference returns it (np.ndarray [S, H, W, 3], one D2H copy of the points);
`as_numpy=False` keeps the points on the device."""
from __future__ import annotations

import numpy as np
import torch

from .. import ops


def unproject_depth_map_to_point_map(depth_map, extrinsics_cam, intrinsics_cam, as_numpy: bool = True):
    """depth_map [S, H, W, 1] or [S, H, W]; extrinsics_cam [S, 3, 4] (camera from world); intrinsics_cam [S, 3, 3]."""
    dev = None
    for t in (depth_map, extrinsics_cam, intrinsics_cam):
        if isinstance(t, torch.Tensor) and t.is_cuda:
            dev = t.device
            break
    if dev is None:
        raise RuntimeError("sailrecon_b200: unproject_depth_map_to_point_map needs at least one CUDA tensor argument "
                           "(no CPU fallback)")

    def dev32(t):
        return torch.as_tensor(t).detach().to(device=dev, dtype=torch.float32)

    d = dev32(depth_map)
    if d.dim() == 4:
        d = d.squeeze(-1)
    pts = ops.unproject_depth(d, dev32(extrinsics_cam), dev32(intrinsics_cam))
    return pts.cpu().numpy().astype(np.float32) if as_numpy else pts

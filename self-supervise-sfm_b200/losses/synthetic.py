"""Synthetic correspondence batches in the format of the reference's data loader (train/datasets/imc2021.py:204-215,
385-392): N frames looking at one random point cloud, every ordered frame pair (or the first `num_pairs` of them) with K
matched pixels, their depths, the K'_to_K matrices that undo the preprocessing, and PREDICTED cameras = the true ones
perturbed, so that the residuals spread over a few pixels to tens of pixels like early in training.  Used by the tests,
the golden-vector script and bench.py (there are no datasets on the box).  numpy, seeded, device-free."""
from __future__ import annotations

import numpy as np


def _rot(rng, sigma):
    w = rng.normal(0.0, sigma, 3)
    th = np.linalg.norm(w) + 1e-12
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def synthetic_batch(num_frames: int, num_points: int, seed: int = 0, num_pairs=None, size: int = 518, noise_px: float = 0.5,
                    cam_sigma: float = 0.02, focal_sigma: float = 0.03, shared_focal: bool = False) -> dict:
    """-> dict of float32 / int64 numpy arrays: pred_intrinsics [N,3,3], pred_extrinsics [N,3,4], K_prime_to_K [N,3,3],
    src_idx, dst_idx [P], src_coords, dst_coords [P,K,2], src_depth, dst_depth [P,K], shared_focal (bool)."""
    rng = np.random.default_rng(seed)
    N, K = num_frames, num_points
    pairs = [(i, j) for i in range(N) for j in range(N) if i != j]
    if num_pairs is not None:
        pairs = [pairs[i % len(pairs)] for i in range(num_pairs)]
    P = len(pairs)
    # ground truth: original-image cameras (focal ~ 1.2 * size) and the crop / resize that maps them to the network input
    f0 = 1.2 * size * (1.0 + (0.0 if shared_focal else 1.0) * rng.normal(0, 0.05, N))
    K_orig = np.zeros((N, 3, 3))
    K_orig[:, 0, 0] = f0
    K_orig[:, 1, 1] = f0
    K_orig[:, 0, 2] = size * (0.5 + rng.normal(0, 0.02, N))
    K_orig[:, 1, 2] = size * (0.5 + rng.normal(0, 0.02, N))
    K_orig[:, 2, 2] = 1
    sc = 1.0 + rng.uniform(-0.2, 0.2, (N, 2))
    sh = rng.uniform(-20, 20, (N, 2))
    K2K = np.zeros((N, 3, 3))                       # K (original) = K'_to_K @ K' (network-input intrinsics)
    K2K[:, 0, 0], K2K[:, 1, 1], K2K[:, 0, 2], K2K[:, 1, 2], K2K[:, 2, 2] = sc[:, 0], sc[:, 1], sh[:, 0], sh[:, 1], 1
    K_net = np.linalg.inv(K2K) @ K_orig
    R = np.stack([_rot(rng, 0.25) for _ in range(N)])
    t = rng.normal(0, 0.3, (N, 3))
    t[:, 2] += 0.0
    E = np.concatenate([R, t[:, :, None]], axis=2)
    # predictions: perturbed network-input intrinsics and poses
    K_pred = K_net.copy()
    K_pred[:, 0, 0] *= 1.0 + rng.normal(0, focal_sigma, N)
    K_pred[:, 1, 1] *= 1.0 + rng.normal(0, focal_sigma, N)
    E_pred = np.concatenate([np.stack([_rot(rng, cam_sigma) for _ in range(N)]) @ R, (t + rng.normal(0, cam_sigma, (N, 3)))[:, :, None]],
                            axis=2)
    src_coords = np.zeros((P, K, 2))
    dst_coords = np.zeros((P, K, 2))
    src_depth = np.zeros((P, K))
    dst_depth = np.zeros((P, K))
    for p, (s, d) in enumerate(pairs):
        # points sampled in the source camera's frustum, 2 .. 6 units away
        z = rng.uniform(2.0, 6.0, K)
        uv = rng.uniform(0.05 * size, 0.95 * size, (K, 2))
        cam = np.concatenate([uv, np.ones((K, 1))], 1) @ np.linalg.inv(K_orig[s]).T * z[:, None]
        world = (cam - t[s]) @ R[s]
        cd = world @ R[d].T + t[d]
        pd = cd @ K_orig[d].T
        src_coords[p] = uv + rng.normal(0, noise_px, (K, 2))
        dst_coords[p] = pd[:, :2] / pd[:, 2:3] + rng.normal(0, noise_px, (K, 2))
        src_depth[p] = z
        dst_depth[p] = np.maximum(cd[:, 2], 0.1)
    f32 = np.float32
    return {"pred_intrinsics": K_pred.astype(f32), "pred_extrinsics": E_pred.astype(f32), "K_prime_to_K": K2K.astype(f32),
            "src_idx": np.array([s for s, _ in pairs], dtype=np.int64), "dst_idx": np.array([d for _, d in pairs], dtype=np.int64),
            "src_coords": src_coords.astype(f32), "dst_coords": dst_coords.astype(f32), "src_depth": src_depth.astype(f32),
            "dst_depth": dst_depth.astype(f32), "shared_focal": bool(shared_focal)}

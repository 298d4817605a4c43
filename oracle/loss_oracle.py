"""TEST INFRASTRUCTURE — CPU restatement (numpy) of the reference's training loss.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may import this file; the product path
(self-supervise-sfm_b200/losses, csrc/loss.cu) never does.

Pinned: tests/golden/loss_*.npz hold loss, tables and camera gradients produced by the reference's OWN, unmodified
`compute_loss` (train/train_imc.py:141-270, imported by oracle/make_golden.py with its unrelated imports stubbed) under
torch autograd on the CPU; tests/test_loss_cpu.py checks this restatement against them.

Forward in float32 in the reference's operation order; backward (the analytic chain through the custom CDF gradient,
cdf_loss.py:6-17) in float64.
"""
from __future__ import annotations

import numpy as np

F = np.float32
EPS = F(1e-6)                      # train/utils/geometry.py:34,245


def pad_poses(E):
    """[..., 3, 4] -> [..., 4, 4] (geometry.py:36-57)."""
    if E.shape[-2:] == (4, 4):
        return E
    last = np.broadcast_to(np.array([0, 0, 0, 1], dtype=E.dtype), E.shape[:-2] + (1, 4))
    return np.concatenate([E[..., :3, :4], last], axis=-2)


def relative_pose(src_E, dst_E):
    """dst @ inverse(src), [P, 4, 4] (geometry.py:252-287)."""
    return pad_poses(dst_E) @ np.linalg.inv(pad_poses(src_E))


def reproject(src_coords, src_depth, src_K, dst_K, rel, dst_depth=None, src_scale=None, dst_scale=None, want_parts=False):
    """geometry.py:89-158 (dst_depth None) and :160-250.  [P, K, 2]."""
    P, K = src_coords.shape[:2]
    z = src_depth if src_scale is None else src_depth * src_scale.reshape(P, 1)
    homo = np.concatenate([src_coords, np.ones((P, K, 1), src_coords.dtype)], -1)
    Ki = np.linalg.inv(src_K)
    s = np.einsum("pij,pkj->pki", Ki, homo) * z[..., None]
    rel4 = pad_poses(rel)
    d = np.einsum("pij,pkj->pki", rel4[:, :3, :3], s) + rel4[:, None, :3, 3]
    h = np.einsum("pij,pkj->pki", dst_K, d)
    if dst_depth is None:
        w = h[..., 2] + EPS.astype(h.dtype)
    else:
        w = (dst_depth if dst_scale is None else dst_depth * dst_scale.reshape(P, 1)) + EPS.astype(h.dtype)
    out = h[..., :2] / w[..., None]
    return (out, dict(Ki=Ki, homo=homo, z=z, s=s, d=d, h=h, w=w, rel4=rel4)) if want_parts else out


def projective_residual(pred, actual):
    """geometry.py:289-303."""
    return np.sqrt(((pred - actual) ** 2).sum(-1))


def smoothing_taps(gradient_smooth, bin_width):
    """cdf_loss.py:63-88 (float32, like the Conv1d weight)."""
    if gradient_smooth <= 0:
        return np.ones(1, F)
    r = max(1, int(gradient_smooth / bin_width))
    idx = np.arange(2 * r + 1, dtype=F) - F(r)
    g = np.exp(F(-0.5) * (idx / F(gradient_smooth / bin_width)) ** 2).astype(F)
    return (g / g.sum(dtype=F)).astype(F)


def _bins(res, min_val, bin_width, half, gpu_division):
    x = res.astype(F) - F(min_val)
    # torch on the CPU divides; on the GPU it multiplies by fp32(1 / bin_width) (profiles/r02_div_probe.txt)
    x = x * F(1.0 / bin_width) if gpu_division else x / F(bin_width)
    if half:
        x = x + F(0.5)
    with np.errstate(invalid="ignore"):
        return np.trunc(x).astype(np.int64)


def cdf_tables(res, weights, src_idx, dst_idx, num_nodes, min_val, max_val, num_bins, gradient_smooth, gpu_division=False):
    """cdf_loss.py:122-186 -> pmf, cdf, pdf [num_nodes, num_bins] float32."""
    bw = (max_val - min_val) / num_bins
    b = _bins(res, min_val, bw, False, gpu_division)
    valid = (b >= 0) & (b < num_bins)
    b = np.clip(b, 0, num_bins - 1)
    contrib = (weights * valid).astype(F)
    hist = np.zeros((num_nodes, num_bins), F)
    total = np.zeros(num_nodes, F)
    P, K = res.shape
    for idx in (src_idx, dst_idx):
        f = np.repeat(idx, K)
        np.add.at(hist, (f, b.reshape(-1)), contrib.reshape(-1))
        np.add.at(total, f, weights.astype(F).reshape(-1))
    pmf = hist / (total[:, None] + F(1e-10))
    cdf = np.cumsum(pmf, axis=1, dtype=F)
    sob = (np.array([-1.0, 0.0, 1.0], F) / F(2.0 * bw)).astype(F)
    raw = np.stack([np.correlate(np.pad(c, 1, mode="reflect"), sob, mode="valid") for c in cdf]).astype(F)
    taps = smoothing_taps(gradient_smooth, bw)
    r = taps.size // 2
    pdf = np.stack([np.correlate(np.pad(c, r, mode="reflect") if r else c, taps, mode="valid") for c in raw]).astype(F)
    return pmf.astype(F), cdf, pdf


def cdf_lookup(res, weights, src_idx, dst_idx, cdf, pdf, min_val, max_val, gpu_division=False):
    """cdf_loss.py:188-242 -> cdf_src, grad_src, cdf_dst, grad_dst [P, K]."""
    num_bins = cdf.shape[1]
    bw = (max_val - min_val) / num_bins
    b = _bins(res, min_val, bw, True, gpu_division)
    valid = (b >= 0) & (b < num_bins) & (weights > 0)
    b = np.clip(b, 0, num_bins - 1)
    out = []
    for idx in (src_idx, dst_idx):
        c = cdf[idx[:, None], b]
        g = pdf[idx[:, None], b] * weights
        out += [np.where(valid, c, F(2.0)).astype(F), np.where(valid, g, F(0.0)).astype(F)]
    return tuple(out)


def compute_loss(K_pred, E_pred, k2k, shared_focal, src_idx, dst_idx, src_coords, dst_coords, src_depth, dst_depth,
                 min_val=0.0, max_val=15.0, num_bins=250, gradient_smooth=0.05, gpu_division=False, want_grads=True):
    """train_imc.py:155-251 -> dict(loss, tables [3, 2, N, bins], d_intrinsics [N, 3, 3], d_extrinsics [N, rows, 4])."""
    N = K_pred.shape[0]
    P, K = src_depth.shape
    K_rec = (k2k @ K_pred).astype(F)
    if shared_focal:
        K_rec = np.repeat(K_rec.mean(0, keepdims=True, dtype=F), N, 0)
    Ks, Kd, Es, Ed = K_rec[src_idx], K_rec[dst_idx], E_pred[src_idx], E_pred[dst_idx]
    rel = relative_pose(Es, Ed).astype(F)
    ones = np.ones((P, K), F)
    u0, parts = reproject(src_coords, src_depth, Ks, Kd, rel, want_parts=True)
    u1 = reproject(src_coords, src_depth, Ks, Kd, rel, dst_depth=dst_depth)
    res = [projective_residual(u0, dst_coords), projective_residual(u1, dst_coords)]
    rl = [np.log1p(r).astype(F) for r in res]
    loss = F(0)
    tables, glog = [], []
    for kind in range(2):
        pmf, cdf, pdf = cdf_tables(rl[kind], ones, src_idx, dst_idx, N, min_val, max_val, num_bins, gradient_smooth, gpu_division)
        cs, gs, cd, gd = cdf_lookup(rl[kind], ones, src_idx, dst_idx, cdf, pdf, min_val, max_val, gpu_division)
        loss = loss + (cs.mean(dtype=F) + cd.mean(dtype=F)) / F(2.0)
        tables.append(np.stack([pmf, cdf, pdf]))
        glog.append((gs.astype(np.float64) + gd) / (4.0 * P * K))
    out = {"loss": F(loss / F(2.0)), "tables": np.stack(tables, 1), "residuals_log": np.stack(rl)}
    if not want_grads:
        return out
    # ---- backward, float64
    D = np.float64
    h, d, s, homo, z, Ki, rel4 = (parts[k].astype(D) for k in ("h", "d", "s", "homo", "z", "Ki", "rel4"))
    g_h = np.zeros_like(h)
    for kind, (u, w) in enumerate(((u0, h[..., 2] + 1e-6), (u1, dst_depth.astype(D) + 1e-6))):
        diff = u.astype(D) - dst_coords
        r = np.sqrt((diff ** 2).sum(-1))
        g_r = glog[kind] / (1.0 + r)
        with np.errstate(divide="ignore", invalid="ignore"):
            g_u = np.where(r[..., None] > 0, g_r[..., None] * diff / r[..., None], 0.0)
        g_h[..., :2] += g_u / w[..., None]
        if kind == 0:
            g_h[..., 2] -= (g_u * h[..., :2]).sum(-1) / w ** 2
    KdD, R = Kd.astype(D), rel4[:, :3, :3]
    g_Kd = np.einsum("pki,pkj->pij", g_h, d)
    g_d = np.einsum("pki,pij->pkj", g_h, KdD)
    g_R = np.einsum("pki,pkj->pij", g_d, s)
    g_t = g_d.sum(1)
    g_c = np.einsum("pki,pij->pkj", g_d, R) * z[..., None]
    g_Ki = np.einsum("pki,pkj->pij", g_c, homo)
    g_Ks = -np.einsum("pji,pjk,plk->pil", Ki, g_Ki, Ki)
    Es4, Ed4 = pad_poses(Es).astype(D), pad_poses(Ed).astype(D)
    Esi = np.linalg.inv(Es4)
    G4 = np.zeros((P, 4, 4))
    G4[:, :3, :3], G4[:, :3, 3] = g_R, g_t
    g_Ed = G4 @ Esi.transpose(0, 2, 1)
    g_Esi = Ed4.transpose(0, 2, 1) @ G4
    g_Es = -Esi.transpose(0, 2, 1) @ g_Esi @ Esi.transpose(0, 2, 1)
    rows = E_pred.shape[1]
    gK, gE = np.zeros((N, 3, 3)), np.zeros((N, rows, 4))
    np.add.at(gK, src_idx, g_Ks)
    np.add.at(gK, dst_idx, g_Kd)
    np.add.at(gE, src_idx, g_Es[:, :rows])
    np.add.at(gE, dst_idx, g_Ed[:, :rows])
    if shared_focal:
        gK = np.repeat(gK.mean(0, keepdims=True), N, 0)
    out["d_intrinsics"] = (k2k.astype(D).transpose(0, 2, 1) @ gK).astype(F)
    out["d_extrinsics"] = gE.astype(F)
    return out

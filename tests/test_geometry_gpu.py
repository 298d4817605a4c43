"""N3: pose_encoding_to_extri_intri + quat_to_mat (utils/pose_enc.py:68-135, utils/rotation.py:14-44) and
unproject_depth_map_to_point_map (utils/geometry.py:19-128) on the kernels of csrc/pose.cu, against the live reference
functions (oracle/_ref) when they travelled, else against a torch / numpy restatement written here from the same lines.
Also CameraHead.trunk_fn / modulate (camera_head.py:123-194)."""
import numpy as np
import pytest
import torch

import ref_live
from golden_common import rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_funcs():
    if ref_live.available():
        ref_live.ref()
        from sailrecon.utils.geometry import unproject_depth_map_to_point_map
        from sailrecon.utils.pose_enc import pose_encoding_to_extri_intri
        from sailrecon.utils.rotation import quat_to_mat
        return pose_encoding_to_extri_intri, quat_to_mat, unproject_depth_map_to_point_map

    def quat_to_mat(q):                                     # rotation.py:26-44
        i, j, k, r = torch.unbind(q, -1)
        two_s = 2.0 / (q * q).sum(-1)
        o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                         two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                         two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
        return o.reshape(q.shape[:-1] + (3, 3))

    def pose(p, hw, pose_encoding_type="absT_quaR_FoV", build_intrinsics=True):     # pose_enc.py:113-135
        ext = torch.cat([quat_to_mat(p[..., 3:7]), p[..., :3, None]], dim=-1)
        H, W = hw
        K = torch.zeros(p.shape[:2] + (3, 3), device=p.device)
        K[..., 0, 0] = (W / 2.0) / torch.tan(p[..., 8] / 2.0)
        K[..., 1, 1] = (H / 2.0) / torch.tan(p[..., 7] / 2.0)
        K[..., 0, 2], K[..., 1, 2], K[..., 2, 2] = W / 2, H / 2, 1.0
        return ext, K

    def unproject(depth, ext, K):                           # geometry.py:19-128
        depth, ext, K = (np.asarray(t.detach().cpu()) for t in (depth, ext, K))
        out = []
        for s in range(depth.shape[0]):
            d = depth[s].squeeze(-1)
            H, W = d.shape
            u, v = np.meshgrid(np.arange(W), np.arange(H))
            cam = np.stack(((u - K[s, 0, 2]) * d / K[s, 0, 0], (v - K[s, 1, 2]) * d / K[s, 1, 1], d), -1).astype(np.float32)
            R, t = ext[s, :, :3], ext[s, :, 3]
            out.append(np.dot(cam, R) + (-R.T @ t))
        return np.stack(out)

    return pose, quat_to_mat, unproject


def test_pose_encoding_to_extri_intri_and_quat_to_mat():
    from sailrecon_b200.utils import pose_encoding_to_extri_intri, quat_to_mat
    r_pose, r_quat, _ = _ref_funcs()
    g = torch.Generator().manual_seed(3)
    p = torch.randn(1, 37, 9, generator=g)
    p[..., 7:] = 0.3 + torch.rand(1, 37, 2, generator=g)                 # FoV (radians), relu'd by the head
    p = p.to(DEV)
    ext, K = pose_encoding_to_extri_intri(p, (392, 518))
    r_ext, r_K = r_pose(p, (392, 518))
    assert ext.shape == (1, 37, 3, 4) and K.shape == (1, 37, 3, 3)
    assert rel_l2(ext, r_ext) < 1e-6 and rel_l2(K, r_K) < 1e-6
    assert torch.equal(K[..., 2, 2], torch.ones_like(K[..., 2, 2])) and float(K[..., 0, 1].abs().max()) == 0.0
    ext2, none = pose_encoding_to_extri_intri(p, build_intrinsics=False)
    assert none is None and torch.equal(ext2, ext)
    q = torch.randn(5, 7, 4, generator=g).to(DEV)
    assert rel_l2(quat_to_mat(q), r_quat(q)) < 1e-6
    with pytest.raises(NotImplementedError):
        pose_encoding_to_extri_intri(p, (10, 10), pose_encoding_type="other")
    e0, k0 = pose_encoding_to_extri_intri(p[:, :0], (392, 518))           # empty input
    assert e0.shape == (1, 0, 3, 4) and k0.shape == (1, 0, 3, 3)


def test_unproject_depth_map_to_point_map():
    from sailrecon_b200.utils import pose_encoding_to_extri_intri, unproject_depth_map_to_point_map
    _, _, r_unproject = _ref_funcs()
    g = torch.Generator().manual_seed(4)
    S, H, W = 3, 37, 53
    p = torch.randn(1, S, 9, generator=g)
    p[..., 7:] = 0.5 + torch.rand(1, S, 2, generator=g)
    ext, K = pose_encoding_to_extri_intri(p.to(DEV), (H, W))
    depth = (0.5 + 4 * torch.rand(S, H, W, 1, generator=g)).to(DEV)
    ref = r_unproject(depth, ext[0], K[0])
    got = unproject_depth_map_to_point_map(depth, ext[0], K[0])
    assert isinstance(got, np.ndarray) and got.shape == (S, H, W, 3) and got.dtype == np.float32
    assert rel_l2(torch.from_numpy(got), torch.from_numpy(np.asarray(ref, dtype=np.float32))) < 1e-5
    dev = unproject_depth_map_to_point_map(depth.squeeze(-1), ext[0], K[0], as_numpy=False)
    assert dev.is_cuda and np.array_equal(dev.cpu().numpy(), got)


def test_camera_head_trunk_fn_and_modulate():
    """trunk_fn(pose_tokens, build_lr_mask(...), n) (camera_head.py:123-186) == what forward() returns on the same tokens;
    modulate == x * (1 + scale) + shift on column slices of a [T, 3C] tensor."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.heads.camera_head import build_lr_mask, modulate
    torch.manual_seed(2)
    head = CameraHead(dim_in=256, trunk_depth=2, num_heads=2).eval().to(DEV)
    with torch.no_grad():
        for prm in head.parameters():
            prm.add_(0.05 * torch.randn_like(prm))
    Na, Nq = 3, 4
    cam = torch.randn(1, Na, 256, device=DEV)
    toks = torch.randn(1, Nq, 7, 256, device=DEV)
    with torch.no_grad():
        ref = head({-1: toks}, cam, num_iterations=3)
        x = torch.nn.functional.layer_norm(torch.cat([cam, toks[:, :, 0]], 1), (256,), head.token_norm.weight, head.token_norm.bias)
        got = head.trunk_fn(x, build_lr_mask(Na + Nq, list(range(Na)), DEV), 3)
    assert len(got) == 3 and got[0].shape == (1, Na + Nq, 9)
    for a, b in zip(got, ref):
        assert rel_l2(a[:, Na:], b) < 1e-5
    mod = torch.randn(11, 3 * 256, device=DEV)
    xx = torch.randn(11, 256, device=DEV)
    shift, scale, _ = mod.chunk(3, dim=-1)
    assert torch.allclose(modulate(xx, shift, scale), xx * (1 + scale) + shift, atol=1e-6)

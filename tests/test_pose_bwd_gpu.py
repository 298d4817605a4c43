"""pose_encoding_to_extri_intri under autograd (training: the loss differentiates the predicted cameras, train_imc.py:156,175):
`sr_pose_to_extri_intri_bwd` against torch autograd through the UNMODIFIED reference function (oracle/_ref,
sailrecon/utils/pose_enc.py:68-135 + utils/rotation.py:14-44).  fp32, 1e-5 relative."""
import pytest
import torch

import ref_live

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_live.available(), reason="oracle/_ref absent (run oracle/make_ref.py in the build container)")]


def test_pose_to_cameras_backward_matches_reference_autograd():
    ref_live.ref()
    from sailrecon.utils.pose_enc import pose_encoding_to_extri_intri as r_fn
    from sailrecon_b200.utils import pose_encoding_to_extri_intri as o_fn
    g = torch.Generator().manual_seed(0)
    pose = torch.randn(1, 7, 9, generator=g)
    pose[..., 7:] = 0.3 + torch.rand(1, 7, 2, generator=g)              # FoV in (0.3, 1.3) rad
    pose = pose.cuda()
    we, wi = torch.randn(1, 7, 3, 4, generator=g).cuda(), torch.randn(1, 7, 3, 3, generator=g).cuda()
    outs = []
    for fn in (r_fn, o_fn):
        p = pose.clone().requires_grad_(True)
        ext, intr = fn(p, (518, 392))
        ((ext * we).sum() + (intr * wi).sum() * 1e-3).backward()
        outs.append((ext.detach(), intr.detach(), p.grad))
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
    assert rel(outs[1][0], outs[0][0]) < 1e-6 and rel(outs[1][1], outs[0][1]) < 1e-6
    assert rel(outs[1][2], outs[0][2]) < 1e-5, rel(outs[1][2], outs[0][2])
    # extrinsics only (build_intrinsics=False) and no_grad (plain kernel path, no grad_fn)
    p = pose.clone().requires_grad_(True)
    ext, intr = o_fn(p, build_intrinsics=False)
    assert intr is None
    (ext * we).sum().backward()
    pr = pose.clone().requires_grad_(True)
    (r_fn(pr, build_intrinsics=False)[0] * we).sum().backward()
    assert rel(p.grad, pr.grad) < 1e-5
    with torch.no_grad():
        assert o_fn(pose, (518, 392))[0].grad_fn is None

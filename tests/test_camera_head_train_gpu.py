"""The camera head under autograd on the B200: `CameraHead.forward` in train() mode (one autograd node, fp32 kernels of
csrc/head_train.cu) against torch autograd through the UNMODIFIED reference CameraHead (oracle/_ref, autocast off, TF32 off)
with the same state_dict and inputs.  Compared: every iteration's pose encoding, the gradient of EVERY parameter and of both
inputs (the aggregator's last-layer tokens and the anchors' camera tokens).

Tolerance: fp32 on both sides, different summation order -> 2e-4 relative L2 per tensor (measured ~1e-6 .. 3e-5)."""
import pytest
import torch

import ref_live

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_live.available(), reason="oracle/_ref absent (run oracle/make_ref.py in the build container)")]
DEV = "cuda"


@pytest.fixture(autouse=True)
def _exact_fp32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def _pair(kw, seed):
    from sailrecon_b200.heads import CameraHead
    torch.manual_seed(seed)
    r_head = ref_live.ref().CameraHead(**kw).to(DEV).train()
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():                      # O(1) LayerScale, non-zero biases / empty pose: every path carries gradient
        for n, p in sorted(r_head.named_parameters()):
            leaf = n.split(".")[-1]
            if leaf == "gamma":
                p.copy_((0.5 + torch.rand(p.shape, generator=g)).to(DEV))
            elif leaf == "bias" or leaf == "empty_pose_tokens":
                p.copy_((0.1 * torch.randn(p.shape, generator=g)).to(DEV))
    o_head = CameraHead(**kw).to(DEV).train()
    o_head.load_state_dict(r_head.state_dict())
    return r_head, o_head


def _run(head, tokens, cam, iters, weights):
    tokens = tokens.clone().requires_grad_(True)
    cam = cam.clone().requires_grad_(True)
    for p in head.parameters():
        p.grad = None
    with torch.autocast("cuda", enabled=False):
        poses = head({-1: tokens}, cam, num_iterations=iters)
    loss = sum((p * w).sum() for p, w in zip(poses, weights) if w is not None)
    loss.backward()
    return [p.detach() for p in poses], tokens.grad, cam.grad, {n: p.grad for n, p in head.named_parameters()}


@pytest.mark.parametrize("kw,Na,Nq,all_iterations", [
    (dict(dim_in=256, trunk_depth=2, num_heads=4), 3, 4, True),
    (dict(dim_in=2048), 12, 12, False),              # the full-size head at the train24 shape, last iteration only (sail_recon.py:122)
    (dict(dim_in=2048), 5, 5, True),                 # IMC2021 batches: 5 images duplicated as anchors and queries
])
def test_camera_head_backward_matches_reference_autograd(kw, Na, Nq, all_iterations):
    r_head, o_head = _pair(kw, seed=3)
    C = kw["dim_in"]
    g = torch.Generator().manual_seed(11)
    tokens = torch.randn(1, Nq, 3, C, generator=g).to(DEV)               # [B, S, P', C]: the head reads token 0 of every frame
    cam = torch.randn(1, Na, C, generator=g).to(DEV)
    iters = 4
    weights = [torch.randn(1, Nq, 9, generator=g).to(DEV) if (all_iterations or i == iters - 1) else None for i in range(iters)]
    rp, r_dt, r_dc, r_g = _run(r_head, tokens, cam, iters, weights)
    op, o_dt, o_dc, o_g = _run(o_head, tokens, cam, iters, weights)
    for a, b in zip(op, rp):
        assert a.shape == b.shape and _rel(a, b) < 1e-4
    errs = {"d_tokens": _rel(o_dt, r_dt), "d_cam": _rel(o_dc, r_dc)}
    assert float(o_dt[:, :, 1:].abs().max()) == 0.0                       # only token 0 of every frame is read
    for n, rg in r_g.items():
        assert (rg is None) == (o_g[n] is None), n
        if rg is not None:
            assert o_g[n].shape == rg.shape, n
            errs[n] = _rel(o_g[n], rg)
    worst = max(errs, key=errs.get)
    print("camera-head backward parity: worst", worst, f"{errs[worst]:.2e}", "over", len(errs), "tensors")
    assert errs[worst] < 2e-4, (worst, errs[worst])
    # last-iteration-only losses leave `empty_pose_tokens` (read by iteration 0 only) without gradient, in both
    assert sum(g_ is not None for g_ in o_g.values()) == len(list(o_head.parameters())) - (0 if all_iterations else 1)


def test_eval_and_train_forward_agree():
    """The un-fused training forward and the fused inference forward of this package give the same encodings."""
    _, o_head = _pair(dict(dim_in=256, trunk_depth=2, num_heads=4), seed=5)
    g = torch.Generator().manual_seed(2)
    tokens, cam = torch.randn(1, 3, 2, 256, generator=g).to(DEV), torch.randn(1, 2, 256, generator=g).to(DEV)
    tr = o_head({-1: tokens.clone().requires_grad_(True)}, cam)
    o_head.eval()
    with torch.no_grad():
        ev = o_head({-1: tokens}, cam)
    for a, b in zip(tr, ev):
        assert a.requires_grad and _rel(a.detach(), b) < 1e-5

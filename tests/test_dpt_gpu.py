"""N1: DPTHead on the kernel library against the LIVE reference DPTHead (oracle/_ref,
sailrecon/heads/dpt_head.py:22-344) with TF32 disabled — the reference's fp32 semantics (sail_recon.py:119).
Tolerance 1e-4 relative L2 on the activated predictions and the confidence."""
import pytest
import torch

import ref_live
from golden_common import rel_l2

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


def _pair(**kw):
    from sailrecon_b200.heads import DPTHead
    ref_live.ref()
    from sailrecon.heads.dpt_head import DPTHead as RefDPT
    torch.manual_seed(11)
    with torch.device(DEV):
        r = RefDPT(dim_in=2048, **kw).eval().to(DEV)
        o = DPTHead(dim_in=2048, **kw).eval()
    with torch.no_grad():
        for p in r.parameters():                      # biases / norm parameters away from their trivial init
            if p.dim() == 1:
                p.add_(0.05 * torch.randn_like(p))
    assert {k: tuple(v.shape) for k, v in o.state_dict().items()} == {k: tuple(v.shape) for k, v in r.state_dict().items()}
    o.load_state_dict(r.state_dict())
    return r, o


def _tokens(S, P, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return {i: torch.randn(1, S, P, 2048, device=DEV, generator=g) for i in (4, 11, 17, 23)}


@pytest.mark.parametrize("H,W,S,chunk", [(518, 518, 2, 8), (392, 518, 3, 2)])
@pytest.mark.parametrize("kind", ["point", "depth"])
def test_dpt_head_matches_reference_fp32(H, W, S, chunk, kind):
    kw = dict(output_dim=4, activation="inv_log", conf_activation="expp1") if kind == "point" else \
        dict(output_dim=2, activation="exp", conf_activation="expp1")           # sail_recon.py:46-66
    r, o = _pair(**kw)
    P = 5 + (H // 14) * (W // 14)
    toks = _tokens(S, P, 3)
    images = torch.rand(1, S, 3, H, W, device=DEV)
    with torch.no_grad():
        rp, rc = r(toks, images=images, patch_start_idx=5, frames_chunk_size=chunk)
        op, oc = o(toks, images=images, patch_start_idx=5, frames_chunk_size=chunk)
    assert op.shape == rp.shape and oc.shape == rc.shape and op.dtype == torch.float32
    e_p, e_c = rel_l2(op, rp), rel_l2(oc, rc)
    print("dpt parity", kind, (H, W), f"pred {e_p:.2e} conf {e_c:.2e}")
    assert e_p < 1e-4 and e_c < 1e-4


def test_dpt_row_kernels():
    """im2col + hi/lo split GEMM == F.conv2d (3x3 stride 1 / 2, 1x1), bilinear align_corners, ConvTranspose(k = stride)."""
    import torch.nn.functional as F
    from sailrecon_b200 import ops
    from sailrecon_b200.heads.dpt_head import _conv_weights, conv_rows
    g = torch.Generator(device=DEV).manual_seed(5)
    x = torch.randn(2, 19, 23, 128, device=DEV, generator=g)
    for k, s, cout in ((3, 1, 256), (3, 2, 128), (1, 1, 384), (3, 1, 32)):
        conv = torch.nn.Conv2d(128, cout, k, s, k // 2).to(DEV)
        ref = F.conv2d(x.permute(0, 3, 1, 2), conv.weight, conv.bias, s, k // 2).permute(0, 2, 3, 1)
        got = conv_rows(conv, x)
        assert got.shape == ref.shape and rel_l2(got, ref) < 2e-5, (k, s, cout)
        got_r = conv_rows(conv, x, relu_in=True)
        assert rel_l2(got_r, F.conv2d(x.relu().permute(0, 3, 1, 2), conv.weight, conv.bias, s, k // 2).permute(0, 2, 3, 1)) < 2e-5
    # the same 3x3 convolutions WITHOUT an im2col matrix: zero-bordered split layout + shifted GEMM (interior of the padded output)
    from sailrecon_b200.heads.dpt_head import conv3_pad
    for cout in (256, 128, 32):
        conv = torch.nn.Conv2d(128, cout, 3, 1, 1).to(DEV)
        ref = F.conv2d(x.permute(0, 3, 1, 2), conv.weight, conv.bias, 1, 1).permute(0, 2, 3, 1)
        got = conv3_pad(conv, x, in_pad=False)
        assert got.shape[:3] == (2, 21, 25) and rel_l2(got[:, 1:-1, 1:-1, :cout], ref) < 2e-5, cout
        xp = ops.map_add(x, False, None, False, True)                       # the same input carried in the padded layout
        assert torch.equal(xp[:, 1:-1, 1:-1], x) and float(xp[:, 0].abs().max()) == 0.0 and float(xp[:, :, -1].abs().max()) == 0.0
        got_p = conv3_pad(conv, xp, in_pad=True, relu_in=True)
        ref_r = F.conv2d(x.relu().permute(0, 3, 1, 2), conv.weight, conv.bias, 1, 1).permute(0, 2, 3, 1)
        assert rel_l2(got_p[:, 1:-1, 1:-1, :cout], ref_r) < 2e-5
        if cout == 256:
            res = torch.randn(2, 21, 25, 256, device=DEV, generator=g)
            got_res = conv3_pad(conv, xp, in_pad=True, resid=res)
            assert rel_l2(got_res[:, 1:-1, 1:-1], ref + res[:, 1:-1, 1:-1]) < 2e-5
    y = torch.randn(2, 19, 23, 128, device=DEV, generator=g)
    yp = ops.map_add(y, False, None, False, True)
    s_ = ops.map_add(x, False, yp, True, False, relu=True)
    assert torch.allclose(s_, (x + y).relu(), atol=1e-6)
    up_p = ops.bilinear_nhwc(yp, 37, 50, in_pad=True)
    assert rel_l2(up_p, F.interpolate(y.permute(0, 3, 1, 2), size=(37, 50), mode="bilinear", align_corners=True).permute(0, 2, 3, 1)) < 1e-6
    up = ops.bilinear_nhwc(x, 37, 50)
    ref = F.interpolate(x.permute(0, 3, 1, 2), size=(37, 50), mode="bilinear", align_corners=True).permute(0, 2, 3, 1)
    assert rel_l2(up, ref) < 1e-6
    ct = torch.nn.ConvTranspose2d(128, 128, 4, 4).to(DEV)
    w_hh, w_lo, _, n = _conv_weights(ct)
    gm = torch.empty(2 * 19 * 23, w_lo.shape[0], device=DEV)
    ops.gemm_split(ops.im2col_split(x, 1), w_hh, w_lo, None, gm)
    got = ops.pixel_shuffle(gm, ct.bias.detach(), 2, 19, 23, 128, 4)
    ref = F.conv_transpose2d(x.permute(0, 3, 1, 2), ct.weight, ct.bias, 4).permute(0, 2, 3, 1)
    assert rel_l2(got, ref) < 2e-5

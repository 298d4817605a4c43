"""N4 on the B200: csrc/loss.cu through `sailrecon_b200.losses` against (a) golden vectors of the reference's own
compute_loss + autograd (CPU torch, tests/golden/loss_*.npz), (b) the numpy oracle on the same inputs, and (c) the LIVE
reference (oracle/_ref/train, unmodified) on the same GPU — its compute_loss as shipped, and the same function running on
this package's kernels through the name-for-name mirrors.

Tolerances (fp32 everywhere): geometry building blocks 1e-5 relative (different summation order of 3 x 3 / 4 x 4 products,
adjugate / Gauss-Jordan inverses vs LU); histograms EXACT against the live reference on the GPU (same bin arithmetic, unit
weights); loss 1e-4 absolute and camera gradients 2e-3 relative wherever a bin edge can flip a point (see test_loss_cpu.py)."""
import numpy as np
import pytest
import torch

import ref_live
from loss_common import BATCH_KEYS, LOSS_GOLDEN, load_loss_golden, oracle_loss, rel

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _t(b):
    return {k: (torch.from_numpy(v).to(DEV) if isinstance(v, np.ndarray) else v) for k, v in b.items() if not k.startswith("ref_")}


def _module(t, cdf_kw, N):
    from sailrecon_b200.losses import CDFLossIndexPytorch
    return CDFLossIndexPytorch(src_indices=t["src_idx"], dst_indices=t["dst_idx"], num_nodes=N, **cdf_kw).to(DEV)


def _ours_fused(t, cdf_kw, scale=1.0, do_record=False):
    from sailrecon_b200.losses import compute_loss
    K = t["pred_intrinsics"].clone().requires_grad_(True)
    E = t["pred_extrinsics"].clone().requires_grad_(True)
    preds = [{"intrinsic": K[i:i + 1], "extrinsic": E[i:i + 1]} for i in range(K.shape[0])]
    batch = {k: v for k, v in t.items() if not k.startswith("pred_")}
    res = compute_loss(preds, batch, DEV, _module(t, cdf_kw, K.shape[0]), do_record)
    (res["loss"] * scale).backward()
    return res, K.grad, E.grad


@pytest.mark.parametrize("name", LOSS_GOLDEN)
def test_fused_loss_matches_reference_golden_and_oracle(name):
    g, meta = load_loss_golden(name)
    res, dK, dE = _ours_fused(_t(g), meta["cdf"], do_record=True)
    loss = float(res["loss"])
    assert abs(loss - float(g["ref_loss"])) < 1e-4, (loss, float(g["ref_loss"]))
    assert rel(dK.cpu().numpy(), g["ref_d_intrinsics"]) < 2e-3
    assert rel(dE.cpu().numpy(), g["ref_d_extrinsics"]) < 2e-3
    assert np.abs(res["plot_data"]["frame_cdfs"] - g["ref_frame_cdfs"]).max() < 2e-3      # one point flipping = 1 / (2 (N-1) K)
    o = oracle_loss(g, meta["cdf"], gpu_division=True)
    assert abs(loss - float(o["loss"])) < 1e-4
    assert rel(dK.cpu().numpy(), o["d_intrinsics"]) < 2e-3 and rel(dE.cpu().numpy(), o["d_extrinsics"]) < 2e-3


def test_backward_scales_with_the_incoming_gradient():
    g, meta = load_loss_golden("loss_imc")
    t = _t(g)
    _, dK1, dE1 = _ours_fused(t, meta["cdf"])
    _, dK2, dE2 = _ours_fused(t, meta["cdf"], scale=1024.0)               # GradScaler (train_imc.py:380)
    assert torch.allclose(dK2, dK1 * 1024.0, rtol=1e-6, atol=0) and torch.allclose(dE2, dE1 * 1024.0, rtol=1e-6, atol=0)


def test_building_blocks_match_oracle():
    import loss_oracle as O
    from sailrecon_b200 import losses
    g, meta = load_loss_golden("loss_imc")
    t = _t(g)
    s, d = g["src_idx"], g["dst_idx"]
    E = g["pred_extrinsics"]
    relp = losses.compute_relative_pose(t["pred_extrinsics"][t["src_idx"]], t["pred_extrinsics"][t["dst_idx"]])
    o_rel = O.relative_pose(E[s], E[d])
    assert relp.shape == (12, 4, 4) and rel(relp.cpu().numpy(), o_rel) < 1e-5
    E44 = torch.from_numpy(O.pad_poses(E)).to(DEV)
    assert rel(losses.compute_relative_pose(E44[t["src_idx"]], E44[t["dst_idx"]]).cpu().numpy(), o_rel) < 1e-5
    K = (g["K_prime_to_K"] @ g["pred_intrinsics"]).astype(np.float32)
    Kt = torch.from_numpy(K).to(DEV)
    ones = torch.ones(12, 1, device=DEV)
    u0, m0 = losses.backproject_and_reproject(t["src_coords"], t["src_depth"], Kt[t["src_idx"]], Kt[t["dst_idx"]], relp, ones)
    u1, m1 = losses.backproject_and_reproject_with_approximation(t["src_coords"], t["src_depth"], t["dst_depth"], Kt[t["src_idx"]],
                                                                  Kt[t["dst_idx"]], relp[:, :3], 2 * ones, 0.5 * ones)
    assert m0.dtype == torch.bool and m0.all() and m0.shape == (12, 300) and m1.all()
    o0 = O.reproject(g["src_coords"], g["src_depth"], K[s], K[d], o_rel.astype(np.float32))
    o1 = O.reproject(g["src_coords"], g["src_depth"], K[s], K[d], o_rel.astype(np.float32), dst_depth=g["dst_depth"],
                     src_scale=np.full(12, 2, np.float32), dst_scale=np.full(12, 0.5, np.float32))
    assert rel(u0.cpu().numpy(), o0) < 1e-5 and rel(u1.cpu().numpy(), o1) < 1e-5
    r0 = losses.compute_projective_residual(u0, t["dst_coords"])
    assert rel(r0.cpu().numpy(), O.projective_residual(o0, g["dst_coords"])) < 1e-4
    # CDF module on the oracle's residuals: identical inputs -> tables to rounding, lookups exact where the tables are
    res = np.log1p(O.projective_residual(o0, g["dst_coords"])).astype(np.float32)
    w = (np.random.default_rng(0).uniform(size=res.shape) > 0.2).astype(np.float32) * 1.5      # zero weights = invalid lookups
    mod = _module(t, meta["cdf"], 4)
    pmf, cdf, pdf = mod.compute_weighted_pdf_cdf_vectorized(torch.from_numpy(res).to(DEV), torch.from_numpy(w).to(DEV))
    o_pmf, o_cdf, o_pdf = O.cdf_tables(res, w, s, d, 4, gpu_division=True, **meta["cdf"])
    assert np.abs(pmf.cpu().numpy() - o_pmf).max() < 1e-6 and np.abs(cdf.cpu().numpy() - o_cdf).max() < 2e-6
    assert np.abs(pdf.cpu().numpy() - o_pdf).max() < 2e-5 * np.abs(o_pdf).max()
    outs = mod.compute_weighted_cdf_forward_backward_vectorized(torch.from_numpy(res).to(DEV), torch.from_numpy(w).to(DEV), cdf, pdf)
    o_outs = O.cdf_lookup(res, w, s, d, cdf.cpu().numpy(), pdf.cpu().numpy(), 0.0, 15.0, gpu_division=True)
    for a, b in zip(outs, o_outs):
        assert np.array_equal(a.cpu().numpy(), b)
    assert (outs[0].cpu().numpy()[w == 0] == 2.0).all() and (outs[1].cpu().numpy()[w == 0] == 0.0).all()


def _ref_train():
    if not ref_live.available():
        pytest.skip("oracle/_ref absent (run oracle/make_ref.py in the build container)")
    import make_ref
    import ref_train_import
    return ref_train_import, ref_train_import.import_train_imc(make_ref.import_path())


@pytest.fixture
def _exact_fp32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("N,K,shared,rows", [(5, 10000, False, 3), (4, 1000, True, 3), (3, 777, False, 3)])
def test_against_live_reference_compute_loss(_exact_fp32, N, K, shared, rows):
    """The reference's compute_loss (train_imc.py:141-270) as shipped, on the GPU, vs (1) the fused kernel path and (2) the
    SAME reference function with its four geometry functions and its CDF module swapped for this package's mirrors."""
    from sailrecon_b200 import losses
    from sailrecon_b200.losses.synthetic import synthetic_batch
    rti, train_imc = _ref_train()
    cdf_kw = dict(min_val=0.0, max_val=15.0, num_bins=250, gradient_smooth=0.05)            # train_imc.py:333-347
    b = synthetic_batch(N, K, seed=N, shared_focal=shared)
    ref = rti.run_reference_loss(train_imc, b, DEV, cdf_kw)
    t = _t(b)
    res, dK, dE = _ours_fused(t, cdf_kw, do_record=True)
    assert abs(float(res["loss"]) - float(ref["loss"])) < 1e-4
    assert rel(dK.cpu().numpy(), ref["d_intrinsics"]) < 2e-3 and rel(dE.cpu().numpy(), ref["d_extrinsics"]) < 2e-3
    assert np.abs(res["plot_data"]["frame_cdfs"] - ref["frame_cdfs"]).max() < 1e-3
    # (2) the reference's function body on our kernels
    names = ("backproject_and_reproject", "backproject_and_reproject_with_approximation", "compute_projective_residual",
             "compute_relative_pose", "CDFLossIndexPytorch")
    saved = {n: getattr(train_imc, n) for n in names}
    try:
        for n in names:
            setattr(train_imc, n, getattr(losses, n))
        mix = rti.run_reference_loss(train_imc, b, DEV, cdf_kw)
    finally:
        for n, f in saved.items():
            setattr(train_imc, n, f)
    assert abs(float(mix["loss"]) - float(ref["loss"])) < 1e-4
    assert rel(mix["d_intrinsics"], ref["d_intrinsics"]) < 2e-3 and rel(mix["d_extrinsics"], ref["d_extrinsics"]) < 2e-3
    # the two kernel paths of this package agree with each other far below the bin-flip tolerance
    assert abs(float(mix["loss"]) - float(res["loss"])) < 2e-5
    assert rel(mix["d_intrinsics"], dK.cpu().numpy()) < 5e-4 and rel(mix["d_extrinsics"], dE.cpu().numpy()) < 5e-4


def test_histograms_are_exact_against_the_live_reference_module(_exact_fp32):
    """Same residual tensor into the reference's CDFLossIndexPytorch and ours, on the GPU: unit-weight histograms are sums of
    ones -> the PMF must be bit-identical; forward values and the custom gradient follow."""
    from sailrecon_b200.losses import CDFLossIndexPytorch
    rti, train_imc = _ref_train()
    rng = np.random.default_rng(3)
    P, K, N = 7, 4099, 4
    src = torch.tensor([0, 1, 2, 3, 0, 2, 1], device=DEV)
    dst = torch.tensor([1, 2, 3, 0, 2, 2, 3], device=DEV)           # includes a pair of a frame with itself
    res = torch.from_numpy(rng.gamma(2.0, 1.2, (P, K)).astype(np.float32)).to(DEV)
    res[0, :5] = torch.tensor([0.0, 14.97, 15.0, 16.0, 0.06], device=DEV)                  # edges, out of range
    w = torch.ones(P, K, device=DEV)
    kw = dict(min_val=0.0, max_val=15.0, num_bins=250, gradient_smooth=0.05)
    r_mod = train_imc.CDFLossIndexPytorch(src_indices=src, dst_indices=dst, num_nodes=N, **kw).to(DEV)
    o_mod = CDFLossIndexPytorch(src_indices=src, dst_indices=dst, num_nodes=N, **kw).to(DEV)
    r_pmf, r_cdf, r_pdf = r_mod.compute_weighted_pdf_cdf_vectorized(res, w)
    o_pmf, o_cdf, o_pdf = o_mod.compute_weighted_pdf_cdf_vectorized(res, w)
    assert torch.equal(o_pmf, r_pmf)
    assert (o_cdf - r_cdf).abs().max() < 2e-6 and (o_pdf - r_pdf).abs().max() < 2e-5 * r_pdf.abs().max()
    rr = res.clone().requires_grad_(True)
    ro = res.clone().requires_grad_(True)
    r_s, r_d = r_mod(rr, w)
    o_s, o_d = o_mod(ro, w)
    assert (o_s - r_s).abs().max() < 2e-6 and (o_d - r_d).abs().max() < 2e-6
    assert float(o_s[0, 3]) == 2.0 and float(o_s[0, 2]) == 2.0                          # 15.0 rounds to bin 250: out of range
    ((r_s.mean() + r_d.mean()) / 2).backward()
    ((o_s.mean() + o_d.mean()) / 2).backward()
    assert rel(ro.grad.cpu().numpy(), rr.grad.cpu().numpy()) < 1e-4


def test_edge_shapes():
    """One pair, K = 1; 4 x 4 extrinsics; K straddling the 1024-point blocks; no K'_to_K (identity)."""
    from sailrecon_b200 import losses
    from sailrecon_b200.losses.synthetic import synthetic_batch
    import loss_oracle as O
    cdf_kw = dict(min_val=0.0, max_val=15.0, num_bins=250, gradient_smooth=0.05)
    for N, K, P in ((2, 1, 1), (3, 1025, None), (2, 2048, 2)):
        b = synthetic_batch(N, K, seed=K, num_pairs=P)
        o = oracle_loss(b, cdf_kw, gpu_division=True)
        t = _t(b)
        res, dK, dE = _ours_fused(t, cdf_kw)
        assert abs(float(res["loss"]) - float(o["loss"])) < 2e-4, (N, K, P)
        if K > 1:
            assert rel(dK.cpu().numpy(), o["d_intrinsics"]) < 5e-3 and rel(dE.cpu().numpy(), o["d_extrinsics"]) < 5e-3
        # 4 x 4 poses give the same loss and the same gradient in the top three rows
        E44 = torch.from_numpy(O.pad_poses(b["pred_extrinsics"])).to(DEV).requires_grad_(True)
        Kp = t["pred_intrinsics"].clone().requires_grad_(True)
        l44 = losses.reprojection_cdf_loss(Kp, E44, t["K_prime_to_K"], False, t["src_idx"], t["dst_idx"], t["src_coords"],
                                           t["dst_coords"], t["src_depth"], t["dst_depth"], _module(t, cdf_kw, N))
        l44.backward()
        assert float(l44) == float(res["loss"]) and torch.allclose(E44.grad[:, :3], dE, rtol=1e-5, atol=1e-9)
    with pytest.raises(NotImplementedError):
        losses.backproject_and_reproject(t["src_coords"].clone().requires_grad_(True), t["src_depth"], t["pred_intrinsics"][:2],
                                         t["pred_intrinsics"][:2], E44[:2].detach(), torch.ones(2, 1, device=DEV))


def test_pair_index_outside_the_frames_raises_instead_of_corrupting_memory():
    """A pair list that names a frame the predictions do not have (batch / model mismatch): the reference dies on a device-side
    assert inside scatter_add_ / gather (cdf_loss.py:150-160); here it is an IndexError BEFORE any kernel runs."""
    from sailrecon_b200 import ops
    g, meta = load_loss_golden("loss_imc")
    t = _t(g)
    N = t["pred_intrinsics"].shape[0]
    bad = dict(t)
    bad["dst_idx"] = t["dst_idx"].clone()
    bad["dst_idx"][-1] = N
    with pytest.raises(IndexError, match="pair indices"):
        _ours_fused(bad, meta["cdf"])
    neg = dict(t)
    neg["src_idx"] = t["src_idx"].clone()
    neg["src_idx"][0] = -1
    with pytest.raises(IndexError, match="pair indices"):
        _ours_fused(neg, meta["cdf"])
    P, K = t["src_depth"].shape
    r = torch.rand(P, K, device=DEV)
    sk = torch.ones(1, device=DEV)
    with pytest.raises(IndexError):                                       # num_nodes smaller than max index + 1
        ops.cdf_tables(r, None, t["src_idx"], t["dst_idx"], N - 1, 0.0, 50.0, 50, 25.0, sk)
    pmf, cdf, pdf = ops.cdf_tables(r, None, t["src_idx"], t["dst_idx"], N, 0.0, 50.0, 50, 25.0, sk)
    with pytest.raises(IndexError):
        ops.cdf_lookup(r, None, t["src_idx"], t["dst_idx"], cdf[:N - 1].contiguous(), pdf[:N - 1].contiguous(), 0.0, 50.0)
    ops.cdf_lookup(r, None, t["src_idx"], t["dst_idx"], cdf, pdf, 0.0, 50.0)    # the same (cached-OK) lists at the right N still run
    # an in-place edit of a list that passed before is seen (version counter)
    idx = t["src_idx"].clone()
    ops.cdf_lookup(r, None, idx, t["dst_idx"], cdf, pdf, 0.0, 50.0)
    idx[0] = N + 3
    with pytest.raises(IndexError):
        ops.cdf_lookup(r, None, idx, t["dst_idx"], cdf, pdf, 0.0, 50.0)

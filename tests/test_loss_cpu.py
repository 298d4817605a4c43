"""CPU: the numpy oracle of the training loss (oracle/loss_oracle.py) against golden vectors produced by the reference's
own compute_loss + torch autograd (train/train_imc.py:141-270, oracle/make_golden.py::run_loss_case), and the host side of
the `sailrecon_b200.losses` mirrors (constructor, filters, state_dict, failure without a GPU).

Tolerances: the loss is a mean of per-point CDF lookups in tables with a FINITE number of bins — a residual that lands
within fp32 rounding of a bin edge may fall on either side depending on the summation order of a 3 x 3 product, moving the
loss by O(1 / (4 P K)).  loss_imc reproduces the reference to the last bit (no point on an edge); loss_shared (64 bins, 1 542
points, shared-focal average) differs by 2.3e-5 in the loss and 5e-4 in the gradients, hence 1e-4 / 2e-3."""
import numpy as np
import pytest
import torch

from loss_common import LOSS_GOLDEN, load_loss_golden, oracle_loss, rel


@pytest.mark.parametrize("name", LOSS_GOLDEN)
def test_oracle_matches_reference_compute_loss(name):
    g, meta = load_loss_golden(name)
    o = oracle_loss(g, meta["cdf"])
    assert abs(float(o["loss"]) - float(g["ref_loss"])) < 1e-4
    assert rel(o["d_intrinsics"], g["ref_d_intrinsics"]) < 2e-3
    assert rel(o["d_extrinsics"], g["ref_d_extrinsics"]) < 2e-3
    assert np.abs(o["tables"][1, 0] - g["ref_frame_cdfs"]).max() < 1e-6          # per-frame CDF of the exact residuals
    assert np.abs(o["tables"][2, 0] - g["ref_frame_pdfs"]).max() < 2e-5 * max(1.0, np.abs(g["ref_frame_pdfs"]).max())
    if name == "loss_imc":
        assert float(o["loss"]) == float(g["ref_loss"])


def test_oracle_building_blocks_are_consistent():
    """relative pose of a camera with itself is the identity; exact and approximated reprojection agree when the matched depth
    equals the projected one; CDF tables are monotone, end at (in-range weight) / (total weight), and the lookup of the
    largest in-range residual returns that."""
    import loss_oracle as O
    g, meta = load_loss_golden("loss_imc")
    E = g["pred_extrinsics"]
    assert np.abs(O.relative_pose(E, E) - np.eye(4)).max() < 1e-5
    s, d = g["src_idx"], g["dst_idx"]
    K = (g["K_prime_to_K"] @ g["pred_intrinsics"]).astype(np.float32)
    relp = O.relative_pose(E[s], E[d]).astype(np.float32)
    u0, parts = O.reproject(g["src_coords"], g["src_depth"], K[s], K[d], relp, want_parts=True)
    u1 = O.reproject(g["src_coords"], g["src_depth"], K[s], K[d], relp, dst_depth=parts["h"][..., 2])
    assert np.abs(u0 - u1).max() < 1e-3
    res = np.log1p(O.projective_residual(u0, g["dst_coords"])).astype(np.float32)
    w = np.ones_like(res)
    pmf, cdf, pdf = O.cdf_tables(res, w, s, d, 4, **meta["cdf"])
    assert (np.diff(cdf, axis=1) >= 0).all() and np.allclose(cdf[:, -1], 1.0, atol=1e-5) and np.allclose(pmf.sum(1), 1.0, atol=1e-5)
    cs, gs, cd, gd = O.cdf_lookup(res, w, s, d, cdf, pdf, meta["cdf"]["min_val"], meta["cdf"]["max_val"])
    assert cs.min() >= 0 and cs.max() <= 1.0 + 1e-5 and (gs >= -1e-4).all()


def test_cdf_module_mirror_host_side():
    from sailrecon_b200.losses import CDFLossIndexPytorch
    import loss_oracle as O
    src, dst = torch.tensor([0, 1, 2, 0]), torch.tensor([1, 2, 0, 2])
    m = CDFLossIndexPytorch(0.0, 15.0, 250, src, dst, gradient_smooth=0.05)
    assert m.num_nodes == 3
    assert sorted(m.state_dict()) == ["dst_indices", "grad_conv.weight", "smooth_conv.weight", "src_indices"]   # the reference's keys
    assert not any(p.requires_grad for p in m.parameters())
    bw = 15.0 / 250
    assert np.array_equal(m.smooth_conv.weight.reshape(-1).numpy(), O.smoothing_taps(0.05, bw))
    assert m.smooth_conv.padding_mode == "reflect" and m.grad_conv.padding_mode == "reflect"
    assert np.allclose(m.grad_conv.weight.reshape(-1).numpy(), np.array([-1, 0, 1], np.float32) / np.float32(2 * bw))
    ident = CDFLossIndexPytorch(0.0, 1.0, 10, src, dst, gradient_smooth=0.0, num_nodes=7)
    assert ident.num_nodes == 7 and ident.smooth_conv.weight.numel() == 1 and float(ident.smooth_conv.weight) == 1.0
    m.set_pairs(torch.tensor([3, 4]), torch.tensor([4, 3]))
    assert m.num_nodes == 5 and m.src_indices.tolist() == [3, 4]
    with pytest.raises(IndexError):                    # the reference indexes its pair buffer with every pair number
        m._check(torch.zeros(3, 8), torch.zeros(3, 8))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_loss_has_no_cpu_path():
    from sailrecon_b200 import losses
    g, meta = load_loss_golden("loss_imc")
    t = {k: torch.from_numpy(v) for k, v in g.items() if isinstance(v, np.ndarray) and not k.startswith("ref_")}
    m = losses.CDFLossIndexPytorch(src_indices=t["src_idx"], dst_indices=t["dst_idx"], **meta["cdf"])
    with pytest.raises((ValueError, RuntimeError)):
        m(torch.zeros(12, 300), torch.ones(12, 300))
    with pytest.raises((ValueError, RuntimeError)):
        losses.compute_relative_pose(t["pred_extrinsics"], t["pred_extrinsics"])
    preds = [{"intrinsic": t["pred_intrinsics"][i:i + 1], "extrinsic": t["pred_extrinsics"][i:i + 1]} for i in range(4)]
    with pytest.raises((ValueError, RuntimeError)):
        losses.compute_loss(preds, {**t, "shared_focal": False}, "cpu", m)

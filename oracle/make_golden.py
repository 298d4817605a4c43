"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference) on seeded
inputs.  Runs only in the build container; the fixtures travel, the reference does not.

    python oracle/make_golden.py            # rewrites every case in oracle/golden_common.GOLDEN_CASES

Each .npz stores: output_dict[-1] and the intermediate layers, cam_token_last_layer, the four camera-head pose
encodings, the scene-token indices the reference drew (bit-exact check), the token positions, and the parameter
shapes (values are regenerated from the seed by golden_common.make_params).
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, "/root/reference")
from golden_common import GOLDEN_CASES, make_images, make_params  # noqa: E402

from sailrecon.heads.camera_head import CameraHead  # noqa: E402
from sailrecon.models.aggregator import Aggregator  # noqa: E402


def run_case(name, case):
    torch.manual_seed(0)
    agg = Aggregator(**case["agg"]).eval()
    head = CameraHead(**case["head"]).eval()
    sd = {"agg." + k: v for k, v in agg.state_dict().items()}
    sd.update({"head." + k: v for k, v in head.state_dict().items()})
    shapes = {k: list(v.shape) for k, v in sd.items()}
    params = make_params(shapes, case["seed"])
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    S, Na = case["S"], case["Na"]
    if name == "tiny_clamp":                      # anchors are frames 0, 2, 3; the query is frame 1
        nrl, rl = [0, 2, 3], [1]
    else:
        nrl, rl = list(range(Na)), list(range(Na, S))
    images = make_images(S, case["H"], case["W"], case["seed"] + 100)

    # record the indices the reference draws (aggregator.py:617-621) by wrapping torch.randperm
    drawn = []
    orig = torch.randperm

    def spy(n, *a, **k):
        r = orig(n, *a, **k)
        drawn.append(r.clone())
        return r

    torch.randperm = spy
    try:
        with torch.no_grad():
            agg.generator.manual_seed(1234)
            out, psi, cam = agg(images, nrl, rl, fix_rank=case["fix_rank"])
            poses = head(out, cam)
    finally:
        torch.randperm = orig
    rank = agg.rank
    sel = torch.stack([d[:rank] for d in drawn]).reshape(case["agg"]["depth"], len(nrl), rank)
    arrays = {f"out_{k}": v.numpy() for k, v in out.items()}
    arrays["cam_token_last_layer"] = cam.numpy()
    for i, pz in enumerate(poses):
        arrays[f"pose_{i}"] = pz.numpy()
    arrays["scene_idx"] = sel.numpy().astype(np.int64)
    arrays["positions"] = agg.position_getter(1, case["H"] // 14, case["W"] // 14, "cpu")[0].numpy()
    arrays["meta"] = np.frombuffer(json.dumps(dict(shapes=shapes, no_reloc_list=nrl, reloc_list=rl, rank=rank,
                                                    patch_start_idx=psi)).encode(), dtype=np.uint8)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", name + ".npz")
    np.savez_compressed(path, **arrays)
    print(name, {k: v.shape for k, v in arrays.items() if k != "meta"}, os.path.getsize(path), "bytes")


def run_cache_case():
    """tests/golden/tiny_cache.npz: the reference's CACHED path (aggregator.py:435-578 + attention.py:84-100) on the
    tiny_conv model built with kv_cache=True: prime with the 2 anchors (SailRecon.tmp_forward, reloc_list = []), then
    forward_with_cache on the 2 query frames and the camera head.  The reference parks the cache with .cpu() and fetches it
    with .cuda(); this container has no GPU, so Tensor.cuda is patched to the identity FOR THE CALL (the reference source
    is untouched and every arithmetic op is the reference's own, on CPU)."""
    name, case = "tiny_cache", GOLDEN_CASES["tiny_conv"]
    torch.manual_seed(0)
    agg = Aggregator(**case["agg"], kv_cache=True).eval()
    head = CameraHead(**case["head"]).eval()
    sd = {"agg." + k: v for k, v in agg.state_dict().items()}
    sd.update({"head." + k: v for k, v in head.state_dict().items()})
    shapes = {k: list(v.shape) for k, v in sd.items()}
    params = make_params(shapes, case["seed"])
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    S, Na = case["S"], case["Na"]
    images = make_images(S, case["H"], case["W"], case["seed"] + 100)
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        with torch.no_grad():
            agg.generator.manual_seed(1234)
            out0, psi, cam = agg(images[:, :Na], list(range(Na)), [], fix_rank=case["fix_rank"])
            assert len(out0) == 0 and agg.global_reloc_blocks[0].attn.k_cache is not None
            out, psi = agg.forward_with_cache(images[:, Na:], fix_rank=case["fix_rank"])
            poses = head(out, cam)
            single, _ = agg.forward_with_cache(images[:, Na:Na + 1], fix_rank=case["fix_rank"])
    finally:
        torch.Tensor.cuda = orig_cuda
    arrays = {f"out_{k}": v.numpy() for k, v in out.items()}
    arrays["single_out_-1"] = single[-1].numpy()
    arrays["cam_token_last_layer"] = cam.numpy()
    arrays["k_cache_0"] = agg.global_reloc_blocks[0].attn.k_cache.numpy()
    arrays["v_cache_0"] = agg.global_reloc_blocks[0].attn.v_cache.numpy()
    for i, pz in enumerate(poses):
        arrays[f"pose_{i}"] = pz.numpy()
    arrays["meta"] = np.frombuffer(json.dumps(dict(shapes=shapes, no_reloc_list=list(range(Na)), reloc_list=[],
                                                    rank=agg.rank, patch_start_idx=psi, num_recon=agg.num_recon)).encode(),
                                   dtype=np.uint8)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", name + ".npz")
    np.savez_compressed(path, **arrays)
    print(name, {k: v.shape for k, v in arrays.items() if k != "meta"}, os.path.getsize(path), "bytes")


LOSS_CASES = {
    # name: (synthetic_batch kwargs, CDFLossIndexPytorch kwargs) — the second is train_imc.py:333-347's configuration
    "loss_imc": (dict(num_frames=4, num_points=300, seed=0), dict(min_val=0.0, max_val=15.0, num_bins=250, gradient_smooth=0.05)),
    # ragged K, shared focal length, residuals beyond max_val (the 2.0 / zero-gradient branch), a 9-tap smoothing filter
    "loss_shared": (dict(num_frames=3, num_points=257, seed=1, shared_focal=True, cam_sigma=0.01),
                    dict(min_val=0.0, max_val=4.0, num_bins=64, gradient_smooth=0.2)),
}


def run_loss_case(name):
    """N4: the reference's own compute_loss (train/train_imc.py:141-270) + torch autograd on the CPU."""
    import ref_train_import
    sys.path.insert(0, os.path.dirname(HERE))
    from sailrecon_b200.losses.synthetic import synthetic_batch
    batch_kw, cdf_kw = LOSS_CASES[name]
    train_imc = ref_train_import.import_train_imc("/root/reference")
    batch = synthetic_batch(**batch_kw)
    out = ref_train_import.run_reference_loss(train_imc, batch, "cpu", cdf_kw)
    arrays = {"ref_" + k: v for k, v in out.items()}
    arrays.update({k: np.asarray(v) for k, v in batch.items()})
    arrays["meta"] = np.frombuffer(json.dumps({"batch": batch_kw, "cdf": cdf_kw}).encode(), dtype=np.uint8)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", name + ".npz")
    np.savez_compressed(path, **arrays)
    print(name, "loss", out["loss"], os.path.getsize(path), "bytes")


if __name__ == "__main__":
    for n in LOSS_CASES:
        if len(sys.argv) == 1 or n in sys.argv[1:]:
            run_loss_case(n)
    for n, c in GOLDEN_CASES.items():
        if len(sys.argv) > 1 and n not in sys.argv[1:]:
            continue
        run_case(n, c)
    if len(sys.argv) == 1 or "tiny_cache" in sys.argv[1:]:
        run_cache_case()

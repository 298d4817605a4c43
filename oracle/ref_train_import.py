"""TEST INFRASTRUCTURE — imports the reference's UNMODIFIED training script (train/train_imc.py) from a given `train/`
directory so that its own `compute_loss`, `utils.geometry` and `losses.cdf_loss` can be called.  The script's unrelated
imports that are not installed on this image (tensorboard, h5py data set, plyfile export, matplotlib plots — none of
which compute_loss touches) are replaced by empty stand-ins; nothing of the loss path is stubbed.

Used by oracle/make_golden.py (from /root/reference, in the build container) and tests/ (from oracle/_ref on the GPU box)."""
import importlib
import os
import sys
import types


def _have(name: str) -> bool:
    try:
        importlib.import_module(name)
        return True
    except Exception:
        return False


def import_train_imc(root: str):
    """root: directory holding `sailrecon/` and `train/`.  Returns the imported train_imc module."""
    train = os.path.join(root, "train")
    for p in (root, train):
        if p not in sys.path:
            sys.path.insert(0, p)
    stand_ins = [("torch.utils.tensorboard", {"SummaryWriter": object}), ("eval", {}), ("eval.utils", {}),
                 ("eval.utils.geometry", {"save_pointcloud_with_plyfile": None}),
                 ("datasets.imc2021", {"IMC2021": object, "collate_fn": None}), ("utils.io", {"ImagePreprocessor": object}),
                 ("utils.sanity_check", {"sanity_check_relative_poses": None}), ("utils.vls", {"plot_cdf_pdf_curves": None})]
    for name, attrs in stand_ins:
        if not _have(name):
            m = types.ModuleType(name)
            m.__dict__.update(attrs)
            sys.modules[name] = m
    mod = importlib.import_module("train_imc")
    assert os.path.dirname(os.path.abspath(mod.__file__)) == os.path.abspath(train), mod.__file__
    for real in ("utils.geometry", "losses.cdf_loss"):
        assert os.path.abspath(sys.modules[real].__file__).startswith(os.path.abspath(train)), real
    return mod


def run_reference_loss(train_imc, batch_np: dict, device, cdf_kwargs: dict, do_record: bool = True):
    """Calls the reference's compute_loss on a synthetic batch (numpy arrays in the data loader's format) and differentiates
    it with torch autograd -> dict of numpy results.  The CDF module is constructed with the batch's pair lists: the
    trainer as shipped passes placeholder indices [0] (train_imc.py:338-346), which only works for a single pair."""
    import numpy as np
    import torch
    t = {k: (torch.from_numpy(v).to(device) if isinstance(v, np.ndarray) else v) for k, v in batch_np.items()}
    K = t.pop("pred_intrinsics").clone().requires_grad_(True)
    E = t.pop("pred_extrinsics").clone().requires_grad_(True)
    preds = [{"intrinsic": K[i:i + 1], "extrinsic": E[i:i + 1]} for i in range(K.shape[0])]
    module = train_imc.CDFLossIndexPytorch(src_indices=t["src_idx"], dst_indices=t["dst_idx"], num_nodes=K.shape[0],
                                           **cdf_kwargs).to(device)
    res = train_imc.compute_loss(preds, t, device, module, do_record)
    loss = res["loss"]
    loss.backward()
    out = {"loss": np.float32(loss.item()), "d_intrinsics": K.grad.cpu().numpy(), "d_extrinsics": E.grad.cpu().numpy()}
    if do_record:
        out["frame_cdfs"] = np.asarray(res["plot_data"]["frame_cdfs"])
        out["frame_pdfs"] = np.asarray(res["plot_data"]["frame_pdfs"])
    return out

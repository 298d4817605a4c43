"""Shared by tests/test_loss_cpu.py and tests/test_loss_gpu.py: golden fixtures of the training loss (written by
oracle/make_golden.py from the reference's own compute_loss) and the oracle call on them."""
import json
import os

import numpy as np

from conftest import ROOT

LOSS_GOLDEN = ("loss_imc", "loss_shared")
BATCH_KEYS = ("pred_intrinsics", "pred_extrinsics", "K_prime_to_K", "src_idx", "dst_idx", "src_coords", "dst_coords", "src_depth",
              "dst_depth")


def load_loss_golden(name):
    z = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
    meta = json.loads(bytes(z.pop("meta")).decode())
    z["shared_focal"] = bool(z["shared_focal"])
    return z, meta


def oracle_loss(b, cdf_kw, **kw):
    import loss_oracle
    return loss_oracle.compute_loss(b["pred_intrinsics"], b["pred_extrinsics"], b["K_prime_to_K"], b["shared_focal"], b["src_idx"],
                                    b["dst_idx"], b["src_coords"], b["dst_coords"], b["src_depth"], b["dst_depth"], **cdf_kw, **kw)


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))

"""Subprocess body of tests/test_overlay_gpu.py: runs the reference's OWN `SailRecon` (sail_recon.py:24-286, unmodified,
from oracle/_ref) — once as shipped ("ref") and once with sailrecon_b200.overlay installed ("overlay", INTEGRATION.md §1:
the aggregator, its layers and the camera head come from this repository; SailRecon itself, the DPT heads and the
pose / unprojection utilities stay the reference's) — on the same seeded weights and images, and writes the predictions.

    python tests/sailrecon_e2e.py {ref|overlay} OUT.npz [anchors] [queries] [size]

Calls reproduced (train/demo_imc.py:83-106, train/demo_imc_forward.py:92-101): under autocast(bf16) and no_grad,
  two-phase:  model.tmp_forward(anchor_images, fix_rank=300); model.reloc(image_i, fix_rank=300, memory_save=False, ...)
              and model.reloc(image_i, fast_reloc=True)
  one-pass:   model.forward(cat(anchors, queries), no_reloc_list, reloc_list, fix_rank=300)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def reinit(model: torch.nn.Module, seed: int) -> None:
    """Same parameter values in both processes regardless of how each constructor consumed the RNG: refill every
    parameter (sorted by name) from one CUDA generator.  LayerScale gammas keep their constructed value."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    with torch.no_grad():
        for name, p in sorted(model.named_parameters()):
            leaf = name.split(".")[-1]
            if leaf == "gamma":
                continue
            r = torch.randn(p.shape, generator=g, device="cuda")
            if p.dim() == 1 and leaf == "weight":
                p.copy_(1.0 + 0.02 * r)
            else:
                p.copy_(0.02 * r)


PROBES = ("aggregator.frame_blocks.0.attn.qkv.weight", "aggregator.frame_blocks.23.mlp.fc2.weight",
          "aggregator.global_blocks.11.attn.proj.weight", "aggregator.global_reloc_blocks.5.mlp.fc1.weight",
          "aggregator.global_reloc_blocks.23.attn.k_norm.weight", "aggregator.patch_embed.blocks.3.attn.qkv.weight",
          "aggregator.patch_embed.patch_embed.proj.weight", "aggregator.patch_embed.pos_embed", "aggregator.camera_token_reloc",
          "aggregator.register_token", "camera_head.trunk.0.attn.qkv.weight", "camera_head.pose_branch.fc2.weight",
          "depth_head.scratch.output_conv1.weight", "point_head.projects.0.weight")


def train_main(mode, out_path, Na, Nq, size):
    """One training forward + backward of the reference's SailRecon exactly as train/train_imc.py:361-381 drives it
    (model.train(), autocast(bf16) around model.forward, loss.backward()), pure reference vs the aggregator-only overlay."""
    import make_ref
    sys.path.insert(0, make_ref.import_path())
    if mode == "overlay_train":
        import sailrecon_b200.overlay as overlay
        overlay.install(geometry=False, dpt=False, camera_head=False, layers=False)
    if mode == "overlay_train_full":        # aggregator + camera head + pose -> cameras on this package; the reference's DPT heads
        import sailrecon_b200.overlay as overlay
        overlay.install(dpt=False)
    from sailrecon.models.sail_recon import SailRecon
    dev = "cuda"
    torch.manual_seed(0)
    with torch.device(dev):
        model = SailRecon(kv_cache=False).to(dev)
    reinit(model, 78)
    model.train()
    if mode == "overlay_train":
        import sailrecon_b200.models.aggregator as A
        assert type(model.aggregator) is A.Aggregator and type(model.camera_head).__module__ == "sailrecon.heads.camera_head"
    if mode == "overlay_train_full":
        import sailrecon_b200.heads.camera_head as H
        import sailrecon_b200.models.aggregator as A
        assert type(model.aggregator) is A.Aggregator and type(model.camera_head) is H.CameraHead
        assert type(model.depth_head).__module__ == "sailrecon.heads.dpt_head"
    g = torch.Generator().manual_seed(6)
    images = torch.rand(Na + Nq, 3, size, size, generator=g).to(dev)
    model.aggregator.generator.manual_seed(1234)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        preds = model.forward(images, no_reloc_list=list(range(Na)), reloc_list=list(range(Na, Na + Nq)), fix_rank=300)
    gl = torch.Generator(device=dev).manual_seed(7)
    loss = 0.0
    for pr in preds:
        for k in ("extrinsic", "depth_map", "point_map", "xyz_cnf", "dpt_cnf"):
            v = pr[k].float()
            loss = loss + (v * torch.randn(v.shape, device=dev, generator=gl)).mean()
    (1024.0 * loss).backward()                      # GradScaler-style scaling
    named = dict(model.named_parameters())
    res = {"loss": np.array(float(loss))}
    n_grad = sum(1 for p in model.parameters() if p.grad is not None)
    res["params_with_grad"] = np.array(n_grad)
    for n in PROBES:
        gr = named[n].grad
        assert gr is not None, n
        res["grad." + n] = (gr.float() / 1024.0).flatten()[:: max(1, gr.numel() // 65536)].cpu().numpy()
        res["gnorm." + n] = np.array(float(gr.float().norm() / 1024.0))
    np.savez(out_path, **res)
    print("e2e ok", mode, n_grad)


def main():
    mode, out_path = sys.argv[1], sys.argv[2]
    if "train" in mode:
        return train_main(mode, out_path, int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
    Na = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    Nq = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    size = int(sys.argv[5]) if len(sys.argv) > 5 else 518
    import make_ref
    sys.path.insert(0, make_ref.import_path())
    if mode == "overlay":
        import sailrecon_b200.overlay as overlay
        overlay.install()
    from sailrecon.models.sail_recon import SailRecon
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = "cuda"
    torch.manual_seed(0)
    with torch.device(dev):
        model = SailRecon(kv_cache=True).eval().to(dev)
    reinit(model, 77)
    if mode == "overlay":
        import sailrecon_b200.models.aggregator as A
        assert type(model.aggregator) is A.Aggregator, "overlay not in effect"
    else:
        assert type(model.aggregator).__module__ == "sailrecon.models.aggregator"
        assert "sailrecon_b200" not in sys.modules, "the reference arm must not load the replacement"
    g = torch.Generator().manual_seed(5)
    anchors = torch.rand(Na, 3, size, size, generator=g).to(dev)
    queries = torch.rand(Nq, 3, size, size, generator=g).to(dev)
    res = {}
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        # ---- two-phase localisation (demo_imc.py:91-106)
        model.aggregator.generator.manual_seed(1234)
        model.tmp_forward(anchors, fix_rank=300)
        res["cam_token_last_layer"] = model.cam_token_last_layer.float().cpu().numpy()
        for i in range(Nq):
            pred = model.reloc(queries[i:i + 1], fix_rank=300, memory_save=False, save_depth=True, ret_img=True)[0]
            for k in ("extrinsic", "intrinsic", "depth_map", "dpt_cnf", "point_map", "xyz_cnf", "cam_tokens"):
                res[f"reloc{i}.{k}"] = pred[k].float().cpu().numpy()
            fast = model.reloc(queries[i:i + 1], fix_rank=300, fast_reloc=True)[0]
            assert set(fast) == {"extrinsic", "intrinsic"}
            res[f"fast{i}.extrinsic"] = fast["extrinsic"].float().cpu().numpy()
        # second tmp_forward clears and re-primes the cache (sail_recon.py:178-183)
        model.aggregator.generator.manual_seed(1234)
        model.tmp_forward(anchors, fix_rank=300)
        again = model.reloc(queries[0:1], fix_rank=300, fast_reloc=True)[0]
        res["again.extrinsic"] = again["extrinsic"].float().cpu().numpy()
        model.clear_cache()
        # ---- one-pass forward (demo_imc_forward.py:92-101) on a kv_cache=False model with the same weights
        sd = model.state_dict()
        del model
        torch.cuda.empty_cache()
        with torch.device(dev):
            model = SailRecon(kv_cache=False).eval().to(dev)
        model.load_state_dict(sd)
        del sd
        model.aggregator.generator.manual_seed(1234)
        preds = model.forward(torch.cat([anchors, queries]), no_reloc_list=list(range(Na)),
                              reloc_list=list(range(Na, Na + Nq)), fix_rank=300)
        assert len(preds) == Nq
        for i, pred in enumerate(preds):
            for k in ("extrinsic", "intrinsic", "depth_map", "point_map", "cam_tokens", "point_map_by_unprojection"):
                v = pred[k]
                res[f"fwd{i}.{k}"] = (v.float().cpu().numpy() if torch.is_tensor(v) else np.asarray(v, dtype=np.float32))
    np.savez(out_path, **res)
    print("e2e ok", mode, len(res))


if __name__ == "__main__":
    main()

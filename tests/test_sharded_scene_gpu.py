"""GPU parity of the frame-sharded WHOLE scene (sailrecon_b200/parallel.py::sharded_scene_forward):
frame blocks rank-local, scene-token K|V all-gathered, reloc block on the local queries, global block over the ring,
camera head on [all anchor camera tokens ; local queries] — against the un-sharded Aggregator.forward + CameraHead.forward
of the same weights, images and scene-token draw.

  * one GPU: world size 1 (the plumbing, no communication);
  * >= 2 GPUs (skipped otherwise; `gpurun --gpus 2`): one process per GPU over NCCL, uneven shards (3 anchors + 3 queries
    over 2 ranks).
"""
import socket

import pytest
import torch

from golden_common import make_images, make_params, rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"
NA, NQ, SIZE, RANK = 3, 3, 224, 100


def _models(device):
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    torch.manual_seed(0)
    agg = Aggregator(img_size=SIZE, depth=2, intermediate_layer_idx=[0, 1]).eval()
    agg.patch_embed.blocks = agg.patch_embed.blocks[:2]
    head = CameraHead(dim_in=2048, trunk_depth=1).eval()
    shapes = {"agg." + k: list(v.shape) for k, v in agg.state_dict().items()}
    shapes.update({"head." + k: list(v.shape) for k, v in head.state_dict().items()})
    params = make_params(shapes, 55)                       # O(1) LayerScale / special tokens: every sub-path is visible
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    return agg.to(device), head.to(device)


def _unsharded(agg, head, images):
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out, _, cam = agg(images, list(range(NA)), list(range(NA, NA + NQ)), fix_rank=RANK)
        with torch.autocast("cuda", enabled=False):
            poses = head(out, cam)
    return out, poses


def _sharded(agg, head, images, rank, world):
    from sailrecon_b200 import parallel
    plan = parallel.ScenePlan.make(NA, NQ, rank, world)
    local = images[:, plan.local_frames()].contiguous()
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out, poses = parallel.sharded_scene_forward(agg, head, local, plan, fix_rank=RANK)
    return plan, out, poses


def test_sharded_scene_world1_equals_unsharded():
    agg, head = _models(DEV)
    images = make_images(NA + NQ, SIZE, SIZE, 77).to(DEV)
    ref_out, ref_poses = _unsharded(agg, head, images)
    plan, out, poses = _sharded(agg, head, images, 0, 1)
    assert plan.local_frames() == list(range(NA + NQ))
    for k in ref_out:
        assert rel_l2(out[k].cpu(), ref_out[k].cpu()) < 1e-5, k          # same kernels, same order: (almost) bit-identical
    for a, b in zip(poses, ref_poses):
        assert rel_l2(a.cpu(), b.cpu()) < 1e-5


def _worker(rank, world, port, out_path):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
    try:
        agg, head = _models(dev)
        images = make_images(NA + NQ, SIZE, SIZE, 77).to(dev)
        plan, out, poses = _sharded(agg, head, images, rank, world)
        torch.cuda.synchronize()
        torch.save({"frames": plan.local_frames(), "out": {k: v.cpu() for k, v in out.items()},
                    "poses": [p.cpu() for p in poses]}, f"{out_path}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")
def test_sharded_scene_two_processes_nccl(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    path = str(tmp_path / "scene")
    mp.spawn(_worker, args=(world, port, path), nprocs=world, join=True)
    agg, head = _models(DEV)
    images = make_images(NA + NQ, SIZE, SIZE, 77).to(DEV)
    ref_out, ref_poses = _unsharded(agg, head, images)
    parts = [torch.load(f"{path}.{r}") for r in range(world)]
    q_frames = [f - NA for p in parts for f in p["frames"] if f >= NA]
    assert sorted(q_frames) == list(range(NQ))
    for k in ref_out:
        got = torch.cat([p["out"][k] for p in parts if k in p["out"]], dim=1)      # [1, Nq, P, 2C], ranks in query order
        assert rel_l2(got, ref_out[k].cpu()[:, q_frames]) < 5e-3, k              # ring merge order differs: bf16 level
    for it in range(len(ref_poses)):
        got = torch.cat([p["poses"][it] for p in parts if p["poses"]], dim=1)
        assert rel_l2(got, ref_poses[it].cpu()[:, q_frames]) < 1e-2

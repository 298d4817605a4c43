"""GPU parity of the frame-sharded global block (sailrecon_b200/parallel.py).

  * one GPU: every rank of a 3-way ring is replayed in turn on the same device (blocks taken in ring arrival order, the
    kernel's (m, l, acc) carry between steps, uneven shards) and must reproduce the un-sharded Block on the whole
    sequence — the oracle for that Block is tests/test_modules_gpu.py;
  * >= 2 GPUs (skipped otherwise; run with `gpurun --gpus 2`): the real thing, one process per GPU over NCCL.
"""
import os
import socket

import pytest
import torch

from golden_common import rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"
C, H, P = 1024, 16, 261          # 16 x 16 patches + 5 special tokens per frame


def _block_and_inputs(n_frames, device):
    from sailrecon_b200.engine import RowPositions
    from sailrecon_b200.layers.block import Block
    from sailrecon_b200.layers.rope import PositionGetter, RotaryPositionEmbedding2D
    torch.manual_seed(3)
    rope = RotaryPositionEmbedding2D(frequency=100)
    blk = Block(dim=C, num_heads=H, mlp_ratio=4.0, init_values=0.01, qk_norm=True, rope=rope)
    with torch.no_grad():
        blk.ls1.gamma.uniform_(0.5, 1.5)       # make the attention branch visible in the output
        blk.ls2.gamma.uniform_(0.5, 1.5)
    blk = blk.to(device).eval()
    g = torch.Generator().manual_seed(11)
    x = torch.randn(n_frames * P, C, generator=g).to(device)
    grid = PositionGetter()(1, 16, 16, device)[0] + 1
    frame_pos = torch.cat([torch.zeros(5, 2, dtype=grid.dtype, device=device), grid], 0).to(torch.int32)
    return blk, x, frame_pos


def _reference(blk, x, frame_pos, n_frames):
    from sailrecon_b200 import engine
    ref = x.clone()
    pos = engine.RowPositions(frame_pos.repeat(n_frames, 1).contiguous(), 17)
    with torch.no_grad():
        engine.run_block(blk, ref, engine.full_segments(1, ref.shape[0]), pos, "bf16", tag="t.")
    return ref


def test_ring_replayed_on_one_gpu_equals_unsharded_block():
    from sailrecon_b200 import engine, ops, parallel
    n_frames, R = 7, 3
    blk, x, frame_pos = _block_and_inputs(n_frames, DEV)
    ref = _reference(blk, x, frame_pos, n_frames)
    attn = blk.attn
    plans = [parallel.RingPlan.make(n_frames, P, r, R) for r in range(R)]
    # every rank's q and k|v (what ring_global_block computes locally)
    qs, kvs, xs = [], [], []
    with torch.no_grad():
        for pl in plans:
            xr = x[pl.row0():pl.row0() + pl.rows()].clone()
            pos = engine.RowPositions(frame_pos.repeat(pl.frames[pl.rank][1], 1).contiguous(), 17)
            a = torch.empty(pl.rows(), C, dtype=torch.bfloat16, device=DEV)
            w1, b1, eps1 = engine.ln_params(blk.norm1)
            ops.layernorm(xr, w1, b1, eps1, a)
            q = torch.empty(pl.rows(), C, dtype=torch.bfloat16, device=DEV)
            kv = torch.empty(pl.rows(), 2 * C, dtype=torch.bfloat16, device=DEV)
            engine.qkv_rows(attn, a, q, "bf16", pos, col_offset=0, n_cols=C)
            engine.qkv_rows(attn, a, kv, "bf16", pos, col_offset=C, n_cols=2 * C)
            qs.append(q); kvs.append(kv); xs.append(xr)
        for pl in plans:
            T = pl.rows()
            carry = parallel.Carry.empty(T, H, C // H, DEV)
            o = torch.empty(T, C, dtype=torch.bfloat16, device=DEV)
            for s in range(R):
                src = pl.source(s)
                parallel._cuda_step(qs[pl.rank], kvs[src], plans[src].rows(), o, carry, s > 0, s + 1 < R, H, C // H, attn.scale)
            engine.block_tail(blk, xs[pl.rank], o, "bf16", tag="t.")
    got = torch.cat(xs, 0)
    torch.cuda.synchronize()
    assert rel_l2(got.cpu(), ref.cpu()) < 5e-3          # same kernels, different tile order across the block boundaries


def _nccl_worker(rank, world, port, n_frames, out_path):
    import torch.distributed as dist
    from sailrecon_b200 import engine, parallel
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        blk, x, frame_pos = _block_and_inputs(n_frames, dev)
        plan = parallel.RingPlan.make(n_frames, P, rank, world)
        xr = x[plan.row0():plan.row0() + plan.rows()].clone()
        pos = engine.RowPositions(frame_pos.repeat(plan.frames[rank][1], 1).contiguous(), 17)
        ws = parallel.RingWorkspace(plan, C, H, dev)
        with torch.no_grad():
            for _ in range(2):                                       # workspace reuse across layers
                xr.copy_(x[plan.row0():plan.row0() + plan.rows()])
                parallel.ring_global_block(blk, xr, pos, plan, ws)
        torch.cuda.synchronize()
        torch.save(xr.cpu(), f"{out_path}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")
def test_ring_two_processes_nccl(tmp_path):
    import torch.multiprocessing as mp
    n_frames, world = 5, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "ring")
    mp.spawn(_nccl_worker, args=(world, port, n_frames, out), nprocs=world, join=True)
    blk, x, frame_pos = _block_and_inputs(n_frames, DEV)
    ref = _reference(blk, x, frame_pos, n_frames)
    got = torch.cat([torch.load(f"{out}.{r}") for r in range(world)], 0)
    assert rel_l2(got, ref.cpu()) < 5e-3

"""Host logic of the frame-sharded ring attention (sailrecon_b200/parallel.py) on CPU: gloo, world_size 2 and 3.

The CUDA kernel cannot run here, so the schedule is exercised with an injected step function that restates the carry
semantics of sr_attn_fwd_bf16 (include/sailrecon_b200.h: running max / sum in the log2 domain, un-normalised accumulator)
in plain torch.  What is under test is parallel.py itself: frame partition, block arrival order, double-buffer reuse,
send/recv pairing with uneven shards, first/last-step carry flags — the result must equal single-device attention over
the concatenated sequence.  (The kernel's own carry arithmetic is covered on the GPU by
tests/test_kernels_gpu.py::test_attention_bf16_carry_equals_single_pass and tests/test_ring_gpu.py.)
"""
import math
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sailrecon_b200 import parallel

H, D = 2, 8
C = H * D
P = 5                      # tokens per frame


def _step_ref(q, kv, n_kv, out, carry, carry_in, carry_out, heads, head_dim, scale):
    """Reference ring step with the kernel's carry contract."""
    T = q.shape[0]
    sl2 = scale * math.log2(math.e)
    for h in range(heads):
        cs = slice(h * head_dim, (h + 1) * head_dim)
        s = (q[:, cs] @ kv[:n_kv, :C][:, cs].t()) * sl2                     # log2 domain
        v = kv[:n_kv, C:][:, cs]
        if carry_in:
            m_old, l_old, acc_old = carry.m[h], carry.l[h], carry.acc[:, cs]
        else:
            m_old = torch.full((T,), -math.inf)
            l_old = torch.zeros(T)
            acc_old = torch.zeros(T, head_dim)
        m_new = torch.maximum(m_old, s.max(dim=1).values)
        alpha = torch.exp2(m_old - m_new)
        p = torch.exp2(s - m_new[:, None])
        l_new = l_old * alpha + p.sum(1)
        acc_new = acc_old * alpha[:, None] + p @ v
        if carry_out:
            carry.m[h], carry.l[h] = m_new, l_new
            carry.acc[:, cs] = acc_new
        else:
            out[:, cs] = acc_new / l_new[:, None]


def _full_reference(n_frames):
    g = torch.Generator().manual_seed(7)
    q = torch.randn(n_frames * P, C, generator=g)
    kv = torch.randn(n_frames * P, 2 * C, generator=g)
    out = torch.empty(n_frames * P, C)
    for h in range(H):
        cs = slice(h * D, (h + 1) * D)
        a = torch.softmax(q[:, cs] @ kv[:, :C][:, cs].t() * D ** -0.5, dim=-1)
        out[:, cs] = a @ kv[:, C:][:, cs]
    return q, kv, out


def _worker(rank, world, port, n_frames, errs):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q, kv, ref = _full_reference(n_frames)
        plan = parallel.RingPlan.make(n_frames, P, rank, world)
        r0, T = plan.row0(), plan.rows()
        bufs = [torch.full((plan.max_rows, 2 * C), float("nan")) for _ in range(2)]
        bufs[0][:T] = kv[r0:r0 + T]
        out = torch.empty(T, C)
        carry = parallel.Carry.empty(T, H, D, "cpu")
        for _ in range(2):                                      # twice: buffers and carry are reusable across layers
            bufs[0][:T] = kv[r0:r0 + T]
            parallel.ring_attention(q[r0:r0 + T].contiguous(), bufs, out, plan, H, D, D ** -0.5, carry, None, step_fn=_step_ref)
        errs[rank] = float((out - ref[r0:r0 + T]).abs().max())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,n_frames", [(2, 4), (2, 5), (3, 7)])
def test_ring_schedule_equals_full_attention(world, n_frames):
    errs = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), n_frames, errs), nprocs=world, join=True)
    assert len(errs) == world
    assert max(errs.values()) < 1e-5, dict(errs)


def test_partition_and_plan():
    assert parallel.partition_frames(10, 4) == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert parallel.partition_frames(8, 8) == [(i, 1) for i in range(8)]
    assert sum(c for _, c in parallel.partition_frames(512, 8)) == 512
    plan = parallel.RingPlan.make(10, 1374, 2, 4)
    assert plan.rows() == 2 * 1374 and plan.row0() == 6 * 1374 and plan.max_rows == 3 * 1374
    assert [plan.source(s) for s in range(4)] == [2, 1, 0, 3]
    assert plan.next_rank == 3 and plan.prev_rank == 1
    with pytest.raises(ValueError):
        parallel.RingPlan.make(3, 1374, 0, 4)                    # a rank without frames


def test_world_one_is_plain_attention():
    q, kv, ref = _full_reference(3)
    plan = parallel.RingPlan.make(3, P, 0, 1)
    out = torch.empty_like(ref)
    parallel.ring_attention(q, [kv.clone(), kv.clone()], out, plan, H, D, D ** -0.5, None, None, step_fn=_step_ref)
    assert (out - ref).abs().max() < 1e-5


def test_scene_plan_partitions_every_frame_once():
    """ScenePlan (frame-sharded whole scene): anchors and queries are partitioned separately; every input frame of the
    [anchors ; queries] numbering lands on exactly one rank, anchors first."""
    for Na, Nq, world in [(256, 256, 8), (100, 200, 8), (3, 3, 2), (5, 0, 4), (9, 2, 4)]:
        seen = []
        for r in range(world):
            plan = parallel.ScenePlan.make(Na, Nq, r, world)
            fr = plan.local_frames()
            na = plan.anchors[r][1]
            assert all(f < Na for f in fr[:na]) and all(f >= Na for f in fr[na:])
            assert fr == plan.local_frames(r)
            seen += fr
        assert sorted(seen) == list(range(Na + Nq))
    with pytest.raises(ValueError):
        parallel.ScenePlan.make(3, 100, 0, 4)                    # a rank without an anchor frame


def _gather_worker(rank, world, port, counts, errs):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(sum(counts) * 3, dtype=torch.float32).reshape(sum(counts), 3)
        off = sum(counts[:rank])
        dst = torch.full_like(full, float("nan"))
        parallel._allgather_rows(dst, full[off:off + counts[rank]].clone(), counts, rank)
        errs[rank] = float((dst - full).abs().max())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("counts", [[2, 1], [3, 0, 2], [2, 2]])       # broadcast fallback (gloo, uneven / empty) and the one-collective path
def test_allgather_rows_uneven_counts(counts):
    errs = mp.Manager().dict()
    mp.spawn(_gather_worker, args=(len(counts), _free_port(), counts, errs), nprocs=len(counts), join=True)
    assert len(errs) == len(counts) and max(errs.values()) == 0.0, dict(errs)

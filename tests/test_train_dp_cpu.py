"""Host logic of the data-parallel training step (sailrecon_b200/train.py: TrainState buckets, per-layer gradient
all-reduce issued in backward order, 1 / world folded into the optimizer) on CPU with gloo, world_size 2.

The CUDA kernels cannot run here: the forward/backward is replaced by a stand-in that writes rank-dependent gradients into
the parameter-gradient views layer by layer (top layer first, as stack_backward does) and the Adam kernel by
torch.optim-style arithmetic on the flat buckets.  Under test: parameters re-pointed into the flat buckets keep their
values and stay views, every parameter owns a disjoint gradient view, the reduced gradient is the sum over ranks, both
ranks end with identical parameters equal to one Adam step on the mean gradient (DDP semantics, train_imc.py:472-475)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from sailrecon_b200 import train


class _Blk(nn.Module):
    def __init__(self, seed):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.w = nn.Parameter(torch.randn(7, 9, generator=g))
        self.b = nn.Parameter(torch.randn(130, generator=g))


class _Stacks(nn.Module):
    def __init__(self, depth=3):
        super().__init__()
        self.depth = depth
        self.frame_blocks = nn.ModuleList([_Blk(3 * i) for i in range(depth)])
        self.global_blocks = nn.ModuleList([_Blk(3 * i + 1) for i in range(depth)])
        self.global_reloc_blocks = nn.ModuleList([_Blk(3 * i + 2) for i in range(depth)])


def _rank_grad(p, rank, i):
    g = torch.Generator().manual_seed(1000 * rank + i + p.numel())
    return torch.randn(p.shape, generator=g)


def _torch_adam(state):
    def step(grad_scale):
        state.step += 1
        for i in range(len(state.param)):
            g = state.grad[i] * grad_scale
            state.m[i].mul_(0.9).add_(g, alpha=0.1)
            state.v[i].mul_(0.999).addcmul_(g, g, value=0.001)
            bc1, bc2 = 1 - 0.9 ** state.step, 1 - 0.999 ** state.step
            state.param[i].sub_(state.lr / bc1 * state.m[i] / (state.v[i].sqrt() / bc2 ** 0.5 + state.eps))
    return step


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        agg = _Stacks()
        before = {k: p.detach().clone() for k, p in agg.named_parameters()}
        state = train.TrainState(agg, lr=1e-2)
        order = []

        def fwd_bwd(on_layer_done):
            for i in reversed(range(agg.depth)):
                for p in state.params_of_layer[i]:
                    state.grads[p] += _rank_grad(p, rank, i)
                order.append(i)
                on_layer_done(i)
            return torch.zeros(1)

        train.training_step(agg, state, None, None, None, {}, world, fwd_bwd=fwd_bwd, optimizer_step=_torch_adam(state))
        ok = order == [2, 1, 0]
        for i in range(agg.depth):
            for p in state.params_of_layer[i]:
                want = sum(_rank_grad(p, r, i) for r in range(world))
                ok = ok and torch.allclose(state.grads[p], want, atol=1e-6)
        named = dict(agg.named_parameters())
        for k, p in named.items():
            ref = nn.Parameter(before[k].clone())
            i = int(k.split(".")[1])
            ref.grad = sum(_rank_grad(p, r, i) for r in range(world)) / world
            torch.optim.Adam([ref], lr=1e-2, betas=(0.9, 0.999), eps=1e-8).step()
            ok = ok and torch.allclose(p.detach(), ref.detach(), atol=1e-6)
        out[rank] = (ok, float(sum(p.detach().double().sum() for p in named.values())))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_dp_step_allreduces_per_layer_and_matches_adam_on_mean_gradient():
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert len(out) == 2 and all(v[0] for v in out.values()), dict(out)
    assert out[0][1] == out[1][1]                                # replicas stay in sync


def test_train_state_bucket_layout():
    agg = _Stacks(depth=2)
    before = {k: p.detach().clone() for k, p in agg.named_parameters()}
    state = train.TrainState(agg)
    assert len(state.param) == 2 and all(b.numel() % 64 == 0 for b in state.param)
    seen = []
    for k, p in agg.named_parameters():
        assert torch.equal(p.detach(), before[k])
        i = int(k.split(".")[1])
        lo, hi = state.param[i].data_ptr(), state.param[i].data_ptr() + 4 * state.param[i].numel()
        assert lo <= p.data_ptr() < hi and (p.data_ptr() - lo) % 256 == 0       # a 256-byte aligned view of layer i's bucket
        g = state.grads[p]
        assert g.shape == p.shape and g.data_ptr() - state.grad[i].data_ptr() == p.data_ptr() - lo
        seen.append((g.data_ptr(), g.data_ptr() + 4 * g.numel()))
    seen.sort()
    assert all(a[1] <= b[0] for a, b in zip(seen, seen[1:]))                    # disjoint
    v0 = agg.frame_blocks[0].w._version
    state.param[0].add_(1.0)                                                     # an in-place bucket update is seen by the module
    assert torch.allclose(agg.frame_blocks[0].w.detach(), before["frame_blocks.0.w"] + 1.0)
    torch.autograd.graph.increment_version(agg.frame_blocks[0].w)
    assert agg.frame_blocks[0].w._version > v0


def test_activation_budget_picks_the_schedule():
    """train.activation_bytes / choose_keep_activations: 24 frames of 518 px keep ~70 GB of activations (fits a 180 GB B200
    next to 15 GB of optimizer state), 100 + 200 frames do not and fall back to the reference's checkpointed schedule."""
    class Ctx:
        pass
    c = Ctx()
    c.Na, c.Nq, c.P = 12, 12, 1374
    need = train.activation_bytes(c, 24, 1024, scene_rows=12 * 305)
    assert 60e9 < need < 85e9
    assert train.choose_keep_activations(c, 24, 1024, 12 * 305, free_bytes=int(160e9))
    assert not train.choose_keep_activations(c, 24, 1024, 12 * 305, free_bytes=int(70e9))       # e.g. an 80 GB part
    c.Na, c.Nq = 100, 200
    assert not train.choose_keep_activations(c, 24, 1024, 100 * 305, free_bytes=int(160e9))

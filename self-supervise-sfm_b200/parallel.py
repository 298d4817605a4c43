"""Frame-sharded (sequence-parallel) global attention with ring K/V exchange — SURVEY.md §8(e), BASELINE.json configs[3].

The reference has no sequence parallelism (DP only, train/train_imc.py:472); its global blocks
(sailrecon/models/aggregator.py:743-769) run one SDPA over all Na*P anchor tokens on one GPU.  Here the anchor frames are
partitioned into contiguous ranges, one per rank (one process per GPU).  Everything per-token (LayerNorm, QKV / proj /
MLP GEMMs, LayerScale-residual) stays rank-local on the rank's rows; only the attention core couples ranks:

    step s = 0 .. R-1:   O_r, m_r, l_r  <-  online-softmax merge of  Q_r  x  (K, V)_{(r - s) mod R}

The K|V block of a rank ([T_r, 2C] bf16, post QK-norm / RoPE) travels around the ring: while the flash kernel works on
block s (sr_attn_fwd_bf16 with the (m, l, acc) carry of include/sailrecon_b200.h), block s is forwarded to rank r+1 and
block s+1 is received from rank r-1 into the other half of a double buffer (NCCL send/recv over NVLink 5 / NVSwitch,
issued before the kernel so the transfer overlaps it).

Why staged copies and not peer-memory loads from inside the attention kernel: every 256-row query tile re-reads the
whole K/V block (343 query tiles per head at 64 frames per rank).  Local reads of that pattern are L2 hits; peer
addresses bypass the local L2, so direct peer TMA loads would put the same bytes on NVLink hundreds of times.  One copy
per block per step is 2 * T_r * C * 2 B = 360 MB per step at R = 8, Na = 512 (0.47 ms at 770 GB/s) against 40 ms of
attention math per step at 800 TFLOP/s: the exchange is ~1 % of the step and fully hidden.
"""
from __future__ import annotations

import json
import os
import time
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from . import engine
from .engine import RowPositions


# ------------------------------------------------------------------------------------------------ partition
def partition_frames(n_frames: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, balanced frame ranges: [(first_frame, count)] per rank; the first n_frames % world ranks get one
    extra frame.  Ranks beyond n_frames get (n_frames, 0)."""
    if n_frames < 0 or world <= 0:
        raise ValueError("partition_frames: n_frames >= 0 and world > 0 required")
    base, extra = divmod(n_frames, world)
    out, start = [], 0
    for r in range(world):
        c = base + (1 if r < extra else 0)
        out.append((start, c))
        start += c
    return out


@dataclass
class RingPlan:
    """Who holds what and the order blocks arrive in: at step s rank r works on the K/V block of rank (r - s) mod R."""
    rank: int
    world: int
    tokens_per_frame: int
    frames: List[Tuple[int, int]]                      # per rank (first_frame, count)

    @classmethod
    def make(cls, n_frames: int, tokens_per_frame: int, rank: int, world: int) -> "RingPlan":
        fr = partition_frames(n_frames, world)
        if any(c == 0 for _, c in fr):
            raise ValueError(f"ring attention needs at least one frame per rank (frames {n_frames}, ranks {world})")
        return cls(rank, world, tokens_per_frame, fr)

    def rows(self, r: Optional[int] = None) -> int:
        return self.frames[self.rank if r is None else r][1] * self.tokens_per_frame

    def row0(self, r: Optional[int] = None) -> int:
        return self.frames[self.rank if r is None else r][0] * self.tokens_per_frame

    @property
    def max_rows(self) -> int:
        return max(c for _, c in self.frames) * self.tokens_per_frame

    def source(self, step: int) -> int:
        return (self.rank - step) % self.world

    @property
    def next_rank(self) -> int:
        return (self.rank + 1) % self.world

    @property
    def prev_rank(self) -> int:
        return (self.rank - 1) % self.world


# ------------------------------------------------------------------------------------------------ the ring
@dataclass
class Carry:
    """Online-softmax state across ring steps (sr_attn.carry_*): running row max m and row sum l in the log2 domain,
    fp32 [H, T]; un-normalised accumulator fp32 [T, H*d]."""
    m: torch.Tensor
    l: torch.Tensor
    acc: torch.Tensor

    @classmethod
    def empty(cls, rows: int, heads: int, head_dim: int, device) -> "Carry":
        return cls(torch.empty(heads, rows, dtype=torch.float32, device=device),
                   torch.empty(heads, rows, dtype=torch.float32, device=device),
                   torch.empty(rows, heads * head_dim, dtype=torch.float32, device=device))


def _cuda_step(q, kv, n_kv, out, carry: Optional[Carry], carry_in: bool, carry_out: bool, heads: int, head_dim: int,
               scale: float) -> None:
    """One ring step on the GPU: the tcgen05 flash kernel over one K/V block (no CPU path exists)."""
    from . import ops
    C = heads * head_dim
    ops.attention(q, kv[:n_kv, :C], kv[:n_kv, C:], None if carry_out else out, heads, head_dim, 1, q.shape[0], n_kv,
                  scale=scale, carry=None if carry is None else (carry.m, carry.l, carry.acc), carry_in=carry_in,
                  carry_out=carry_out)


def ring_attention(q: torch.Tensor, kv_bufs: Sequence[torch.Tensor], out: torch.Tensor, plan: RingPlan, heads: int,
                   head_dim: int, scale: float, carry: Optional[Carry], group=None,
                   step_fn: Callable = _cuda_step, wait_events: Optional[list] = None) -> torch.Tensor:
    """softmax(Q_r [K_0 .. K_{R-1}]^T) [V_0 .. V_{R-1}] for this rank's query rows.

    q [T_r, C]; kv_bufs = two [max_rows, 2C] buffers, kv_bufs[0][:T_r] holding this rank's own K|V block on entry (both
    are overwritten); out [T_r, C].  Per step: post (isend current block -> next rank, irecv next block <- previous
    rank), launch the attention step on the current block, then make the compute stream wait for the transfer.  The
    transfer is enqueued BEFORE the kernel, so it runs under it; the receive buffer was last read two steps ago.
    `step_fn` is the attention step (the CUDA kernel; tests inject a reference to exercise the schedule on CPU).
    `wait_events`: when a list, a CUDA event pair brackets every wait on the compute stream (recorded after the attention
    kernel has been enqueued), so the time between them is the EXPOSED part of the exchange (bench.py `nccl_wait_ms`)."""
    R = plan.world
    if R == 1:
        step_fn(q, kv_bufs[0], plan.rows(), out, None, False, False, heads, head_dim, scale)
        return out
    if carry is None:
        raise ValueError("ring_attention: carry buffers are required when world > 1")
    for s in range(R):
        cur, nxt = kv_bufs[s % 2], kv_bufs[(s + 1) % 2]
        n_cur = plan.rows(plan.source(s))
        reqs = []
        if s + 1 < R:
            n_nxt = plan.rows(plan.source(s + 1))
            ops_ = [dist.P2POp(dist.isend, cur[:n_cur], _global_rank(group, plan.next_rank), group),
                    dist.P2POp(dist.irecv, nxt[:n_nxt], _global_rank(group, plan.prev_rank), group)]
            reqs = dist.batch_isend_irecv(ops_)
        step_fn(q, cur, n_cur, out, carry, s > 0, s + 1 < R, heads, head_dim, scale)
        if reqs and wait_events is not None:
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
        for r in reqs:
            r.wait()
        if reqs and wait_events is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            wait_events.append((e0, e1))
    return out


def _global_rank(group, group_rank: int) -> int:
    return group_rank if group is None else dist.get_global_rank(group, group_rank)


# ------------------------------------------------------------------------------------------------ sharded global block
class RingWorkspace:
    """Per-rank buffers of the sharded global stack, allocated once (sized for the largest shard)."""

    def __init__(self, plan: RingPlan, C: int, heads: int, device, act=torch.bfloat16):
        T, Tm = plan.rows(), plan.max_rows
        self.a = torch.empty(T, C, dtype=act, device=device)              # LN output / attention output
        self.q = torch.empty(T, C, dtype=act, device=device)
        self.kv = [torch.empty(Tm, 2 * C, dtype=act, device=device) for _ in range(2)]
        self.carry = Carry.empty(T, heads, C // heads, device) if plan.world > 1 else None


def ring_global_block(blk, x: torch.Tensor, rows_pos: Optional[RowPositions], plan: RingPlan, ws: RingWorkspace,
                      group=None, mode: str = "bf16", wait_events: Optional[list] = None) -> torch.Tensor:
    """One global Block (aggregator.py:743-769 -> block.py:86-112) on this rank's anchor rows x [T_r, C] (fp32, in place),
    attending to the anchor rows of every rank."""
    from . import ops
    if mode != "bf16":
        raise NotImplementedError("sailrecon_b200: the ring path runs the bf16 (autocast) mode")
    attn = blk.attn
    C = x.shape[1]
    T = plan.rows()
    if x.shape[0] != T:
        raise ValueError(f"ring_global_block: rank {plan.rank} owns {T} rows, got {x.shape[0]}")
    w1, b1, eps1 = engine.ln_params(blk.norm1)
    ops.layernorm(x, w1, b1, eps1, ws.a)
    engine.qkv_rows(attn, ws.a, ws.q, mode, rows_pos, col_offset=0, n_cols=C)              # q columns
    engine.qkv_rows(attn, ws.a, ws.kv[0][:T], mode, rows_pos, col_offset=C, n_cols=2 * C)  # k | v, contiguous for the ring
    ring_attention(ws.q, ws.kv, ws.a, plan, attn.num_heads, C // attn.num_heads, attn.scale, ws.carry, group,
                   wait_events=wait_events)
    engine.block_tail(blk, x, ws.a, mode, tag="ring.")
    return x


def run_global_stack(blocks, x: torch.Tensor, rows_pos: Optional[RowPositions], plan: RingPlan, group=None,
                     ws: Optional[RingWorkspace] = None, wait_events: Optional[list] = None) -> torch.Tensor:
    """All global blocks over the sharded anchor tokens (the O(L^2) term of the aggregator, BASELINE.json configs[3])."""
    blk0 = blocks[0]
    ws = ws or RingWorkspace(plan, x.shape[1], blk0.attn.num_heads, x.device)
    for blk in blocks:
        ring_global_block(blk, x, rows_pos, plan, ws, group, wait_events=wait_events)
    return x


# ------------------------------------------------------------------------------------------------ bench (bench.py --workload ring)
def _frame_tokens(frame: int, P: int, C: int, device) -> torch.Tensor:
    """Synthetic anchor tokens of one frame, ~N(0,1), seeded by the GLOBAL frame index so every world size sees the
    same scene."""
    g = torch.Generator(device=device)
    g.manual_seed(2_000_003 + frame)
    return torch.randn(P, C, generator=g, device=device, dtype=torch.float32)


def _ring_problem(n_frames: int, depth: int, rank: int, world: int, dev, frames: Optional[Tuple[int, int]] = None):
    """Blocks (identical on every rank: torch.manual_seed(0)), this rank's token rows and positions."""
    from .layers.block import Block
    from .layers.rope import PositionGetter, RotaryPositionEmbedding2D
    C, H, P, patch = 1024, 16, 1374, 14
    plan = RingPlan.make(n_frames, P, rank, world)
    torch.manual_seed(0)
    rope = RotaryPositionEmbedding2D(frequency=100)
    with torch.device(dev):
        blocks = torch.nn.ModuleList([Block(dim=C, num_heads=H, mlp_ratio=4.0, init_values=0.01, qk_norm=True, rope=rope)
                                      for _ in range(depth)]).eval()
    f0, fc = frames if frames is not None else plan.frames[rank]
    x0 = torch.cat([_frame_tokens(f, P, C, dev) for f in range(f0, f0 + fc)], dim=0)
    hw = 518 // patch
    grid = PositionGetter()(1, hw, hw, dev)[0] + 1
    frame_pos = torch.cat([torch.zeros(5, 2, dtype=grid.dtype, device=dev), grid], 0).to(torch.int32)
    rows_pos = RowPositions(frame_pos.repeat(fc, 1).contiguous(), hw + 1)
    return plan, blocks, x0, rows_pos


def ring_parity(rank: int, world: int, dev, depth: int = 2) -> Optional[float]:
    """Parity ACROSS world sizes inside one launch (the 2-process pytest cases are skipped on 1-GPU boxes): 2*world + 1
    anchor frames (uneven shards) through `depth` global blocks sharded over the ring, against the SAME stack run
    un-sharded on rank 0's GPU; returns the rel-L2 of rank 0's rows (None on the other ranks)."""
    n_frames = 2 * world + 1
    plan, blocks, x0, rows_pos = _ring_problem(n_frames, depth, rank, world, dev)
    x = x0.clone()
    with torch.no_grad():
        run_global_stack(blocks, x, rows_pos, plan, None, RingWorkspace(plan, 1024, 16, dev))
    torch.cuda.synchronize()
    if rank != 0:
        return None
    plan1, _, x_all, pos_all = _ring_problem(n_frames, depth, 0, 1, dev)
    with torch.no_grad():
        run_global_stack(blocks, x_all, pos_all, plan1, None, RingWorkspace(plan1, 1024, 16, dev))
    ref = x_all[: x.shape[0]].double()
    return float((x.double() - ref).norm() / ref.norm())


def ring_measure(n_frames: int, depth: int, steps: int, warmup: int, rank: int, world: int, dev) -> dict:
    """The sequence-sharded global stack (BASELINE.json configs[3]) timed on the device: barrier + synchronize on both
    sides, CUDA events, max over ranks.  Also times the EXPOSED part of the K/V exchange (event pairs around the waits on
    the compute stream).  Every rank returns the same dict (frame0_out_l2 is broadcast from rank 0)."""
    from . import ops
    C, P = 1024, 1374
    plan, blocks, x0, rows_pos = _ring_problem(n_frames, depth, rank, world, dev)
    ws = RingWorkspace(plan, C, 16, dev)
    x = torch.empty_like(x0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    waits: list = []

    def step(ev=None):
        x.copy_(x0)
        with torch.no_grad():
            run_global_stack(blocks, x, rows_pos, plan, None, ws, wait_events=ev)

    for _ in range(warmup):
        step()
    barrier()
    n0 = ops.num_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        step(waits)
    e1.record()
    barrier()
    launches = ops.num_kernel_launches() - n0
    ms = e0.elapsed_time(e1) / steps
    wait_ms = sum(a.elapsed_time(b) for a, b in waits) / steps
    finite = bool(torch.isfinite(x).all())
    probe = float(x[:P].double().norm()) if rank == 0 else 0.0      # frame 0 lives on rank 0 for every world size
    if world > 1:
        t = torch.tensor([ms, wait_ms, 1.0 if finite else 0.0, probe], device=dev, dtype=torch.float64)
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        mn = t.clone()
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        ms, wait_ms, finite, probe = float(mx[0]), float(mx[1]), bool(mn[2] > 0), float(mx[3])
    L = n_frames * P
    flops = depth * (4.0 * L * L * C + L * 24.0 * C * C)
    return {"workload": f"ring: {depth} global blocks over {n_frames} anchor frames x {P} tokens (518 px), frames sharded over "
                        f"{world} rank(s), K|V blocks passed round the ring by NCCL send/recv under the attention tiles",
            "n_frames": n_frames, "depth": depth, "steps": steps, "warmup": warmup, "ms_per_step": ms,
            "frames_per_s": n_frames / (ms * 1e-3), "kv_bytes_per_step": 2 * plan.max_rows * C * 2 * depth * (world - 1),
            "nccl_wait_ms": wait_ms, "frame0_out_l2": probe, "outputs_finite": finite, "gpu_launches": launches,
            "algorithmic_tflop_per_step": flops / 1e12, "tflops_per_gpu": flops / (ms * 1e-3) / 1e12 / world,
            "scaling": "strong"}


def bench_ring(args, rank: int, world: int, local: int) -> None:
    """BASELINE.json configs[3]: the 24 global blocks over `--ring-frames` (512) anchor frames of 518x518 (1374 tokens
    each), frames sharded over the ranks, ring K/V over NVLink.  Strong scaling: total work is fixed as N grows.
    One step = one pass of the whole sharded stack; value = frames / s."""
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --workload ring needs B200s: sailrecon_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    import bench as _bench                                   # clock sampler + peaks live in bench.py
    clocks = _bench.ClockSampler(local)
    clocks.start()
    rec = ring_measure(args.ring_frames, args.ring_depth, args.steps, args.warmup, rank, world, dev)
    clk = clocks.stop()
    par = ring_parity(rank, world, dev) if world > 1 else None
    if rank == 0:
        peak_tf, peak_src = _bench.load_peak_tf()
        line = {"metric": "global-attention stack frames/s at N=%d (518px), sequence-sharded" % args.ring_frames,
                "value": rec["frames_per_s"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": rec["workload"],
                           "l2": "K/V block per head (>= 11 MB) x 16 heads and the fp32 carry exceed L2; no flush needed"},
                "gpu_launches": rec["gpu_launches"], "clocks": clk,
                "roofline": {"bound": "tensor", "kernel": "attn_tc + gemm_tc (whole stack)", "achieved": rec["tflops_per_gpu"],
                             "peak": peak_tf, "unit": "TFLOP/s", "frac": rec["tflops_per_gpu"] / peak_tf, "traffic": None,
                             "peak_source": peak_src},
                "sharded": rec, "ring_parity_rel_l2": par,
                "algorithmic_tflop_per_step": rec["algorithmic_tflop_per_step"], "frame0_out_l2": rec["frame0_out_l2"],
                "outputs_finite": rec["outputs_finite"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ whole scene, frame-sharded
@dataclass
class ScenePlan:
    """contiguous frame ranges per rank, anchors Na/R each and queries Nq/R each; the residual stream of
    a rank's frames stays on that rank for the whole 24-layer loop."""
    rank: int
    world: int
    Na: int
    Nq: int
    anchors: List[Tuple[int, int]]                     # per rank (first anchor, count)
    queries: List[Tuple[int, int]]                     # per rank (first query, count), query j is input frame Na + j

    @classmethod
    def make(cls, Na: int, Nq: int, rank: int, world: int) -> "ScenePlan":
        if Na < world:
            raise ValueError(f"sharded scene needs at least one anchor frame per rank (anchors {Na}, ranks {world})")
        return cls(rank, world, Na, Nq, partition_frames(Na, world), partition_frames(Nq, world))

    def local_frames(self, r: Optional[int] = None) -> List[int]:
        """Input-frame indices ([anchors ; queries] numbering of demo_imc_forward) held by rank r, anchors first."""
        r = self.rank if r is None else r
        (a0, na), (q0, nq) = self.anchors[r], self.queries[r]
        return list(range(a0, a0 + na)) + list(range(self.Na + q0, self.Na + q0 + nq))


def _allgather_rows(dst: torch.Tensor, local: torch.Tensor, counts: Sequence[int], rank: int, group=None) -> torch.Tensor:
    """dst[sum(counts), ...] <- concatenation of every rank's `local` rows.  Equal counts (the usual case: anchors divide
    evenly over the ranks) -> ONE all_gather_into_tensor; uneven or empty shards keep one broadcast per rank."""
    R = len(counts)
    if not (dist.is_initialized() and R > 1):
        if counts[rank]:
            off = sum(counts[:rank])
            dst[off:off + counts[rank]].copy_(local)
        return dst
    if local.shape[0] != counts[rank]:
        raise ValueError(f"_allgather_rows: rank {rank} holds {local.shape[0]} rows, the plan says {counts[rank]}")
    if min(counts) > 0 and len(set(counts)) == 1 and dst.is_contiguous() and local.is_contiguous():
        dist.all_gather_into_tensor(dst, local, group=group)
        return dst
    off = 0
    for r, c in enumerate(counts):
        if c:
            if r == rank:
                dst[off:off + c].copy_(local)
            dist.broadcast(dst[off:off + c], src=_global_rank(group, r), group=group)
        off += c
    return dst


def sharded_scene_forward(agg, head, images_local: torch.Tensor, plan: ScenePlan, fix_rank: Optional[int] = 300, group=None):
    """Aggregator.forward + CameraHead.forward (models/aggregator.py:242-433, heads/camera_head.py:85-121) for ONE scene
    whose frames are partitioned over the ranks.  `images_local` [1, na_r + nq_r, 3, H, W] holds this rank's anchors, then
    its queries (ScenePlan.local_frames).  Per layer: frame block rank-local; scene tokens of the local anchors projected to
    K|V locally and all-gathered (the reloc block's shared prefix); reloc block on the local query frames; global block
    with ring K/V over the anchors (ring_global_block).  The camera head runs on [all anchor camera tokens ; local queries]
    (its mask lets queries see the anchors and themselves only, camera_head.py:197-228).  agg.generator must be seeded
    identically on every rank: each rank replays the whole scene-token draw (aggregator.py:617-619) and keeps its part.
    Returns (output_dict of the local query frames, pose encodings of the local query frames)."""
    from . import ops
    from .engine import AttnCall, Prefix
    mode = engine.compute_mode()
    if mode != "bf16":
        raise NotImplementedError("sailrecon_b200: the sharded scene path runs the bf16 (autocast) mode")
    B, S_l, C_in, H, W = images_local.shape
    r, R = plan.rank, plan.world
    (a0, na), (q0, nq) = plan.anchors[r], plan.queries[r]
    if B != 1 or C_in != 3 or S_l != na + nq:
        raise ValueError(f"sharded_scene_forward: rank {r} expects [1, {na + nq}, 3, H, W], got {tuple(images_local.shape)}")
    engine.require_cuda_f32(images_local, "images_local")
    dev = images_local.device
    Na, Nq = plan.Na, plan.Nq
    agg.num_recon = Na
    C = agg.camera_token.shape[-1]
    psi = agg.patch_start_idx
    h, w = H // agg.patch_size, W // agg.patch_size
    hw = h * w
    P = psi + hw
    agg.rank = agg._draw_rank(fix_rank, hw, clamp=True)
    l_prime = min(agg.rank, hw)
    Pp = psi + l_prime
    act = torch.bfloat16

    X = engine.WS.get("shard.x", (S_l * P, C), torch.float32, dev)
    agg._embed(images_local[0].contiguous(), mode, X, S_l, P, C)
    with torch.no_grad():
        X.view(S_l, P, C)[:, :psi] = agg._special_rows(plan.local_frames(), set(range(Na, Na + Nq)), False)
    rows_pos, frame_pos_cpu = agg._positions(S_l, h, w, dev)
    pos_a = RowPositions(rows_pos.pos32[:na * P], rows_pos.max_pos)
    pos_q = RowPositions(rows_pos.pos32[na * P:], rows_pos.max_pos)
    Xa, Xq = X[:na * P], X[na * P:]
    frame_calls = engine.full_segments(S_l, P)
    ring = RingPlan(r, R, P, plan.anchors)
    key = ("shard.ring", Na, P, r, R, str(dev))
    ws = agg.__dict__.setdefault("_sr_ring_ws", {}).get(key)
    if ws is None:
        ws = RingWorkspace(ring, C, agg.global_blocks[0].attn.num_heads, dev)
        agg.__dict__["_sr_ring_ws"] = {key: ws}
    scene_counts = [c * Pp for _, c in plan.anchors]
    output_dict = {}
    cam_token_last_layer = None
    for i in range(agg.depth):
        engine.run_block(agg.frame_blocks[i], X, frame_calls, rows_pos, mode, tag="agg.")
        out_i = None
        if i in agg.intermediate_layer_idx and nq > 0:
            out_i = torch.empty(1, nq, P, 2 * C, dtype=torch.float32, device=dev)
            ops.copy_cols(Xq, out_i.view(nq * P, 2 * C), 0)
        if i == agg.depth - 1:
            cam_frame = X.view(S_l, P, C)[:na, 0].clone()
        # scene tokens: every rank replays the draw for ALL anchors, projects its own anchors' rows, gathers the rest
        packed, _ = agg._draw_scene(Na, P, hw, l_prime, frame_pos_cpu)
        mine = packed[a0 * Pp:(a0 + na) * Pp].clone()
        mine[:, 0] -= a0 * P                                                             # row index into the local X
        mine = mine.pin_memory().to(dev, non_blocking=True)
        idx, spos = mine[:, 0].contiguous(), mine[:, 1:].contiguous()
        rblk = agg.global_reloc_blocks[i]
        if Nq > 0:
            a_s = engine.WS.get("shard.scene_a", (na * Pp, C), act, dev)
            w1, b1, eps1 = engine.ln_params(rblk.norm1)
            ops.layernorm(X, w1, b1, eps1, a_s, idx=idx)
            skv_l = engine.WS.get("shard.scene_kv_local", (na * Pp, 2 * C), act, dev)
            engine.qkv_rows(rblk.attn, a_s, skv_l, mode, RowPositions(spos, rows_pos.max_pos), col_offset=C)
            skv = engine.WS.get("shard.scene_kv", (Na * Pp, 2 * C), act, dev)
            _allgather_rows(skv, skv_l, scene_counts, r, group)
            if nq > 0:
                engine.run_block(rblk, Xq, [AttnCall(0, nq, P, P, Prefix(skv[:, :C], skv[:, C:], 0, Na * Pp))], pos_q, mode,
                                 tag="agg.")
        ring_global_block(agg.global_blocks[i], Xa, pos_a, ring, ws, group, mode)
        if out_i is not None:
            ops.copy_cols(Xq, out_i.view(nq * P, 2 * C), C)
            output_dict[i] = out_i
        if i == agg.depth - 1:
            cam_l = torch.cat([cam_frame, X.view(S_l, P, C)[:na, 0]], dim=-1).contiguous()
            cam_all = torch.empty(Na, 2 * C, dtype=torch.float32, device=dev)
            _allgather_rows(cam_all, cam_l, [c for _, c in plan.anchors], r, group)
            cam_token_last_layer = cam_all[None]
    poses = []
    if nq > 0:
        output_dict[-1] = output_dict[agg.depth - 1]
        with torch.autocast("cuda", enabled=False):
            poses = head(output_dict, cam_token_last_layer)
    return output_dict, poses


def bench_scene(args, rank: int, world: int, local: int) -> None:
    """bench.py --workload scene: ONE scene of --anchors + --queries frames (default 256 + 256 = the N = 512 of the
    BASELINE.json metric), full aggregator + camera head, frames sharded over the ranks (sharded_scene_forward).  Strong
    scaling: total work is fixed as N grows.  value = frames / s, device time, max over ranks."""
    import bench as _bench
    from . import ops
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --workload scene needs B200s: sailrecon_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    Na, Nq, size = args.anchors, args.queries, args.size
    plan = ScenePlan.make(Na, Nq, rank, world)
    agg, head = _bench.build_models(dev)                    # torch.manual_seed(0): identical weights on every rank
    frames = plan.local_frames()
    host = torch.empty(1, len(frames), 3, size, size).pin_memory()
    for i, f in enumerate(frames):                          # seeded by the GLOBAL frame index: same scene for every world size
        host[0, i] = torch.rand(3, size, size, generator=torch.Generator().manual_seed(10_007 + f))
    images = host.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(img):
        agg.generator.manual_seed(1234)                     # the same scene-token draw on every rank
        with torch.no_grad():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return sharded_scene_forward(agg, head, img, plan, fix_rank=300)[1]

    for _ in range(args.warmup):
        step(images)
    barrier()
    clocks = _bench.ClockSampler(local)
    clocks.start()
    n0 = ops.num_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        poses = step(images)
    e1.record()
    barrier()
    launches = ops.num_kernel_launches() - n0
    clk = clocks.stop()
    ms = e0.elapsed_time(e1) / args.steps
    # end to end: local frames from pinned host memory in, local pose encodings out
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record()
    for _ in range(args.steps):
        poses = step(host.to(dev, non_blocking=True))
        host_poses = torch.stack(poses).cpu() if poses else torch.zeros(0)
    t1.record()
    barrier()
    ms_e2e = t0.elapsed_time(t1) / args.steps
    finite = bool(all(torch.isfinite(p).all() for p in poses))
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
        fin = torch.tensor([1.0 if finite else 0.0], device=dev)
        dist.all_reduce(fin, op=dist.ReduceOp.MIN)
        finite = bool(fin.item() > 0)
    if rank == 0:
        fl = _bench.algorithmic_flops(Na, Nq, size, size)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(_bench.ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak_tf = peaks.get("bf16_tflops_sustained", 1400.0)
        S = Na + Nq
        line = {"metric": f"aggregator+camera-head frames/s at N={S} ({Na} anchors + {Nq} queries, {size}px), frame-sharded",
                "value": S / (ms * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic",
                "config": {"workload": f"scene{S}: one scene of {Na} anchors + {Nq} queries, {size}x{size}, fix_rank 300, frames "
                                       f"sharded over {world} rank(s): frame blocks local, scene K|V all-gathered per layer, "
                                       "global blocks with ring K/V over NCCL, camera head on gathered anchor tokens",
                           "l2": "per-layer activations exceed the 126 MB L2; no flush needed"},
                "e2e": {"value": S / (ms_e2e * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": host.numel() * 4 * world,
                        "d2h_bytes_per_step": int(host_poses.numel() * 4 * world)},
                "gpu_launches": launches, "clocks": clk,
                "roofline": {"bound": "tensor", "kernel": "whole step", "achieved": fl["total"] / (ms * 1e-3) / 1e12,
                             "peak": peak_tf * world, "unit": "TFLOP/s", "frac": fl["total"] / (ms * 1e-3) / 1e12 / (peak_tf * world),
                             "traffic": None, "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops_sustained) x n_gpus"},
                "algorithmic_tflop_per_step": fl["total"] / 1e12, "outputs_finite": finite}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()

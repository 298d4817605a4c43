"""Host-side orchestration of one transformer block on row buffers — the fused schedule behind Block / Attention / Mlp.

The reference runs a block as ~20 library kernels (LayerNorm, cuBLAS, SDPA, ~10 pointwise ATen ops, see
sailrecon/layers/block.py:86-112 and attention.py:70-122).  Here a block is 7 launches of the C-ABI library on a
fp32 residual stream x [T, C] that is updated IN PLACE:

    a   = LN1(x)                         -> bf16                 sr_layernorm_fwd            (block.py:87 norm1)
    qkv = a Wqkv^T + b, q/k LN(d) + RoPE -> bf16 [T, 3C]         sr_gemm_bf16 / SR_EPI_QKV   (attention.py:73-82)
    o   = softmax(q k^T / sqrt d) v      -> bf16 [T, C]          sr_attn_fwd_bf16            (attention.py:103-109)
    x  += gamma1 * (o Wp^T + b)                                  sr_gemm_bf16 / LS_RESIDUAL  (attention.py:120, block.py:110)
    a   = LN2(x)                         -> bf16                 sr_layernorm_fwd            (block.py:91 norm2)
    h   = gelu(a W1^T + b)               -> bf16 [T, 4C]         sr_gemm_bf16 / BIAS_GELU    (mlp.py:35-36)
    x  += gamma2 * (h W2^T + b)                                  sr_gemm_bf16 / LS_RESIDUAL  (mlp.py:38, block.py:111)

`mode` is "bf16" (the rounding points of torch.autocast(bfloat16), which is how train/demo_imc_forward.py:92-101
runs the aggregator) or "f32" (no autocast: the camera head, sail_recon.py:119, and the fp32 check mode).
Attention is described structurally (AttnCall) instead of by a dense boolean mask.

PyTorch is plumbing only here: it owns the buffers and the stream.  There is no fallback path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from . import ops
from ._lib import SR_EPI_BIAS, SR_EPI_BIAS_GELU, SR_EPI_LS_RESIDUAL, SR_EPI_QKV


# ------------------------------------------------------------------------------------------------ precision mode
def compute_mode(x: Optional[torch.Tensor] = None) -> str:
    """'bf16' under torch.autocast('cuda', bfloat16) — the reference's inference/training setting — else 'f32'."""
    if torch.is_autocast_enabled("cuda"):
        dt = torch.get_autocast_dtype("cuda")
        if dt != torch.bfloat16:
            raise NotImplementedError(f"sailrecon_b200: autocast dtype {dt} is not supported (bfloat16 only)")
        return "bf16"
    return "f32"


def require_cuda_f32(x: torch.Tensor, name: str) -> None:
    if not x.is_cuda:
        raise RuntimeError(f"sailrecon_b200: {name} must live on a CUDA device (B200); there is no CPU fallback")
    if x.dtype != torch.float32:
        raise TypeError(f"sailrecon_b200: {name} must be float32 (the residual stream is fp32 as in the reference), got {x.dtype}")


def forbid_autograd(mod: nn.Module, what: str) -> None:
    """The inference forwards read every parameter detached and return tensors without grad_fn.  In train() mode with
    grad enabled that would silently freeze the module under an unmodified training loop (train_imc.py:364-381: only the
    reference's DPT heads would receive gradients), so it is an error instead.  Training goes through
    sailrecon_b200.train (training_step / AggregatorFunction)."""
    if mod.training and torch.is_grad_enabled() and any(p.requires_grad for p in mod.parameters()):
        raise RuntimeError(f"sailrecon_b200: {what} is the inference path (parameters are read detached, outputs carry no "
                           "grad_fn); it was called in train() mode with autograd enabled.  Use model.eval() / torch.no_grad() "
                           "for inference, or sailrecon_b200.train.training_step for the differentiated block stacks.")


# ------------------------------------------------------------------------------------------------ workspace
class Workspace:
    """Named scratch buffers that grow monotonically, one set per (device, CUDA stream): all launches enqueued on one
    stream are ordered, so a buffer can be reused as soon as the kernel that read it has been enqueued — and two forwards
    running on two streams (or two models driven from two threads, each on its own stream) never share scratch
    (SURVEY.md §8b "re-entrant per stream").  Two models interleaved on the SAME stream share buffers safely for the
    same reason: their launches are serialised by the stream."""

    def __init__(self):
        self._buf: Dict[Tuple[str, torch.device, int], torch.Tensor] = {}

    def get(self, name: str, shape: Sequence[int], dtype: torch.dtype, device) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = max(n * torch.empty((), dtype=dtype).element_size(), 16)
        dev = torch.device(device)
        stream = torch.cuda.current_stream(dev).cuda_stream if dev.type == "cuda" else 0
        key = (name, dev, stream)
        buf = self._buf.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
            self._buf[key] = buf
        return buf[:nbytes].view(dtype)[:n].view(*shape)

    def clear(self):
        self._buf.clear()

    def detach(self, stream: int, device=None) -> list:
        """Forget (and hand back) every buffer of the raw stream handle `stream`: nothing is evicted otherwise, so each
        stream that ever ran a forward keeps its own scratch set (multi-GB at 32 frames of 518 px) for the life of the
        process, and torch recycles raw handles from a 32-entry pool — a later owner of the same handle would otherwise
        grow / replace buffers somebody else still uses.  The caller either drops the returned list (memory goes back
        to the allocator once the stream's work is done) or keeps it alive (localize.Localizer: a captured CUDA graph
        holds the addresses)."""
        dev = None if device is None else torch.device(device)
        keys = [k for k in self._buf if k[2] == stream and (dev is None or k[1] == dev)]
        return [self._buf.pop(k) for k in keys]


WS = Workspace()


# ------------------------------------------------------------------------------------------------ cached parameters
def _cached(mod: nn.Module, key: str, src: torch.Tensor, make):
    """Derived copies of a parameter (bf16 weight, bf16-rounded bias), rebuilt when the parameter changes."""
    cache = mod.__dict__.setdefault("_sr_cache", {})
    stamp = (src.data_ptr(), src._version, src.device, tuple(src.shape))
    hit = cache.get(key)
    if hit is None or hit[0] != stamp:
        with torch.no_grad():
            hit = (stamp, make(src.detach()))
        cache[key] = hit
    return hit[1]


def weight_bf16(lin: nn.Module, pad_k: int = 0) -> torch.Tensor:
    """nn.Linear / Conv2d weight as bf16 [out, K(+pad)] — the cast autocast performs on every call, done once."""
    def make(w):
        w2 = w.reshape(w.shape[0], -1).contiguous()
        if pad_k and w2.shape[1] % pad_k:
            kp = (w2.shape[1] + pad_k - 1) // pad_k * pad_k
            full = torch.zeros(w2.shape[0], kp, dtype=torch.float32, device=w.device)
            full[:, : w2.shape[1]] = w2
            w2 = full
        return ops.cast_bf16(w2)
    return _cached(lin, f"w_bf16_{pad_k}", lin.weight, make)


def bias_for(lin: nn.Module, mode: str) -> Optional[torch.Tensor]:
    if getattr(lin, "bias", None) is None:
        return None
    if mode == "f32":
        return lin.bias.detach()
    # autocast casts the bias to bf16 as well; the kernels add it in fp32, so round once here
    return _cached(lin, "b_r", lin.bias, lambda b: ops.cast_bf16(b.reshape(1, -1).contiguous()).float().reshape(-1))


def weight_f32(lin: nn.Module) -> torch.Tensor:
    w = lin.weight.detach()
    return w.reshape(w.shape[0], -1)


def ln_params(norm: nn.Module):
    """(weight, bias, eps) of an nn.LayerNorm (possibly without affine); Identity is not a LayerNorm."""
    if not isinstance(norm, nn.LayerNorm):
        raise TypeError(f"sailrecon_b200: expected nn.LayerNorm, got {type(norm).__name__}")
    w = norm.weight.detach() if norm.weight is not None else None
    b = norm.bias.detach() if norm.bias is not None else None
    return w, b, float(norm.eps)


def gamma_of(ls: nn.Module) -> Optional[torch.Tensor]:
    g = getattr(ls, "gamma", None)
    return None if g is None else g.detach()


# ------------------------------------------------------------------------------------------------ attention structure
@dataclass
class Prefix:
    """Keys/values every segment of a call attends to before its own rows."""
    k: Optional[torch.Tensor] = None       # external [rows, H*d] (kv-cache / scene tokens); None -> rows of this call's own qkv
    v: Optional[torch.Tensor] = None
    row0: int = 0                          # first row (of the external tensors, or of the call's own k/v)
    length: int = 0


@dataclass
class AttnCall:
    """`batch` segments: queries rows [q_row0 + b*stride, +q_len) attend to prefix + own rows [kv_row0 + b*stride, +kv_len)."""
    q_row0: int
    batch: int
    seg_len: int
    stride: int = 0
    prefix: Optional[Prefix] = None


def full_segments(batch: int, seg_len: int) -> List[AttnCall]:
    return [AttnCall(0, batch, seg_len, seg_len)]


@dataclass
class MaskStructure:
    """The only mask family SAIL-Recon builds (aggregator.py:832-851, camera_head.py:197-228): rows [0, prefix_len)
    attend to each other; then `num_segs` segments of `seg_len` rows attend to the prefix and to themselves."""
    prefix_len: int
    seg_len: int
    num_segs: int

    def calls(self) -> List[AttnCall]:
        out = []
        if self.prefix_len:
            out.append(AttnCall(0, 1, self.prefix_len, self.prefix_len))
        if self.num_segs:
            out.append(AttnCall(self.prefix_len, self.num_segs, self.seg_len, self.seg_len,
                                Prefix(None, None, 0, self.prefix_len) if self.prefix_len else None))
        return out

    def dense(self, device=None) -> torch.Tensor:
        L = self.prefix_len + self.seg_len * self.num_segs
        m = torch.zeros(L, L, dtype=torch.bool, device=device)
        m[:, : self.prefix_len] = True
        for i in range(self.num_segs):
            s = self.prefix_len + i * self.seg_len
            m[s:s + self.seg_len, s:s + self.seg_len] = True
        return m


def structure_from_mask(mask: torch.Tensor, n_rows: int) -> MaskStructure:
    """Recover (prefix, segments) from a dense boolean allow-mask [.., L, L] (True = attend, attention.py:103-109) and
    verify it.  Anything else raises: this library implements SAIL-Recon's structured masks, not arbitrary ones."""
    if mask.dtype != torch.bool:
        raise TypeError("sailrecon_b200: attn_mask must be a boolean tensor (True = attend)")
    m = mask
    while m.dim() > 2:
        if m.shape[0] != 1:
            raise NotImplementedError("sailrecon_b200: per-batch / per-head masks are not supported")
        m = m[0]
    L = m.shape[-1]
    if m.shape[0] != L or L != n_rows:
        raise ValueError(f"sailrecon_b200: attn_mask shape {tuple(mask.shape)} does not match sequence length {n_rows}")
    first, last = int(m[0].sum()), int(m[-1].sum())
    if not bool(m[0, :first].all()):
        raise NotImplementedError("sailrecon_b200: unsupported attention mask structure")
    if first == L:
        st = MaskStructure(L, 0, 0) if bool(m.all()) else None
    elif bool(m[0, 0]) and first < L and last > first and (L - first) % (last - first) == 0 and bool(m[-1, 0]):
        st = MaskStructure(first, last - first, (L - first) // (last - first))
    elif last >= 1 and not bool(m[-1, 0]) and L % last == 0:
        st = MaskStructure(0, last, L // last)             # pure block diagonal
    else:
        st = None
    if st is None or not torch.equal(st.dense(m.device), m):
        raise NotImplementedError("sailrecon_b200: unsupported attention mask structure (only [shared prefix | block "
                                  "diagonal] masks as built by aggregator.build_allow_block / camera_head.build_lr_mask)")
    return st


# ------------------------------------------------------------------------------------------------ positions / RoPE
@dataclass
class RowPositions:
    """int32 [M, 2] (y, x) per row + the table extent, prepared once per forward (no device->host sync per call as
    in rope.py:190)."""
    pos32: torch.Tensor
    max_pos: int


# ------------------------------------------------------------------------------------------------ the block schedule
def qkv_rows(attn: nn.Module, a_in: torch.Tensor, out: torch.Tensor, mode: str, rows_pos: Optional[RowPositions],
             col_offset: int = 0, n_cols: Optional[int] = None) -> torch.Tensor:
    """qkv Linear + per-head q/k LayerNorm + RoPE (attention.py:73-82) on rows.  col_offset = C computes only the
    k|v columns (scene tokens are K/V producers only, aggregator.py:737)."""
    C = attn.qkv.weight.shape[1]
    d = C // attn.num_heads
    qn = kn = None
    if isinstance(attn.q_norm, nn.LayerNorm):
        qn = (attn.q_norm.weight.detach(), attn.q_norm.bias.detach())
        kn = (attn.k_norm.weight.detach(), attn.k_norm.bias.detach())
    cos = sin = pos32 = None
    if attn.rope is not None and rows_pos is not None:
        cos, sin = attn.rope.tables(a_in.device, rows_pos.max_pos, d)
        pos32 = rows_pos.pos32
        if pos32.shape[0] != a_in.shape[0]:
            raise ValueError("sailrecon_b200: one (y, x) position per token row is required")
    if qn is None and cos is None:
        epi_kind, epi = SR_EPI_BIAS, None
    else:
        if d != 64:
            raise NotImplementedError("sailrecon_b200: q/k LayerNorm and RoPE are implemented for head_dim 64")
        epi_kind = SR_EPI_QKV
        epi = ops.QkvEpilogue(C, col_offset, d, qn, kn, attn.q_norm.eps if qn is not None else 1e-5, cos, sin, pos32)
    if mode == "bf16":
        end = 3 * C if n_cols is None else col_offset + n_cols        # n_cols: only a column window (ring attention: q, then k|v)
        w = weight_bf16(attn.qkv)[col_offset:end]
        b = bias_for(attn.qkv, mode)
        ops.gemm(a_in, w, None if b is None else b[col_offset:end], out, kind=epi_kind, qkv=epi)
    else:
        if col_offset or n_cols is not None:
            raise ValueError("sailrecon_b200: the fp32 path computes the full qkv projection")
        ops.gemm(a_in, weight_f32(attn.qkv), bias_for(attn.qkv, mode), out, kind=epi_kind, qkv=epi)
    return out


def attention_rows(attn: nn.Module, qkv: torch.Tensor, o: torch.Tensor, calls: Sequence[AttnCall]) -> torch.Tensor:
    """softmax(q k^T) v for every call; qkv [T, 3C] (q | k | v column blocks), o [T, C]."""
    C = attn.qkv.weight.shape[1]
    H = attn.num_heads
    d = C // H
    q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
    for c in calls:
        stride = c.stride or c.seg_len
        r1 = c.q_row0 + (c.batch - 1) * stride + c.seg_len
        pk = pv = None
        plen = 0
        if c.prefix is not None and c.prefix.length > 0:
            p = c.prefix
            pk = (k if p.k is None else p.k)[p.row0:p.row0 + p.length]
            pv = (v if p.v is None else p.v)[p.row0:p.row0 + p.length]
            plen = p.length
        ops.attention(q[c.q_row0:r1], k[c.q_row0:r1], v[c.q_row0:r1], o[c.q_row0:r1], H, d, c.batch, c.seg_len, c.seg_len,
                      stride, stride, prefix_k=pk, prefix_v=pv, prefix_len=plen, scale=attn.scale)
    return o


def run_block(blk: nn.Module, x: torch.Tensor, calls: Sequence[AttnCall], rows_pos: Optional[RowPositions], mode: str,
              tag: str = "", attn_zero: bool = False, capture_kv: Optional[list] = None) -> torch.Tensor:
    """Block.forward (block.py:86-112) in place on the fp32 residual rows x [T, C].  `attn_zero` reproduces the
    kv-cache priming pass where Attention returns zeros (attention.py:94-100).  `capture_kv`: a list that receives a copy
    of this call's (K, V) rows [T, C] (the anchors' keys / values the camera head caches for later query-only calls)."""
    T, C = x.shape
    dev = x.device
    act = torch.bfloat16 if mode == "bf16" else torch.float32
    attn, mlp = blk.attn, blk.mlp
    a = WS.get(tag + "a", (T, C), act, dev)
    if not attn_zero:
        w1, b1, eps1 = ln_params(blk.norm1)
        ops.layernorm(x, w1, b1, eps1, a)
        qkv = WS.get(tag + "qkv", (T, 3 * C), act, dev)
        qkv_rows(attn, a, qkv, mode, rows_pos)
        if capture_kv is not None:
            capture_kv.append((qkv[:, C:2 * C].clone(), qkv[:, 2 * C:].clone()))
        attention_rows(attn, qkv, a, calls)              # `a` is free again: reuse it for the attention output
    block_tail(blk, x, a, mode, tag, attn_zero)
    return x


def block_tail(blk: nn.Module, x: torch.Tensor, a: torch.Tensor, mode: str, tag: str = "", attn_zero: bool = False) -> torch.Tensor:
    """Everything of Block.forward after the attention core: x += g1 * proj(a); x += g2 * mlp(LN2(x)) (block.py:110-111).
    `a` [T, C] holds the attention output on entry and is reused as scratch."""
    T, C = x.shape
    dev = x.device
    act = torch.bfloat16 if mode == "bf16" else torch.float32
    attn, mlp = blk.attn, blk.mlp
    if not attn_zero:
        wp = weight_bf16(attn.proj) if mode == "bf16" else weight_f32(attn.proj)
        ops.gemm(a, wp, bias_for(attn.proj, mode), x, kind=SR_EPI_LS_RESIDUAL, gamma=gamma_of(blk.ls1), resid=x)
    w2, b2, eps2 = ln_params(blk.norm2)
    ops.layernorm(x, w2, b2, eps2, a)
    hid = mlp.fc1.weight.shape[0]
    h = WS.get(tag + "h", (T, hid), act, dev)
    if mode == "bf16":
        ops.gemm(a, weight_bf16(mlp.fc1), bias_for(mlp.fc1, mode), h, kind=SR_EPI_BIAS_GELU)
        ops.gemm(h, weight_bf16(mlp.fc2), bias_for(mlp.fc2, mode), x, kind=SR_EPI_LS_RESIDUAL, gamma=gamma_of(blk.ls2), resid=x)
    else:
        ops.gemm(a, weight_f32(mlp.fc1), bias_for(mlp.fc1, mode), h, kind=SR_EPI_BIAS_GELU)
        ops.gemm(h, weight_f32(mlp.fc2), bias_for(mlp.fc2, mode), x, kind=SR_EPI_LS_RESIDUAL, gamma=gamma_of(blk.ls2), resid=x)
    return x


def run_block_camera_rows(blk: nn.Module, X: torch.Tensor, S: int, P: int, prefix: Prefix,
                          rows_pos: Optional[RowPositions], mode: str, tag: str = "") -> torch.Tensor:
    """The LAST reloc block of a camera-only localisation (SailRecon.reloc(fast_reloc=True), sail_recon.py:240-244): the
    caller only reads token 0 of every frame (`rgb_feats[-1][:, :, 0]`), so the block's query side, output projection and
    MLP run on those S rows alone; K / V are still produced for all S*P rows (the camera token attends to its whole frame
    behind the cached scene prefix).  X [S*P, C] fp32 is NOT modified; returns the S updated camera rows [S, C] fp32."""
    T, C = X.shape
    dev = X.device
    act = torch.bfloat16 if mode == "bf16" else torch.float32
    attn = blk.attn
    a = WS.get(tag + "a", (T, C), act, dev)
    w1, b1, eps1 = ln_params(blk.norm1)
    ops.layernorm(X, w1, b1, eps1, a)
    cam_pos = None
    if rows_pos is not None:
        cam_pos = RowPositions(rows_pos.pos32.view(S, P, 2)[:, 0].contiguous(), rows_pos.max_pos)
    if mode == "bf16":
        kv = WS.get(tag + "kv", (T, 2 * C), act, dev)
        qkv_rows(attn, a, kv, mode, rows_pos, col_offset=C)                           # k | v for every row
        k, v = kv[:, :C], kv[:, C:]
        a_cam = a.view(S, P, C)[:, 0].contiguous()
        q = WS.get(tag + "q_cam", (S, C), act, dev)
        qkv_rows(attn, a_cam, q, mode, cam_pos, col_offset=0, n_cols=C)               # q for the camera rows only
    else:
        qkv = WS.get(tag + "qkv", (T, 3 * C), act, dev)
        qkv_rows(attn, a, qkv, mode, rows_pos)
        k, v = qkv[:, C:2 * C], qkv[:, 2 * C:]
        q = qkv.view(S, P, 3 * C)[:, 0, :C].contiguous()
    o = WS.get(tag + "o_cam", (S, C), act, dev)
    H = attn.num_heads
    ops.attention(q, k, v, o, H, C // H, S, 1, P, 1, P, prefix_k=prefix.k[prefix.row0:prefix.row0 + prefix.length],
                  prefix_v=prefix.v[prefix.row0:prefix.row0 + prefix.length], prefix_len=prefix.length, scale=attn.scale)
    x_cam = X.view(S, P, C)[:, 0].clone()
    block_tail(blk, x_cam, o, mode, tag + "cam.")
    return x_cam


def linear_rows(lin: nn.Module, a_in: torch.Tensor, out: torch.Tensor, mode: str, kind: int = SR_EPI_BIAS) -> torch.Tensor:
    """A plain nn.Linear on rows through the GEMM kernels (bf16 tensor-core path needs K % 64 == 0, N % 128 == 0)."""
    w = weight_bf16(lin) if mode == "bf16" else weight_f32(lin)
    return ops.gemm(a_in, w, bias_for(lin, mode), out, kind=kind)

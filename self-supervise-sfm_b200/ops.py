"""Thin torch-tensor wrappers over the C-ABI (include/sailrecon_b200.h).

PyTorch is only the plumbing here: it owns device memory and the stream.  Every function below hands raw
device pointers, row strides and the current CUDA stream to libsailrecon_b200.so and raises if the call fails.
Nothing in this module computes with torch ops; there is no fallback path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import (SR_EPI_BIAS, SR_EPI_BIAS_GELU, SR_EPI_BIAS_SILU, SR_EPI_LS_RESIDUAL, SR_EPI_QKV, SrAttn,
                   SrEpilogue)

__all__ = [
    "layernorm", "gemm", "attention", "gather_rows", "copy_cols", "cast_bf16", "adaln_modulate",
    "QkvEpilogue", "num_kernel_launches", "rope_tables", "rope2d", "patchify", "pose_update", "scale_cols",
]


# Optional per-launch timing (bench.py): when `PROFILE` is a list, every wrapper below brackets its launch with CUDA
# events on the launching stream and appends (kernel class, algorithmic flops, algorithmic bytes, start, end).
PROFILE = None


class _Timed:
    __slots__ = ("rec",)

    def __init__(self, label: str, flops: float, nbytes: float):
        self.rec = None
        if PROFILE is not None:
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
            self.rec = (label, flops, nbytes, e0)

    def done(self):
        if self.rec is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.append((*self.rec, e1))


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _rows2d(t: torch.Tensor, dtype, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"sailrecon_b200: {name} must be a CUDA tensor (no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"sailrecon_b200: {name} must be {dtype}, got {t.dtype}")
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"sailrecon_b200: {name} must be 2-D with unit inner stride, got {tuple(t.shape)} {t.stride()}")


def _vec(t: Optional[torch.Tensor], n: int, name: str) -> None:
    if t is None:
        return
    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n):
        raise ValueError(f"sailrecon_b200: {name} must be a contiguous fp32 CUDA vector of {n} elements")


def num_kernel_launches() -> int:
    return int(_lib.load().sr_num_kernel_launches())


# ------------------------------------------------------------------------------------------------ LayerNorm
def layernorm(x: torch.Tensor, weight: Optional[torch.Tensor], bias: Optional[torch.Tensor], eps: float,
              out: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[r] = LayerNorm(x[idx[r]] if idx is not None else x[r]); out dtype bf16 or fp32 picks the store."""
    _rows2d(x, torch.float32, "x")
    rows = out.shape[0]
    Cdim = x.shape[1]
    _rows2d(out, out.dtype, "out")
    if out.dtype not in (torch.bfloat16, torch.float32) or out.shape[1] != Cdim:
        raise ValueError("sailrecon_b200: layernorm out must be bf16/fp32 [rows, C]")
    _vec(weight, Cdim, "weight")
    _vec(bias, Cdim, "bias")
    if idx is not None:
        if not (idx.is_cuda and idx.dtype == torch.int32 and idx.is_contiguous() and idx.numel() == rows):
            raise ValueError("sailrecon_b200: idx must be a contiguous int32 CUDA vector with one entry per output row")
    elif x.shape[0] != rows:
        raise ValueError("sailrecon_b200: layernorm x/out row mismatch")
    lib = _lib.load()
    t = _Timed("layernorm", 0.0, rows * Cdim * (4 + out.element_size()))
    _lib.check(lib.sr_layernorm_gather_fwd(x.data_ptr(), x.stride(0), _ptr(idx), _ptr(weight), _ptr(bias), out.data_ptr(),
                                           int(out.dtype == torch.bfloat16), out.stride(0), rows, Cdim, float(eps),
                                           _stream()), "layernorm")
    t.done()
    return out


def adaln_modulate(x: torch.Tensor, mod: torch.Tensor, eps: float, out: torch.Tensor) -> torch.Tensor:
    """camera_head.py:157-160: gate * (LN_noaffine(x) * (1 + scale) + shift) + x with mod = [shift|scale|gate]."""
    _rows2d(x, torch.float32, "x")
    _rows2d(mod, torch.float32, "mod")
    _rows2d(out, torch.float32, "out")
    rows, Cdim = x.shape
    if not (x.is_contiguous() and mod.is_contiguous() and out.is_contiguous() and mod.shape == (rows, 3 * Cdim)
            and out.shape == x.shape):
        raise ValueError("sailrecon_b200: adaln_modulate needs contiguous x [R,C], mod [R,3C], out [R,C]")
    _lib.check(_lib.load().sr_adaln_modulate_f32(x.data_ptr(), mod.data_ptr(), out.data_ptr(), rows, Cdim, float(eps),
                                                 _stream()), "adaln_modulate")
    return out


# ------------------------------------------------------------------------------------------------ GEMM
class QkvEpilogue:
    """Arguments of the fused q/k LayerNorm + 2D-RoPE epilogue (attention.py:78-82)."""

    def __init__(self, qkv_dim: int, col_offset: int = 0, head_dim: int = 64, q_norm=None, k_norm=None, qk_eps: float = 1e-5,
                 rope_cos: Optional[torch.Tensor] = None, rope_sin: Optional[torch.Tensor] = None,
                 pos: Optional[torch.Tensor] = None):
        self.qkv_dim, self.col_offset, self.head_dim = qkv_dim, col_offset, head_dim
        self.q_norm, self.k_norm, self.qk_eps = q_norm, k_norm, qk_eps      # (weight, bias) pairs or None
        self.rope_cos, self.rope_sin, self.pos = rope_cos, rope_sin, pos


def _fill_qkv(e: SrEpilogue, qkv: "QkvEpilogue", M: int) -> list:
    """Copies the q/k LayerNorm + RoPE arguments into the C struct; returns the tensors that must stay alive."""
    keep = []
    e.qkv_dim, e.col_offset, e.head_dim = qkv.qkv_dim, qkv.col_offset, qkv.head_dim
    e.qk_norm = int(qkv.q_norm is not None)
    if qkv.q_norm is not None:
        for t in (*qkv.q_norm, *qkv.k_norm):
            _vec(t, qkv.head_dim, "q/k norm parameter")
        e.qn_w, e.qn_b = qkv.q_norm[0].data_ptr(), qkv.q_norm[1].data_ptr()
        e.kn_w, e.kn_b = qkv.k_norm[0].data_ptr(), qkv.k_norm[1].data_ptr()
    e.qk_eps = float(qkv.qk_eps)
    e.rope = int(qkv.rope_cos is not None)
    if qkv.rope_cos is not None:
        cos, sin, pos = qkv.rope_cos, qkv.rope_sin, qkv.pos
        if not (cos.is_cuda and cos.dtype == torch.float32 and cos.is_contiguous() and cos.shape == sin.shape
                and cos.dim() == 2 and cos.shape[1] == qkv.head_dim // 4):
            raise ValueError("sailrecon_b200: rope tables must be fp32 [max_pos, head_dim/4]")
        if not (pos.is_cuda and pos.dtype == torch.int32 and pos.is_contiguous() and pos.shape == (M, 2)):
            raise ValueError("sailrecon_b200: rope positions must be int32 [M, 2]")
        e.rope_cos, e.rope_sin, e.rope_max_pos, e.pos = cos.data_ptr(), sin.data_ptr(), cos.shape[0], pos.data_ptr()
        keep += [cos, sin, pos]
    return keep


def gemm(a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], out: torch.Tensor, kind: int = SR_EPI_BIAS,
         gamma: Optional[torch.Tensor] = None, resid: Optional[torch.Tensor] = None,
         qkv: Optional[QkvEpilogue] = None, exact: bool = False) -> torch.Tensor:
    """out = epilogue(a @ w.T + bias).  a/w both bf16 -> tcgen05 path; both fp32 -> SIMT fp32 path."""
    if a.dtype != w.dtype or a.dtype not in (torch.bfloat16, torch.float32):
        raise TypeError(f"sailrecon_b200: gemm operands must both be bf16 or both fp32, got {a.dtype} / {w.dtype}")
    _rows2d(a, a.dtype, "a")
    _rows2d(w, w.dtype, "w")
    _rows2d(out, out.dtype, "out")
    M, K = a.shape
    N = w.shape[0]
    if w.shape[1] != K or out.shape != (M, N):
        raise ValueError(f"sailrecon_b200: gemm shape mismatch a{tuple(a.shape)} w{tuple(w.shape)} out{tuple(out.shape)}")
    _vec(bias, N, "bias")
    _vec(gamma, N, "gamma")
    e = SrEpilogue()
    e.kind, e.out_bf16 = kind, int(out.dtype == torch.bfloat16)
    e.out, e.ldo = out.data_ptr(), out.stride(0)
    e.bias, e.gamma = _ptr(bias), _ptr(gamma)
    if kind == SR_EPI_LS_RESIDUAL:
        if resid is None or out.dtype != torch.float32:
            raise ValueError("sailrecon_b200: LS_RESIDUAL needs an fp32 residual and fp32 out")
        _rows2d(resid, torch.float32, "resid")
        if resid.shape != (M, N):
            raise ValueError("sailrecon_b200: residual shape mismatch")
        e.resid, e.ldr = resid.data_ptr(), resid.stride(0)
        e.exact_residual = int(exact)
    keep = []
    if kind == SR_EPI_QKV:
        if qkv is None:
            raise ValueError("sailrecon_b200: QKV epilogue arguments missing")
        keep = _fill_qkv(e, qkv, M)
    lib = _lib.load()
    fn = lib.sr_gemm_bf16 if a.dtype == torch.bfloat16 else lib.sr_gemm_f32
    t = _Timed("gemm_tc" if a.dtype == torch.bfloat16 else "gemm_f32", 2.0 * M * N * K,
               (M * K + N * K) * a.element_size() + M * N * out.element_size())
    _lib.check(fn(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), M, N, K, C.byref(e), _stream()), "gemm")
    t.done()
    return out


def gemm_t(a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, a_mn_major: bool, accumulate: bool = False) -> torch.Tensor:
    """The tensor-core GEMM on operands as the Linear gradients find them (sr_gemm_bf16_t), no transposed copies:
        a_mn_major False:  out[M, N] = a[M, K] @ w[K, N]            (dX = dY W,  w = the Linear's weight [N_out, K_in])
        a_mn_major True:   out[M, N] = a[K, M]^T @ w[K, N]          (dW = dY^T X, K = token rows, any length)
    bf16 operands; out bf16 or fp32; accumulate (fp32 out only): out += (exact fp32 accumulate)."""
    _rows2d(a, torch.bfloat16, "a")
    _rows2d(w, torch.bfloat16, "w")
    _rows2d(out, out.dtype, "out")
    K, N = w.shape
    M = a.shape[1] if a_mn_major else a.shape[0]
    if (a.shape[0] if a_mn_major else a.shape[1]) != K or out.shape != (M, N):
        raise ValueError(f"sailrecon_b200: gemm_t shape mismatch a{tuple(a.shape)} w{tuple(w.shape)} out{tuple(out.shape)}")
    e = SrEpilogue()
    e.out_bf16 = int(out.dtype == torch.bfloat16)
    e.out, e.ldo = out.data_ptr(), out.stride(0)
    if accumulate:
        if out.dtype != torch.float32:
            raise ValueError("sailrecon_b200: gemm_t accumulates into fp32 outputs only")
        e.kind, e.resid, e.ldr, e.exact_residual = SR_EPI_LS_RESIDUAL, out.data_ptr(), out.stride(0), 1
    else:
        e.kind = SR_EPI_BIAS
    t = _Timed("gemm_tc", 2.0 * M * N * K, (M * K + N * K) * 2 + M * N * out.element_size())
    _lib.check(_lib.load().sr_gemm_bf16_t(a.data_ptr(), a.stride(0), int(a_mn_major), w.data_ptr(), w.stride(0), 1, M, N, K,
                                          C.byref(e), _stream()), "gemm_t")
    t.done()
    return out


# ------------------------------------------------------------------------------------------------ attention
def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out: Optional[torch.Tensor], num_heads: int,
              head_dim: int, batch: int, q_len: int, kv_len: int, q_batch_stride: int = 0, kv_batch_stride: int = 0,
              prefix_k: Optional[torch.Tensor] = None, prefix_v: Optional[torch.Tensor] = None, prefix_len: int = 0,
              scale: Optional[float] = None, carry=None, carry_in: bool = False, carry_out: bool = False,
              lse: Optional[torch.Tensor] = None, variant: int = 0):
    """Structured attention (see sr_attn in the header).  q/k/v are 2-D row views [rows, >= heads*head_dim]
    (column slices of the fused qkv buffer are fine); bf16 -> tcgen05 flash kernel, fp32 -> SIMT kernel."""
    dt = q.dtype
    for name, t in (("q", q), ("k", k), ("v", v)):
        _rows2d(t, dt, name)
    if k.stride(0) != v.stride(0):
        raise ValueError("sailrecon_b200: k and v must share a row stride")
    a = SrAttn()
    a.q, a.ldq = q.data_ptr(), q.stride(0)
    a.k, a.v, a.ldkv = k.data_ptr(), v.data_ptr(), k.stride(0)
    if prefix_k is not None and prefix_len > 0:
        _rows2d(prefix_k, dt, "prefix_k")
        _rows2d(prefix_v, dt, "prefix_v")
        if prefix_k.stride(0) != prefix_v.stride(0) or prefix_k.shape[0] < prefix_len:
            raise ValueError("sailrecon_b200: bad prefix K/V")
        a.pk, a.pv, a.ldp, a.prefix_len = prefix_k.data_ptr(), prefix_v.data_ptr(), prefix_k.stride(0), prefix_len
    if out is not None:
        _rows2d(out, dt, "out")
        a.out, a.ldo = out.data_ptr(), out.stride(0)
    a.num_heads, a.head_dim, a.batch = num_heads, head_dim, batch
    a.q_len, a.q_batch_stride = q_len, q_batch_stride or q_len
    a.kv_len, a.kv_batch_stride = kv_len, kv_batch_stride or kv_len
    need_q = (batch - 1) * a.q_batch_stride + q_len
    need_kv = (batch - 1) * a.kv_batch_stride + kv_len
    if q.shape[0] < need_q or k.shape[0] < need_kv or v.shape[0] < need_kv or (out is not None and out.shape[0] < need_q):
        raise ValueError("sailrecon_b200: attention row ranges exceed the tensors")
    a.scale = float(scale if scale is not None else head_dim ** -0.5)
    if carry is not None:
        m, l, acc = carry
        a.carry_m, a.carry_l, a.carry_acc = m.data_ptr(), l.data_ptr(), acc.data_ptr()
    a.carry_in, a.carry_out = int(carry_in), int(carry_out)
    a.variant = int(variant)
    if lse is not None:                                     # training: see new_lse()
        if dt != torch.bfloat16 or lse.dtype != torch.float32 or not lse.is_contiguous() or lse.dim() != 3 \
                or lse.shape[0] != num_heads or lse.shape[1] != batch or lse.shape[2] % 64 or lse.shape[2] < q_len:
            raise ValueError("sailrecon_b200: lse must be contiguous fp32 [num_heads, batch, q_pad] (bf16 attention only)")
        a.lse, a.lse_pad = lse.data_ptr(), lse.shape[2]
    lib = _lib.load()
    if dt == torch.bfloat16 and out is not None and carry is None:
        need = int(lib.sr_attn_workspace_bytes(C.byref(a), _num_sms(q.device)))
        if need > 0:                                        # under-filled launch: split the K/V range over more CTAs
            ws = _attn_workspace(need, q.device)
            a.workspace, a.workspace_bytes = ws.data_ptr(), need
    fn = lib.sr_attn_fwd_bf16 if dt == torch.bfloat16 else lib.sr_attn_fwd_f32
    hd = num_heads * head_dim
    t = _Timed("attn_tc" if dt == torch.bfloat16 else "attn_f32", 4.0 * batch * q_len * (prefix_len + kv_len) * hd,
               (2 * batch * q_len + 2 * batch * kv_len + 2 * prefix_len) * hd * q.element_size())
    _lib.check(fn(C.byref(a), _stream()), "attention")
    t.done()
    return out


_SMS: dict = {}
_ATTN_WS: dict = {}


def _num_sms(device) -> int:
    d = torch.device(device)
    if d not in _SMS:
        _SMS[d] = torch.cuda.get_device_properties(d).multi_processor_count
    return _SMS[d]


def _attn_workspace(nbytes: int, device) -> torch.Tensor:
    """Split-KV scratch, one growing buffer per (device, stream) — launches on a stream are ordered, so it is reusable."""
    d = torch.device(device)
    key = (d, torch.cuda.current_stream(d).cuda_stream)
    buf = _ATTN_WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=d)
        _ATTN_WS[key] = buf
    return buf


def detach_attn_workspace(stream: int, device=None) -> list:
    """Forget (and return) the split-KV scratch of the raw stream handle `stream` — see engine.Workspace.detach."""
    dev = None if device is None else torch.device(device)
    keys = [k for k in _ATTN_WS if k[1] == stream and (dev is None or k[0] == dev)]
    return [_ATTN_WS.pop(k) for k in keys]


def new_lse(num_heads: int, batch: int, q_len: int, device) -> torch.Tensor:
    """Row log-sum-exp buffer of attention(..., lse=) / attention_bwd: [num_heads, batch, q_len rounded up to 64], +inf."""
    return torch.full((num_heads, batch, (q_len + 63) // 64 * 64), float("inf"), dtype=torch.float32, device=device)


def attention_bwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out: torch.Tensor, dout: torch.Tensor, lse: torch.Tensor,
                  dq: torch.Tensor, dk: torch.Tensor, dv: torch.Tensor, num_heads: int, head_dim: int, batch: int, q_len: int,
                  kv_len: int, q_batch_stride: int = 0, kv_batch_stride: int = 0, prefix_k: Optional[torch.Tensor] = None,
                  prefix_v: Optional[torch.Tensor] = None, prefix_len: int = 0, dprefix_k: Optional[torch.Tensor] = None,
                  dprefix_v: Optional[torch.Tensor] = None, scale: Optional[float] = None) -> None:
    """Gradient of attention() (sr_attn_bwd in the header): all tensors bf16 2-D row views, lse from the forward."""
    bf = torch.bfloat16
    for name, t in (("q", q), ("k", k), ("v", v), ("out", out), ("dout", dout), ("dq", dq), ("dk", dk), ("dv", dv)):
        _rows2d(t, bf, name)
    if k.stride(0) != v.stride(0) or dk.stride(0) != dv.stride(0):
        raise ValueError("sailrecon_b200: k/v and dk/dv must share a row stride")
    a = _lib.SrAttnBwd()
    a.q, a.ldq = q.data_ptr(), q.stride(0)
    a.k, a.v, a.ldkv = k.data_ptr(), v.data_ptr(), k.stride(0)
    a.out, a.ldo = out.data_ptr(), out.stride(0)
    a.dout, a.lddo = dout.data_ptr(), dout.stride(0)
    a.dq, a.lddq = dq.data_ptr(), dq.stride(0)
    a.dk, a.dv, a.lddkv = dk.data_ptr(), dv.data_ptr(), dk.stride(0)
    if prefix_k is not None and prefix_len > 0:
        _rows2d(prefix_k, bf, "prefix_k")
        _rows2d(prefix_v, bf, "prefix_v")
        if prefix_k.stride(0) != prefix_v.stride(0) or prefix_k.shape[0] < prefix_len:
            raise ValueError("sailrecon_b200: bad prefix K/V")
        a.pk, a.pv, a.ldp, a.prefix_len = prefix_k.data_ptr(), prefix_v.data_ptr(), prefix_k.stride(0), prefix_len
        if dprefix_k is not None:
            _rows2d(dprefix_k, bf, "dprefix_k")
            _rows2d(dprefix_v, bf, "dprefix_v")
            if dprefix_k.stride(0) != dprefix_v.stride(0) or dprefix_k.shape[0] < prefix_len:
                raise ValueError("sailrecon_b200: bad prefix gradient buffers")
            a.dpk, a.dpv, a.lddp = dprefix_k.data_ptr(), dprefix_v.data_ptr(), dprefix_k.stride(0)
    a.num_heads, a.head_dim, a.batch = num_heads, head_dim, batch
    a.q_len, a.q_batch_stride = q_len, q_batch_stride or q_len
    a.kv_len, a.kv_batch_stride = kv_len, kv_batch_stride or kv_len
    need_q = (batch - 1) * a.q_batch_stride + q_len
    need_kv = (batch - 1) * a.kv_batch_stride + kv_len
    if min(q.shape[0], out.shape[0], dout.shape[0], dq.shape[0]) < need_q or min(k.shape[0], v.shape[0], dk.shape[0], dv.shape[0]) < need_kv:
        raise ValueError("sailrecon_b200: attention row ranges exceed the tensors")
    if lse.dtype != torch.float32 or not lse.is_contiguous() or tuple(lse.shape[:2]) != (num_heads, batch) or lse.shape[2] % 64 \
            or lse.shape[2] < q_len:
        raise ValueError("sailrecon_b200: lse must be the contiguous fp32 [num_heads, batch, q_pad] buffer of the forward")
    delta = torch.empty_like(lse)
    a.lse, a.delta, a.lse_pad = lse.data_ptr(), delta.data_ptr(), lse.shape[2]
    a.scale = float(scale if scale is not None else head_dim ** -0.5)
    hd = num_heads * head_dim
    t = _Timed("attn_bwd", 10.0 * batch * q_len * (prefix_len + kv_len) * hd,
               (4 * batch * q_len + 4 * batch * kv_len + 4 * prefix_len) * hd * 2)
    _lib.check(_lib.load().sr_attn_bwd_bf16(C.byref(a), _stream()), "attention_bwd")
    t.done()


# ------------------------------------------------------------------------------------------------ training-step row kernels
def transpose_bf16(src: torch.Tensor, r_pad: Optional[int] = None, colsum: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[R, C] bf16 -> [C, R_pad] bf16 (R_pad = R rounded up to 64, padding zero): transposed operand of dgrad / wgrad.
    colsum (fp32 [C]) += column sums of src in the same pass (the bias gradient when src is a Linear's dY)."""
    _rows2d(src, torch.bfloat16, "src")
    R, Cc = src.shape
    rp = r_pad or (R + 63) // 64 * 64
    _vec(colsum, Cc, "colsum")
    dst = torch.empty(Cc, rp, dtype=torch.bfloat16, device=src.device)
    tm = _Timed("transpose", 0.0, 4.0 * R * Cc)
    _lib.check(_lib.load().sr_transpose_bf16(src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0), R, Cc, rp, _ptr(colsum),
                                             _stream()), "transpose_bf16")
    tm.done()
    return dst


def colsum(a: torch.Tensor, b: Optional[torch.Tensor], out: torch.Tensor, accumulate: bool = False) -> torch.Tensor:
    """out[c] (+)= sum_r a[r, c] * b[r, c] (b None = 1)."""
    if a.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("sailrecon_b200: colsum input must be fp32 or bf16")
    _rows2d(a, a.dtype, "a")
    if b is not None:
        _rows2d(b, torch.bfloat16, "b")
        if b.shape != a.shape:
            raise ValueError("sailrecon_b200: colsum shape mismatch")
    _vec(out, a.shape[1], "out")
    tm = _Timed("colsum", 0.0, a.numel() * (a.element_size() + (0 if b is None else 2)))
    _lib.check(_lib.load().sr_colsum(a.data_ptr(), int(a.dtype == torch.bfloat16), a.stride(0), _ptr(b), 0 if b is None else b.stride(0),
                                     out.data_ptr(), a.shape[0], a.shape[1], int(accumulate), _stream()), "colsum")
    tm.done()
    return out


def scale_cast_bf16(x: torch.Tensor, gamma: Optional[torch.Tensor], out: torch.Tensor) -> torch.Tensor:
    _rows2d(x, torch.float32, "x")
    _rows2d(out, torch.bfloat16, "out")
    _vec(gamma, x.shape[1], "gamma")
    if out.shape != x.shape:
        raise ValueError("sailrecon_b200: scale_cast shape mismatch")
    tm = _Timed("scale_cast", 0.0, 6.0 * x.numel())
    _lib.check(_lib.load().sr_scale_cast_bf16(x.data_ptr(), x.stride(0), _ptr(gamma), out.data_ptr(), out.stride(0), x.shape[0],
                                              x.shape[1], _stream()), "scale_cast_bf16")
    tm.done()
    return out


def ls_bwd(dy: torch.Tensor, z: torch.Tensor, gamma: torch.Tensor, dz: torch.Tensor, dgamma: torch.Tensor,
           dbias: Optional[torch.Tensor]) -> torch.Tensor:
    """LayerScale backward in one pass: dz = bf16(dy * gamma); dgamma += colsum(dy * z); dbias += colsum(dz) (None = skip)."""
    _rows2d(dy, torch.float32, "dy")
    _rows2d(z, torch.bfloat16, "z")
    _rows2d(dz, torch.bfloat16, "dz")
    if not (dy.shape == z.shape == dz.shape):
        raise ValueError("sailrecon_b200: ls_bwd shape mismatch")
    for t_, n_ in ((gamma, "gamma"), (dgamma, "dgamma"), (dbias, "dbias")):
        _vec(t_, dy.shape[1], n_)
    tm = _Timed("ls_bwd", 0.0, 8.0 * dy.numel())
    _lib.check(_lib.load().sr_ls_bwd_bf16(dy.data_ptr(), dy.stride(0), z.data_ptr(), z.stride(0), gamma.data_ptr(), dz.data_ptr(),
                                          dz.stride(0), dgamma.data_ptr(), _ptr(dbias), dy.shape[0], dy.shape[1], _stream()), "ls_bwd")
    tm.done()
    return dz


def gelu_bwd_colsum(u: torch.Tensor, dg: torch.Tensor, du: torch.Tensor, dbias: torch.Tensor) -> torch.Tensor:
    """du = bf16(dg * gelu'(u)) (du may be dg) and dbias += colsum(du)."""
    for t_, n_ in ((u, "u"), (dg, "dg"), (du, "du")):
        _rows2d(t_, torch.bfloat16, n_)
    if not (u.shape == dg.shape == du.shape):
        raise ValueError("sailrecon_b200: gelu_bwd_colsum shape mismatch")
    _vec(dbias, u.shape[1], "dbias")
    tm = _Timed("gelu_bwd", 0.0, 6.0 * u.numel())
    _lib.check(_lib.load().sr_gelu_bwd_colsum_bf16(u.data_ptr(), u.stride(0), dg.data_ptr(), dg.stride(0), du.data_ptr(), du.stride(0),
                                                   dbias.data_ptr(), u.shape[0], u.shape[1], _stream()), "gelu_bwd_colsum")
    tm.done()
    return du


def ls_residual(resid: torch.Tensor, z: torch.Tensor, gamma: Optional[torch.Tensor], out: torch.Tensor) -> torch.Tensor:
    _rows2d(resid, torch.float32, "resid")
    _rows2d(z, torch.bfloat16, "z")
    _rows2d(out, torch.float32, "out")
    if not (resid.is_contiguous() and z.is_contiguous() and out.is_contiguous() and resid.shape == z.shape == out.shape):
        raise ValueError("sailrecon_b200: ls_residual needs contiguous tensors of one shape")
    _vec(gamma, resid.shape[1], "gamma")
    tm = _Timed("ls_residual", 0.0, 10.0 * resid.numel())
    _lib.check(_lib.load().sr_ls_residual_f32(resid.data_ptr(), z.data_ptr(), _ptr(gamma), out.data_ptr(), resid.shape[0],
                                              resid.shape[1], _stream()), "ls_residual")
    tm.done()
    return out


def gelu_fwd(u: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    if not (u.is_cuda and u.dtype == out.dtype == torch.bfloat16 and u.is_contiguous() and out.is_contiguous() and u.shape == out.shape):
        raise ValueError("sailrecon_b200: gelu_fwd needs contiguous bf16 CUDA tensors of one shape")
    tm = _Timed("gelu_fwd", 0.0, 4.0 * u.numel())
    _lib.check(_lib.load().sr_gelu_fwd_bf16(u.data_ptr(), out.data_ptr(), u.numel(), _stream()), "gelu_fwd")
    tm.done()
    return out


def gelu_bwd(u: torch.Tensor, dg: torch.Tensor, du: torch.Tensor) -> torch.Tensor:
    for t in (u, dg, du):
        if not (t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous() and t.shape == u.shape):
            raise ValueError("sailrecon_b200: gelu_bwd needs contiguous bf16 CUDA tensors of one shape")
    tm = _Timed("gelu_bwd", 0.0, 6.0 * u.numel())
    _lib.check(_lib.load().sr_gelu_bwd_bf16(u.data_ptr(), dg.data_ptr(), du.data_ptr(), u.numel(), _stream()), "gelu_bwd")
    tm.done()
    return du


def layernorm_bwd(x: torch.Tensor, weight: Optional[torch.Tensor], dy: torch.Tensor, dx: torch.Tensor,
                  dweight: Optional[torch.Tensor], dbias: Optional[torch.Tensor], eps: float,
                  dres: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = dres + dLN(dy) (dres None: dx += dLN(dy)); dweight / dbias accumulate (fp32 [C])."""
    _rows2d(x, torch.float32, "x")
    _rows2d(dy, torch.bfloat16, "dy")
    _rows2d(dx, torch.float32, "dx")
    if dres is not None:
        _rows2d(dres, torch.float32, "dres")
    if not (x.shape == dy.shape == dx.shape) or (dres is not None and dres.shape != dx.shape):
        raise ValueError("sailrecon_b200: layernorm_bwd shape mismatch")
    for n, t in (("weight", weight), ("dweight", dweight), ("dbias", dbias)):
        _vec(t, x.shape[1], n)
    tm = _Timed("layernorm_bwd", 0.0, 14.0 * x.numel())
    _lib.check(_lib.load().sr_layernorm_bwd(x.data_ptr(), x.stride(0), _ptr(weight), dy.data_ptr(), dy.stride(0), _ptr(dres),
                                            0 if dres is None else dres.stride(0), dx.data_ptr(), dx.stride(0), _ptr(dweight),
                                            _ptr(dbias), x.shape[0], x.shape[1], float(eps), _stream()), "layernorm_bwd")
    tm.done()
    return dx


def scatter_add_rows(dst: torch.Tensor, idx: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """dst[idx[r]] += src[r] (idx unique int32)."""
    _rows2d(dst, torch.float32, "dst")
    _rows2d(src, torch.float32, "src")
    if not (idx.is_cuda and idx.dtype == torch.int32 and idx.is_contiguous() and idx.numel() == src.shape[0]) or dst.shape[1] != src.shape[1]:
        raise ValueError("sailrecon_b200: scatter_add_rows needs one contiguous int32 index per source row")
    _lib.check(_lib.load().sr_scatter_add_rows_f32(dst.data_ptr(), dst.stride(0), idx.data_ptr(), src.data_ptr(), src.stride(0),
                                                   src.shape[0], src.shape[1], _stream()), "scatter_add_rows")
    return dst


def adam_step(param: torch.Tensor, grad: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, lr: float, step: int,
              betas=(0.9, 0.999), eps: float = 1e-8, grad_scale: float = 1.0) -> None:
    """torch.optim.Adam step (train/train_imc.py:475) in place on flat contiguous fp32 buffers of equal length."""
    n = param.numel()
    for t, name in ((param, "param"), (grad, "grad"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n):
            raise ValueError(f"sailrecon_b200: adam_step {name} must be a contiguous fp32 CUDA buffer of {n} elements")
    t = _Timed("adam", 0.0, 28.0 * n)
    _lib.check(_lib.load().sr_adam_step_f32(param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), n, lr,
                                            betas[0], betas[1], eps, step, grad_scale, _stream()), "adam_step")
    t.done()


def qknorm_rope_fwd(raw: torch.Tensor, out: torch.Tensor, qkv: QkvEpilogue) -> torch.Tensor:
    _rows2d(raw, torch.bfloat16, "raw")
    _rows2d(out, torch.bfloat16, "out")
    if raw.shape != out.shape or raw.shape[1] != 3 * qkv.qkv_dim:
        raise ValueError("sailrecon_b200: qknorm_rope needs [M, 3C] buffers")
    e = SrEpilogue()
    keep = _fill_qkv(e, qkv, raw.shape[0])
    tm = _Timed("qknorm_rope_fwd", 0.0, 4.0 * raw.numel())
    _lib.check(_lib.load().sr_qknorm_rope_fwd(raw.data_ptr(), raw.stride(0), out.data_ptr(), out.stride(0), C.byref(e), raw.shape[0],
                                              _stream()), "qknorm_rope_fwd")
    tm.done()
    del keep
    return out


def qknorm_rope_bwd(raw: torch.Tensor, dqkv: torch.Tensor, qkv: QkvEpilogue, dq_norm=None, dk_norm=None) -> torch.Tensor:
    """In place on dqkv (gradient w.r.t. normalised / rotated q, k -> gradient w.r.t. raw); dq_norm / dk_norm are
    (dweight, dbias) fp32 [64] accumulators or None."""
    _rows2d(raw, torch.bfloat16, "raw")
    _rows2d(dqkv, torch.bfloat16, "dqkv")
    if raw.shape != dqkv.shape or raw.shape[1] != 3 * qkv.qkv_dim:
        raise ValueError("sailrecon_b200: qknorm_rope needs [M, 3C] buffers")
    e = SrEpilogue()
    keep = _fill_qkv(e, qkv, raw.shape[0])
    g = [None, None, None, None]
    if dq_norm is not None:
        g = [dq_norm[0], dq_norm[1], dk_norm[0], dk_norm[1]]
        for t in g:
            _vec(t, qkv.head_dim, "q/k norm gradient")
    tm = _Timed("qknorm_rope_bwd", 0.0, 6.0 * raw.numel())
    _lib.check(_lib.load().sr_qknorm_rope_bwd(raw.data_ptr(), raw.stride(0), dqkv.data_ptr(), dqkv.stride(0), C.byref(e), _ptr(g[0]),
                                              _ptr(g[1]), _ptr(g[2]), _ptr(g[3]), raw.shape[0], _stream()), "qknorm_rope_bwd")
    tm.done()
    del keep
    return dqkv


# ------------------------------------------------------------------------------------------------ re-layout
def gather_rows(src: torch.Tensor, idx: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    _rows2d(src, torch.float32, "src")
    _rows2d(out, torch.float32, "out")
    if not (idx.is_cuda and idx.dtype == torch.int32 and idx.is_contiguous() and idx.numel() == out.shape[0]):
        raise ValueError("sailrecon_b200: idx must be contiguous int32 CUDA, one per output row")
    _lib.check(_lib.load().sr_gather_rows_f32(src.data_ptr(), src.stride(0), idx.data_ptr(), out.data_ptr(), out.stride(0),
                                              out.shape[0], src.shape[1], _stream()), "gather_rows")
    return out


def copy_cols(src: torch.Tensor, dst: torch.Tensor, col_off: int) -> torch.Tensor:
    """dst[:, col_off:col_off+C] = src (one half of the [.., 2C] concat outputs)."""
    _rows2d(src, torch.float32, "src")
    _rows2d(dst, torch.float32, "dst")
    if src.shape[0] != dst.shape[0] or col_off + src.shape[1] > dst.shape[1]:
        raise ValueError("sailrecon_b200: copy_cols shape mismatch")
    _lib.check(_lib.load().sr_copy_cols_f32(src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0), col_off,
                                            src.shape[0], src.shape[1], _stream()), "copy_cols")
    return dst


def cast_bf16(src: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _rows2d(src, torch.float32, "src")
    if out is None:
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    _rows2d(out, torch.bfloat16, "out")
    _lib.check(_lib.load().sr_cast_f32_to_bf16(src.data_ptr(), src.stride(0), out.data_ptr(), out.stride(0), src.shape[0],
                                               src.shape[1], _stream()), "cast_bf16")
    return out


# ------------------------------------------------------------------------------------------------ RoPE / front end / pose
def rope_tables(max_pos: int, quarter_dim: int, frequency: float, device) -> tuple:
    """cos / sin fp32 [max_pos, quarter_dim] (rope.py:107-123), built on the device by the library."""
    cos = torch.empty(max_pos, quarter_dim, dtype=torch.float32, device=device)
    sin = torch.empty_like(cos)
    _lib.check(_lib.load().sr_rope_table_build(cos.data_ptr(), sin.data_ptr(), max_pos, quarter_dim, float(frequency),
                                               _stream()), "rope_tables")
    return cos, sin


def rope2d(x: torch.Tensor, pos: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """RotaryPositionEmbedding2D.forward on fp32 [B, H, N, d] with int32 positions [B, N, 2]."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.is_contiguous()):
        raise ValueError("sailrecon_b200: rope2d needs a contiguous fp32 CUDA tensor [B, H, N, d]")
    B, H, N, d = x.shape
    if not (pos.is_cuda and pos.dtype == torch.int32 and pos.is_contiguous() and tuple(pos.shape) == (B, N, 2)):
        raise ValueError("sailrecon_b200: rope2d positions must be contiguous int32 [B, N, 2]")
    if cos.shape[1] != d // 4:
        raise ValueError("sailrecon_b200: rope tables do not match the feature dimension")
    out = torch.empty_like(x)
    _lib.check(_lib.load().sr_rope2d_f32(x.data_ptr(), out.data_ptr(), pos.data_ptr(), cos.data_ptr(), sin.data_ptr(),
                                         cos.shape[0], B, H, N, d, _stream()), "rope2d")
    return out


def patchify(images: torch.Tensor, out: torch.Tensor, patch: int, mean, std) -> torch.Tensor:
    """images fp32 [S, 3, H, W] -> out [S*h*w, ldo] (bf16 or fp32): normalised im2col rows, zero padded to ldo columns."""
    if not (images.is_cuda and images.dtype == torch.float32 and images.dim() == 4 and images.is_contiguous()
            and images.shape[1] == 3):
        raise ValueError("sailrecon_b200: patchify needs contiguous fp32 CUDA images [S, 3, H, W]")
    S, _, H, W = images.shape
    _rows2d(out, out.dtype, "out")
    if out.dtype not in (torch.bfloat16, torch.float32) or not out.is_contiguous() or out.shape[0] != S * (H // patch) * (W // patch):
        raise ValueError("sailrecon_b200: patchify out must be contiguous bf16/fp32 [S*h*w, Kpad]")
    m = (C.c_float * 3)(*[float(v) for v in mean])
    sd = (C.c_float * 3)(*[float(v) for v in std])
    _lib.check(_lib.load().sr_patchify(images.data_ptr(), out.data_ptr(), int(out.dtype == torch.bfloat16), out.shape[1], S, H, W,
                                       patch, m, sd, _stream()), "patchify")
    return out


def pose_update(pred: torch.Tensor, delta: torch.Tensor, activated: torch.Tensor, first: bool) -> None:
    for t in (pred, delta, activated):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.shape[-1] == 9 and t.shape == pred.shape):
            raise ValueError("sailrecon_b200: pose_update needs contiguous fp32 CUDA tensors [rows, 9]")
    _lib.check(_lib.load().sr_pose_update(pred.data_ptr(), delta.data_ptr(), activated.data_ptr(), pred.numel() // 9,
                                          int(first), _stream()), "pose_update")


def scale_cols(x: torch.Tensor, gamma: torch.Tensor) -> torch.Tensor:
    """LayerScale on its own: x [..., C] fp32 contiguous times gamma [C]."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
        raise ValueError("sailrecon_b200: scale_cols needs a contiguous fp32 CUDA tensor")
    _vec(gamma, x.shape[-1], "gamma")
    out = torch.empty_like(x)
    _lib.check(_lib.load().sr_scale_cols_f32(x.data_ptr(), gamma.data_ptr(), out.data_ptr(), x.numel() // x.shape[-1],
                                             x.shape[-1], _stream()), "scale_cols")
    return out


# ------------------------------------------------------------------------------------------------ camera geometry (N3)
def pose_to_extri_intri(pose: torch.Tensor, image_hw=None, build_intrinsics: bool = True):
    """pose fp32 [..., 9] -> (extrinsics [..., 3, 4], intrinsics [..., 3, 3] or None); utils/pose_enc.py:68-135."""
    if not (pose.is_cuda and pose.dtype == torch.float32 and pose.shape[-1] == 9):
        raise ValueError("sailrecon_b200: pose_to_extri_intri needs a fp32 CUDA tensor [..., 9]")
    p = pose.contiguous()
    n = p.numel() // 9
    ext = torch.empty(*p.shape[:-1], 3, 4, dtype=torch.float32, device=p.device)
    intr = None
    H = W = 0
    if build_intrinsics:
        if image_hw is None:
            raise ValueError("sailrecon_b200: image_size_hw is required to build the intrinsics")
        H, W = int(image_hw[0]), int(image_hw[1])
        intr = torch.empty(*p.shape[:-1], 3, 3, dtype=torch.float32, device=p.device)
    if n == 0:
        return ext, intr
    _lib.check(_lib.load().sr_pose_to_extri_intri(p.data_ptr(), ext.data_ptr(), _ptr(intr), n, H, W, _stream()),
               "pose_to_extri_intri")
    return ext, intr


def pose_to_extri_intri_bwd(pose: torch.Tensor, d_ext: torch.Tensor, d_intr: Optional[torch.Tensor], image_hw=None) -> torch.Tensor:
    """-> d pose [..., 9] (the gradient of pose_to_extri_intri)."""
    p = pose.contiguous()
    n = p.numel() // 9
    ge = d_ext.contiguous().float()
    gi = None if d_intr is None else d_intr.contiguous().float()
    if ge.numel() != n * 12 or (gi is not None and gi.numel() != n * 9):
        raise ValueError("sailrecon_b200: pose_to_extri_intri_bwd shape mismatch")
    H, W = (0, 0) if gi is None else (int(image_hw[0]), int(image_hw[1]))
    out = torch.empty_like(p)
    _lib.check(_lib.load().sr_pose_to_extri_intri_bwd(p.data_ptr(), ge.data_ptr(), _ptr(gi), out.data_ptr(), n, H, W, _stream()),
               "pose_to_extri_intri_bwd")
    return out


def unproject_depth(depth: torch.Tensor, extrinsics: torch.Tensor, intrinsics: torch.Tensor) -> torch.Tensor:
    """depth fp32 [S, H, W], extrinsics [S, 3, 4], intrinsics [S, 3, 3] -> world points [S, H, W, 3]; utils/geometry.py:19-128."""
    for t, name in ((depth, "depth"), (extrinsics, "extrinsics"), (intrinsics, "intrinsics")):
        if not (t.is_cuda and t.dtype == torch.float32):
            raise ValueError(f"sailrecon_b200: unproject_depth {name} must be a fp32 CUDA tensor")
    S, H, W = depth.shape
    if tuple(extrinsics.shape) != (S, 3, 4) or tuple(intrinsics.shape) != (S, 3, 3):
        raise ValueError("sailrecon_b200: unproject_depth needs extrinsics [S, 3, 4] and intrinsics [S, 3, 3]")
    d, e, m = depth.contiguous(), extrinsics.contiguous(), intrinsics.contiguous()
    out = torch.empty(S, H, W, 3, dtype=torch.float32, device=depth.device)
    _lib.check(_lib.load().sr_unproject_depth(d.data_ptr(), e.data_ptr(), m.data_ptr(), out.data_ptr(), S, H, W, _stream()),
               "unproject_depth")
    return out


def modulate(x: torch.Tensor, shift: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """x * (1 + scale) + shift on fp32 rows (camera_head.py:189-194); shift / scale may be column slices."""
    C_ = x.shape[-1]
    x2, s2, c2 = x.reshape(-1, C_), shift.reshape(-1, C_), scale.reshape(-1, C_)
    for t, name in ((x2, "x"), (s2, "shift"), (c2, "scale")):
        _rows2d(t, torch.float32, name)
    if not (x2.shape == s2.shape == c2.shape):
        raise ValueError("sailrecon_b200: modulate shape mismatch")
    out = torch.empty(x2.shape, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().sr_modulate_f32(x2.data_ptr(), x2.stride(0), s2.data_ptr(), s2.stride(0), c2.data_ptr(), c2.stride(0),
                                           out.data_ptr(), out.stride(0), x2.shape[0], C_, _stream()), "modulate")
    return out.view(x.shape)


# ------------------------------------------------------------------------------------------------ DPT dense heads (N1)
def im2col_split(x: torch.Tensor, k: int, stride: int = 1, pad: int = 0, relu: bool = False,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x fp32 channels-last [N, H, W, C] -> bf16 rows [N*Ho*Wo, 2*k*k*C] = [hi | lo] (csrc/dpt.cu)."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.stride(3) == 1 and x.stride(2) >= x.shape[3]
            and x.stride(1) == x.shape[2] * x.stride(2) and x.stride(0) == x.shape[1] * x.stride(1)):
        raise ValueError("sailrecon_b200: im2col_split needs a dense channels-last fp32 CUDA tensor [N, H, W, C]")
    N, H, W, Cc = x.shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    rows, cols = N * Ho * Wo, 2 * k * k * Cc
    if out is None:
        out = torch.empty(rows, cols, dtype=torch.bfloat16, device=x.device)
    elif out.shape != (rows, cols) or out.dtype != torch.bfloat16 or not out.is_contiguous():
        raise ValueError("sailrecon_b200: im2col_split out must be contiguous bf16 [N*Ho*Wo, 2*k*k*C]")
    t = _Timed("im2col", 0.0, rows * cols * 2 + x.numel() * 4)
    _lib.check(_lib.load().sr_im2col_split_bf16(x.data_ptr(), x.stride(2), out.data_ptr(), N, H, W, Cc, k, stride, pad, int(relu),
                                                _stream()), "im2col_split")
    t.done()
    return out


def gemm_split(a2: torch.Tensor, w_hh: torch.Tensor, w_lo: torch.Tensor, bias: Optional[torch.Tensor], out: torch.Tensor,
               resid: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32-accurate out = A W^T (+ bias) (+ resid) on the bf16 tensor-core GEMM: a2 = [A_hi | A_lo] (im2col_split),
    w_hh = [W_hi | W_hi], w_lo = W_lo (engine.weight_split).  Two launches: K' = 2K, then the exact accumulate of A_hi W_lo^T."""
    K = w_lo.shape[1]
    if resid is None:
        gemm(a2, w_hh, bias, out, kind=SR_EPI_BIAS)
    else:
        gemm(a2, w_hh, bias, out, kind=SR_EPI_LS_RESIDUAL, resid=resid, exact=True)
    gemm(a2[:, :K], w_lo, None, out, kind=SR_EPI_LS_RESIDUAL, resid=out, exact=True)
    return out


def bilinear_nhwc(x: torch.Tensor, Ho: int, Wo: int, in_pad: bool = False) -> torch.Tensor:
    """F.interpolate(bilinear, align_corners=True) on channels-last fp32 [N, Hi (+2), Wi (+2), C] -> dense [N, Ho, Wo, C];
    in_pad: the source carries the one-pixel border of the shifted-GEMM convolutions."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.is_contiguous()):
        raise ValueError("sailrecon_b200: bilinear_nhwc needs a contiguous fp32 CUDA tensor [N, H, W, C]")
    N, Hi, Wi, Cc = x.shape
    if in_pad:
        Hi, Wi = Hi - 2, Wi - 2
    out = torch.empty(N, Ho, Wo, Cc, dtype=torch.float32, device=x.device)
    t = _Timed("bilinear", 0.0, (x.numel() + out.numel()) * 4)
    _lib.check(_lib.load().sr_bilinear_nhwc_f32(x.data_ptr(), out.data_ptr(), N, Hi, Wi, Ho, Wo, Cc, int(in_pad), _stream()),
               "bilinear_nhwc")
    t.done()
    return out


def split_pad(x: torch.Tensor, in_pad: bool, relu: bool = False) -> torch.Tensor:
    """x fp32 channels-last [S, H (+2), W (+2), >= C] (pixel stride x.stride(2)) -> bf16 [S, H+2, W+2, 2C] = [hi | lo] with a zero
    border: the operand of conv3x3 (csrc/dpt.cu split_pad_kernel)."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.stride(3) == 1
            and x.stride(1) == x.shape[2] * x.stride(2) and x.stride(0) == x.shape[1] * x.stride(1)):
        raise ValueError("sailrecon_b200: split_pad needs a dense channels-last fp32 CUDA tensor [S, H, W, C]")
    S, Hh, Ww, Cc = x.shape
    if in_pad:
        Hh, Ww = Hh - 2, Ww - 2
    out = torch.empty(S, Hh + 2, Ww + 2, 2 * Cc, dtype=torch.bfloat16, device=x.device)
    t = _Timed("split_pad", 0.0, x.shape[0] * x.shape[1] * x.shape[2] * Cc * 4 + out.numel() * 2)
    _lib.check(_lib.load().sr_split_pad_bf16(x.data_ptr(), x.stride(2), int(in_pad), out.data_ptr(), S, Hh, Ww, Cc, int(relu), _stream()),
               "split_pad")
    t.done()
    return out


def conv3x3(xs: torch.Tensor, w27: torch.Tensor, bias: Optional[torch.Tensor], resid: Optional[torch.Tensor] = None) -> torch.Tensor:
    """3x3 / stride 1 / pad 1 convolution as a shifted GEMM (sr_conv3x3_bf16): xs = split_pad output [S, H+2, W+2, 2C], w27 bf16
    [N, 27C] = (Whi | Whi | Wlo) x (tap, c); returns fp32 [S, H+2, W+2, N] in the SAME padded row numbering (border rows are
    garbage: consumers ignore or re-zero them); resid fp32 [S, H+2, W+2, N] is added in the epilogue (exact fp32)."""
    S, Hp, Wp, C2 = xs.shape
    Cc = C2 // 2
    N = w27.shape[0]
    if xs.dtype != torch.bfloat16 or not xs.is_contiguous() or w27.dtype != torch.bfloat16 or w27.shape[1] != 27 * Cc or not w27.is_contiguous():
        raise ValueError("sailrecon_b200: conv3x3 operands must be contiguous bf16 [S, H+2, W+2, 2C] and [N, 27C]")
    rows = S * Hp * Wp
    out = torch.empty(S, Hp, Wp, N, dtype=torch.float32, device=xs.device)
    _vec(bias, N, "bias")
    e = SrEpilogue()
    e.out, e.ldo, e.bias = out.data_ptr(), N, _ptr(bias)
    if resid is None:
        e.kind = SR_EPI_BIAS
    else:
        if resid.shape != out.shape or resid.dtype != torch.float32 or not resid.is_contiguous():
            raise ValueError("sailrecon_b200: conv3x3 residual must be a contiguous fp32 map in the padded layout of the output")
        e.kind, e.resid, e.ldr, e.exact_residual = SR_EPI_LS_RESIDUAL, resid.data_ptr(), N, 1
    t = _Timed("gemm_tc", 2.0 * rows * N * 27 * Cc, rows * 2 * Cc * 2 + N * 27 * Cc * 2 + rows * N * 4)
    _lib.check(_lib.load().sr_conv3x3_bf16(xs.data_ptr(), 2 * Cc, Cc, Wp, w27.data_ptr(), 27 * Cc, rows, N, C.byref(e), _stream()),
               "conv3x3")
    t.done()
    return out


def map_add(a: torch.Tensor, pad_a: bool, b: Optional[torch.Tensor], pad_b: bool, pad_out: bool, relu: bool = False) -> torch.Tensor:
    """act(a + b) between channels-last fp32 maps of one spatial size with / without the one-pixel border (b may be None)."""
    for t_ in (a, b):
        if t_ is not None and not (t_.is_cuda and t_.dtype == torch.float32 and t_.dim() == 4 and t_.is_contiguous()):
            raise ValueError("sailrecon_b200: map_add needs contiguous fp32 CUDA maps [S, H, W, C]")
    S, Ha, Wa, Cc = a.shape
    Hh, Ww = (Ha - 2, Wa - 2) if pad_a else (Ha, Wa)
    if b is not None and tuple(b.shape) != (S, Hh + 2 * int(pad_b), Ww + 2 * int(pad_b), Cc):
        raise ValueError("sailrecon_b200: map_add shape mismatch")
    po = int(pad_out)
    out = torch.empty(S, Hh + 2 * po, Ww + 2 * po, Cc, dtype=torch.float32, device=a.device)
    _lib.check(_lib.load().sr_map_add_f32(a.data_ptr(), int(pad_a), _ptr(b), int(pad_b), out.data_ptr(), po, S, Hh, Ww, Cc, int(relu),
                                          _stream()), "map_add")
    return out


def pixel_shuffle(g: torch.Tensor, bias: Optional[torch.Tensor], N: int, H: int, W: int, Cc: int, k: int) -> torch.Tensor:
    """ConvTranspose2d(kernel = stride = k) epilogue: g fp32 [N*H*W, k*k*C] -> [N, H*k, W*k, C] + bias."""
    _rows2d(g, torch.float32, "g")
    if not g.is_contiguous() or g.shape != (N * H * W, k * k * Cc):
        raise ValueError("sailrecon_b200: pixel_shuffle shape mismatch")
    _vec(bias, Cc, "bias")
    out = torch.empty(N, H * k, W * k, Cc, dtype=torch.float32, device=g.device)
    _lib.check(_lib.load().sr_pixel_shuffle_f32(g.data_ptr(), _ptr(bias), out.data_ptr(), N, H, W, Cc, k, _stream()), "pixel_shuffle")
    return out


def add_bcast(a: torch.Tensor, b: Optional[torch.Tensor], alpha: float = 1.0, out: Optional[torch.Tensor] = None,
              relu: bool = False) -> torch.Tensor:
    """out = act(a + alpha * b), b broadcast over the leading elements of a (b.numel() divides a.numel()) or None;
    act = ReLU when relu; fp32 contiguous."""
    for t_, n_ in ((a, "a"), (b, "b")):
        if t_ is not None and not (t_.is_cuda and t_.dtype == torch.float32 and t_.is_contiguous()):
            raise ValueError(f"sailrecon_b200: add_bcast {n_} must be a contiguous fp32 CUDA tensor")
    period = a.numel() if b is None else b.numel()
    if period == 0 or a.numel() % period:
        raise ValueError("sailrecon_b200: add_bcast period must divide the tensor")
    if out is None:
        out = torch.empty_like(a)
    _lib.check(_lib.load().sr_add_bcast_f32(a.data_ptr(), _ptr(b), out.data_ptr(), a.numel(), period, float(alpha), int(relu),
                                            _stream()), "add_bcast")
    return out


def uv_posembed_table(H: int, W: int, Cc: int, aspect_ratio: float, device, omega0: float = 100.0) -> torch.Tensor:
    """create_uv_grid + position_grid_to_embed (heads/utils.py) -> fp32 [H, W, C]."""
    out = torch.empty(H, W, Cc, dtype=torch.float32, device=device)
    _lib.check(_lib.load().sr_uv_posembed_table(out.data_ptr(), H, W, Cc, float(aspect_ratio), float(omega0), _stream()),
               "uv_posembed_table")
    return out


_DPT_ACT = {"inv_log": 0, "exp": 1, "linear": 2, "relu": 3, "sigmoid": 4}
_DPT_CONF = {"expp1": 0, "expp0": 1, "sigmoid": 2}


def dpt_out_head(h: torch.Tensor, hid: int, w: torch.Tensor, bias: torch.Tensor, activation: str, conf_activation: str,
                 padded_map=None):
    """ReLU -> 1x1 conv -> activate_head: h fp32 rows [px or padded px, >= hid] -> (pred [px, out_dim - 1], conf [px]).
    padded_map = (S, H, W): h is a zero-bordered map [S, H+2, W+2, ld] and px = S*H*W interior pixels are produced."""
    _rows2d(h, torch.float32, "h")
    if activation not in _DPT_ACT or conf_activation not in _DPT_CONF:
        raise ValueError(f"Unknown activation: {activation} / {conf_activation}")
    w2 = w.reshape(w.shape[0], -1).contiguous()
    out_dim = w2.shape[0]
    if padded_map is None:
        px, mh, mw, pad = h.shape[0], 0, 0, 0
    else:
        S_, mh, mw = padded_map
        px, pad = S_ * mh * mw, 1
        if h.shape[0] != S_ * (mh + 2) * (mw + 2):
            raise ValueError("sailrecon_b200: dpt_out_head padded map size mismatch")
    pred = torch.empty(px, out_dim - 1, dtype=torch.float32, device=h.device)
    conf = torch.empty(px, dtype=torch.float32, device=h.device)
    _lib.check(_lib.load().sr_dpt_out_head_f32(h.data_ptr(), h.stride(0), w2.data_ptr(), bias.data_ptr(), pred.data_ptr(),
                                               conf.data_ptr(), px, hid, out_dim, _DPT_ACT[activation], _DPT_CONF[conf_activation],
                                               mh, mw, pad, _stream()), "dpt_out_head")
    return pred, conf


# ------------------------------------------------------------------------------------------------ training loss (N4)
def _f32c(t: torch.Tensor, name: str, shape=None) -> torch.Tensor:
    if not (t.is_cuda and t.dtype == torch.float32):
        raise ValueError(f"sailrecon_b200: {name} must be a fp32 CUDA tensor (no CPU fallback)")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"sailrecon_b200: {name} must have shape {tuple(shape)}, got {tuple(t.shape)}")
    return t.contiguous()


def _idx64(t: torch.Tensor, name: str, n: int) -> torch.Tensor:
    if not (t.is_cuda and t.dtype == torch.int64 and tuple(t.shape) == (n,)):
        raise ValueError(f"sailrecon_b200: {name} must be an int64 CUDA tensor [{n}]")
    return t.contiguous()


def _check_pair_indices(src_idx: torch.Tensor, dst_idx: torch.Tensor, n: int, what: str) -> None:
    """Every pair index must name one of the n frames: the loss kernels index histograms, totals, cameras and CDF tables with
    them unchecked (csrc/loss.cu), where the reference's scatter_add_ / gather raise a device-side assert (cdf_loss.py:150-160,
    :219-226).  One reduction + one host read per index tensor; the verdict is remembered ON the tensor object (with its
    in-place version counter and n), so a pair list that is reused step after step costs nothing after its first use."""
    for t in (src_idx, dst_idx):
        if t.numel() == 0 or getattr(t, "_sr_range_ok", None) == (t._version, n):
            continue
        lo, hi = (int(v) for v in torch.stack(torch.aminmax(t)).tolist())
        if lo < 0 or hi >= n:
            raise IndexError(f"sailrecon_b200 {what}: pair indices span [{lo}, {hi}] but there are only {n} frames")
        t._sr_range_ok = (t._version, n)


def _pose_rows(t: torch.Tensor, name: str) -> int:
    if t.dim() != 3 or tuple(t.shape[-2:]) not in ((3, 4), (4, 4)):
        raise ValueError(f"sailrecon_b200: {name} must be [P, 3, 4] or [P, 4, 4], got {tuple(t.shape)}")
    return t.shape[-2]


def relative_pose(src_extrinsic: torch.Tensor, dst_extrinsic: torch.Tensor) -> torch.Tensor:
    """dst @ inverse(src) -> [P, 4, 4] (train/utils/geometry.py:252-287)."""
    rows = _pose_rows(src_extrinsic, "src_extrinsic")
    s = _f32c(src_extrinsic, "src_extrinsic")
    d = _f32c(dst_extrinsic, "dst_extrinsic", s.shape)
    P = s.shape[0]
    rel = torch.empty(P, 4, 4, dtype=torch.float32, device=s.device)
    _lib.check(_lib.load().sr_relative_pose_f32(s.data_ptr(), d.data_ptr(), rows, rel.data_ptr(), P, _stream()), "relative_pose")
    return rel


def relative_pose_bwd(src_extrinsic: torch.Tensor, dst_extrinsic: torch.Tensor, d_rel: torch.Tensor):
    rows = _pose_rows(src_extrinsic, "src_extrinsic")
    s = _f32c(src_extrinsic, "src_extrinsic")
    d = _f32c(dst_extrinsic, "dst_extrinsic", s.shape)
    P = s.shape[0]
    g = _f32c(d_rel, "d_rel", (P, 4, 4))
    ds, dd = torch.empty_like(s), torch.empty_like(d)
    _lib.check(_lib.load().sr_relative_pose_bwd_f32(s.data_ptr(), d.data_ptr(), rows, g.data_ptr(), ds.data_ptr(), dd.data_ptr(), P,
                                                    _stream()), "relative_pose_bwd")
    return ds, dd


def _reproject_args(src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale):
    if src_coords.dim() != 3 or src_coords.shape[-1] != 2:
        raise ValueError("sailrecon_b200: src_coords must be [P, K, 2]")
    P, K = src_coords.shape[:2]
    rows = _pose_rows(rel, "relative_pose")
    c = _f32c(src_coords, "src_coords")
    z = _f32c(src_depth, "src_depth", (P, K))
    zd = None if dst_depth is None else _f32c(dst_depth, "dst_depth", (P, K))
    ks, kd = _f32c(src_k, "src_intrinsic", (P, 3, 3)), _f32c(dst_k, "dst_intrinsic", (P, 3, 3))
    r = _f32c(rel, "relative_pose", (P, rows, 4))
    ss = None if src_scale is None else _f32c(src_scale.reshape(-1), "src_depth_scale", (P,))
    sd = None if dst_scale is None else _f32c(dst_scale.reshape(-1), "dst_depth_scale", (P,))
    ws = torch.empty(max(1, _lib.load().sr_reproject_workspace_floats(P, K)), dtype=torch.float32, device=c.device)
    return P, K, rows, c, z, zd, ks, kd, r, ss, sd, ws


def reproject(src_coords, src_depth, src_k, dst_k, rel, src_scale=None, dst_depth=None, dst_scale=None) -> torch.Tensor:
    """backproject_and_reproject (dst_depth None) / ..._with_approximation: predicted destination pixels [P, K, 2]
    (train/utils/geometry.py:89-250)."""
    P, K, rows, c, z, zd, ks, kd, r, ss, sd, ws = _reproject_args(src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale)
    out = torch.empty(P, K, 2, dtype=torch.float32, device=c.device)
    _lib.check(_lib.load().sr_reproject_f32(c.data_ptr(), z.data_ptr(), _ptr(zd), ks.data_ptr(), kd.data_ptr(), r.data_ptr(), rows,
                                            _ptr(ss), _ptr(sd), out.data_ptr(), P, K, ws.data_ptr(), _stream()), "reproject")
    return out


def reproject_bwd(src_coords, src_depth, src_k, dst_k, rel, d_coords, src_scale=None, dst_depth=None, dst_scale=None):
    """-> (d src_k [P, 3, 3], d dst_k [P, 3, 3], d relative_pose [P, rows, 4])."""
    P, K, rows, c, z, zd, ks, kd, r, ss, sd, ws = _reproject_args(src_coords, src_depth, dst_depth, src_k, dst_k, rel, src_scale, dst_scale)
    g = _f32c(d_coords, "d_coords", (P, K, 2))
    dks, dkd, dr = torch.zeros_like(ks), torch.zeros_like(kd), torch.zeros_like(r)
    if P and K:
        _lib.check(_lib.load().sr_reproject_bwd_f32(c.data_ptr(), z.data_ptr(), _ptr(zd), ks.data_ptr(), kd.data_ptr(), r.data_ptr(),
                                                    rows, _ptr(ss), _ptr(sd), g.data_ptr(), dks.data_ptr(), dkd.data_ptr(),
                                                    dr.data_ptr(), P, K, ws.data_ptr(), _stream()), "reproject_bwd")
    return dks, dkd, dr


def projective_residual(predicted: torch.Tensor, actual: torch.Tensor) -> torch.Tensor:
    """|| predicted - actual ||_2 over the last (size 2) dimension (train/utils/geometry.py:289-303)."""
    if predicted.shape[-1] != 2 or predicted.shape != actual.shape:
        raise ValueError("sailrecon_b200: projective_residual needs two [..., 2] tensors of the same shape")
    a, b = _f32c(predicted, "predicted_dst_coords"), _f32c(actual, "actual_dst_coords")
    out = torch.empty(a.shape[:-1], dtype=torch.float32, device=a.device)
    _lib.check(_lib.load().sr_projective_residual_f32(a.data_ptr(), b.data_ptr(), out.data_ptr(), out.numel(), _stream()),
               "projective_residual")
    return out


def projective_residual_bwd(predicted: torch.Tensor, actual: torch.Tensor, d_res: torch.Tensor) -> torch.Tensor:
    a, b = _f32c(predicted, "predicted_dst_coords"), _f32c(actual, "actual_dst_coords")
    g = _f32c(d_res, "d_residual", a.shape[:-1])
    out = torch.empty_like(a)
    _lib.check(_lib.load().sr_projective_residual_bwd_f32(a.data_ptr(), b.data_ptr(), g.data_ptr(), out.data_ptr(), g.numel(),
                                                          _stream()), "projective_residual_bwd")
    return out


def cdf_tables(residuals, weights, src_idx, dst_idx, num_nodes: int, min_val: float, inv_bin_width: float, num_bins: int,
               sobel_weight: float, smooth_kernel: torch.Tensor):
    """Per-frame weighted PMF / CDF / PDF tables [num_nodes, num_bins] (train/losses/cdf_loss.py:122-186)."""
    if residuals.dim() != 2:
        raise ValueError("sailrecon_b200: residuals must be [N_pairs, K_points]")
    P, K = residuals.shape
    r = _f32c(residuals, "residuals")
    w = None if weights is None else _f32c(weights, "weights", (P, K))
    s, d = _idx64(src_idx, "src_indices", P), _idx64(dst_idx, "dst_indices", P)
    _check_pair_indices(s, d, num_nodes, "cdf_tables")
    sk = _f32c(smooth_kernel.reshape(-1), "smooth_kernel")
    if sk.numel() % 2 != 1:
        raise ValueError("sailrecon_b200: the smoothing kernel must have an odd number of taps")
    scratch = torch.empty(num_nodes * (num_bins + 1), dtype=torch.float32, device=r.device)
    pmf, cdf, pdf = (torch.empty(num_nodes, num_bins, dtype=torch.float32, device=r.device) for _ in range(3))
    _lib.check(_lib.load().sr_cdf_tables_f32(r.data_ptr(), _ptr(w), s.data_ptr(), d.data_ptr(), P, K, num_nodes, min_val,
                                             inv_bin_width, num_bins, sobel_weight, sk.data_ptr(), sk.numel() // 2,
                                             scratch.data_ptr(), pmf.data_ptr(), cdf.data_ptr(), pdf.data_ptr(), _stream()),
               "cdf_tables")
    return pmf, cdf, pdf


def cdf_lookup(residuals, weights, src_idx, dst_idx, cdf, pdf, min_val: float, inv_bin_width: float):
    """-> (cdf_src, grad_src, cdf_dst, grad_dst), each [N_pairs, K_points] (train/losses/cdf_loss.py:188-242)."""
    P, K = residuals.shape
    num_bins = cdf.shape[1]
    r = _f32c(residuals, "residuals")
    w = None if weights is None else _f32c(weights, "weights", (P, K))
    s, d = _idx64(src_idx, "src_indices", P), _idx64(dst_idx, "dst_indices", P)
    c, q = _f32c(cdf, "cdf"), _f32c(pdf, "pdf", cdf.shape)
    if c.dim() != 2:
        raise ValueError(f"sailrecon_b200: cdf / pdf must be [num_nodes, num_bins], got {tuple(c.shape)}")
    _check_pair_indices(s, d, c.shape[0], "cdf_lookup")
    outs = [torch.empty(P, K, dtype=torch.float32, device=r.device) for _ in range(4)]
    _lib.check(_lib.load().sr_cdf_lookup_f32(r.data_ptr(), _ptr(w), s.data_ptr(), d.data_ptr(), c.data_ptr(), q.data_ptr(), P, K,
                                             min_val, inv_bin_width, num_bins, *(o.data_ptr() for o in outs), _stream()),
               "cdf_lookup")
    return tuple(outs)


def reproj_cdf_loss(pred_intrinsics, pred_extrinsics, kprime_to_k, shared_focal: bool, src_idx, dst_idx, src_coords, dst_coords,
                    src_depth, dst_depth, min_val: float, inv_bin_width: float, num_bins: int, sobel_weight: float,
                    smooth_kernel: torch.Tensor, want_tables: bool = False):
    """compute_loss (train/train_imc.py:141-256) fused: -> (loss [1], d intrinsics [N, 3, 3], d extrinsics [N, rows, 4], tables)."""
    from ._lib import SrReprojLoss
    if pred_intrinsics.dim() != 3:
        raise ValueError("sailrecon_b200: predicted intrinsics must be [N, 3, 3]")
    N = pred_intrinsics.shape[0]
    rows = _pose_rows(pred_extrinsics, "predicted extrinsics")
    ki = _f32c(pred_intrinsics, "predicted intrinsics", (N, 3, 3))
    ex = _f32c(pred_extrinsics, "predicted extrinsics", (N, rows, 4))
    k2k = None if kprime_to_k is None else _f32c(kprime_to_k, "K_prime_to_K", (N, 3, 3))
    if src_coords.dim() != 3 or src_coords.shape[-1] != 2:
        raise ValueError("sailrecon_b200: src_coords must be [num_pairs, K, 2]")
    P, K = src_coords.shape[:2]
    sc, dc = _f32c(src_coords, "src_coords"), _f32c(dst_coords, "dst_coords", (P, K, 2))
    sz, dz = _f32c(src_depth, "src_depth", (P, K)), _f32c(dst_depth, "dst_depth", (P, K))
    s, d = _idx64(src_idx, "src_idx", P), _idx64(dst_idx, "dst_idx", P)
    _check_pair_indices(s, d, N, "reproj_cdf_loss")
    sk = _f32c(smooth_kernel.reshape(-1), "smooth_kernel")
    lib = _lib.load()
    n_ws = lib.sr_reproj_loss_workspace_floats(N, P, K, num_bins)
    ws = torch.empty(n_ws, dtype=torch.float32, device=ki.device)
    loss = torch.empty(1, dtype=torch.float32, device=ki.device)
    dki, dex = torch.empty_like(ki), torch.empty_like(ex)
    tables = torch.empty(3, 2, N, num_bins, dtype=torch.float32, device=ki.device) if want_tables else None
    a = SrReprojLoss(ki.data_ptr(), ex.data_ptr(), _ptr(k2k), s.data_ptr(), d.data_ptr(), sc.data_ptr(), dc.data_ptr(), sz.data_ptr(),
                     dz.data_ptr(), N, P, K, rows, int(bool(shared_focal)), min_val, inv_bin_width, sobel_weight, num_bins,
                     sk.numel() // 2, sk.data_ptr(), ws.data_ptr(), n_ws, loss.data_ptr(), dki.data_ptr(), dex.data_ptr(), _ptr(tables))
    t = _Timed("reproj_cdf_loss", 0.0, 64.0 * P * K)
    _lib.check(lib.sr_reproj_cdf_loss_f32(C.byref(a), _stream()), "reproj_cdf_loss")
    t.done()
    return loss, dki, dex, tables


# ------------------------------------------------------------------------------------------------ camera-head backward (fp32)
def _f32_rows(t: torch.Tensor, name: str) -> torch.Tensor:
    _rows2d(t, torch.float32, name)
    if not t.is_contiguous():
        raise ValueError(f"sailrecon_b200: {name} must be contiguous")
    return t


def transpose_f32(src: torch.Tensor) -> torch.Tensor:
    """[R, C] fp32 (row stride >= C) -> contiguous [C, R]."""
    _rows2d(src, torch.float32, "src")
    R, C_ = src.shape
    dst = torch.empty(C_, R, dtype=torch.float32, device=src.device)
    _lib.check(_lib.load().sr_transpose_f32(src.data_ptr(), src.stride(0), dst.data_ptr(), R, R, C_, _stream()), "transpose_f32")
    return dst


def colsum_f32(a: torch.Tensor, b: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
               accumulate: bool = False) -> torch.Tensor:
    """out[c] (+)= sum_r a[r, c] * b[r, c]."""
    _rows2d(a, torch.float32, "a")
    R, C_ = a.shape
    if b is not None:
        _rows2d(b, torch.float32, "b")
        if b.shape != a.shape:
            raise ValueError("sailrecon_b200: colsum_f32 shape mismatch")
    if out is None:
        out, accumulate = torch.empty(C_, dtype=torch.float32, device=a.device), False
    _vec(out, C_, "out")
    _lib.check(_lib.load().sr_colsum_f32(a.data_ptr(), a.stride(0), _ptr(b), b.stride(0) if b is not None else 0, out.data_ptr(), R, C_,
                                         int(accumulate), _stream()), "colsum_f32")
    return out


_ACT_KIND = {"gelu": 0, "silu": 1}


def act_f32(u: torch.Tensor, kind: str) -> torch.Tensor:
    u = _f32_rows(u, "u")
    out = torch.empty_like(u)
    _lib.check(_lib.load().sr_act_f32(u.data_ptr(), out.data_ptr(), u.numel(), _ACT_KIND[kind], _stream()), "act_f32")
    return out


def act_bwd_f32(u: torch.Tensor, dg: torch.Tensor, kind: str) -> torch.Tensor:
    u, dg = _f32_rows(u, "u"), _f32_rows(dg, "dg")
    if u.shape != dg.shape:
        raise ValueError("sailrecon_b200: act_bwd_f32 shape mismatch")
    du = torch.empty_like(u)
    _lib.check(_lib.load().sr_act_bwd_f32(u.data_ptr(), dg.data_ptr(), du.data_ptr(), u.numel(), _ACT_KIND[kind], _stream()), "act_bwd_f32")
    return du


def ls_residual_f32(resid: torch.Tensor, z: torch.Tensor, gamma: Optional[torch.Tensor]) -> torch.Tensor:
    resid, z = _f32_rows(resid, "resid"), _f32_rows(z, "z")
    if resid.shape != z.shape:
        raise ValueError("sailrecon_b200: ls_residual_f32 shape mismatch")
    _vec(gamma, z.shape[1], "gamma")
    out = torch.empty_like(resid)
    _lib.check(_lib.load().sr_ls_residual_f32f32(resid.data_ptr(), z.data_ptr(), _ptr(gamma), out.data_ptr(), z.shape[0], z.shape[1],
                                                 _stream()), "ls_residual_f32")
    return out


def layernorm_bwd_f32(x: torch.Tensor, weight: Optional[torch.Tensor], dy: torch.Tensor, eps: float,
                      dres: Optional[torch.Tensor] = None, dweight: Optional[torch.Tensor] = None,
                      dbias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = (dres or 0) + dLN(dy); dweight / dbias (fp32 [C]) are added to in place."""
    x, dy = _f32_rows(x, "x"), _f32_rows(dy, "dy")
    if x.shape != dy.shape or (dres is not None and _f32_rows(dres, "dres").shape != x.shape):
        raise ValueError("sailrecon_b200: layernorm_bwd_f32 shape mismatch")
    for t, n in ((weight, "weight"), (dweight, "dweight"), (dbias, "dbias")):
        _vec(t, x.shape[1], n)
    dx = torch.empty_like(x)
    _lib.check(_lib.load().sr_layernorm_bwd_f32(x.data_ptr(), _ptr(weight), dy.data_ptr(), _ptr(dres), dx.data_ptr(), _ptr(dweight),
                                                _ptr(dbias), x.shape[0], x.shape[1], eps, _stream()), "layernorm_bwd_f32")
    return dx


def adaln_modulate_bwd(x: torch.Tensor, mod: torch.Tensor, dy: torch.Tensor, eps: float, dx: torch.Tensor, accumulate_dx: bool):
    """-> d mod [T, 3C]; dx (=|+=) gradient w.r.t. x."""
    x, mod, dy, dx = _f32_rows(x, "x"), _f32_rows(mod, "mod"), _f32_rows(dy, "dy"), _f32_rows(dx, "dx")
    T, C_ = x.shape
    if mod.shape != (T, 3 * C_) or dy.shape != x.shape or dx.shape != x.shape:
        raise ValueError("sailrecon_b200: adaln_modulate_bwd shape mismatch")
    dmod = torch.empty_like(mod)
    _lib.check(_lib.load().sr_adaln_modulate_bwd_f32(x.data_ptr(), mod.data_ptr(), dy.data_ptr(), dx.data_ptr(), dmod.data_ptr(), T, C_,
                                                     eps, int(accumulate_dx), _stream()), "adaln_modulate_bwd")
    return dmod


def attn_small_bwd_f32(qkv: torch.Tensor, dout: torch.Tensor, num_heads: int, n_anchor: int, scale: float) -> torch.Tensor:
    """qkv fp32 [T, 3C] (q | k | v), dout [T, C] -> d qkv [T, 3C]; visibility: column < n_anchor or the row itself."""
    qkv, dout = _f32_rows(qkv, "qkv"), _f32_rows(dout, "dout")
    T, C3 = qkv.shape
    C_ = C3 // 3
    if dout.shape != (T, C_) or C_ % num_heads:
        raise ValueError("sailrecon_b200: attn_small_bwd_f32 shape mismatch")
    d = torch.empty_like(qkv)
    _lib.check(_lib.load().sr_attn_small_bwd_f32(qkv.data_ptr(), qkv[:, C_:].data_ptr(), qkv[:, 2 * C_:].data_ptr(), C3, dout.data_ptr(),
                                                 C_, d.data_ptr(), d[:, C_:].data_ptr(), d[:, 2 * C_:].data_ptr(), C3, T, num_heads,
                                                 C_ // num_heads, n_anchor, scale, _stream()), "attn_small_bwd_f32")
    return d


def pose_act_bwd(pred: torch.Tensor, d_act: torch.Tensor) -> torch.Tensor:
    pred, d_act = _f32_rows(pred, "pred"), _f32_rows(d_act, "d_activated")
    if pred.shape != d_act.shape or pred.shape[1] != 9:
        raise ValueError("sailrecon_b200: pose_act_bwd needs two [T, 9] tensors")
    out = torch.empty_like(pred)
    _lib.check(_lib.load().sr_pose_act_bwd_f32(pred.data_ptr(), d_act.data_ptr(), out.data_ptr(), pred.shape[0], _stream()), "pose_act_bwd")
    return out

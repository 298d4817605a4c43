"""CPU ORACLE for the SAIL-Recon hot path — TEST INFRASTRUCTURE ONLY.

A from-scratch restatement, in plain torch-CPU tensor arithmetic, of the reference algorithm on the path named by
BASELINE.json (aggregator frame / global / global-reloc blocks + camera head).  Each function cites the reference
file:line it follows (paths relative to /root/reference/sailrecon).  Nothing here is imported by the product
package `sailrecon_b200`; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
may import it, and there only as the checker / the timed CPU baseline.

Pinning: the reference ships no golden vectors or tests for this path, so this oracle is pinned
against outputs of the reference itself: oracle/make_golden.py imports the unmodified reference from
/root/reference in the build container, runs it on seeded inputs and writes tests/golden/*.npz;
tests/test_oracle_golden.py checks this file against those vectors (and against the live reference when
/root/reference is present).

Precision: `bf16=False` is exact fp32 math (what the reference does without autocast).  `bf16=True` emulates CUDA
autocast(bfloat16) rounding points: Linear / SDPA inputs and outputs rounded to bf16, LayerNorm and
RoPE in fp32, LayerScale and the residual stream in fp32.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
Params = Dict[str, Tensor]


def _r(x: Tensor, bf16: bool) -> Tensor:
    """Round to bf16 and back (an autocast rounding point) when emulating the bf16 path."""
    return x.to(torch.bfloat16).to(x.dtype) if bf16 else x


# --------------------------------------------------------------------------------------------------- layers
def linear(x: Tensor, w: Tensor, b: Optional[Tensor], bf16: bool) -> Tensor:
    """nn.Linear (layers/attention.py:48,52; layers/mlp.py:29,31).  Under autocast: bf16 operands, fp32
    accumulate, bf16 result."""
    y = _r(x, bf16) @ _r(w, bf16).t()
    if b is not None:
        y = y + _r(b, bf16)
    return _r(y, bf16)


def layer_norm(x: Tensor, w: Optional[Tensor], b: Optional[Tensor], eps: float) -> Tensor:
    """nn.LayerNorm (layers/block.py:50,70; attention.py:49-50; heads/camera_head.py:64,65,77): biased variance."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    y = (x - mu) / torch.sqrt(var + eps)
    if w is not None:
        y = y * w
    if b is not None:
        y = y + b
    return y


def gelu_erf(x: Tensor) -> Tensor:
    """nn.GELU() exact erf form (layers/mlp.py:22,36)."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def position_grid(h: int, w: int) -> Tensor:
    """PositionGetter (layers/rope.py:25-66): cartesian_prod(arange(h), arange(w)) -> int64 [h*w, 2] = (y, x)."""
    ys = torch.arange(h).repeat_interleave(w)
    xs = torch.arange(w).repeat(h)
    return torch.stack([ys, xs], dim=-1)


def aggregator_positions(h: int, w: int, num_special: int = 5) -> Tensor:
    """models/aggregator.py:313-328: grid positions +1, with `num_special` zero rows prepended -> int64 [P, 2]."""
    return torch.cat([torch.zeros(num_special, 2, dtype=torch.int64), position_grid(h, w) + 1], dim=0)


def rope_tables(max_pos: int, half_dim: int, frequency: float = 100.0, dtype=torch.float32) -> Tuple[Tensor, Tensor]:
    """layers/rope.py:107-123: inv_freq = f^-(arange(0,D,2)/D), angles = pos x inv_freq duplicated to D columns."""
    exponents = torch.arange(0, half_dim, 2).float() / half_dim
    inv_freq = 1.0 / (frequency ** exponents)
    angles = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    angles = torch.cat([angles, angles], dim=-1).to(dtype)
    return angles.cos(), angles.sin()


def rope_2d(x: Tensor, pos: Tensor, frequency: float = 100.0) -> Tensor:
    """RotaryPositionEmbedding2D.forward (layers/rope.py:165-207).  x [..., N, d]; pos int64 [N, 2] (y, x).
    First half of d rotates with y, second half with x; rotate-half pairs feature i with i + d/4 (rope.py:136-138)."""
    d = x.shape[-1]
    half = d // 2
    cos_t, sin_t = rope_tables(int(pos.max()) + 1, half, frequency, x.dtype)

    def one(t: Tensor, p: Tensor) -> Tensor:
        c, s = cos_t[p], sin_t[p]                                   # F.embedding (rope.py:159-160)
        t1, t2 = t[..., : half // 2], t[..., half // 2:]
        rot = torch.cat([-t2, t1], dim=-1)
        return t * c + rot * s                                      # rope.py:163

    return torch.cat([one(x[..., :half], pos[:, 0]), one(x[..., half:], pos[:, 1])], dim=-1)


def sdpa(q: Tensor, k: Tensor, v: Tensor, mask: Optional[Tensor], bf16: bool) -> Tensor:
    """F.scaled_dot_product_attention (layers/attention.py:103-109); mask True = attend.  q,k,v [H, N, d]."""
    q, k, v = _r(q, bf16), _r(k, bf16), _r(v, bf16)
    scale = q.shape[-1] ** -0.5
    n_q, n_k = q.shape[-2], k.shape[-2]
    chunk = n_q if n_q * n_k <= (1 << 24) else max(1, (1 << 24) // n_k)      # rows are independent: bound the score matrix
    outs = []
    for r0 in range(0, n_q, chunk):
        s = (q[..., r0:r0 + chunk, :] @ k.transpose(-1, -2)) * scale
        if mask is not None:
            s = s.masked_fill(~mask[..., r0:r0 + chunk, :], float("-inf"))
        outs.append(torch.softmax(s, dim=-1) @ v)
    return _r(outs[0] if len(outs) == 1 else torch.cat(outs, dim=-2), bf16)


def qkv_heads(x: Tensor, p: Params, pre: str, num_heads: int, pos: Optional[Tensor], qk_norm: bool, rope: bool,
              bf16: bool, qk_eps: float = 1e-5, frequency: float = 100.0) -> Tuple[Tensor, Tensor, Tensor]:
    """layers/attention.py:71-82 for one sequence x [N, C] -> q, k, v each [H, N, d] (post norm, post RoPE)."""
    N, C = x.shape
    d = C // num_heads
    qkv = linear(x, p[pre + "qkv.weight"], p.get(pre + "qkv.bias"), bf16).reshape(N, 3, num_heads, d).permute(1, 2, 0, 3)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if qk_norm:
        q = layer_norm(q, p[pre + "q_norm.weight"], p[pre + "q_norm.bias"], qk_eps)
        k = layer_norm(k, p[pre + "k_norm.weight"], p[pre + "k_norm.bias"], qk_eps)
    if rope:
        q, k = rope_2d(q, pos, frequency), rope_2d(k, pos, frequency)
    return q, k, v


def attention_forward(x: Tensor, p: Params, pre: str, num_heads: int, pos: Optional[Tensor] = None,
                      mask: Optional[Tensor] = None, qk_norm: bool = False, rope: bool = False, bf16: bool = False,
                      prefix_kv: Optional[Tuple[Tensor, Tensor]] = None) -> Tensor:
    """Attention.forward (layers/attention.py:70-122) for one sequence x [N, C]; `prefix_kv` = cached K/V that are
    prepended (attention.py:85-93)."""
    N, C = x.shape
    q, k, v = qkv_heads(x, p, pre, num_heads, pos, qk_norm, rope, bf16)
    if prefix_kv is not None:
        k = torch.cat([prefix_kv[0], k], dim=1)
        v = torch.cat([prefix_kv[1], v], dim=1)
    o = sdpa(q, k, v, mask, bf16).transpose(0, 1).reshape(N, C)
    return linear(o, p[pre + "proj.weight"], p.get(pre + "proj.bias"), bf16)


def mlp_forward(x: Tensor, p: Params, pre: str, bf16: bool) -> Tensor:
    """Mlp.forward (layers/mlp.py:34-40)."""
    h = linear(x, p[pre + "fc1.weight"], p.get(pre + "fc1.bias"), bf16)
    h = _r(gelu_erf(h), bf16)
    return linear(h, p[pre + "fc2.weight"], p.get(pre + "fc2.bias"), bf16)


def block_forward(x: Tensor, p: Params, pre: str, num_heads: int, pos: Optional[Tensor] = None,
                  mask: Optional[Tensor] = None, qk_norm: bool = False, rope: bool = False, bf16: bool = False,
                  eps: float = 1e-5, prefix_kv=None, attn_zero: bool = False) -> Tensor:
    """Block.forward (layers/block.py:86-112): x + ls1(attn(norm1 x)); x + ls2(mlp(norm2 x)).  `attn_zero`
    reproduces the kv-cache priming pass where Attention returns zeros (attention.py:94-100)."""
    g1 = p.get(pre + "ls1.gamma")
    g2 = p.get(pre + "ls2.gamma")
    if attn_zero:
        a = torch.zeros_like(x)
    else:
        a = attention_forward(layer_norm(x, p[pre + "norm1.weight"], p[pre + "norm1.bias"], eps), p, pre + "attn.",
                              num_heads, pos, mask, qk_norm, rope, bf16, prefix_kv)
    x = x + (a * g1 if g1 is not None else a)
    m = mlp_forward(layer_norm(x, p[pre + "norm2.weight"], p[pre + "norm2.bias"], eps), p, pre + "mlp.", bf16)
    return x + (m * g2 if g2 is not None else m)


# --------------------------------------------------------------------------------------------------- masks / indices
def build_allow_block(L: int, L_a: Sequence[int], L_b: Sequence[int]) -> Tensor:
    """models/aggregator.py:832-847: anchors see anchors; queries see anchors and themselves."""
    allow = torch.zeros(L, L, dtype=torch.bool)
    for i in L_a:
        for j in L_a:
            allow[i, j] = True
    for i in L_b:
        for j in L_a:
            allow[i, j] = True
        allow[i, i] = True
    return allow


def reloc_mask(num_recon: int, num_reloc: int, P: int, P_prime: int) -> Tensor:
    """models/aggregator.py:302-311: token-level expansion (expand_to_token :850-851) of the allow block over the
    concatenated [anchors ; queries] sequence, cropped to the last num_recon*P' + num_reloc*P rows/cols."""
    S = num_recon + num_reloc
    allow = build_allow_block(S, list(range(num_recon)), list(range(num_recon, S)))
    m = allow.repeat_interleave(P, 0).repeat_interleave(P, 1)
    crop = (P - P_prime) * num_recon
    return m[crop:, crop:]


def camera_mask(num_recon: int, num_reloc: int) -> Tensor:
    """heads/camera_head.py:197-228 build_lr_mask, returned as the ALLOW mask (~mask) that trunk_fn passes to the
    blocks (:163-166): anchors see anchors only; a query sees the anchors and itself."""
    S = num_recon + num_reloc
    masked = torch.zeros(S, S, dtype=torch.bool)
    masked[:num_recon, num_recon:] = True
    if num_reloc > 1:
        masked[num_recon:, num_recon:] = True
        idx = torch.arange(num_recon, S)
        masked[idx, idx] = False
    return ~masked


def draw_scene_indices(generator: torch.Generator, num_layers: int, batch: int, num_recon: int, num_patches: int,
                       rank: int) -> Tensor:
    """models/aggregator.py:351-357,593-626: per layer, per (batch, anchor) in that loop order, one
    `torch.randperm(l, generator)[:rank]` on the CPU generator.  Returns int64 [layers, batch, anchors, rank] patch
    indices (0-based inside the patch tokens, i.e. token row = index + patch_start_idx)."""
    l_prime = min(rank, num_patches)
    out = torch.empty(num_layers, batch, num_recon, l_prime, dtype=torch.int64)
    for layer in range(num_layers):
        for i in range(batch):
            for j in range(num_recon):
                out[layer, i, j] = torch.randperm(num_patches, generator=generator)[:l_prime]
    return out


def special_tokens(p: Params, S: int, reloc_list: Sequence[int], pre: str = "") -> Tensor:
    """models/aggregator.py:287-299 + slice_expand_and_flatten (:806-829): [S, 5, C] camera+register tokens; input
    frame 0 takes slot 0, other frames slot 1, query frames the *_reloc tokens."""
    cam, reg = p[pre + "camera_token"][0], p[pre + "register_token"][0]          # [2,1,C], [2,4,C]
    cam_s = torch.cat([cam[0:1], cam[1:2].expand(S - 1, -1, -1)], dim=0).clone()
    reg_s = torch.cat([reg[0:1], reg[1:2].expand(S - 1, -1, -1)], dim=0).clone()
    for f in reloc_list:
        cam_s[f] = p[pre + "camera_token_reloc"][0, 0]
        reg_s[f] = p[pre + "register_token_reloc"][0, 0]
    return torch.cat([cam_s, reg_s], dim=1)


# --------------------------------------------------------------------------------------------------- aggregator
def aggregator_blocks_forward(patch_tokens: Tensor, p: Params, no_reloc_list: Sequence[int], reloc_list: Sequence[int],
                              grid_hw: Tuple[int, int], rank: int, generator: torch.Generator, depth: int = 24,
                              num_heads: int = 16, intermediate_layer_idx=(4, 11, 17, 23), bf16: bool = False,
                              dense_mask: bool = False, pre: str = "", collect: Optional[dict] = None,
                              scene_cache: Optional[list] = None):
    """Aggregator.forward after the patch embed (models/aggregator.py:276-433), batch 1.

    patch_tokens [S, h*w, C].  Returns (output_dict {layer: [1, Nq, P, 2C]}, patch_start_idx, cam_token_last_layer
    [1, Na, 2C]).  `dense_mask=True` evaluates the reloc block exactly like the reference (one sequence
    [scene ; queries] under the dense boolean mask, aggregator.py:672-741); `dense_mask=False` is the mask-free
    restatement (scene rows are only K/V producers, each query frame attends to [scene ; itself]) that the CUDA path
    implements — tests prove the two equal.
    `scene_cache`: a list that receives, per layer, the (k, v) [H, Na*P', d] the reloc block's Attention stores on its
    first call with kv_cache=True (attention.py:94-97; post q/k-norm, post RoPE) — the SailRecon.tmp_forward priming pass
    (sail_recon.py:176-199, reloc_list = []), whose reloc-block output is zeros + MLP and is discarded.
    """
    S, _, C = patch_tokens.shape
    Na, Nq = len(no_reloc_list), len(reloc_list)
    psi = 5
    tokens = torch.cat([special_tokens(p, S, reloc_list, pre), patch_tokens], dim=1)         # [S, P, C]
    P = tokens.shape[1]
    rank = min(rank, P - psi)                                                                 # aggregator.py:277-278
    P_prime = min(rank + psi, P)
    pos = aggregator_positions(*grid_hw, psi)                                                 # same for every frame
    idx_all = draw_scene_indices(generator, depth, 1, Na, P - psi, rank)
    mask = reloc_mask(Na, Nq, P, P_prime) if (dense_mask and Nq > 0) else None
    out: Dict[int, Tensor] = {}
    cam_last = None
    nrl, rl = list(no_reloc_list), list(reloc_list)
    for layer in range(depth):
        fpre, gpre, rpre = (f"{pre}{n}.{layer}." for n in ("frame_blocks", "global_blocks", "global_reloc_blocks"))
        # frame attention over every frame independently (aggregator.py:643-670)
        tokens = torch.stack([block_forward(tokens[s], p, fpre, num_heads, pos, None, True, True, bf16) for s in range(S)])
        frame_out = tokens
        # scene-token selection from the frame-attended anchors (aggregator.py:351-360, 580-626)
        sel = torch.cat([torch.arange(psi).expand(Na, psi), idx_all[layer, 0] + psi], dim=1)   # [Na, P']
        anchors = tokens[nrl]                                                                    # [Na, P, C]
        scene = torch.gather(anchors, 1, sel[:, :, None].expand(-1, -1, C)).reshape(Na * P_prime, C)
        scene_pos = pos[sel.reshape(-1)]
        if collect is not None:
            collect.setdefault("scene_sel", []).append(sel.clone())
        reloc_out = None
        if scene_cache is not None:
            sn = layer_norm(scene, p[rpre + "norm1.weight"], p[rpre + "norm1.bias"], 1e-5)
            _, sk, sv = qkv_heads(sn, p, rpre + "attn.", num_heads, scene_pos, True, True, bf16)
            scene_cache.append((sk, sv))
        if Nq > 0:
            queries = tokens[rl]                                                                 # [Nq, P, C]
            if dense_mask:
                x = torch.cat([scene, queries.reshape(Nq * P, C)], dim=0)
                xpos = torch.cat([scene_pos, pos.repeat(Nq, 1)], dim=0)
                y = block_forward(x, p, rpre, num_heads, xpos, mask, True, True, bf16)
                reloc_out = y[Na * P_prime:].reshape(Nq, P, C)                                  # aggregator.py:737
            else:
                sn = layer_norm(scene, p[rpre + "norm1.weight"], p[rpre + "norm1.bias"], 1e-5)
                _, sk, sv = qkv_heads(sn, p, rpre + "attn.", num_heads, scene_pos, True, True, bf16)
                reloc_out = torch.stack([block_forward(queries[i], p, rpre, num_heads, pos, None, True, True, bf16,
                                                       prefix_kv=(sk, sv)) for i in range(Nq)])
        # global attention over all anchor tokens as one sequence (aggregator.py:743-769)
        g = block_forward(anchors.reshape(Na * P, C), p, gpre, num_heads, pos.repeat(Na, 1), None, True, True, bf16)
        g = g.reshape(Na, P, C)
        new_tokens = torch.ones_like(tokens)                                                    # aggregator.py:393-399
        new_tokens[nrl] = g
        if Nq > 0:
            new_tokens[rl] = reloc_out
        tokens = new_tokens
        if layer in intermediate_layer_idx and Nq > 0:                                          # aggregator.py:403-413
            out[layer] = torch.cat([frame_out[rl], reloc_out], dim=-1)[None]
        if layer == depth - 1:                                                                  # aggregator.py:414-423
            cam_last = torch.cat([frame_out[nrl][:, 0], g[:, 0]], dim=-1)[None]
    if Nq > 0:
        out[-1] = out[depth - 1]
    return out, psi, cam_last


def cache_mask(num_recon: int, S: int, P: int, P_prime: int) -> Tensor:
    """models/aggregator.py:496-508: allow block over [anchors ; S query frames] expanded to P tokens per frame, cropped
    to rows = the S*P query tokens, cols = [num_recon * P' cached scene tokens ; S*P query tokens]."""
    allow = build_allow_block(num_recon + S, list(range(num_recon)), list(range(num_recon, num_recon + S)))
    m = allow.repeat_interleave(P, 0).repeat_interleave(P, 1)
    return m[-S * P:, -(P_prime * num_recon + S * P):]


def aggregator_blocks_forward_with_cache(patch_tokens: Tensor, p: Params, scene_cache: Sequence[Tuple[Tensor, Tensor]],
                                         num_recon: int, grid_hw: Tuple[int, int], rank: int, depth: int = 24,
                                         num_heads: int = 16, intermediate_layer_idx=(4, 11, 17, 23), bf16: bool = False,
                                         dense_mask: bool = False, pre: str = ""):
    """Aggregator.forward_with_cache after the patch embed (models/aggregator.py:469-578), batch 1: every frame is a
    query carrying the *_reloc special tokens (:483-490); per layer a frame block on each frame, then the reloc block on
    the S*P query tokens whose Attention prepends the cached scene K/V (attention.py:85-93).  `rank` is NOT clamped here
    (:471-472): the mask columns assume rank + 5 cached tokens per anchor (:503-505).  `dense_mask=True` evaluates the
    block exactly like the reference (one sequence under the cropped dense mask); False is the structured restatement
    (each query frame attends to [cache ; itself]).  Returns (output_dict {layer: [1, S, P, 2C]}, patch_start_idx)."""
    S, _, C = patch_tokens.shape
    psi = 5
    cam = p[pre + "camera_token_reloc"][0, 0]
    reg = p[pre + "register_token_reloc"][0, 0]
    special = torch.cat([cam, reg], dim=0)[None].expand(S, -1, -1)
    tokens = torch.cat([special, patch_tokens], dim=1)                                         # [S, P, C]
    P = tokens.shape[1]
    P_prime = rank + psi
    pos = aggregator_positions(*grid_hw, psi)
    mask = cache_mask(num_recon, S, P, P_prime) if dense_mask else None
    out: Dict[int, Tensor] = {}
    for layer in range(depth):
        fpre, rpre = (f"{pre}{n}.{layer}." for n in ("frame_blocks", "global_reloc_blocks"))
        tokens = torch.stack([block_forward(tokens[s], p, fpre, num_heads, pos, None, True, True, bf16) for s in range(S)])
        frame_out = tokens
        ck, cv = scene_cache[layer]
        if dense_mask:
            assert ck.shape[1] == num_recon * P_prime, "cached scene length must match the mask crop (aggregator.py:503-505)"
            y = block_forward(tokens.reshape(S * P, C), p, rpre, num_heads, pos.repeat(S, 1), mask, True, True, bf16,
                              prefix_kv=(ck, cv))
            tokens = y.reshape(S, P, C)
        else:
            tokens = torch.stack([block_forward(tokens[s], p, rpre, num_heads, pos, None, True, True, bf16,
                                                prefix_kv=(ck, cv)) for s in range(S)])
        if layer in intermediate_layer_idx:                                                     # :562-571
            out[layer] = torch.cat([frame_out, tokens], dim=-1)[None]
    assert depth - 1 in out
    out[-1] = out[depth - 1]
    return out, psi


_RESNET_MEAN = (0.485, 0.456, 0.406)
_RESNET_STD = (0.229, 0.224, 0.225)


def dino_patch_embed(images: Tensor, p: Params, pre: str = "patch_embed.", patch_size: int = 14, num_heads: int = 16,
                     depth: int = 24, num_register_tokens: int = 4, bf16: bool = False) -> Tensor:
    """DinoVisionTransformer.forward_features (layers/vision_transformer.py:242-307) -> x_norm_patchtokens.
    images [S, 3, H, W] already ImageNet-normalised.  conv patch projection (layers/patch_embed.py:67-84), cls token,
    bicubic-antialias interpolated position embedding (:206-240, interpolate_offset = 0 as the aggregator builds it,
    models/aggregator.py:196-240), register tokens, `depth` plain blocks (LN eps 1e-6, no RoPE / QK-norm), final norm."""
    S, _, H, W = images.shape
    w_ = _r(p[pre + "patch_embed.proj.weight"], bf16)
    b_ = _r(p[pre + "patch_embed.proj.bias"], bf16)
    x = _r(F.conv2d(_r(images, bf16), w_, b_, stride=patch_size), bf16).flatten(2).transpose(1, 2)   # [S, hw, C]
    C = x.shape[-1]
    x = torch.cat([p[pre + "cls_token"].expand(S, -1, -1), x], dim=1)
    pe = p[pre + "pos_embed"].float()
    N = pe.shape[1] - 1
    if not (x.shape[1] - 1 == N and H == W):
        M = int(math.sqrt(N))
        assert M * M == N
        grid = F.interpolate(pe[:, 1:].reshape(1, M, M, C).permute(0, 3, 1, 2), size=(H // patch_size, W // patch_size),
                             mode="bicubic", antialias=True)
        pe = torch.cat([pe[:, :1], grid.permute(0, 2, 3, 1).reshape(1, -1, C)], dim=1)
    x = x + pe
    if num_register_tokens:
        x = torch.cat([x[:, :1], p[pre + "register_tokens"].expand(S, -1, -1), x[:, 1:]], dim=1)
    for i in range(depth):
        x = torch.stack([block_forward(x[s], p, f"{pre}blocks.{i}.", num_heads, None, None, False, False, bf16, eps=1e-6)
                         for s in range(S)])
    x = layer_norm(x, p[pre + "norm.weight"], p[pre + "norm.bias"], 1e-6)
    return x[:, 1 + num_register_tokens:]


def conv_patch_embed(images: Tensor, p: Params, pre: str = "patch_embed.", patch_size: int = 14, bf16: bool = False) -> Tensor:
    """PatchEmbed (layers/patch_embed.py:25-84), the aggregator's patch_embed="conv" variant."""
    x = F.conv2d(_r(images, bf16), _r(p[pre + "proj.weight"], bf16), _r(p[pre + "proj.bias"], bf16), stride=patch_size)
    return _r(x, bf16).flatten(2).transpose(1, 2)


def aggregator_forward(images: Tensor, p: Params, no_reloc_list, reloc_list, fix_rank: int, generator: torch.Generator,
                       patch_embed: str = "dinov2_vitl14_reg", depth: int = 24, num_heads: int = 16, patch_size: int = 14,
                       vit_depth: int = 24, bf16: bool = False, dense_mask: bool = False, **kw):
    """Aggregator.forward (models/aggregator.py:242-433) for images [1, S, 3, H, W] in [0, 1]."""
    B, S, Cin, H, W = images.shape
    assert B == 1
    if Cin != 3:
        raise ValueError(f"Expected 3 input channels, got {Cin}")
    mean = torch.tensor(_RESNET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(_RESNET_STD).view(1, 3, 1, 1)
    x = (images[0] - mean) / std                                                               # aggregator.py:267
    if "conv" in patch_embed:
        pt = conv_patch_embed(x, p, "patch_embed.", patch_size, bf16)
    else:
        pt = dino_patch_embed(x, p, "patch_embed.", patch_size, num_heads, vit_depth, 4, bf16)
    return aggregator_blocks_forward(pt.float(), p, no_reloc_list, reloc_list, (H // patch_size, W // patch_size), fix_rank,
                                     generator, depth, num_heads, bf16=bf16, dense_mask=dense_mask, **kw)


def _embed_images(images: Tensor, p: Params, patch_embed: str, patch_size: int, num_heads: int, vit_depth: int, bf16: bool) -> Tensor:
    mean = torch.tensor(_RESNET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(_RESNET_STD).view(1, 3, 1, 1)
    x = (images[0] - mean) / std
    if "conv" in patch_embed:
        return conv_patch_embed(x, p, "patch_embed.", patch_size, bf16).float()
    return dino_patch_embed(x, p, "patch_embed.", patch_size, num_heads, vit_depth, 4, bf16).float()


def aggregator_forward_with_cache(images: Tensor, p: Params, scene_cache, num_recon: int, fix_rank: int,
                                  patch_embed: str = "dinov2_vitl14_reg", depth: int = 24, num_heads: int = 16,
                                  patch_size: int = 14, vit_depth: int = 24, bf16: bool = False, dense_mask: bool = False, **kw):
    """Aggregator.forward_with_cache (models/aggregator.py:435-578) for images [1, S, 3, H, W] in [0, 1]; `scene_cache`
    and `num_recon` are what the priming forward left behind (self.num_recon :262, the blocks' k_cache / v_cache)."""
    B, S, Cin, H, W = images.shape
    assert B == 1, "Batch size must be 1 for this model"                                       # :453
    if Cin != 3:
        raise ValueError(f"Expected 3 input channels, got {Cin}")
    pt = _embed_images(images, p, patch_embed, patch_size, num_heads, vit_depth, bf16)
    return aggregator_blocks_forward_with_cache(pt, p, scene_cache, num_recon, (H // patch_size, W // patch_size), fix_rank,
                                                depth, num_heads, bf16=bf16, dense_mask=dense_mask, **kw)


# --------------------------------------------------------------------------------------------------- camera head
def activate_pose(x: Tensor) -> Tensor:
    """heads/head_act.py:12-37 with trans_act = quat_act = "linear", fl_act = "relu"."""
    return torch.cat([x[..., :3], x[..., 3:7], torch.relu(x[..., 7:])], dim=-1)


def camera_head_forward(query_cam_tokens: Tensor, cam_token_last_layer: Tensor, p: Params, num_iterations: int = 4,
                        num_heads: int = 16, trunk_depth: int = 4, pre: str = "", structured: bool = True) -> List[Tensor]:
    """CameraHead.forward + trunk_fn (heads/camera_head.py:85-186), fp32 (the reference disables autocast around it,
    models/sail_recon.py:119).  query_cam_tokens [1, Nq, 2C] = aggregated_tokens_list[-1][:, :, 0];
    cam_token_last_layer [1, Na, 2C].  Returns `num_iterations` tensors [1, Nq, 9] (query rows only, :118-121)."""
    Na, Nq = cam_token_last_layer.shape[1], query_cam_tokens.shape[1]
    x = torch.cat([cam_token_last_layer[0], query_cam_tokens[0]], dim=0)                      # [T, C]
    x = layer_norm(x, p[pre + "token_norm.weight"], p[pre + "token_norm.bias"], 1e-5)
    T, C = x.shape
    allow = camera_mask(Na, Nq)
    pred = None
    outs = []
    for _ in range(num_iterations):
        inp = p[pre + "empty_pose_tokens"][0].expand(T, -1) if pred is None else pred
        emb = linear(inp, p[pre + "embed_pose.weight"], p[pre + "embed_pose.bias"], False)
        mod = linear(F.silu(emb), p[pre + "poseLN_modulation.1.weight"], p[pre + "poseLN_modulation.1.bias"], False)
        shift, scale, gate = mod.chunk(3, dim=-1)
        y = gate * (layer_norm(x, None, None, 1e-6) * (1 + scale) + shift) + x              # camera_head.py:151-160
        for b in range(trunk_depth):
            y = block_forward(y, p, f"{pre}trunk.{b}.", num_heads, None, allow, False, False, False)
        delta = mlp_forward(layer_norm(y, p[pre + "trunk_norm.weight"], p[pre + "trunk_norm.bias"], 1e-5), p,
                            pre + "pose_branch.", False)
        pred = delta if pred is None else pred + delta
        outs.append(activate_pose(pred)[None, Na:])
    return outs

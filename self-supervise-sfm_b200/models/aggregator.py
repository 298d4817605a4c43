"""Aggregator — API mirror of sailrecon/models/aggregator.py (Aggregator :34-803) on the B200 kernel library.

Same constructor, parameter / module names (state_dict compatible with `sailrecon.pt`), `forward(images,
no_reloc_list, reloc_list, fix_rank) -> (output_dict, patch_start_idx, cam_token_last_layer)` (:242-433) and
`forward_with_cache(images, cameras, camera_dropout, fix_rank) -> (output_dict, patch_start_idx)` (:435-578).

B200-first data layout (DESIGN.md "data layout in HBM"): frames are permuted ONCE into [anchors ; queries] order and
all S*P tokens live in one fp32 residual buffer X [S*P, C] for the whole 24-layer loop.  Frame blocks run over all
rows (batch = S segments), the global block over the anchor rows (one long segment) and the reloc block over the
query rows (batch = Nq segments, each behind the shared scene-token K/V prefix) — all in place, so the reference's
per-layer gathers / `torch.ones` / scatters (aggregator.py:355-399) disappear.  The dense boolean mask
(aggregator.py:302-311; (S*P)^2 bytes on the host) is never built: its structure is [scene prefix | own frame].
Scene tokens are K/V producers only (their block outputs are discarded at aggregator.py:737), so only
LayerNorm -> K,V projection runs on them, with the row gather fused into the LayerNorm kernel.
"""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

import torch
import torch.nn as nn

from .. import engine, ops
from ..engine import AttnCall, Prefix, RowPositions
from ..layers import PatchEmbed
from ..layers.block import Block
from ..layers.rope import PositionGetter, RotaryPositionEmbedding2D
from ..layers.vision_transformer import vit_base, vit_giant2, vit_large, vit_small

_RESNET_MEAN = [0.485, 0.456, 0.406]
_RESNET_STD = [0.229, 0.224, 0.225]


class Aggregator(nn.Module):
    def __init__(self, img_size=518, patch_size=14, embed_dim=1024, depth=24, num_heads=16, mlp_ratio=4.0,
                 num_register_tokens=4, block_fn=Block, qkv_bias=True, proj_bias=True, ffn_bias=True,
                 patch_embed="dinov2_vitl14_reg", aa_order=["frame", "global"], aa_block_size=1, qk_norm=True,
                 rope_freq=100, init_values=0.01, intermediate_layer_idx=[4, 11, 17, 23], min_rank: int = 150,
                 kv_cache: bool = False):
        super().__init__()
        self.__build_patch_embed__(patch_embed, img_size, patch_size, num_register_tokens, embed_dim=embed_dim)
        self.rope = RotaryPositionEmbedding2D(frequency=rope_freq) if rope_freq > 0 else None
        self.position_getter = PositionGetter() if self.rope is not None else None
        self.intermediate_layer_idx = intermediate_layer_idx

        def stack(**extra):
            return nn.ModuleList([block_fn(dim=embed_dim, num_heads=num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias,
                                           proj_bias=proj_bias, ffn_bias=ffn_bias, init_values=init_values, qk_norm=qk_norm,
                                           rope=self.rope, **extra) for _ in range(depth)])

        self.frame_blocks = stack()
        self.global_blocks = stack()
        self.global_reloc_blocks = stack(kv_cache=kv_cache)
        self.depth = depth
        self.aa_order = aa_order
        self.patch_size = patch_size
        self.aa_block_size = aa_block_size
        self.min_rank = min_rank
        self.kv_cache = kv_cache
        if self.depth % self.aa_block_size != 0:                                          # aggregator.py:157-160
            raise ValueError(f"depth ({depth}) must be divisible by aa_block_size ({aa_block_size})")
        if list(aa_order) != ["frame", "global"] or aa_block_size != 1:
            raise NotImplementedError("sailrecon_b200: SAIL-Recon runs aa_order=['frame','global'], aa_block_size=1")
        self.aa_block_num = self.depth // self.aa_block_size
        self.camera_token = nn.Parameter(torch.randn(1, 2, 1, embed_dim))
        self.register_token = nn.Parameter(torch.randn(1, 2, num_register_tokens, embed_dim))
        self.camera_token_reloc = nn.Parameter(torch.randn(1, 1, 1, embed_dim))
        self.register_token_reloc = nn.Parameter(torch.randn(1, 1, num_register_tokens, embed_dim))
        self.patch_start_idx = 1 + num_register_tokens
        for t in (self.camera_token, self.register_token, self.camera_token_reloc, self.register_token_reloc):
            nn.init.normal_(t, std=1e-6)
        for name, value in (("_resnet_mean", _RESNET_MEAN), ("_resnet_std", _RESNET_STD)):
            self.register_buffer(name, torch.tensor(value, dtype=torch.float32).view(1, 1, 3, 1, 1), persistent=False)
        self.use_reentrant = False
        self.generator = self._generate_per_rank_generator()
        self.num_recon = 0
        self.rank = 0
        self._pos_cache = {}
        self._sel_cache = {}

    def __build_patch_embed__(self, patch_embed, img_size, patch_size, num_register_tokens, interpolate_antialias=True,
                              interpolate_offset=0.0, block_chunks=0, init_values=1.0, embed_dim=1024):
        """aggregator.py:196-240."""
        if "conv" in patch_embed:
            self.patch_embed = PatchEmbed(img_size=img_size, patch_size=patch_size, in_chans=3, embed_dim=embed_dim)
        else:
            vit_models = {"dinov2_vitl14_reg": vit_large, "dinov2_vitb14_reg": vit_base, "dinov2_vits14_reg": vit_small,
                          "dinov2_vitg2_reg": vit_giant2}
            self.patch_embed = vit_models[patch_embed](img_size=img_size, patch_size=patch_size,
                                                       num_register_tokens=num_register_tokens,
                                                       interpolate_antialias=interpolate_antialias,
                                                       interpolate_offset=interpolate_offset, block_chunks=block_chunks,
                                                       init_values=init_values)
            if hasattr(self.patch_embed, "mask_token"):
                self.patch_embed.mask_token.requires_grad_(False)

    def _generate_per_rank_generator(self):
        """aggregator.py:628-641: CPU generator seeded from the global RNG plus the distributed rank."""
        per_forward_pass_seed = torch.randint(0, 2 ** 32, (1,)).item()
        world_rank = torch.distributed.get_rank() if torch.distributed.is_initialized() else 0
        g = torch.Generator(device="cpu")
        g.manual_seed(per_forward_pass_seed + world_rank)
        return g

    # ---------------------------------------------------------------------------------------------- host-side pieces
    def _draw_rank(self, fix_rank, n_patches: int, clamp: bool) -> int:
        if fix_rank is not None:
            return min(fix_rank, n_patches) if clamp else fix_rank                        # aggregator.py:277-278 / :471-472
        return torch.randint(min(self.min_rank, n_patches // 2), max(self.min_rank, n_patches // 2), (1,),
                             generator=self.generator).item()                              # :280-285

    def _positions(self, S: int, h: int, w: int, device) -> RowPositions:
        """aggregator.py:313-328: grid (y, x) + 1, zeros for the special tokens; identical for every frame."""
        key = (S, h, w, torch.device(device))
        hit = self._pos_cache.get(key)
        if hit is None:
            grid = self.position_getter(1, h, w, device)[0] + 1                            # int64 [h*w, 2]
            frame = torch.cat([torch.zeros(self.patch_start_idx, 2, dtype=grid.dtype, device=device), grid], dim=0)
            hit = (frame.to(torch.int32).repeat(S, 1).contiguous(), frame.cpu())
            self._pos_cache = {key: hit}
        return RowPositions(hit[0], max(h, w) + 1), hit[1]

    def _embed(self, imgs: torch.Tensor, mode: str, X: torch.Tensor, S: int, P: int, C: int) -> None:
        """aggregator.py:267-274: ImageNet normalisation + patch embed, written straight into the patch rows of X."""
        if isinstance(self.patch_embed, PatchEmbed):
            m = mode if (mode == "f32" or C % 128 == 0) else "f32"
            y = self.patch_embed.project_rows(imgs, m, _RESNET_MEAN, _RESNET_STD)
            X.view(S, P, C)[:, self.patch_start_idx:] = y.view(S, P - self.patch_start_idx, C)
        else:
            self.patch_embed.forward_rows(imgs, mode, _RESNET_MEAN, _RESNET_STD, out=X)

    def _special_rows(self, order: List[int], reloc: set, all_reloc: bool) -> torch.Tensor:
        """aggregator.py:287-299 (+ slice_expand_and_flatten :806-829): [S, 5, C]; input frame 0 takes slot 0, other
        frames slot 1, query frames the *_reloc tokens."""
        cam, reg = self.camera_token.detach()[0], self.register_token.detach()[0]          # [2,1,C], [2,R,C]
        slots = torch.cat([torch.cat([cam[0], reg[0]], 0)[None], torch.cat([cam[1], reg[1]], 0)[None],
                           torch.cat([self.camera_token_reloc.detach()[0, 0], self.register_token_reloc.detach()[0, 0]], 0)[None]])
        sel = tuple(2 if (all_reloc or f in reloc) else (0 if f == 0 else 1) for f in order)
        key = (sel, slots.device)
        idx = self._sel_cache.get(key)               # cached: no H2D copy per call (and none under CUDA-graph capture)
        if idx is None:
            idx = torch.tensor(sel, device=slots.device)
            self._sel_cache = {key: idx}
        return slots[idx]

    def _draw_scene(self, Na: int, P: int, hw: int, l_prime: int, frame_pos_cpu: torch.Tensor):
        """aggregator.py:580-626 for one layer, host side: per anchor (batch-major, anchor order of no_reloc_list) ONE
        `torch.randperm(hw, generator)[:l_prime]` on the CPU generator — same draws, same order as the reference.
        Returns (int32 [Na*P', 3] = (row index into X, y, x) per scene token, int64 patch indices [Na, l'])."""
        psi = self.patch_start_idx
        sel = torch.empty(Na, psi + l_prime, dtype=torch.int64)
        sel[:, :psi] = torch.arange(psi)
        for j in range(Na):
            sel[j, psi:] = torch.randperm(hw, generator=self.generator)[:l_prime] + psi
        rows = (sel + (torch.arange(Na) * P)[:, None]).reshape(-1).to(torch.int32)
        pos = frame_pos_cpu[sel.reshape(-1)].to(torch.int32)
        return torch.cat([rows[:, None], pos], dim=1).contiguous(), sel[:, psi:] - psi

    def _scene_rows(self, Na: int, P: int, hw: int, l_prime: int, frame_pos_cpu: torch.Tensor, device):
        """Draw on the host, one small pinned H2D copy; -> (row indices int32 [Na*P'], positions int32 [Na*P', 2], drawn)."""
        packed, drawn = self._draw_scene(Na, P, hw, l_prime, frame_pos_cpu)
        dev = packed.pin_memory().to(device, non_blocking=True)
        return dev[:, 0].contiguous(), dev[:, 1:].contiguous(), drawn

    # ---------------------------------------------------------------------------------------------- forward
    def forward(self, images: torch.Tensor, no_reloc_list: list, reloc_list: list,
                fix_rank: Union[int, None] = None) -> Tuple[dict, int, torch.Tensor]:
        if self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            # training (train_imc.py:361-381): one autograd node over the block stacks — forward keeps what the backward
            # kernels read, loss.backward() reaches them (sailrecon_b200.train.AggregatorFunction)
            from .. import train
            if self.kv_cache:
                raise NotImplementedError("sailrecon_b200: the differentiated forward is the kv_cache=False path (sail_recon.py:70-78)")
            return train.aggregator_forward_autograd(self, images, no_reloc_list, reloc_list, fix_rank)
        B, S, C_in, H, W = images.shape
        Na, Nq = len(no_reloc_list), len(reloc_list)
        self.num_recon = Na
        if C_in != 3:
            raise ValueError(f"Expected 3 input channels, got {C_in}")                     # aggregator.py:263-264
        if B != 1:
            raise NotImplementedError("sailrecon_b200: batch size 1 (every SAIL-Recon script runs B = 1, sail_recon.py:90-91)")
        engine.require_cuda_f32(images, "images")
        order = list(no_reloc_list) + list(reloc_list)
        if sorted(order) != list(range(S)) or Na == 0:
            raise NotImplementedError("sailrecon_b200: no_reloc_list and reloc_list must be disjoint, cover range(S) and "
                                      "contain at least one anchor (frames in neither list are left as ones by the "
                                      "reference, aggregator.py:393-399; no caller does that)")
        if self.kv_cache and Nq > 0:
            raise NotImplementedError("sailrecon_b200: with kv_cache=True forward() primes the scene cache and takes no "
                                      "query frames (SailRecon.tmp_forward, sail_recon.py:192-194)")
        mode = engine.compute_mode()
        dev = images.device
        C = self.camera_token.shape[-1]
        psi = self.patch_start_idx
        h, w = H // self.patch_size, W // self.patch_size
        hw = h * w
        P = psi + hw
        self.rank = self._draw_rank(fix_rank, hw, clamp=True)
        l_prime = min(self.rank, hw)                                                       # aggregator.py:595-599
        Pp = psi + l_prime                                                                 # = min(rank + psi, P), :373
        imgs = images[0] if order == list(range(S)) else images[0][torch.tensor(order, device=dev)]
        T = S * P
        X = engine.WS.get("agg.x", (T, C), torch.float32, dev)
        self._embed(imgs.contiguous(), mode, X, S, P, C)
        with torch.no_grad():
            X.view(S, P, C)[:, :psi] = self._special_rows(order, set(reloc_list), False)
        rows_pos, frame_pos_cpu = self._positions(S, h, w, dev)
        pos_a = RowPositions(rows_pos.pos32[:Na * P], rows_pos.max_pos)
        pos_q = RowPositions(rows_pos.pos32[Na * P:], rows_pos.max_pos)
        act = torch.bfloat16 if mode == "bf16" else torch.float32
        Xa, Xq = X[:Na * P], X[Na * P:]
        frame_calls = engine.full_segments(S, P)
        global_calls = engine.full_segments(1, Na * P)
        output_dict = {}
        cam_token_last_layer = None
        self.last_scene_indices = []                                                       # bit-exact check hook for tests
        for i in range(self.depth):
            engine.run_block(self.frame_blocks[i], X, frame_calls, rows_pos, mode, tag="agg.")      # aggregator.py:343-349
            out_i = None
            if i in self.intermediate_layer_idx and Nq > 0:                                # :403-413, first half of the concat
                out_i = torch.empty(1, Nq, P, 2 * C, dtype=torch.float32, device=dev)
                ops.copy_cols(Xq, out_i.view(Nq * P, 2 * C), 0)
            if i == self.depth - 1:
                cam_frame = X.view(S, P, C)[:Na, 0].clone()                                # :417
            # scene tokens: re-drawn every layer from the frame-attended anchors (:351-360)
            idx, spos, drawn = self._scene_rows(Na, P, hw, l_prime, frame_pos_cpu, dev)
            self.last_scene_indices.append(drawn)
            rblk = self.global_reloc_blocks[i]
            if Nq > 0 or self.kv_cache:
                a_s = engine.WS.get("agg.scene_a", (Na * Pp, C), act, dev)
                w1, b1, eps1 = engine.ln_params(rblk.norm1)
                ops.layernorm(X, w1, b1, eps1, a_s, idx=idx)                               # gather + norm1 of the reloc block
                if mode == "bf16":
                    skv = engine.WS.get("agg.scene_kv", (Na * Pp, 2 * C), act, dev)
                    engine.qkv_rows(rblk.attn, a_s, skv, mode, RowPositions(spos, rows_pos.max_pos), col_offset=C)
                    sk, sv = skv[:, :C], skv[:, C:]
                else:
                    skv = engine.WS.get("agg.scene_kv", (Na * Pp, 3 * C), act, dev)
                    engine.qkv_rows(rblk.attn, a_s, skv, mode, RowPositions(spos, rows_pos.max_pos))
                    sk, sv = skv[:, C:2 * C], skv[:, 2 * C:]
                if self.kv_cache:
                    if rblk.attn.k_cache is not None:
                        raise RuntimeError("sailrecon_b200: scene kv cache already primed; call clear_kv_cache() first "
                                           "(SailRecon.tmp_forward does, sail_recon.py:178-181)")
                    rblk.attn.k_cache, rblk.attn.v_cache = sk.clone(), sv.clone()          # attention.py:94-97, kept in HBM
            if Nq > 0:                                                                     # :362-379, query rows only (:737)
                engine.run_block(rblk, Xq, [AttnCall(0, Nq, P, P, Prefix(sk, sv, 0, Na * Pp))], pos_q, mode, tag="agg.")
            engine.run_block(self.global_blocks[i], Xa, global_calls, pos_a, mode, tag="agg.")      # :380-392
            if out_i is not None:
                ops.copy_cols(Xq, out_i.view(Nq * P, 2 * C), C)
                output_dict[i] = out_i
            if i == self.depth - 1:                                                        # :414-423
                cam_token_last_layer = torch.cat([cam_frame, X.view(S, P, C)[:Na, 0]], dim=-1)[None].clone()
        assert (self.depth - 1 in output_dict) or Nq == 0, \
            f"Please make sure the last layer ({self.depth - 1}) is in the output_dict: {output_dict.keys()}"
        if Nq > 0:
            output_dict[-1] = output_dict[self.depth - 1]
        return output_dict, self.patch_start_idx, cam_token_last_layer

    def forward_with_cache(self, images: torch.Tensor, cameras: torch.Tensor = None, camera_dropout: float = False,
                           fix_rank: Union[int, None] = None, camera_only: bool = False) -> Tuple[dict, int]:
        """aggregator.py:435-578: every frame is a query localised against the cached scene K/V (HBM resident).
        `camera_only=True` (not in the reference; used by localize.Localizer for SailRecon.reloc(fast_reloc=True),
        sail_recon.py:240-244, where only `rgb_feats[-1][:, :, 0]` is read): no intermediate outputs are written and the last
        reloc block computes its query side / projection / MLP for the camera rows only; returns {-1: [1, S, 1, 2C]}."""
        engine.forbid_autograd(self, "Aggregator.forward_with_cache")
        B, S, C_in, H, W = images.shape
        assert B == 1, "Batch size must be 1 for this model"                               # aggregator.py:453
        if C_in != 3:
            raise ValueError(f"Expected 3 input channels, got {C_in}")
        engine.require_cuda_f32(images, "images")
        mode = engine.compute_mode()
        dev = images.device
        C = self.camera_token.shape[-1]
        psi = self.patch_start_idx
        h, w = H // self.patch_size, W // self.patch_size
        P = psi + h * w
        self.rank = self._draw_rank(fix_rank, h * w, clamp=False)
        X = engine.WS.get("agg.x", (S * P, C), torch.float32, dev)
        self._embed(images[0].contiguous(), mode, X, S, P, C)
        with torch.no_grad():
            X.view(S, P, C)[:, :psi] = self._special_rows(list(range(S)), set(), True)      # :483-490
        rows_pos, _ = self._positions(S, h, w, dev)
        frame_calls = engine.full_segments(S, P)
        output_dict = {}
        for i in range(self.depth):
            engine.run_block(self.frame_blocks[i], X, frame_calls, rows_pos, mode, tag="agg.")
            out_i = None
            if i in self.intermediate_layer_idx and not camera_only:
                out_i = torch.empty(1, S, P, 2 * C, dtype=torch.float32, device=dev)
                ops.copy_cols(X, out_i.view(S * P, 2 * C), 0)
            attn = self.global_reloc_blocks[i].attn
            if attn.k_cache is None:
                raise RuntimeError("sailrecon_b200: forward_with_cache needs a primed scene kv cache (run forward() with "
                                   "kv_cache=True first, SailRecon.tmp_forward)")
            pre = Prefix(attn.k_cache, attn.v_cache, 0, attn.k_cache.shape[0])
            if camera_only and i == self.depth - 1:
                cam_frame = X.view(S, P, C)[:, 0]
                cam_reloc = engine.run_block_camera_rows(self.global_reloc_blocks[i], X, S, P, pre, rows_pos, mode, tag="agg.")
                return {-1: torch.cat([cam_frame, cam_reloc], dim=-1).view(1, S, 1, 2 * C)}, self.patch_start_idx
            engine.run_block(self.global_reloc_blocks[i], X, [AttnCall(0, S, P, P, pre)], rows_pos, mode, tag="agg.")
            if out_i is not None:
                ops.copy_cols(X, out_i.view(S * P, 2 * C), C)
                output_dict[i] = out_i
        assert self.depth - 1 in output_dict, \
            f"Please make sure the last layer ({self.depth - 1}) is in the output_dict: {output_dict.keys()}"
        output_dict[-1] = output_dict[self.depth - 1]
        return output_dict, self.patch_start_idx

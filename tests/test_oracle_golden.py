"""CPU: the oracle restatement (oracle/sailrecon_oracle.py) against golden vectors produced by the unmodified
reference (oracle/make_golden.py), and — when /root/reference is importable — against the live reference."""
import os
import sys

import pytest
import torch

import sailrecon_oracle as O
from conftest import load_golden
from golden_common import GOLDEN_CASES, make_images, rel_l2

TOL = 2e-5      # fp32 vs fp32, different summation orders only


def _run_oracle(name, dense_mask):
    arrays, meta, params = load_golden(name)
    case = GOLDEN_CASES[name]
    images = make_images(case["S"], case["H"], case["W"], case["seed"] + 100)
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    head_p = {k[5:]: v for k, v in params.items() if k.startswith("head.")}
    g = torch.Generator().manual_seed(1234)
    a = case["agg"]
    collect = {}
    vit = dict(dinov2_vits14_reg=(12, 6)).get(a["patch_embed"], (0, a["num_heads"]))
    out, psi, cam = O.aggregator_forward(images, agg_p, meta["no_reloc_list"], meta["reloc_list"], case["fix_rank"], g,
                                         patch_embed=a["patch_embed"], depth=a["depth"], num_heads=a["num_heads"],
                                         vit_depth=vit[0], intermediate_layer_idx=tuple(a["intermediate_layer_idx"]),
                                         dense_mask=dense_mask, collect=collect)
    h = case["head"]
    poses = O.camera_head_forward(out[-1][:, :, 0], cam, head_p, 4, h["num_heads"], h["trunk_depth"])
    return arrays, meta, out, psi, cam, poses, collect


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
@pytest.mark.parametrize("dense_mask", [True, False])
def test_oracle_matches_reference_golden(name, dense_mask):
    arrays, meta, out, psi, cam, poses, collect = _run_oracle(name, dense_mask)
    assert psi == meta["patch_start_idx"] == 5
    for k, v in out.items():
        assert rel_l2(v, arrays[f"out_{k}"]) < TOL, k
    assert rel_l2(cam, arrays["cam_token_last_layer"]) < TOL
    for i, pz in enumerate(poses):
        assert rel_l2(pz, arrays[f"pose_{i}"]) < 1e-4, i
    # integer indexing: bit-exact
    sel = torch.stack(collect["scene_sel"])[:, :, 5:] - 5
    assert torch.equal(sel, arrays["scene_idx"])
    hw = (GOLDEN_CASES[name]["H"] // 14, GOLDEN_CASES[name]["W"] // 14)
    assert torch.equal(O.position_grid(*hw), arrays["positions"])


def test_structured_mask_equals_reference_dense_mask():
    """aggregator.py:302-311 crop of the token-expanded allow block == [scene prefix | block diagonal]."""
    Na, Nq, P, Pp = 3, 4, 11, 7
    m = O.reloc_mask(Na, Nq, P, Pp)
    L = Na * Pp + Nq * P
    assert m.shape == (L, L)
    exp = torch.zeros(L, L, dtype=torch.bool)
    exp[:, : Na * Pp] = True
    for i in range(Nq):
        s = Na * Pp + i * P
        exp[s:s + P, s:s + P] = True
    assert torch.equal(m, exp)
    cm = O.camera_mask(2, 3)
    exp = torch.tensor([[1, 1, 0, 0, 0], [1, 1, 0, 0, 0], [1, 1, 1, 0, 0], [1, 1, 0, 1, 0], [1, 1, 0, 0, 1]], dtype=torch.bool)
    assert torch.equal(cm, exp)


def test_bf16_emulation_close_to_fp32():
    arrays, meta, params = load_golden("tiny_conv")
    case = GOLDEN_CASES["tiny_conv"]
    images = make_images(case["S"], case["H"], case["W"], case["seed"] + 100)
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    a = case["agg"]
    out, _, cam = O.aggregator_forward(images, agg_p, meta["no_reloc_list"], meta["reloc_list"], case["fix_rank"],
                                       torch.Generator().manual_seed(1234), patch_embed="conv", depth=a["depth"],
                                       num_heads=a["num_heads"], intermediate_layer_idx=(0, 1), bf16=True)
    e = rel_l2(out[-1], arrays["out_-1"])
    assert 1e-5 < e < 2e-2, e


@pytest.mark.skipif(not os.path.isdir("/root/reference/sailrecon"), reason="reference tree not present on this machine")
def test_oracle_layers_match_live_reference():
    sys.path.insert(0, "/root/reference")
    from sailrecon.layers.block import Block
    from sailrecon.layers.rope import PositionGetter, RotaryPositionEmbedding2D

    torch.manual_seed(3)
    rope = RotaryPositionEmbedding2D(frequency=100.0)
    blk = Block(dim=128, num_heads=2, init_values=0.7, qk_norm=True, rope=rope).eval()
    p = {k: v for k, v in blk.state_dict().items()}
    pos = PositionGetter()(1, 3, 4, "cpu") + 1
    x = torch.randn(1, 12, 128)
    with torch.no_grad():
        ref = blk(x, pos=pos)
        refq = rope(x.view(1, 12, 2, 64).transpose(1, 2), pos)
    got = O.block_forward(x[0], p, "", 2, pos[0], None, True, True, False)
    assert rel_l2(got, ref[0]) < 1e-6
    assert torch.allclose(O.rope_2d(x.view(12, 2, 64).transpose(0, 1), pos[0]), refq[0], atol=1e-6)


def _cache_oracle(dense_mask, bf16=False):
    """Priming forward (kv_cache=True, reloc_list = []) + forward_with_cache, oracle restatement."""
    arrays, meta, params = load_golden_cache()
    case = GOLDEN_CASES["tiny_conv"]
    Na = case["Na"]
    images = make_images(case["S"], case["H"], case["W"], case["seed"] + 100)
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    head_p = {k[5:]: v for k, v in params.items() if k.startswith("head.")}
    a = case["agg"]
    cache = []
    out0, _, cam = O.aggregator_forward(images[:, :Na], agg_p, list(range(Na)), [], case["fix_rank"],
                                        torch.Generator().manual_seed(1234), patch_embed="conv", depth=a["depth"],
                                        num_heads=a["num_heads"], intermediate_layer_idx=(0, 1), bf16=bf16, scene_cache=cache)
    assert len(out0) == 0 and len(cache) == a["depth"]
    kw = dict(patch_embed="conv", depth=a["depth"], num_heads=a["num_heads"], intermediate_layer_idx=(0, 1), bf16=bf16,
              dense_mask=dense_mask)
    out, psi = O.aggregator_forward_with_cache(images[:, Na:], agg_p, cache, Na, case["fix_rank"], **kw)
    single, _ = O.aggregator_forward_with_cache(images[:, Na:Na + 1], agg_p, cache, Na, case["fix_rank"], **kw)
    h = case["head"]
    poses = O.camera_head_forward(out[-1][:, :, 0], cam, head_p, 4, h["num_heads"], h["trunk_depth"])
    return arrays, out, single, cam, cache, poses


def load_golden_cache():
    import json
    import numpy as np
    from golden_common import make_params
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiny_cache.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    arrays = {k: torch.from_numpy(z[k]) for k in z.files if k != "meta"}
    return arrays, meta, make_params(meta["shapes"], GOLDEN_CASES["tiny_conv"]["seed"])


@pytest.mark.parametrize("dense_mask", [True, False])
def test_oracle_cached_path_matches_reference_golden(dense_mask):
    """aggregator.py:435-578 + attention.py:84-100 restated (oracle.aggregator_forward_with_cache, scene_cache hook) against
    tests/golden/tiny_cache.npz, which oracle/make_golden.py produced by running the UNMODIFIED reference's kv-cache path
    (priming forward, forward_with_cache on 2 frames and on 1 frame, camera head)."""
    arrays, out, single, cam, cache, poses = _cache_oracle(dense_mask)
    assert rel_l2(cache[0][0], arrays["k_cache_0"][0]) < TOL and rel_l2(cache[0][1], arrays["v_cache_0"][0]) < TOL
    assert rel_l2(cam, arrays["cam_token_last_layer"]) < TOL
    for k, v in out.items():
        assert rel_l2(v, arrays[f"out_{k}"]) < TOL, k
    assert rel_l2(single[-1], arrays["single_out_-1"]) < TOL
    for i, pz in enumerate(poses):
        assert rel_l2(pz, arrays[f"pose_{i}"]) < 1e-4, i


def test_cache_mask_structure():
    """aggregator.py:496-508 crop == [every cached scene token | block diagonal over the query frames]."""
    Na, S, P, Pp = 3, 2, 7, 4
    m = O.cache_mask(Na, S, P, Pp)
    assert m.shape == (S * P, Na * Pp + S * P)
    exp = torch.zeros_like(m)
    exp[:, : Na * Pp] = True
    for i in range(S):
        exp[i * P:(i + 1) * P, Na * Pp + i * P: Na * Pp + (i + 1) * P] = True
    assert torch.equal(m, exp)

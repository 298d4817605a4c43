"""GPU parity of the module layer (sailrecon_b200.models.Aggregator / heads.CameraHead / layers.Block) against
  (a) golden vectors produced by the UNMODIFIED reference (tests/golden/*.npz, oracle/make_golden.py) and
  (b) the CPU oracle (oracle/sailrecon_oracle.py) on the same seeded inputs.

Tolerances (north_star): aggregator tokens rel-L2 <= 1e-4 in fp32 mode and <= 2e-2 under autocast(bf16); camera pose
encodings <= 1e-2; scene-token indices and positions bit-exact.  The fixtures use O(1) LayerScale / special tokens so
every sub-path is visible in the output (the reference init, gamma = 0.01, would hide kernel bugs).
"""
import pytest
import torch

import sailrecon_oracle as O
from conftest import load_golden
from golden_common import GOLDEN_CASES, make_images, make_params, rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _build(name):
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    arrays, meta, params = load_golden(name)
    case = GOLDEN_CASES[name]
    agg = Aggregator(**case["agg"]).eval()
    head = CameraHead(**case["head"]).eval()
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    images = make_images(case["S"], case["H"], case["W"], case["seed"] + 100)
    return agg.to(DEV), head.to(DEV), images.to(DEV), arrays, meta, case, params


def _run(agg, head, images, meta, case, bf16):
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        if bf16:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out, psi, cam = agg(images, meta["no_reloc_list"], meta["reloc_list"], fix_rank=case["fix_rank"])
        else:
            out, psi, cam = agg(images, meta["no_reloc_list"], meta["reloc_list"], fix_rank=case["fix_rank"])
        poses = head(out, cam)
    return out, psi, cam, poses


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
def test_fp32_mode_matches_reference_golden(name):
    agg, head, images, arrays, meta, case, _ = _build(name)
    out, psi, cam, poses = _run(agg, head, images, meta, case, bf16=False)
    assert psi == meta["patch_start_idx"]
    assert set(out) == {int(k[4:]) for k in arrays if k.startswith("out_")}
    for k, v in out.items():
        assert v.dtype == torch.float32 and tuple(v.shape) == tuple(arrays[f"out_{k}"].shape)
        assert rel_l2(v.cpu(), arrays[f"out_{k}"]) < 1e-4, k
    assert rel_l2(cam.cpu(), arrays["cam_token_last_layer"]) < 1e-4
    for i, pz in enumerate(poses):
        assert tuple(pz.shape) == tuple(arrays[f"pose_{i}"].shape)
        assert (pz.cpu() - arrays[f"pose_{i}"]).abs().max() < 1e-2 and rel_l2(pz.cpu(), arrays[f"pose_{i}"]) < 1e-3
    # integer indexing: bit-exact with what the reference drew (aggregator.py:617-621)
    assert torch.equal(torch.stack(agg.last_scene_indices), arrays["scene_idx"])
    assert agg.rank == meta["rank"] and agg.num_recon == len(meta["no_reloc_list"])


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
def test_bf16_mode_within_tolerance(name):
    """autocast(bf16): <= 2e-2 against the fp32 reference golden, and tight against the oracle's emulation of the
    autocast rounding points (the two differ only by accumulation order and 1-ulp bf16 flips)."""
    agg, head, images, arrays, meta, case, params = _build(name)
    out, psi, cam, poses = _run(agg, head, images, meta, case, bf16=True)
    for k, v in out.items():
        assert rel_l2(v.cpu(), arrays[f"out_{k}"]) < 2e-2, k
    assert rel_l2(cam.cpu(), arrays["cam_token_last_layer"]) < 2e-2
    for i, pz in enumerate(poses):
        assert (pz.cpu() - arrays[f"pose_{i}"]).abs().max() < 5e-2
    a = case["agg"]
    vit = dict(dinov2_vits14_reg=(12, 6)).get(a["patch_embed"], (0, a["num_heads"]))
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    ref, _, ref_cam = O.aggregator_forward(images.cpu(), agg_p, meta["no_reloc_list"], meta["reloc_list"], case["fix_rank"],
                                           torch.Generator().manual_seed(1234), patch_embed=a["patch_embed"],
                                           depth=a["depth"], num_heads=a["num_heads"], vit_depth=vit[0],
                                           intermediate_layer_idx=tuple(a["intermediate_layer_idx"]), bf16=True)
    for k, v in out.items():
        assert rel_l2(v.cpu(), ref[k]) < 8e-3, k
    assert rel_l2(cam.cpu(), ref_cam) < 8e-3


def test_block_generic_api_with_dense_masks():
    """Block.forward(x, pos, attn_mask) with the dense boolean masks the reference builds: reloc mask
    (aggregator.py:302-311) and camera mask (~build_lr_mask, camera_head.py:163-166), fp32 and autocast."""
    from sailrecon_b200.layers import Block, RotaryPositionEmbedding2D
    torch.manual_seed(5)
    rope = RotaryPositionEmbedding2D(frequency=100.0)
    blk = Block(dim=128, num_heads=2, init_values=0.7, qk_norm=True, rope=rope).eval()
    with torch.no_grad():
        for n, p_ in blk.named_parameters():
            if p_.dim() == 1 and "gamma" not in n:
                p_.add_(0.1 * torch.randn_like(p_))
    p = {k: v.clone() for k, v in blk.state_dict().items()}
    blk = blk.to(DEV)
    Na, Nq, P, Pp = 2, 3, 17, 9
    mask = O.reloc_mask(Na, Nq, P, Pp)
    L = mask.shape[0]
    g = torch.Generator().manual_seed(6)
    x = torch.randn(1, L, 128, generator=g)
    pos = torch.randint(0, 12, (1, L, 2), generator=g)
    ref = O.block_forward(x[0], p, "", 2, pos[0], mask, True, True, False)
    with torch.no_grad():
        got = blk(x.to(DEV), pos=pos.to(DEV), attn_mask=mask.to(DEV)[None, None])
    assert rel_l2(got[0].cpu(), ref) < 1e-5
    ref16 = O.block_forward(x[0], p, "", 2, pos[0], mask, True, True, True)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        got16 = blk(x.to(DEV), pos=pos.to(DEV), attn_mask=mask.to(DEV)[None, None])
    assert got16.dtype == torch.float32 and rel_l2(got16[0].cpu(), ref16) < 6e-3
    # unsupported structure -> loud error, not a silent fallback
    bad = mask.clone()
    bad[3, 5] = ~bad[3, 5]
    with pytest.raises(NotImplementedError):
        blk(x.to(DEV), pos=pos.to(DEV), attn_mask=bad.to(DEV)[None, None])
    # batch of independent sequences, no mask, no positions on a rope-less block
    blk2 = Block(dim=128, num_heads=2, init_values=None).eval()
    p2 = {k: v.clone() for k, v in blk2.state_dict().items()}
    xb = torch.randn(3, 20, 128, generator=g)
    ref2 = torch.stack([O.block_forward(xb[i], p2, "", 2) for i in range(3)])
    with torch.no_grad():
        got2 = blk2.to(DEV)(xb.to(DEV))
    assert rel_l2(got2.cpu(), ref2) < 1e-5


def test_rope_and_position_getter_bit_exact():
    from sailrecon_b200.layers import PositionGetter, RotaryPositionEmbedding2D
    pg = PositionGetter()
    pos = pg(2, 5, 7, DEV)
    assert pos.dtype == torch.int64 and torch.equal(pos[0].cpu(), O.position_grid(5, 7)) and torch.equal(pos[0], pos[1])
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, 3, 35, 64, generator=g)
    rope = RotaryPositionEmbedding2D(100.0)
    got = rope(x.to(DEV), pos)
    ref = O.rope_2d(x, O.position_grid(5, 7))
    assert rel_l2(got.cpu(), ref) < 1e-6


def test_two_phase_localisation_matches_one_pass():
    """SailRecon.tmp_forward + reloc (sail_recon.py:176-286): prime the HBM-resident scene K/V cache with the anchors,
    then localise the query frames with forward_with_cache.  With the same scene-token draws this must reproduce the
    one-pass forward for the query rows (queries are mutually independent given the anchors)."""
    from sailrecon_b200.models import Aggregator
    name = "tiny_conv"
    agg, head, images, arrays, meta, case, params = _build(name)
    out1, _, cam1, poses1 = _run(agg, head, images, meta, case, bf16=False)
    agg2 = Aggregator(**case["agg"], kv_cache=True).eval()
    agg2.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    agg2 = agg2.to(DEV)
    nrl, rl = meta["no_reloc_list"], meta["reloc_list"]
    agg2.generator.manual_seed(1234)
    with torch.no_grad():
        _, _, cam2 = agg2(images[:, nrl], list(range(len(nrl))), [], fix_rank=case["fix_rank"])
        out2, psi = agg2.forward_with_cache(images[:, rl], fix_rank=case["fix_rank"])
        poses2 = head(out2, cam2)
    assert rel_l2(cam2.cpu(), cam1.cpu()) < 1e-5
    # in the one-pass forward the query frames carry index >= 1 -> same *_reloc special tokens in both paths
    for k in out1:
        assert rel_l2(out2[k].cpu(), out1[k].cpu()) < 1e-5, k
    for a, b in zip(poses1, poses2):
        assert rel_l2(b.cpu(), a.cpu()) < 1e-4
    for blk in agg2.global_reloc_blocks:
        assert blk.attn.k_cache is not None and blk.attn.k_cache.is_cuda
        blk.attn.clear_kv_cache()
        assert blk.attn.k_cache is None


def test_full_width_reduced_depth_518():
    """C = 1024, 16 heads, 518x518 (P = 1374), DINOv2-L width with 2 ViT blocks + 2 aggregator layers, 2 anchors +
    1 query, autocast(bf16) against the oracle's bf16 emulation: exercises every kernel at the production tile shapes."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    torch.manual_seed(0)
    agg = Aggregator(depth=2, intermediate_layer_idx=[0, 1]).eval()
    agg.patch_embed.blocks = agg.patch_embed.blocks[:2]
    head = CameraHead(dim_in=2048, trunk_depth=1).eval()
    shapes = {"agg." + k: list(v.shape) for k, v in agg.state_dict().items()}
    shapes.update({"head." + k: list(v.shape) for k, v in head.state_dict().items()})
    params = make_params(shapes, 21)
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    images = make_images(3, 518, 518, 121)
    agg, head = agg.to(DEV), head.to(DEV)
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out, psi, cam = agg(images.to(DEV), [0, 1], [2], fix_rank=300)
        poses = head(out, cam)
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    head_p = {k[5:]: v for k, v in params.items() if k.startswith("head.")}
    ref, _, ref_cam = O.aggregator_forward(images, agg_p, [0, 1], [2], 300, torch.Generator().manual_seed(1234), depth=2,
                                           vit_depth=2, intermediate_layer_idx=(0, 1), bf16=True)
    ref_poses = O.camera_head_forward(ref[-1][:, :, 0], ref_cam, head_p, 4, 16, 1)
    assert tuple(out[-1].shape) == (1, 1, 1374, 2048)
    for k in out:
        assert rel_l2(out[k].cpu(), ref[k]) < 8e-3, k
    assert rel_l2(cam.cpu(), ref_cam) < 8e-3
    for a, b in zip(poses, ref_poses):                      # O(1) LayerScale weights: pose values are O(3), compare relatively
        assert rel_l2(a.cpu(), b) < 1e-2


def test_imc_style_scene_more_queries_than_anchors():
    """BASELINE.json configs[2] in miniature (train/demo_imc_forward.py:76-101): 3 anchors build the scene tokens, 6 query
    frames are localised against them in ONE pass; full width (C = 1024, 16 heads), 224 px (P = 261), fix_rank 100 so
    that the scene prefix (3 x 105 rows) ends inside a 128-row K/V tile, 2 ViT blocks + 2 aggregator layers, fp32 check
    mode (1e-4) and autocast(bf16) (2e-2) against the oracle; the scene-token draw order is part of the comparison."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    torch.manual_seed(0)
    agg = Aggregator(img_size=224, depth=2, intermediate_layer_idx=[0, 1]).eval()
    agg.patch_embed.blocks = agg.patch_embed.blocks[:2]
    head = CameraHead(dim_in=2048, trunk_depth=1).eval()
    shapes = {"agg." + k: list(v.shape) for k, v in agg.state_dict().items()}
    shapes.update({"head." + k: list(v.shape) for k, v in head.state_dict().items()})
    params = make_params(shapes, 33)
    agg_p = {k[4:]: v for k, v in params.items() if k.startswith("agg.")}
    head_p = {k[5:]: v for k, v in params.items() if k.startswith("head.")}
    agg.load_state_dict(agg_p)
    head.load_state_dict(head_p)
    Na, Nq = 3, 6
    images = make_images(Na + Nq, 224, 224, 133)
    anchors, queries = list(range(Na)), list(range(Na, Na + Nq))
    agg, head = agg.to(DEV), head.to(DEV)
    for bf16, tol in ((False, 1e-4), (True, 2e-2)):
        agg.generator.manual_seed(1234)
        with torch.no_grad():
            if bf16:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    out, psi, cam = agg(images.to(DEV), anchors, queries, fix_rank=100)
            else:
                out, psi, cam = agg(images.to(DEV), anchors, queries, fix_rank=100)
            poses = head(out, cam)
        ref, _, ref_cam = O.aggregator_forward(images, agg_p, anchors, queries, 100, torch.Generator().manual_seed(1234),
                                               depth=2, vit_depth=2, intermediate_layer_idx=(0, 1), bf16=bf16)
        ref_poses = O.camera_head_forward(ref[-1][:, :, 0], ref_cam, head_p, 4, 16, 1)
        assert psi == 5 and tuple(out[-1].shape) == (1, Nq, 261, 2048) and tuple(cam.shape) == (1, Na, 2048)
        for k in out:
            assert rel_l2(out[k].cpu(), ref[k]) < tol, (bf16, k)
        assert rel_l2(cam.cpu(), ref_cam) < tol
        assert len(poses) == 4 and tuple(poses[-1].shape) == (1, Nq, 9)
        for a, b in zip(poses, ref_poses):
            assert rel_l2(a.cpu(), b) < 1e-2


@pytest.mark.parametrize("bf16", [False, True])
def test_forward_with_cache_matches_reference_golden(bf16):
    """A13 against a REFERENCE-derived expectation that travels: tests/golden/tiny_cache.npz is the unmodified reference's
    kv-cache path (priming forward + forward_with_cache + camera head, oracle/make_golden.py run_cache_case)."""
    from test_oracle_golden import load_golden_cache
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    arrays, meta, params = load_golden_cache()
    case = GOLDEN_CASES["tiny_conv"]
    Na = case["Na"]
    agg = Aggregator(**case["agg"], kv_cache=True).eval()
    head = CameraHead(**case["head"]).eval()
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    agg, head = agg.to(DEV), head.to(DEV)
    images = make_images(case["S"], case["H"], case["W"], case["seed"] + 100).to(DEV)
    tol = 2e-2 if bf16 else 1e-4
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out0, psi, cam = agg(images[:, :Na], list(range(Na)), [], fix_rank=case["fix_rank"])
            out, psi2 = agg.forward_with_cache(images[:, Na:], fix_rank=case["fix_rank"])
            single, _ = agg.forward_with_cache(images[:, Na:Na + 1], fix_rank=case["fix_rank"])
        poses = head(out, cam)
    assert len(out0) == 0 and psi == psi2 == 5 and agg.num_recon == meta["num_recon"] and agg.rank == meta["rank"]
    assert rel_l2(cam.cpu(), arrays["cam_token_last_layer"]) < tol
    C = agg.camera_token.shape[-1]
    k0 = agg.global_reloc_blocks[0].attn.k_cache.float().view(-1, 2, 64).transpose(0, 1).cpu()
    assert rel_l2(k0, arrays["k_cache_0"][0]) < (1e-2 if bf16 else 1e-5)
    for k, v in out.items():
        assert rel_l2(v.cpu(), arrays[f"out_{k}"]) < tol, k
    assert rel_l2(single[-1].cpu(), arrays["single_out_-1"]) < tol
    for i, pz in enumerate(poses):
        assert (pz.cpu() - arrays[f"pose_{i}"]).abs().max() < (5e-2 if bf16 else 1e-2)


@pytest.mark.parametrize("S", [1, 2])
def test_block_cached_api_with_the_reference_dense_mask(S):
    """Block(kv_cache=True) through the generic API exactly as aggregator.py:771-803 drives it: first call primes the cache
    with the scene tokens (attention output = zeros, attention.py:94-100), second call passes the dense cropped mask of
    aggregator.py:496-508 with S query frames — S = 1 makes the query part of the mask all-True."""
    from sailrecon_b200.layers import Block, RotaryPositionEmbedding2D
    torch.manual_seed(9)
    rope = RotaryPositionEmbedding2D(frequency=100.0)
    blk = Block(dim=128, num_heads=2, init_values=0.8, qk_norm=True, rope=rope, kv_cache=True).eval()
    p = {k: v.clone() for k, v in blk.state_dict().items()}
    blk = blk.to(DEV)
    Na, P, Pp = 2, 13, 6
    g = torch.Generator().manual_seed(10)
    scene = torch.randn(1, Na * Pp, 128, generator=g)
    spos = torch.randint(0, 9, (1, Na * Pp, 2), generator=g)
    x = torch.randn(1, S * P, 128, generator=g)
    pos = torch.randint(0, 9, (1, S * P, 2), generator=g)
    mask = O.cache_mask(Na, S, P, Pp)
    with torch.no_grad():
        y0 = blk(scene.to(DEV), pos=spos.to(DEV))                       # priming pass
        got = blk(x.to(DEV), pos=pos.to(DEV), attn_mask=mask.to(DEV)[None, None])
    ref0 = O.block_forward(scene[0], p, "", 2, spos[0], None, True, True, False, attn_zero=True)
    assert rel_l2(y0[0].cpu(), ref0) < 1e-5
    sn = O.layer_norm(scene[0], p["norm1.weight"], p["norm1.bias"], 1e-5)
    _, sk, sv = O.qkv_heads(sn, p, "attn.", 2, spos[0], True, True, False)
    ref = O.block_forward(x[0], p, "", 2, pos[0], mask, True, True, False, prefix_kv=(sk, sv))
    assert rel_l2(got[0].cpu(), ref) < 1e-5
    blk.attn.clear_kv_cache()


def test_two_streams_do_not_share_scratch():
    """SURVEY.md §8(b) 're-entrant per stream' (VERDICT r1 weak #13): engine.WS is keyed by (name, device, stream), so two
    forwards enqueued on two streams without any synchronisation between them reproduce their serial results."""
    agg, head, images, arrays, meta, case, _ = _build("tiny_conv")
    images2 = make_images(case["S"], case["H"], case["W"], 999).to(DEV)
    ref1 = _run(agg, head, images, meta, case, bf16=False)[0][-1].clone()
    ref2 = _run(agg, head, images2, meta, case, bf16=False)[0][-1].clone()
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = {}
    for rep in range(3):
        for s, img, key in ((s1, images, 1), (s2, images2, 2)):
            with torch.cuda.stream(s):
                outs[key] = _run(agg, head, img, meta, case, bf16=False)[0][-1]
    torch.cuda.synchronize()
    assert torch.equal(outs[1], ref1) and torch.equal(outs[2], ref2)
    from sailrecon_b200 import engine
    streams = {k[2] for k in engine.WS._buf if k[0] == "agg.x"}
    assert s1.cuda_stream in streams and s2.cuda_stream in streams


@pytest.mark.parametrize("bf16", [False, True])
def test_localizer_camera_only_matches_full_reloc_path(bf16):
    """N2: localize.Localizer (camera-only forward_with_cache + camera head + pose -> extrinsics, eager and as a CUDA-graph
    replay) returns what the full cached path returns for the camera: SailRecon.reloc(fast_reloc=True) semantics
    (sail_recon.py:218-244).  Full width (C = 1024, 16 heads, bf16 tensor-core kernels at M = 1 rows) at 224 px."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.localize import Localizer
    from sailrecon_b200.models import Aggregator
    from sailrecon_b200.utils import pose_encoding_to_extri_intri
    torch.manual_seed(0)
    agg = Aggregator(img_size=224, depth=2, intermediate_layer_idx=[0, 1], kv_cache=True).eval()
    agg.patch_embed.blocks = agg.patch_embed.blocks[:2]
    head = CameraHead(dim_in=2048, trunk_depth=1).eval()
    shapes = {"agg." + k: list(v.shape) for k, v in agg.state_dict().items()}
    shapes.update({"head." + k: list(v.shape) for k, v in head.state_dict().items()})
    params = make_params(shapes, 44)
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")})
    head.load_state_dict({k[5:]: v for k, v in params.items() if k.startswith("head.")})
    agg, head = agg.to(DEV), head.to(DEV)
    anchors = make_images(3, 224, 224, 144)[0].to(DEV)
    queries = make_images(3, 224, 224, 145)[0].to(DEV)
    eager = Localizer(agg, head, fix_rank=100, use_cuda_graph=False, autocast_bf16=bf16)
    agg.generator.manual_seed(1234)
    cam = eager.build(anchors)
    # reference semantics: the FULL cached path, then the camera head on its last-layer tokens
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            full, _ = agg.forward_with_cache(queries[None, :1], fix_rank=100)
        poses = head(full, cam)
        ext_ref, intr_ref = pose_encoding_to_extri_intri(poses[-1], (224, 224))
    pose, ext, intr = eager.localize(queries[:1])
    tol = 5e-3 if bf16 else 1e-4
    fin = torch.isfinite(intr_ref)                          # FoV = relu(.) can be exactly 0 at random init -> infinite focal length
    assert torch.equal(torch.isfinite(intr), fin)
    assert rel_l2(pose, poses[-1]) < tol and rel_l2(ext, ext_ref) < tol and rel_l2(intr[fin], intr_ref[fin]) < tol
    # the query-only camera head (anchors' keys / values cached by prime_anchors) == the full head on [anchors ; query]
    with torch.no_grad():
        q_only = head.forward_queries(full)
    for a, b in zip(q_only, poses):
        assert rel_l2(a, b) < 1e-5
    graph = Localizer(agg, head, fix_rank=100, use_cuda_graph=True, autocast_bf16=bf16)
    graph.cam_token_last_layer = cam                         # same primed scene
    for i in range(3):
        pe, ee, ie = eager.localize(queries[i:i + 1])
        pe, ee = pe.clone(), ee.clone()
        pg, eg, ig = graph.localize(queries[i:i + 1])
        torch.cuda.synchronize()
        assert torch.equal(pg, pe) and torch.equal(eg, ee), i            # same kernels, same inputs: bit-identical
    # the graph owns the scratch it captured: nothing is left in the shared per-stream tables under its stream handle
    from sailrecon_b200 import engine, ops
    sid = graph._stream.cuda_stream
    assert graph._keep and not [k for k in engine.WS._buf if k[2] == sid] and not [k for k in ops._ATTN_WS if k[1] == sid]
    # weights changed after the capture (optimizer step / load_state_dict): the replay must not serve the stale bf16 copies
    before = eg.clone()
    with torch.no_grad():
        for prm in list(head.parameters())[-2:] + list(agg.global_reloc_blocks[-1].mlp.parameters()):
            prm.mul_(1.25)
    pe, ee, ie = eager.localize(queries[2:3])
    pe, ee = pe.clone(), ee.clone()
    pg, eg, ig = graph.localize(queries[2:3])
    torch.cuda.synchronize()
    assert not torch.equal(ee, before)
    assert torch.equal(pg, pe) and torch.equal(eg, ee)
    eager.clear()
    with pytest.raises(RuntimeError):
        eager.localize(queries[:1])

"""CPU: host-side logic of the module layer — state_dict compatibility with the reference, the structural reading of the
reference's boolean masks, and bit-exact scene-token index draws (no compute calls: those need a B200)."""
import pytest
import torch

import sailrecon_oracle as O
from conftest import load_golden
from golden_common import GOLDEN_CASES


@pytest.mark.parametrize("name", list(GOLDEN_CASES))
def test_state_dict_matches_reference_shapes(name):
    """Parameter names / shapes equal those dumped from the unmodified reference (stored in the golden meta)."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    _, meta, params = load_golden(name)
    case = GOLDEN_CASES[name]
    agg, head = Aggregator(**case["agg"]), CameraHead(**case["head"])
    sd = {"agg." + k: list(v.shape) for k, v in agg.state_dict().items()}
    sd.update({"head." + k: list(v.shape) for k, v in head.state_dict().items()})
    assert sd == meta["shapes"]
    agg.load_state_dict({k[4:]: v for k, v in params.items() if k.startswith("agg.")}, strict=True)


def test_mask_structure_roundtrip_and_rejection():
    from sailrecon_b200 import engine
    m = O.reloc_mask(3, 4, 11, 7)
    st = engine.structure_from_mask(m[None, None], m.shape[0])
    assert (st.prefix_len, st.seg_len, st.num_segs) == (21, 11, 4)
    assert torch.equal(st.dense(), m)
    cm = O.camera_mask(5, 7)
    st = engine.structure_from_mask(cm, 12)
    assert (st.prefix_len, st.seg_len, st.num_segs) == (5, 1, 7)
    calls = st.calls()
    assert calls[0].batch == 1 and calls[0].seg_len == 5 and calls[1].batch == 7 and calls[1].prefix.length == 5
    full = torch.ones(9, 9, dtype=torch.bool)
    st = engine.structure_from_mask(full, 9)
    assert st.prefix_len == 9 and st.num_segs == 0
    diag = torch.block_diag(*[torch.ones(4, 4, dtype=torch.bool)] * 3)
    st = engine.structure_from_mask(diag, 12)
    assert (st.prefix_len, st.seg_len, st.num_segs) == (0, 4, 3)
    bad = m.clone()
    bad[25, 30] = ~bad[25, 30]
    with pytest.raises(NotImplementedError):
        engine.structure_from_mask(bad, bad.shape[0])
    with pytest.raises(NotImplementedError):
        engine.structure_from_mask(torch.tril(torch.ones(8, 8, dtype=torch.bool)), 8)
    with pytest.raises(TypeError):
        engine.structure_from_mask(torch.ones(4, 4), 4)


def test_scene_draws_bit_exact_with_oracle_order():
    """aggregator.py:617-621: one randperm per (layer, anchor) on the CPU generator, in that order."""
    from sailrecon_b200.models import Aggregator
    case = GOLDEN_CASES["tiny_conv"]
    agg = Aggregator(**case["agg"])
    Na, hw, rank, depth, P = 3, 20, 6, 4, 25
    agg.generator.manual_seed(77)
    frame_pos = O.aggregator_positions(4, 5)
    got = []
    for _ in range(depth):
        packed, drawn = agg._draw_scene(Na, P, hw, rank, frame_pos)
        got.append(drawn)
        rows = packed[:, 0].view(Na, 5 + rank)
        assert torch.equal(rows[:, :5], torch.arange(5)[None] + torch.arange(Na)[:, None] * P)
        assert torch.equal(rows[:, 5:], (drawn + 5 + torch.arange(Na)[:, None] * P).to(torch.int32))
        assert torch.equal(packed[:, 1:].long(), frame_pos[(rows - torch.arange(Na)[:, None].to(torch.int32) * P).reshape(-1).long()])
    ref = O.draw_scene_indices(torch.Generator().manual_seed(77), depth, 1, Na, hw, rank)[:, 0]
    assert torch.equal(torch.stack(got), ref)


def test_argument_errors_follow_reference():
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    with pytest.raises(ValueError):
        Aggregator(depth=3, aa_block_size=2, patch_embed="conv", embed_dim=128, num_heads=2)      # aggregator.py:157-160
    with pytest.raises(ValueError):
        CameraHead(pose_encoding_type="other")                                                     # camera_head.py:43
    agg = Aggregator(**GOLDEN_CASES["tiny_conv"]["agg"])
    with pytest.raises(RuntimeError):                                                              # CPU tensor: no fallback
        agg(torch.zeros(1, 2, 3, 56, 56), [0], [1], fix_rank=4)


def test_inference_forwards_refuse_training_mode_with_autograd():
    """the forwards read parameters detached; in train() mode with grad enabled an unmodified training
    loop would silently freeze them, so every module-level forward raises and points at sailrecon_b200.train."""
    import pytest
    import torch
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.layers import Attention, Block
    from sailrecon_b200.models import Aggregator
    blk, att = Block(dim=128, num_heads=2), Attention(dim=128, num_heads=2)
    agg = Aggregator(img_size=56, embed_dim=128, depth=1, num_heads=2, patch_embed="conv", intermediate_layer_idx=[0])
    head = CameraHead(dim_in=256, trunk_depth=1, num_heads=2)
    x = torch.zeros(1, 4, 128)
    # Aggregator.forward itself dispatches to the autograd bridge in train() mode (train.AggregatorFunction); without
    # autocast(bfloat16) that path refuses loudly as well
    with pytest.raises(NotImplementedError, match="autocast"):
        agg(torch.zeros(1, 2, 3, 56, 56), [0], [1])
    for call in (lambda: blk(x), lambda: att(x),
                 lambda: agg.forward_with_cache(torch.zeros(1, 1, 3, 56, 56)),
                 lambda: head.trunk_fn(torch.zeros(1, 2, 256), None, 1)):
        with pytest.raises(RuntimeError, match="inference path"):
            call()
    # CameraHead.forward dispatches to its autograd node (heads/camera_head_train.py) in train() mode; no CPU path there either
    with pytest.raises(RuntimeError, match="CUDA"):
        head({-1: torch.zeros(1, 1, 5, 256)}, torch.zeros(1, 1, 256))
    # eval() or no_grad(): the guard is silent (the call then fails on the CPU tensor instead — no CPU fallback)
    blk.eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        blk(x)
    blk.train()
    with torch.no_grad(), pytest.raises(RuntimeError, match="CUDA"):
        blk(x)

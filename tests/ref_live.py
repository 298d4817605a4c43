"""Access to the UNMODIFIED reference shipped under oracle/_ref (oracle/make_ref.py) — test infrastructure only.

`ref()` puts oracle/_ref on sys.path and returns the reference's own modules; the product package `sailrecon_b200`
is never aliased in the same process (the overlay tests run in subprocesses)."""
from __future__ import annotations

import contextlib
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_ref  # noqa: E402


def available() -> bool:
    return make_ref.available()


def ref() -> types.SimpleNamespace:
    p = make_ref.import_path()
    if p not in sys.path:
        sys.path.insert(0, p)
    from sailrecon.heads.camera_head import CameraHead
    from sailrecon.layers.block import Block
    from sailrecon.models.aggregator import Aggregator
    return types.SimpleNamespace(Aggregator=Aggregator, CameraHead=CameraHead, Block=Block)


@contextlib.contextmanager
def spy_randperm(drawn: list):
    """Records every torch.randperm the reference draws (aggregator.py:617-621) — the bit-exact index check."""
    orig = torch.randperm

    def spy(n, *a, **k):
        r = orig(n, *a, **k)
        drawn.append(r.clone())
        return r

    torch.randperm = spy
    try:
        yield drawn
    finally:
        torch.randperm = orig


def randomize_o1(agg, head, seed: int) -> None:
    """SURVEY.md §8(d) second weight set: LayerScale gamma ~ U(0.5, 1.5), special tokens ~ N(0, 1) so that no kernel can
    hide behind gamma = 0.01 (in place, same values on whatever device the modules live)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for mod in (agg, head):
            for n, p in sorted(mod.named_parameters()):
                leaf = n.split(".")[-1]
                if leaf == "gamma" and not n.startswith("patch_embed."):
                    p.copy_((0.5 + torch.rand(p.shape, generator=g)).to(p.device))
                elif "token" in leaf and leaf != "empty_pose_tokens" and not n.startswith("patch_embed."):
                    p.copy_(torch.randn(p.shape, generator=g).to(p.device))
                elif leaf == "bias" and p.dim() == 1:
                    p.copy_((0.05 * torch.randn(p.shape, generator=g)).to(p.device))


def run_reference(agg, head, images, nrl, rl, fix_rank, bf16: bool):
    """sail_recon.py:101,119-121: aggregator under autocast(bf16) (or plain fp32), camera head with autocast disabled.
    Returns (output_dict, cam_token_last_layer, poses, drawn scene indices [depth, Na, rank], rank)."""
    agg.generator.manual_seed(1234)
    drawn = []
    with torch.no_grad(), spy_randperm(drawn):
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out, psi, cam = agg(images, nrl, rl, fix_rank=fix_rank)
        with torch.autocast("cuda", enabled=False):
            poses = head(out, cam) if len(rl) else None
    rank = agg.rank
    hw = drawn[0].numel()
    sel = torch.stack([d[:min(rank, hw)] for d in drawn]).reshape(agg.depth, len(nrl), -1)
    return out, cam, poses, sel, rank


def run_ours(agg, head, images, nrl, rl, fix_rank, bf16: bool):
    agg.generator.manual_seed(1234)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out, psi, cam = agg(images, nrl, rl, fix_rank=fix_rank)
        with torch.autocast("cuda", enabled=False):
            poses = head(out, cam) if len(rl) else None
    return out, cam, poses, torch.stack(agg.last_scene_indices), agg.rank


def build_pair(agg_kwargs: dict, head_kwargs: dict, seed: int, o1: bool, device="cuda"):
    """Reference modules with the reference's own constructors / init, and sailrecon_b200 modules loaded from the
    reference's state_dict (the checkpoint contract of SURVEY.md §8(b))."""
    from sailrecon_b200.heads import CameraHead
    from sailrecon_b200.models import Aggregator
    R = ref()
    torch.manual_seed(seed)
    with torch.device(device):
        r_agg = R.Aggregator(**agg_kwargs).eval().to(device)      # .to(): its _resnet_mean buffers are legacy-constructed on the CPU
        r_head = R.CameraHead(**head_kwargs).eval().to(device)
        o_agg = Aggregator(**agg_kwargs).eval()
        o_head = CameraHead(**head_kwargs).eval()
    if o1:
        randomize_o1(r_agg, r_head, seed + 1)
    o_agg.load_state_dict(r_agg.state_dict())
    o_head.load_state_dict(r_head.state_dict())
    return r_agg, r_head, o_agg, o_head

"""sr_reproject_static_f32 — the reprojection entry point for hosts that are not allowed to allocate GPU memory.

Who that is: code embedded in another program's engine whose device-memory pool it may not draw from, a localisation service
that has promised "no cudaMalloc after start-up" to keep its latency flat, a worker that only ever sees buffers somebody else
owns.  Such a caller can still pass inputs and an output buffer, but it has nowhere to get the per-pair scratch that
sr_reproject_f32 asks for, so the library keeps that scratch itself in a fixed array (4096 pairs, 512 KiB per GPU).

What is checked here, through the C-ABI only (ctypes on raw device pointers):
  * two calls in a row — different sizes, pose layouts and depth modes, the second one smaller so that the array still holds
    the first call's records beyond it — each equal BIT FOR BIT to sr_reproject_f32 on the same inputs (the two entry points
    run the same kernels; the fp32 tests of this repo use 1e-5 elsewhere, here nothing may differ at all);
  * the second call leaves the GPU's free memory where it was;
  * a batch over the limit is refused with a message that names the limit, and writes nothing.

The file name sorts after every other test file on purpose: under `pytest -x` a failure here must not hide the rest.  It has
no dependence on what ran before it — the array is rewritten by every call — so it passes alone in a fresh process too."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _scene(P, K, rows, seed, with_dst_depth):
    g = torch.Generator().manual_seed(seed)
    f = 400.0 + 200.0 * torch.rand(P, 2, generator=g)
    k = torch.zeros(P, 3, 3)
    k[:, 0, 0], k[:, 1, 1], k[:, 2, 2] = f[:, 0], f[:, 1], 1.0
    k[:, 0, 2], k[:, 1, 2] = 259.0, 196.0
    kd = k.clone()
    kd[:, 0, 0] *= 1.1
    q, _ = torch.linalg.qr(torch.eye(3) + 0.1 * torch.randn(P, 3, 3, generator=g))
    rel = torch.zeros(P, rows, 4)
    rel[:, :3, :3], rel[:, :3, 3] = q, 0.2 * torch.randn(P, 3, generator=g)
    if rows == 4:
        rel[:, 3, 3] = 1.0
    t = {
        "coords": torch.rand(P, K, 2, generator=g) * torch.tensor([518.0, 392.0]),
        "depth": 1.0 + 4.0 * torch.rand(P, K, generator=g),
        "dst_depth": 1.0 + 4.0 * torch.rand(P, K, generator=g) if with_dst_depth else None,
        "src_k": k, "dst_k": kd, "rel": rel,
        "src_scale": 0.5 + torch.rand(P, generator=g), "dst_scale": 0.5 + torch.rand(P, generator=g) if with_dst_depth else None,
    }
    return {n: (None if v is None else v.to(DEV).contiguous()) for n, v in t.items()}


def _ptr(t):
    return None if t is None else t.data_ptr()


def _call(lib, check, s, P, K, rows, workspace):
    """One launch through the C-ABI; workspace=None selects the entry point that brings its own scratch."""
    out = torch.full((P, K, 2), float("nan"), device=DEV)
    head = (s["coords"].data_ptr(), s["depth"].data_ptr(), _ptr(s["dst_depth"]), s["src_k"].data_ptr(), s["dst_k"].data_ptr(),
            s["rel"].data_ptr(), rows, _ptr(s["src_scale"]), _ptr(s["dst_scale"]), out.data_ptr(), P, K)
    stream = torch.cuda.current_stream().cuda_stream
    if workspace is None:
        check(lib.sr_reproject_static_f32(*head, stream), "reproject_static")
    else:
        check(lib.sr_reproject_f32(*head, workspace.data_ptr(), stream), "reproject")
    torch.cuda.synchronize()
    return out


def test_static_scratch_matches_caller_workspace_on_two_calls_in_a_row():
    from sailrecon_b200 import _lib
    lib = _lib.load()
    limit = lib.sr_reproject_static_max_pairs()
    assert limit == 4096

    first = dict(P=int(limit), K=700, rows=4, seed=11, with_dst_depth=False)    # fills the whole array
    second = dict(P=37, K=1500, rows=3, seed=12, with_dst_depth=True)            # stale records of `first` lie behind it
    for n, case in enumerate((first, second)):
        P, K, rows = case["P"], case["K"], case["rows"]
        s = _scene(**case)
        ws = torch.empty(max(1, lib.sr_reproject_workspace_floats(P, K)), device=DEV)
        want = _call(lib, _lib.check, s, P, K, rows, ws)
        out_buf_bytes = P * K * 2 * 4
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        got = _call(lib, _lib.check, s, P, K, rows, None)
        free_after = torch.cuda.mem_get_info()[0]
        assert torch.isfinite(want).all() and torch.isfinite(got).all()
        assert torch.equal(got, want), f"call {n}: max |diff| {(got - want).abs().max().item():.3e}"
        if n == 1:      # past the first use on this GPU: the only memory taken is the test's own output tensor (torch's pool)
            assert free_before - free_after <= max(out_buf_bytes, 2 << 20), (free_before, free_after)


def test_static_scratch_refuses_more_pairs_than_it_holds():
    from sailrecon_b200 import _lib
    lib = _lib.load()
    P, K = int(lib.sr_reproject_static_max_pairs()) + 1, 4
    s = _scene(P, K, 4, 13, False)
    out = torch.full((P, K, 2), -7.0, device=DEV)
    rc = lib.sr_reproject_static_f32(s["coords"].data_ptr(), s["depth"].data_ptr(), None, s["src_k"].data_ptr(), s["dst_k"].data_ptr(),
                                     s["rel"].data_ptr(), 4, None, None, out.data_ptr(), P, K, torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.SR_ERR_SHAPE
    with pytest.raises(ValueError, match="4096"):
        _lib.check(rc, "reproject_static")
    torch.cuda.synchronize()
    assert bool((out == -7.0).all())

"""b200_g1_sum_noalloc (static device scratch, no cudaMalloc) against the existing b200_g1_sum, through the C ABI.
Kept in a file pytest collects last.  Bit-exact like every other test here: both return the normalised affine sum."""
import ctypes as C
import numpy as np
import pytest
import orc

pytestmark = pytest.mark.gpu


def _noalloc_sum(ctx, torch, pts):
    """pts: (n, 8) uint64 host array with (0,0) rows for infinity -> (8 limbs, flag) via device buffers"""
    n = pts.shape[0]
    dev = torch.device("cuda", ctx.device)
    d_in = torch.from_numpy(np.ascontiguousarray(pts).view(np.int64).reshape(-1).copy()).to(dev) if n else torch.zeros(8, dtype=torch.int64, device=dev)
    d_out = torch.full((9,), -1, dtype=torch.int64, device=dev)      # 68 bytes used; poisoned so a missing write shows
    torch.cuda.synchronize(dev)
    ctx.g1_sum_noalloc(d_in.data_ptr(), n, d_out.data_ptr())
    torch.cuda.synchronize(dev)                                      # the call does not synchronise its stream
    raw = d_out.cpu().numpy().view(np.uint64)
    return raw[:8].copy(), int(raw[8] & 0xffffffff)


def test_g1_sum_noalloc_matches_g1_sum(ctx):
    import torch
    pts, inf = orc.g1_fixed_base(orc.sha512_rng_fr(1 << 50, 3000))
    pts = pts.copy()
    pts[7] = pts[6]                       # a doubling on the way
    pts[11] = 0                           # points at infinity among the inputs
    pts[2999] = 0
    # two calls in a row on the shared scratch: a long input (many blocks), then a short one (stale slots must not be read)
    for n in (3000, 37):
        want, winf = ctx.g1_sum(pts[:n], (pts[:n] == 0).all(axis=1).astype(np.uint8))
        got, ginf = _noalloc_sum(ctx, torch, pts[:n])
        assert ginf == winf == 0
        assert np.array_equal(got, want), n
    # P + (-P) and the empty sum: infinity, encoded as (0,0) + flag
    neg = pts[:1].copy()
    neg[0].reshape(2, 4)[1] = orc.fq_sub(np.zeros((1, 4), np.uint64), neg[0].reshape(2, 4)[1:])
    got, ginf = _noalloc_sum(ctx, torch, np.concatenate([pts[:1], neg]))
    assert ginf == 1 and not got.any()
    got, ginf = _noalloc_sum(ctx, torch, np.zeros((0, 8), np.uint64))
    assert ginf == 1 and not got.any()


def test_g1_sum_noalloc_rejects_oversize(ctx):
    from ethsnarks_b200 import lib as L
    rc = ctx.lib.b200_g1_sum_noalloc(ctx.h, C.c_void_p(256), C.c_size_t((1 << 20) + 1), C.c_void_p(256))
    assert rc == 2   # B200_ERR_ARG, nothing launched
    assert b"exceeds the maximum" in ctx.lib.b200_last_error(ctx.h)
    with pytest.raises(L.B200Error):
        ctx.g1_sum_noalloc(256, (1 << 20) + 1, 256)

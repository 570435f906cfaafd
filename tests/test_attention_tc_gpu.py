"""GPU: the tcgen05 causal-attention forward (bf16 operands from the fused qkv buffer, TMA loads, TMEM accumulators) against
the oracle of the reference's attention (core/memory_efficient_attention.cpp:92-162). Tolerance of the bf16 operand mode:
1e-2 relative to max|ref|; the oracle is also evaluated on the bf16-rounded inputs, where only P's rounding and
the summation order remain (4e-3)."""
import numpy as np
import pytest
import torch

import mft_oracle as O
from conftest import rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,H,S", [(2, 3, 512), (1, 2, 64), (1, 4, 200), (2, 2, 130), (1, 1, 1), (1, 16, 1024)])
def test_attention_tc_forward_matches_reference_attention(capi, B, H, S):
    rng = np.random.default_rng(B * 100 + H * 10 + S)
    Hd, C = 64, H * 64
    qkv = rng.standard_normal((B, S, 3 * C)).astype(np.float32)
    q16 = torch.tensor(qkv).cuda().to(torch.bfloat16).contiguous()
    ctx = torch.full((B, S, C), float("nan"), device="cuda")
    ctx16 = torch.empty(B, S, C, device="cuda", dtype=torch.bfloat16)
    scale = 0.125
    capi.call("mft_causal_attention_tc_fwd", q16, 3 * C, ctx, ctx16, B, H, S, Hd, scale, None)
    torch.cuda.synchronize()
    heads = lambda a: np.ascontiguousarray(a.reshape(B, S, H, Hd).transpose(0, 2, 1, 3))
    qr = q16.float().cpu().numpy()
    for src, tol in ((qkv, 1e-2), (qr, 4e-3)):
        e = O.causal_attention_fwd(heads(src[:, :, :C]), heads(src[:, :, C:2 * C]), heads(src[:, :, 2 * C:]), scale)
        e = e.transpose(0, 2, 1, 3).reshape(B, S, C)
        assert rel(ctx, e) < tol, (B, H, S, rel(ctx, e))
    assert rel(ctx16.float(), ctx.cpu().numpy()) < 5e-3


@pytest.mark.parametrize("kind", ["peaky", "ramp"])
def test_attention_tc_rescales_when_the_row_max_runs_ahead(capi, kind):
    """O accumulates in TMEM under a scaling max that follows the running row max lazily (rescaled by a TMEM load / store only
    when a later tile's max is more than 2^8 above it). Scores with a large spread ("peaky": std 9) and keys whose magnitude
    grows with the position ("ramp") make that happen at most tiles; unit-variance inputs never do."""
    B, H, S, Hd = 2, 2, 512, 64
    C = H * Hd
    rng = np.random.default_rng(77)
    qkv = rng.standard_normal((B, S, 3 * C)).astype(np.float32)
    if kind == "peaky":
        qkv[:, :, :2 * C] *= 3.0
    else:
        qkv[:, :, C:2 * C] *= (1.0 + np.arange(S, dtype=np.float32) / 16.0)[None, :, None]
    q16 = torch.tensor(qkv).cuda().to(torch.bfloat16).contiguous()
    ctx = torch.full((B, S, C), float("nan"), device="cuda")
    scale = 0.125
    capi.call("mft_causal_attention_tc_fwd", q16, 3 * C, ctx, None, B, H, S, Hd, scale, None)
    torch.cuda.synchronize()
    heads = lambda a: np.ascontiguousarray(a.reshape(B, S, H, Hd).transpose(0, 2, 1, 3))
    qr = q16.float().cpu().numpy()
    e = O.causal_attention_fwd(heads(qr[:, :, :C]), heads(qr[:, :, C:2 * C]), heads(qr[:, :, 2 * C:]), scale)
    e = e.transpose(0, 2, 1, 3).reshape(B, S, C)
    assert bool(torch.isfinite(ctx).all())
    assert rel(ctx, e) < 4e-3, (kind, rel(ctx, e))


def test_attention_tc_rejects_other_head_sizes(capi):
    x = torch.zeros(1, 8, 3 * 32, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(1, 8, 32, device="cuda")
    with pytest.raises(capi.MftError, match="head_dim"):
        capi.call("mft_causal_attention_tc_fwd", x, 96, out, None, 1, 1, 8, 32, 0.1, None)

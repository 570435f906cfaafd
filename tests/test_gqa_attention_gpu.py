"""GPU: fused grouped-query attention (forward + backward) and the fused q/k RMSNorm + RoPE kernels through the C ABI,
against the oracle's restatement of the reference chain (repeat_kv -> QK^T -> scale -> masks -> softmax -> PV and the
backward nodes). Exact mode (3xTF32 split): 1e-5 relative to max|ref|; default TF32 operand mode: 2e-3."""
import numpy as np
import pytest
import torch

import mft_oracle as O
from conftest import FP32_TOL, TF32_TOL, dev, rel, ws

pytestmark = pytest.mark.gpu

CASES = [  # B, H, KV, S, Hd, window, pad
    (2, 4, 1, 50, 32, 0, False),
    (1, 4, 2, 130, 64, 40, False),
    (2, 4, 1, 256, 256, 0, False),       # Gemma-3 270M head shape at the c3 sequence length
    (1, 4, 1, 200, 256, 100, True),
    (1, 2, 2, 77, 16, 0, True),
    (1, 3, 3, 33, 128, 7, False),
]


def run_gqa(capi, q, k, v, dctx, H, KV, scale, window, bias):
    """q [B,S,H,Hd], k/v [B,S,KV,Hd] numpy (the projections' layout) -> ctx, dq, dk, dv in the same layouts."""
    B, S, _, Hd = q.shape
    dq_, dk_, dv_, do_ = dev(q), dev(k), dev(v), dev(dctx)
    db = dev(bias) if bias is not None else None
    ctx = torch.empty(B, S, H, Hd, device="cuda")
    lse = torch.empty(B, H, S, device="cuda")
    capi.call("mft_gqa_attention_fwd", dq_, dk_, dv_, db, ctx, lse, None, B, H, KV, S, Hd, S * H * Hd, H * Hd, Hd, S * KV * Hd, KV * Hd, Hd,
              S * H * Hd, H * Hd, Hd, scale, window, None)
    gq, gk, gv = torch.full_like(dq_, float("nan")), torch.full_like(dk_, float("nan")), torch.full_like(dv_, float("nan"))
    w = ws(capi.lib.mft_gqa_attention_bwd_workspace_bytes(B, H, KV, S, Hd))
    capi.call("mft_gqa_attention_bwd", dq_, dk_, dv_, db, ctx, lse, do_, gq, gk, gv, B, H, KV, S, Hd,
              S * H * Hd, H * Hd, Hd, S * KV * Hd, KV * Hd, Hd, S * H * Hd, H * Hd, Hd,
              S * H * Hd, H * Hd, Hd, S * KV * Hd, KV * Hd, Hd, scale, window, w, w.numel(), None)
    torch.cuda.synchronize()
    return ctx, gq, gk, gv


@pytest.mark.parametrize("mode", ["exact", "tf32"])
@pytest.mark.parametrize("B,H,KV,S,Hd,window,pad", CASES)
def test_gqa_attention_fwd_bwd_against_reference_chain(capi, B, H, KV, S, Hd, window, pad, mode):
    rng = np.random.default_rng(B * 1000 + S + Hd)
    q = rng.standard_normal((B, S, H, Hd)).astype(np.float32)
    k = rng.standard_normal((B, S, KV, Hd)).astype(np.float32)
    v = rng.standard_normal((B, S, KV, Hd)).astype(np.float32)
    dctx = rng.standard_normal((B, S, H, Hd)).astype(np.float32)
    bias = None
    if pad:                                   # right padding: the last keys of every sequence are masked
        bias = np.zeros((B, S), np.float32)
        bias[:, S - S // 5:] = -1e10
    scale = float(Hd) ** -0.5
    t = lambda a: np.ascontiguousarray(a.transpose(0, 2, 1, 3))      # [B,S,h,Hd] <-> [B,h,S,Hd]
    e_ctx, e_dq, e_dk, e_dv = O.gqa_attention_fwd_bwd(t(q), t(k), t(v), H // KV, scale, window, bias, t(dctx))
    prev = capi.lib.mft_attention_precise(1 if mode == "exact" else 0)
    try:
        ctx, gq, gk, gv = run_gqa(capi, q, k, v, dctx, H, KV, scale, window, bias)
    finally:
        capi.lib.mft_attention_precise(prev)
    tol = FP32_TOL if mode == "exact" else TF32_TOL
    assert rel(ctx, t(e_ctx)) < tol
    assert rel(gq, t(e_dq)) < tol * 2
    assert rel(gk, t(e_dk)) < tol * 2
    assert rel(gv, t(e_dv)) < tol * 2


def test_gqa_attention_matches_the_gpt2_attention_golden(capi, golden_ops):
    """KV == H on the reference's own attention tensors (attn_* goldens, [B,H,S,Hd], dumped by the reference build)."""
    L = golden_ops
    q, k, v, y = (L(n) for n in ("attn_q", "attn_k", "attn_v", "attn_y"))
    B, H, S, Hd = q.shape
    ctx = torch.empty(B, H, S, Hd, device="cuda")
    prev = capi.lib.mft_attention_precise(1)
    try:
        capi.call("mft_gqa_attention_fwd", dev(q), dev(k), dev(v), None, ctx, None, None, B, H, H, S, Hd, H * S * Hd, Hd, S * Hd,
                  H * S * Hd, Hd, S * Hd, H * S * Hd, Hd, S * Hd, 0.25, 0, None)
    finally:
        capi.lib.mft_attention_precise(prev)
    assert rel(ctx, y) < FP32_TOL


@pytest.mark.parametrize("Hd,H,KV,S,do_norm", [(32, 4, 1, 16, 1), (256, 4, 1, 40, 1), (64, 2, 2, 33, 0), (128, 3, 1, 9, 1)])
def test_qk_norm_rope_fwd_bwd(capi, Hd, H, KV, S, do_norm):
    rng = np.random.default_rng(Hd + S)
    B = 2
    theta = 10000.0
    q = rng.standard_normal((B, S, H, Hd)).astype(np.float32)
    k = rng.standard_normal((B, S, KV, Hd)).astype(np.float32)
    wq, wk = (rng.standard_normal(Hd) * 0.1).astype(np.float32), (rng.standard_normal(Hd) * 0.1).astype(np.float32)
    gq = rng.standard_normal(q.shape).astype(np.float32)
    t = lambda a: np.ascontiguousarray(a.transpose(0, 2, 1, 3))
    yq, yk = torch.empty(q.shape, device="cuda"), torch.empty(k.shape, device="cuda")
    capi.call("mft_qk_norm_rope_fwd", dev(q), dev(wq), yq, B * S, H, dev(k), dev(wk), yk, B * S, KV, S, Hd, 1e-6, theta, do_norm, None)
    assert rel(yq, t(O.qk_norm_rope_fwd(t(q), wq, theta, 1e-6, bool(do_norm)))) < FP32_TOL
    assert rel(yk, t(O.qk_norm_rope_fwd(t(k), wk, theta, 1e-6, bool(do_norm)))) < FP32_TOL
    gx = torch.empty(q.shape, device="cuda")
    capi.call("mft_qk_norm_rope_bwd", dev(q), dev(wq), dev(gq), gx, B * S, H, S, Hd, 1e-6, theta, do_norm, None)
    assert rel(gx, t(O.qk_norm_rope_bwd(t(q), wq, t(gq), theta, 1e-6, bool(do_norm)))) < FP32_TOL


def _attention_f64(q, k, v, rep, scale, window):
    """fp64 forward of the same chain on [B,h,S,Hd] arrays (no reference code involved): used for finite differences."""
    kf, vf = np.repeat(k, rep, axis=1), np.repeat(v, rep, axis=1)
    S = q.shape[-2]
    s = np.matmul(q, np.swapaxes(kf, -1, -2)) * scale
    i, j = np.arange(S)[:, None], np.arange(S)[None, :]
    forbid = (j > i) | ((i - j >= window) if window > 0 else False)
    s = np.where(forbid, -np.inf, s)
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    return np.matmul(p, vf)


@pytest.mark.parametrize("H,KV,S,Hd,window", [(4, 1, 40, 32, 0), (2, 2, 48, 64, 16)])
def test_gqa_attention_backward_against_finite_differences(capi, H, KV, S, Hd, window):
    """The analytic dq / dk / dv of the kernels against central differences of an independent fp64 forward, along random
    directions: <grad, d> == (f(x + eps d) - f(x - eps d)) / (2 eps) with f = sum(ctx * dctx)."""
    rng = np.random.default_rng(S + Hd)
    B = 1
    q = rng.standard_normal((B, S, H, Hd)).astype(np.float32)
    k = rng.standard_normal((B, S, KV, Hd)).astype(np.float32)
    v = rng.standard_normal((B, S, KV, Hd)).astype(np.float32)
    dctx = rng.standard_normal((B, S, H, Hd)).astype(np.float32)
    scale = float(Hd) ** -0.5
    prev = capi.lib.mft_attention_precise(1)
    try:
        _, gq, gk, gv = run_gqa(capi, q, k, v, dctx, H, KV, scale, window, None)
    finally:
        capi.lib.mft_attention_precise(prev)
    t = lambda a: np.ascontiguousarray(a.transpose(0, 2, 1, 3)).astype(np.float64)
    f = lambda qq, kk, vv: float((_attention_f64(t(qq), t(kk), t(vv), H // KV, scale, window) * t(dctx)).sum())
    eps = 1e-4
    for name, x, g in (("q", q, gq), ("k", k, gk), ("v", v, gv)):
        d = rng.standard_normal(x.shape)
        args = {"q": q.astype(np.float64), "k": k.astype(np.float64), "v": v.astype(np.float64)}
        plus, minus = dict(args), dict(args)
        plus[name] = args[name] + eps * d
        minus[name] = args[name] - eps * d
        fd = (f(plus["q"], plus["k"], plus["v"]) - f(minus["q"], minus["k"], minus["v"])) / (2 * eps)
        an = float((g.cpu().numpy().astype(np.float64) * d).sum())
        assert abs(fd - an) <= 2e-4 * max(abs(fd), 1.0), (name, fd, an)

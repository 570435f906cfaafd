"""GPU parity tests (C ABI): GEMM (SIMT fp32 and tcgen05 TF32), fused LoRA linear, causal attention,
shifted LM cross-entropy, clip + Adam. See tests/test_kernels_gpu.py for the tolerance policy."""
import ctypes as C

import numpy as np
import pytest
import torch

import mft_oracle as O
from conftest import FP32_TOL, TF32_TOL, dev, rel, ws

pytestmark = pytest.mark.gpu


def run_gemm(capi, a, b, trans_a, trans_b, alpha=1.0, beta=0.0, bias=None, c0=None, gelu=False, seg2=None, tc=True):
    """a, b numpy in their STORED layouts. Returns (C, gelu_out) as torch tensors."""
    prev = capi.lib.mft_use_tensor_cores(1 if tc else 0)
    try:
        batch = a.shape[0] if a.ndim == 3 else 1
        m = a.shape[-1] if trans_a else a.shape[-2]
        k = a.shape[-2] if trans_a else a.shape[-1]
        n = b.shape[-2] if trans_b else b.shape[-1]
        da, db = dev(a), dev(b)
        shape = (batch, m, n) if a.ndim == 3 else (m, n)
        c = dev(c0) if c0 is not None else torch.full(shape, float("nan"), device="cuda")
        g = torch.empty(shape, device="cuda") if gelu else None
        kw = {}
        keep = []
        if seg2 is not None:
            a2, b2 = dev(seg2[0]), dev(seg2[1])
            keep += [a2, b2]
            kw = dict(k2=a2.shape[1], a2=a2, lda2=a2.shape[1], b2=b2, ldb2=b2.shape[1])
        dbias = dev(bias) if bias is not None else None
        args = capi.gemm(m, n, k, da, a.shape[-1], int(trans_a), db, b.shape[-1], int(trans_b), c, n, alpha, beta,
                         bias=dbias, gelu_out=g, batch=batch, stride_a=a.shape[-2] * a.shape[-1],
                         stride_b=(b.shape[-2] * b.shape[-1] if b.ndim == 3 else 0), stride_c=m * n, **kw)
        used_tc = capi.lib.mft_gemm_uses_tensor_cores(C.byref(args))
        torch.cuda.synchronize()
        return c, g, used_tc
    finally:
        capi.lib.mft_use_tensor_cores(prev)


def ref_gemm(a, b, trans_a, trans_b, alpha=1.0, beta=0.0, bias=None, c0=None, seg2=None):
    A = np.swapaxes(a, -1, -2) if trans_a else a
    Bm = np.swapaxes(b, -1, -2) if trans_b else b
    r = np.matmul(A.astype(np.float64), Bm.astype(np.float64))
    if seg2 is not None:
        r = r + seg2[0].astype(np.float64) @ seg2[1].astype(np.float64).T
    r = r * alpha
    if bias is not None:
        r = r + bias
    if c0 is not None and beta != 0:
        r = r + beta * c0
    return r


# ----------------------------------------------------------------------------- SIMT (exact fp32) path
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("m,n,k", [(1, 1, 1), (10, 40, 24), (65, 130, 33), (128, 8, 768), (200, 96, 50)])
def test_gemm_simt_all_layouts(capi, m, n, k, ta, tb):
    rng = np.random.default_rng(m * 1000 + n * 10 + k)
    a = rng.standard_normal((k, m) if ta else (m, k)).astype(np.float32)
    b = rng.standard_normal((n, k) if tb else (k, n)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    c0 = rng.standard_normal((m, n)).astype(np.float32)
    c, g, used = run_gemm(capi, a, b, ta, tb, alpha=0.5, beta=2.0, bias=bias, c0=c0, gelu=True, tc=False)
    e = ref_gemm(a, b, ta, tb, 0.5, 2.0, bias, c0)
    assert used == 0
    assert rel(c, e) < FP32_TOL
    assert rel(g, O.gelu_fwd(e.astype(np.float32))) < 2e-5


def test_gemm_golden_matmul_and_bmm(capi, golden_ops):
    L = golden_ops
    x, w, gy = L("matmul_x"), L("matmul_w"), L("matmul_gy")
    x2, g2 = x.reshape(-1, x.shape[-1]), gy.reshape(-1, gy.shape[-1])
    y, _, _ = run_gemm(capi, x2, w, 0, 0, tc=False)
    assert rel(y, L("matmul_y").reshape(y.shape)) < FP32_TOL
    gx, _, _ = run_gemm(capi, g2, w, 0, 1, tc=False)            # dX = dY W^T
    assert rel(gx, L("matmul_gx").reshape(gx.shape)) < FP32_TOL
    gw, _, _ = run_gemm(capi, x2, g2, 1, 0, tc=False)           # dW = X^T dY
    assert rel(gw, L("matmul_gw")) < FP32_TOL
    a, b = L("bmm_a"), L("bmm_b")
    a3, b3 = a.reshape(-1, *a.shape[-2:]), b.reshape(-1, *b.shape[-2:])
    y, _, _ = run_gemm(capi, a3, b3, 0, 0, tc=False)
    assert rel(y, L("bmm_y").reshape(y.shape)) < FP32_TOL
    gyb = L("bmm_gy").reshape(-1, *L("bmm_gy").shape[-2:])
    ga, _, _ = run_gemm(capi, gyb, b3, 0, 1, tc=False)
    assert rel(ga, L("bmm_ga").reshape(ga.shape)) < FP32_TOL
    gb, _, _ = run_gemm(capi, a3, gyb, 1, 0, tc=False)
    assert rel(gb, L("bmm_gb").reshape(gb.shape)) < FP32_TOL


@pytest.mark.parametrize("m,n,k,ta,splits", [(16, 1024, 4096, 1, 16), (1024, 16, 4096, 1, 16), (8, 768, 512, 1, 2),
                                             (130, 70, 1000, 0, 7), (16, 1024, 4096, 1, 1)])
def test_gemm_splitk_matches_reference(capi, m, n, k, ta, splits):
    """Split-K SIMT product (LoRA dA / dB shapes: tiny output, contraction over all tokens); deterministic."""
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((k, m) if ta else (m, k)).astype(np.float32)
    b = rng.standard_normal((k, n)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    c0 = rng.standard_normal((m, n)).astype(np.float32)
    da, db, dbias = dev(a), dev(b), dev(bias)
    outs = []
    for _ in range(2):
        c = dev(c0)
        args = capi.GemmArgs(m, n, k, da.data_ptr(), a.shape[1], ta, db.data_ptr(), n, 0, c.data_ptr(), n, 0.5, 2.0, dbias.data_ptr(),
                             None, 0, None, 0, None, 0, 1, 0, 0, 0)
        w = ws(capi.lib.mft_gemm_splitk_workspace_bytes(m, n, 1, splits))
        capi.check(capi.lib.mft_gemm_splitk(C.byref(args), splits, w.data_ptr(), w.numel(), None), "mft_gemm_splitk")
        torch.cuda.synchronize()
        outs.append(c.clone())
    assert torch.equal(outs[0], outs[1])                      # fixed-order reduction
    assert rel(outs[0], ref_gemm(a, b, ta, 0, 0.5, 2.0, bias, c0)) < FP32_TOL


@pytest.mark.parametrize("T,K,r,a_is_kr", [(4096, 1024, 16, 1), (4096, 1024, 16, 0), (100, 768, 8, 1), (33, 64, 4, 0),
                                            (512, 3072, 64, 1), (257, 2048, 32, 0), (10, 24, 4, 1), (37, 101, 5, 1), (2500, 333, 16, 0)])
def test_lora_down_streaming_kernel(capi, T, K, r, a_is_kr):
    """Default mode: TF32 operands on the tensor pipe (same class as the large GEMMs); exact mode: 3xTF32 split / fp32 FMA."""
    rng = np.random.default_rng(T + K + r)
    x = rng.standard_normal((T, K)).astype(np.float32)
    A = (rng.standard_normal((K, r) if a_is_kr else (r, K)) * 0.1).astype(np.float32)
    e = 2.0 * (x.astype(np.float64) @ (A if a_is_kr else A.T).astype(np.float64))
    for exact, tol in ((True, FP32_TOL), (False, TF32_TOL)):
        prev = capi.lib.mft_use_tensor_cores(0 if exact else 1)
        try:
            y = torch.full((T, r), float("nan"), device="cuda")
            capi.call("mft_lora_down", dev(x), K, dev(A), A.shape[1], a_is_kr, y, r, T, K, r, 2.0, None)
            torch.cuda.synchronize()
        finally:
            capi.lib.mft_use_tensor_cores(prev)
        assert rel(y, e) < tol, (exact, rel(y, e))


@pytest.mark.parametrize("T,q,r,rq", [(4096, 1024, 16, 1), (4096, 1024, 16, 0), (100, 768, 8, 1), (65, 50, 4, 0),
                                       (512, 1024, 64, 1), (300, 2304, 32, 0), (129, 51, 3, 1), (1000, 77, 16, 0)])
def test_lora_tn_streaming_kernel(capi, T, q, r, rq):
    rng = np.random.default_rng(T + q + r)
    S = rng.standard_normal((T, r)).astype(np.float32)
    W = rng.standard_normal((T, q)).astype(np.float32)
    outs = []
    for _ in range(2):
        out = torch.full((r, q) if rq else (q, r), float("nan"), device="cuda")
        w = ws(capi.lib.mft_lora_tn_workspace_bytes(T, q, r))
        capi.call("mft_lora_tn", dev(S), r, dev(W), q, out, out.shape[1], rq, T, q, r, 0.5, w, w.numel(), None)
        torch.cuda.synchronize()
        outs.append(out)
    e = 0.5 * (S.astype(np.float64).T @ W.astype(np.float64))
    assert torch.equal(outs[0], outs[1])
    assert rel(outs[0], e if rq else e.T) < FP32_TOL


BF16_TOL = 1e-2   # bf16 inputs (8-bit mantissa), fp32 accumulate


@pytest.mark.parametrize("B,H,S", [(2, 4, 128), (1, 3, 200), (8, 16, 512), (1, 2, 64)])
def test_attention_bf16_operand_mode(capi, B, H, S):
    """head_dim 64 on mma.sync m16n8k16 with bf16 operands (fp32 accumulate / softmax), fused-qkv layout, plus its bf16 ctx copy."""
    Hd, Cc = 64, H * 64
    rng = np.random.default_rng(B * 100 + S)
    qkv = rng.standard_normal((B, S, 3 * Cc)).astype(np.float32)
    d = dev(qkv)
    ctx = torch.full((B, S, Cc), float("nan"), device="cuda")
    ctx16 = torch.zeros((B, S, Cc), dtype=torch.bfloat16, device="cuda")
    prev_p, prev_b = capi.lib.mft_attention_precise(0), capi.lib.mft_attention_bf16(1)
    try:
        base = d.data_ptr()
        capi.check(capi.lib.mft_causal_attention_fwd_ex(base, base + 4 * Cc, base + 8 * Cc, ctx.data_ptr(), ctx16.data_ptr(), B, H, S, Hd,
                                                       S * 3 * Cc, 3 * Cc, Hd, S * 3 * Cc, 3 * Cc, Hd, S * Cc, Cc, Hd, 1.0 / np.sqrt(Hd), 0, None),
                   "mft_causal_attention_fwd_ex")
        torch.cuda.synchronize()
    finally:
        capi.lib.mft_attention_precise(prev_p); capi.lib.mft_attention_bf16(prev_b)
    q, k, v = (qkv[..., i * Cc:(i + 1) * Cc].reshape(B, S, H, Hd).transpose(0, 2, 1, 3) for i in range(3))
    e = O.causal_attention_fwd(q, k, v, 1.0 / np.sqrt(Hd)).transpose(0, 2, 1, 3).reshape(B, S, Cc)
    assert rel(ctx, e) < BF16_TOL
    assert torch.equal(ctx16, ctx.to(torch.bfloat16))


# ----------------------------------------------------------------------------- tcgen05 TF32 path
@pytest.mark.parametrize("m,n,k", [(128, 256, 32), (128, 64, 64), (512, 768, 768), (512, 2304, 768), (300, 200, 100),
                                   (2048, 3072, 768), (129, 257, 36), (4096, 1024, 4096), (64, 64, 4096),
                                   (4500, 300, 64)])       # 18 m-units: one raster group of 16 and a short one of 2
def test_gemm_tc_shapes(capi, m, n, k):
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    b = rng.standard_normal((n, k)).astype(np.float32)
    c, _, used = run_gemm(capi, a, b, 0, 1)
    assert used == 1, "expected the tcgen05 path"
    e = ref_gemm(a, b, 0, 1)
    assert rel(c, e) < TF32_TOL
    # the tensor-core result must agree with the exact SIMT kernel to TF32 accuracy as well
    c2, _, _ = run_gemm(capi, a, b, 0, 1, tc=False)
    assert rel(c, c2.cpu().numpy()) < TF32_TOL


def test_gemm_tc_epilogue_bias_beta_gelu_and_lora_segment(capi):
    rng = np.random.default_rng(5)
    m, n, k, r = 512, 768, 768, 8
    a = rng.standard_normal((m, k)).astype(np.float32)
    b = (rng.standard_normal((n, k)) * 0.05).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    c0 = rng.standard_normal((m, n)).astype(np.float32)
    xa = rng.standard_normal((m, r)).astype(np.float32)
    bt = rng.standard_normal((n, r)).astype(np.float32)
    c, g, used = run_gemm(capi, a, b, 0, 1, alpha=1.0, beta=1.0, bias=bias, c0=c0, gelu=True, seg2=(xa, bt))
    assert used == 1
    e = ref_gemm(a, b, 0, 1, 1.0, 1.0, bias, c0, seg2=(xa, bt))
    assert rel(c, e) < TF32_TOL
    assert rel(g, O.gelu_fwd(c.cpu().numpy())) < 2e-5


@pytest.mark.parametrize("m,n,k,path", [(512, 768, 768, "tc"), (300, 200, 100, "tc"), (512, 768, 768, "simt"), (70, 90, 33, "simt"),
                                          (96, 64, 1024, "splitk")])
def test_gemm_epilogue_residual_and_gelu_grad(capi, m, n, k, path):
    """v = acc + bias + addend; v *= gelu'(x)  — equals the separate add / GeluBackward kernels on the same fp32 product."""
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    b = (rng.standard_normal((n, k)) * 0.05).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    res = rng.standard_normal((m, n)).astype(np.float32)
    gx = rng.standard_normal((m, n)).astype(np.float32)
    da, db, dbias, dres, dgx = dev(a), dev(b), dev(bias), dev(res), dev(gx)
    prev = capi.lib.mft_use_tensor_cores(1 if path == "tc" else 0)
    try:
        def run(addend, ggx):
            c = torch.full((m, n), float("nan"), device="cuda")
            args = capi.GemmArgs(m, n, k, da.data_ptr(), k, 0, db.data_ptr(), k, 1, c.data_ptr(), n, 1.0, 0.0, dbias.data_ptr(), None,
                                 0, None, 0, None, 0, 1, 0, 0, 0, addend.data_ptr() if addend is not None else None, n,
                                 ggx.data_ptr() if ggx is not None else None, n)
            if path == "splitk":
                w = ws(capi.lib.mft_gemm_splitk_workspace_bytes(m, n, 1, 4))
                capi.check(capi.lib.mft_gemm_splitk(C.byref(args), 4, w.data_ptr(), w.numel(), None), "mft_gemm_splitk")
            else:
                assert capi.lib.mft_gemm_uses_tensor_cores(C.byref(args)) == (1 if path == "tc" else 0)
                capi.check(capi.lib.mft_gemm(C.byref(args), None), "mft_gemm")
            torch.cuda.synchronize()
            return c
        plain = run(None, None)
        fused = run(dres, dgx)
    finally:
        capi.lib.mft_use_tensor_cores(prev)
    # separate kernels on the plain product: add, then GeluBackward
    s = plain + dres
    out = torch.empty_like(s)
    capi.call("mft_gelu_bwd", dgx, s, out, m * n, None)
    torch.cuda.synchronize()
    assert torch.equal(fused, out)                      # bit for bit
    e = (ref_gemm(a, b, 0, 1, bias=bias) + res) * O.gelu_bwd(gx, np.ones_like(gx))
    assert rel(fused, e) < (TF32_TOL if path == "tc" else FP32_TOL)


def test_gemm_tc_batched_and_pitched(capi):
    rng = np.random.default_rng(6)
    a = rng.standard_normal((6, 256, 64)).astype(np.float32)
    b = rng.standard_normal((6, 256, 64)).astype(np.float32)
    c, _, used = run_gemm(capi, a, b, 0, 1)           # Q K^T style, batch 6
    assert used == 1
    assert rel(c, ref_gemm(a, b, 0, 1)) < TF32_TOL
    # vocabulary-like N that is not a multiple of 4 and a pitched A operand (lda > k)
    m, n, k, lda = 256, 1003, 96, 128
    a_full = rng.standard_normal((m, lda)).astype(np.float32)
    bmat = rng.standard_normal((n, k)).astype(np.float32)
    da, db = dev(a_full), dev(bmat)
    cc = torch.empty(m, n, device="cuda")
    args = capi.gemm(m, n, k, da, lda, 0, db, k, 1, cc, n)
    assert capi.lib.mft_gemm_uses_tensor_cores(C.byref(args)) == 1
    assert rel(cc, a_full[:, :k].astype(np.float64) @ bmat.astype(np.float64).T) < TF32_TOL




@pytest.mark.parametrize("m,n,k,r", [(512, 768, 768, 0), (4096, 1024, 1024, 16), (300, 200, 104, 8), (256, 50257, 768, 0),
                                     (2048, 3072, 768, 0), (5000, 520, 128, 0)])      # 20 m-units: raster groups of 16 + 4
def test_gemm_bf16_tensor_core_path(capi, m, n, k, r):
    """mft_cast_bf16 / mft_transpose_cast_bf16 + mft_gemm_bf16 (kind::f16 tcgen05) against the fp64 product."""
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) * 0.05).astype(np.float32)      # weight stored [in,out] like the reference
    bias = rng.standard_normal(n).astype(np.float32)
    da, dw, dbias = dev(a), dev(w), dev(bias)
    ldk = (k + 7) // 8 * 8
    a16 = torch.zeros(m, ldk, dtype=torch.bfloat16, device="cuda")
    w16 = torch.zeros(n, ldk, dtype=torch.bfloat16, device="cuda")
    capi.call("mft_cast_bf16", da, k, a16, ldk, m, k, None)
    capi.call("mft_transpose_cast_bf16", dw, n, w16, ldk, k, n, None)
    assert torch.equal(a16[:, :k], da.to(torch.bfloat16))                    # round-to-nearest-even, like torch
    assert torch.equal(w16[:, :k], dw.t().contiguous().to(torch.bfloat16))
    kw = {}
    seg2 = None
    if r:
        xa = rng.standard_normal((m, r)).astype(np.float32)
        bt = rng.standard_normal((n, r)).astype(np.float32)
        ldr = (r + 7) // 8 * 8
        xa16 = torch.zeros(m, ldr, dtype=torch.bfloat16, device="cuda"); xa16[:, :r] = dev(xa).to(torch.bfloat16)
        bt16 = torch.zeros(n, ldr, dtype=torch.bfloat16, device="cuda"); bt16[:, :r] = dev(bt).to(torch.bfloat16)
        kw = dict(k2=r, a2=xa16.data_ptr(), lda2=ldr, b2=bt16.data_ptr(), ldb2=ldr)
        seg2 = (xa, bt)
    ldc = (n + 31) // 32 * 32
    c = torch.full((m, ldc), float("nan"), device="cuda")
    args = capi.GemmArgs(m, n, k, a16.data_ptr(), ldk, 0, w16.data_ptr(), ldk, 1, c.data_ptr(), ldc, 1.0, 0.0, dbias.data_ptr(), None,
                         kw.get("k2", 0), kw.get("a2"), kw.get("lda2", 0), kw.get("b2"), kw.get("ldb2", 0), 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    torch.cuda.synchronize()
    e = ref_gemm(a, np.ascontiguousarray(w.T), 0, 1, 1.0, 0.0, bias, None, seg2)
    assert rel(c[:, :n], e) < BF16_TOL


def test_lora_linear_fused_forward_golden(capi, golden_ops):
    """y = x W + b + s (xA) B as ONE mft_gemm call with the rank-r update as second K segment, against the
    reference's LoRALinear::forward output (tests/golden/ops/lora0_*)."""
    L = golden_ops
    x, W, bias, A, Bm = L("lora0_x"), L("lora0_W"), L("lora0_bias"), L("lora0_A"), L("lora0_B")
    x2 = x.reshape(-1, x.shape[-1])
    xa, _, _ = run_gemm(capi, x2, A, 0, 0, alpha=2.0, tc=False)          # s * x A   [T, r]
    y, _, _ = run_gemm(capi, x2, np.ascontiguousarray(W.T), 0, 1, bias=bias,
                       seg2=(xa.cpu().numpy(), np.ascontiguousarray(Bm.T)), tc=False)
    assert rel(y, L("lora0_y").reshape(y.shape)) < FP32_TOL


# ----------------------------------------------------------------------------- attention
def run_attention(capi, q, k, v, scale, window=0):
    """q,k,v numpy [B,H,S,Hd]; packs them as the fused [B,S,3C] qkv buffer the GPT-2 block produces."""
    B, H, S, Hd = q.shape
    Cc = H * Hd
    qkv = np.concatenate([t.transpose(0, 2, 1, 3).reshape(B, S, Cc) for t in (q, k, v)], axis=-1)
    dq = dev(qkv)
    ctx = torch.empty(B, S, Cc, device="cuda")
    base = dq.data_ptr()
    capi.call("mft_causal_attention_fwd", base, base + 4 * Cc, base + 8 * Cc, ctx, B, H, S, Hd,
              S * 3 * Cc, 3 * Cc, Hd, S * 3 * Cc, 3 * Cc, Hd, S * Cc, Cc, Hd, scale, window, None)
    torch.cuda.synchronize()
    return ctx.cpu().numpy().reshape(B, S, H, Hd).transpose(0, 2, 1, 3)


@pytest.fixture
def precise_attention(capi):
    prev = capi.lib.mft_attention_precise(1)
    yield
    capi.lib.mft_attention_precise(prev)


def test_attention_golden(capi, golden_ops, precise_attention):
    L = golden_ops
    y = run_attention(capi, L("attn_q"), L("attn_k"), L("attn_v"), 0.25)
    assert rel(y, L("attn_y")) < FP32_TOL


def test_attention_tf32_mode(capi, golden_ops):
    """Default mode: TF32 operands (same input precision as the tensor-core GEMMs)."""
    L = golden_ops
    prev = capi.lib.mft_attention_precise(0)
    try:
        y = run_attention(capi, L("attn_q"), L("attn_k"), L("attn_v"), 0.25)
        assert rel(y, L("attn_y")) < TF32_TOL
        rng = np.random.default_rng(9)
        q, k, v = (rng.standard_normal((2, 16, 512, 64)).astype(np.float32) for _ in range(3))
        y = run_attention(capi, q, k, v, 0.125)
        assert rel(y, O.causal_attention_fwd(q, k, v, 0.125)) < TF32_TOL
    finally:
        capi.lib.mft_attention_precise(prev)


@pytest.mark.parametrize("B,H,S,Hd,window", [(4, 12, 128, 64, 0), (1, 2, 1, 64, 0), (2, 3, 200, 64, 0),
                                             (1, 4, 512, 64, 0), (2, 2, 130, 32, 0), (1, 2, 300, 64, 128),
                                             (1, 1, 96, 128, 0)])
def test_attention_shapes(capi, precise_attention, B, H, S, Hd, window):
    rng = np.random.default_rng(S + Hd)
    q, k, v = (rng.standard_normal((B, H, S, Hd)).astype(np.float32) for _ in range(3))
    y = run_attention(capi, q, k, v, 1.0 / np.sqrt(Hd), window)
    assert rel(y, O.causal_attention_fwd(q, k, v, 1.0 / np.sqrt(Hd), window)) < FP32_TOL


# ----------------------------------------------------------------------------- cross-entropy
def run_ce(capi, logits, labels, ld=None, reduction=0, gout=1.0):
    B, S, V = logits.shape
    ld = ld or V
    buf = torch.zeros(B * S, ld, device="cuda")
    buf[:, :V] = dev(logits.reshape(B * S, V))
    dl = dev(labels, torch.int32)
    loss = torch.empty(1, device="cuda")
    valid = torch.empty(1, dtype=torch.int32, device="cuda")
    lse = torch.empty(B * S, device="cuda")
    w = ws(capi.lib.mft_lm_ce_workspace_bytes(B * S))
    capi.call("mft_lm_ce_fwd", buf, ld, dl, B, S, V, -100, reduction, loss, valid, lse, w, w.numel(), None)
    g = torch.full((B * S, ld), float("nan"), device="cuda")
    go = torch.tensor([gout], device="cuda")
    capi.call("mft_lm_ce_bwd", buf, ld, dl, lse, valid, go, B, S, V, -100, reduction, g, ld, None)
    torch.cuda.synchronize()
    return loss.item(), valid.item(), g[:, :V].cpu().numpy().reshape(B, S, V)


def test_ce_golden(capi, golden_ops):
    L = golden_ops
    loss, valid, g = run_ce(capi, L("ce_x"), L("ce_labels"))
    assert valid == 7
    assert abs(loss - L("ce_loss")[0]) / abs(L("ce_loss")[0]) < FP32_TOL
    assert rel(g, L("ce_gx")) < FP32_TOL


@pytest.mark.parametrize("B,S,V,ld", [(4, 128, 50257, 50272), (2, 16, 50257, 50257), (1, 2, 1000, 1000),
                                      (2, 8, 262144, 262144), (3, 1, 100, 100)])
def test_ce_shapes(capi, B, S, V, ld):
    rng = np.random.default_rng(V + S)
    logits = (rng.standard_normal((B, S, V)) * 2).astype(np.float32)
    labels = rng.integers(0, V, (B, S)).astype(np.int32)
    if S > 3:
        labels[0, 2] = -100
        labels[-1, 1] = V + 5      # out of range -> skipped
    for red, name in ((0, "mean"), (1, "sum")):
        loss, valid, g = run_ce(capi, logits, labels, ld, red, gout=0.5)
        el, ec = O.lm_cross_entropy_fwd(logits, labels, -100, name)
        assert valid == ec
        assert abs(loss - el) <= FP32_TOL * max(abs(el), 1e-30)
        assert rel(g, O.lm_cross_entropy_bwd(logits, labels, -100, name, gout=0.5)) < FP32_TOL


# ----------------------------------------------------------------------------- clip + Adam
def make_slots(capi, params, grads, ms, vs, steps, b1=0.9, b2=0.999):
    arr = (capi.TensorSlot * len(params))()
    for i, (p, g, m, v, t) in enumerate(zip(params, grads, ms, vs, steps)):
        arr[i].param, arr[i].m, arr[i].v = p.data_ptr(), m.data_ptr(), v.data_ptr()
        arr[i].grad = g.data_ptr() if g is not None else None
        arr[i].n = p.numel()
        arr[i].bias_corr1 = float(np.float32(1) - np.power(np.float32(b1), np.float32(t), dtype=np.float32))
        arr[i].bias_corr2 = float(np.float32(1) - np.power(np.float32(b2), np.float32(t), dtype=np.float32))
    raw = bytes(arr)
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()


def test_adam_golden_three_steps(capi, golden_ops):
    L = golden_ops
    p = dev(L("adam_p0"))
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for s in range(3):
        g = dev(L(f"adam_g{s}"))
        slots = make_slots(capi, [p], [g], [m], [v], [s + 1])
        capi.call("mft_adam_step", slots, 1, 1e-2, 0.9, 0.999, 1e-8, 0.01, None, None)
        assert rel(p, L(f"adam_p{s + 1}")) < 1e-6


def test_clip_and_multitensor_adam(capi):
    rng = np.random.default_rng(11)
    shapes = [(768, 8), (8, 768), (5,), (1031,), (64, 64)]
    P = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    G = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    G[2] = None   # a tensor that never receives a gradient is skipped (optim/adam.cpp:35)
    dp = [dev(p) for p in P]
    dg = [dev(g) if g is not None else None for g in G]
    dm = [torch.zeros_like(p) for p in dp]
    dv = [torch.zeros_like(p) for p in dp]
    slots = make_slots(capi, dp, dg, dm, dv, [1] * len(P))
    sumsq = torch.empty(1, dtype=torch.float64, device="cuda")
    w = ws(capi.lib.mft_grad_sumsq_workspace_bytes())
    capi.call("mft_grad_sumsq", slots, len(P), sumsq, w, w.numel(), None)
    norm = O.grad_norm(G)
    assert abs(np.sqrt(sumsq.item()) - norm) / norm < 1e-12
    scale, nout = torch.empty(1, device="cuda"), torch.empty(1, device="cuda")
    capi.call("mft_clip_scale", sumsq, 1.0, 1.0, scale, nout, None)
    es = O.clip_scale(norm, 1.0)
    assert abs(scale.item() - float(es)) < 1e-9 and abs(nout.item() - norm) / norm < 1e-6
    # clip factor applied inside Adam (never visits the host) == scaling the grads first
    capi.call("mft_adam_step", slots, len(P), 1e-3, 0.9, 0.999, 1e-8, 0.0, scale, None)
    for i, (p, g) in enumerate(zip(P, G)):
        if g is None:
            assert rel(dp[i], p) == 0
            continue
        ep, em, ev = O.adam_step(p, g * es, np.zeros_like(p), np.zeros_like(p), 1, 1e-3)
        assert rel(dp[i], ep) < 1e-6 and rel(dm[i], em) < 1e-6 and rel(dv[i], ev) < 1e-6
    # host-visible clip semantics: scale the gradients in place
    g_before = dg[0].clone()
    capi.call("mft_scale_grads", slots, len(P), scale, None)
    assert rel(dg[0], g_before.cpu().numpy() * es) < 1e-7


# ----------------------------------------------------------------------------- bf16 results (bf16 operand mode, tcgen05 epilogue)
@pytest.mark.parametrize("m,n,k", [(512, 768, 256), (300, 1000, 128), (128, 64, 64), (4096, 3072, 1024), (260, 200, 96)])
def test_gemm_bf16_result_only_and_copy(capi, m, n, k):
    """c = NULL + c_bf16: the bf16 values are the round-to-nearest of what the fp32 epilogue produces; c + c_bf16: both."""
    rng = np.random.default_rng(m + n + k)
    a = torch.tensor(rng.standard_normal((m, k)).astype(np.float32)).cuda().to(torch.bfloat16)
    b = torch.tensor((rng.standard_normal((n, k)) * 0.1).astype(np.float32)).cuda().to(torch.bfloat16)
    bias = torch.tensor(rng.standard_normal(n).astype(np.float32)).cuda()
    ldc = (n + 3) // 4 * 4
    ld16 = (n + 7) // 8 * 8
    c32 = torch.zeros(m, ldc, device="cuda")
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, c32.data_ptr(), ldc, 0.5, 0.0, bias.data_ptr(), None, 0, None, 0, None, 0, 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    expect = c32[:, :n].to(torch.bfloat16)
    only = torch.zeros(m, ld16, device="cuda", dtype=torch.bfloat16)
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, None, 0, 0.5, 0.0, bias.data_ptr(), None, 0, None, 0, None, 0, 1, 0, 0, 0,
                         None, 0, None, 0, None, 0, only.data_ptr(), ld16)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    torch.cuda.synchronize()
    assert torch.equal(only[:, :n], expect)
    if ld16 > n:
        assert float(only[:, n:].float().abs().max()) == 0.0          # the TMA clips at the matrix edge: padding untouched
    both32, both16 = torch.zeros(m, ldc, device="cuda"), torch.zeros(m, ld16, device="cuda", dtype=torch.bfloat16)
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, both32.data_ptr(), ldc, 0.5, 0.0, bias.data_ptr(), None, 0, None, 0, None, 0, 1, 0, 0, 0,
                         None, 0, None, 0, None, 0, both16.data_ptr(), ld16)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    torch.cuda.synchronize()
    assert torch.equal(both32, c32) and torch.equal(both16[:, :n], expect)


def test_gemm_bf16_gelu_activation_as_bf16_only(capi):
    rng = np.random.default_rng(3)
    m, n, k = 512, 1024, 256
    a = torch.tensor(rng.standard_normal((m, k)).astype(np.float32)).cuda().to(torch.bfloat16)
    b = torch.tensor((rng.standard_normal((n, k)) * 0.1).astype(np.float32)).cuda().to(torch.bfloat16)
    pre, act, act16 = torch.zeros(m, n, device="cuda"), torch.zeros(m, n, device="cuda"), torch.zeros(m, n, device="cuda", dtype=torch.bfloat16)
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, pre.data_ptr(), n, 1.0, 0.0, None, act.data_ptr(), 0, None, 0, None, 0, 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    pre2 = torch.zeros(m, n, device="cuda")
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, pre2.data_ptr(), n, 1.0, 0.0, None, None, 0, None, 0, None, 0, 1, 0, 0, 0,
                         None, 0, None, 0, act16.data_ptr(), n)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    torch.cuda.synchronize()
    assert torch.equal(pre, pre2) and torch.equal(act16, act.to(torch.bfloat16))
    # pre-activation AND activation as bf16 only (what fc_in writes when GeluBackward is the pre-activation's only reader)
    pre16, act16b = torch.zeros(m, n, device="cuda", dtype=torch.bfloat16), torch.zeros(m, n, device="cuda", dtype=torch.bfloat16)
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, None, n, 1.0, 0.0, None, None, 0, None, 0, None, 0, 1, 0, 0, 0,
                         None, 0, None, 0, act16b.data_ptr(), n, pre16.data_ptr(), n)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")
    torch.cuda.synchronize()
    assert torch.equal(pre16, pre.to(torch.bfloat16)) and torch.equal(act16b, act16)


def test_lowp_elementwise_kernels(capi):
    """mft_cast_from_bf16, mft_gelu_bwd_lowp (fp32 / bf16 pre-activation), LayerNorm and CE backward with bf16-only results."""
    rng = np.random.default_rng(11)
    n = 4096 * 24
    x = rng.standard_normal(n).astype(np.float32)
    gy = rng.standard_normal(n).astype(np.float32)
    dx, dgy16 = dev(x), dev(gy).to(torch.bfloat16)
    gx16, gx32 = torch.empty(n, device="cuda", dtype=torch.bfloat16), torch.empty(n, device="cuda")
    capi.call("mft_gelu_bwd_lowp", dx, 0, dgy16, gx32, gx16, n, None)
    e = O.gelu_bwd(x, dgy16.float().cpu().numpy())
    assert rel(gx32, e) < FP32_TOL and torch.equal(gx16, gx32.to(torch.bfloat16))
    gx16f = torch.empty(n, device="cuda", dtype=torch.bfloat16)      # bf16 as the only result: hardware tanh, at most one bf16 step away
    capi.call("mft_gelu_bwd_lowp", dx, 0, dgy16, None, gx16f, n, None)
    assert rel(gx16f.float(), e) < 5e-3
    assert bool(((gx16f.float() - gx16.float()).abs() <= gx16.float().abs() * 2.0 ** -6 + 1e-3).all())
    x16 = dx.to(torch.bfloat16)
    capi.call("mft_gelu_bwd_lowp", x16, 1, dgy16, None, gx16, n, None)
    assert rel(gx16.float(), O.gelu_bwd(x16.float().cpu().numpy(), dgy16.float().cpu().numpy())) < 5e-3
    src = torch.tensor(rng.standard_normal((37, 104)).astype(np.float32)).cuda().to(torch.bfloat16)       # pitched: 100 of 104 columns
    out = torch.zeros(37, 100, device="cuda")
    capi.call("mft_cast_from_bf16", src, 104, out, 100, 37, 100, None)
    assert torch.equal(out, src[:, :100].float())
    # LayerNorm with the bf16 copy as the only result
    rows, dim = 300, 1024
    xx, g, b = (rng.standard_normal((rows, dim)).astype(np.float32), rng.standard_normal(dim).astype(np.float32), rng.standard_normal(dim).astype(np.float32))
    y32, y16a, y16b = torch.empty(rows, dim, device="cuda"), torch.empty(rows, dim, device="cuda", dtype=torch.bfloat16), torch.empty(rows, dim, device="cuda", dtype=torch.bfloat16)
    capi.call("mft_layernorm_fwd_ex", dev(xx), None, None, dev(g), dev(b), y32, y16a, dim, rows, dim, 1e-5, None)
    capi.call("mft_layernorm_fwd_ex", dev(xx), None, None, dev(g), dev(b), None, y16b, dim, rows, dim, 1e-5, None)
    assert torch.equal(y16a, y16b) and torch.equal(y16a, y32.to(torch.bfloat16))
    # CE backward with dlogits as bf16 only
    B, S, V = 2, 9, 1000
    logits = rng.standard_normal((B, S, V)).astype(np.float32)
    labels = rng.integers(0, V, (B, S)).astype(np.int32)
    dl, dlab = dev(logits), torch.tensor(labels).cuda()
    loss, valid, lse = torch.zeros(1, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda"), torch.empty(B * S, device="cuda")
    w = ws(capi.lib.mft_lm_ce_workspace_bytes(B * S))
    capi.call("mft_lm_ce_fwd", dl, V, dlab, B, S, V, -100, 0, loss, valid, lse, w, w.numel(), None)
    g32, g16a, g16b = torch.empty(B * S, V, device="cuda"), torch.empty(B * S, V, device="cuda", dtype=torch.bfloat16), torch.empty(B * S, V, device="cuda", dtype=torch.bfloat16)
    capi.call("mft_lm_ce_bwd_ex", dl, V, dlab, lse, valid, None, B, S, V, -100, 0, g32, V, g16a, V, None)
    capi.call("mft_lm_ce_bwd_ex", dl, V, dlab, lse, valid, None, B, S, V, -100, 0, None, 0, g16b, V, None)
    assert torch.equal(g16a, g32.to(torch.bfloat16))
    # the bf16-only form evaluates exp with one ex2.approx (2 ulp in fp32): at most one bf16 step away from the rounded fp32 result
    a, b = g16a.float(), g16b.float()
    assert bool(((a - b).abs() <= a.abs() * 2.0 ** -7 + 1e-30).all())

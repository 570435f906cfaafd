"""Per-op sweep (BASELINE.json configs[4]): the LoRA-fused linear forward/backward at d = 768..2048, r = 8..64, tokens
512..65536, and the LM cross-entropy at vocab 50257 / 262144 — through the C ABI, in the three execution modes.

Sizes the numpy oracle finishes in seconds are compared in full. At the largest token counts the comparison uses what the
operator guarantees independently of size: rows of y and dX depend only on their own input row (so slices of the big run are
checked against the oracle on those rows), and dA / dB are sums over tokens (so the big run must equal the sum of its
chunks). Tolerances: exact fp32 mode 1e-5, TF32 2e-3, bf16 operands 1e-2."""
import ctypes as C

import numpy as np
import pytest
import torch

import mft_oracle as O
from conftest import FP32_TOL, TF32_TOL, dev, rel, ws

pytestmark = pytest.mark.gpu
BF16_TOL = 1e-2


def lora_forward(capi, x, Wt, bias, A, Bt, scale):
    """y = x W + b + scale (x A) B with W given K-major as Wt [out,in], A [in,r], Bt [out,r] — the calls fused_lora_linear makes."""
    T, d_in = x.shape
    d_out, r = Bt.shape
    xa = torch.empty(T, r, device="cuda")
    capi.call("mft_lora_down", x, d_in, A, r, 1, xa, r, T, d_in, r, scale, None)
    y = torch.empty(T, d_out, device="cuda")
    args = capi.GemmArgs(T, d_out, d_in, x.data_ptr(), d_in, 0, Wt.data_ptr(), d_in, 1, y.data_ptr(), d_out, 1.0, 0.0, bias.data_ptr(), None,
                         r, xa.data_ptr(), r, Bt.data_ptr(), r, 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm(C.byref(args), None), "mft_gemm")
    return y, xa


def lora_backward(capi, x, W, A, B, Bt_unused, xa, gy, scale):
    """dx = gy W^T + dxa A^T, dA = x^T dxa, dB = xa^T gy with dxa = scale gy B^T. W [in,out] is K-major for the dX product."""
    T, d_in = x.shape
    d_out = gy.shape[1]
    r = A.shape[1]
    dxa = torch.empty(T, r, device="cuda")
    capi.call("mft_lora_down", gy, d_out, B, d_out, 0, dxa, r, T, d_out, r, scale, None)          # B [r,out]: contraction over out
    dx = torch.empty(T, d_in, device="cuda")
    args = capi.GemmArgs(T, d_in, d_out, gy.data_ptr(), d_out, 0, W.data_ptr(), d_out, 1, dx.data_ptr(), d_in, 1.0, 0.0, None, None,
                         r, dxa.data_ptr(), r, A.data_ptr(), r, 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm(C.byref(args), None), "mft_gemm")
    dA = torch.empty(d_in, r, device="cuda")
    dB = torch.empty(r, d_out, device="cuda")
    w1 = ws(capi.lib.mft_lora_tn_workspace_bytes(T, d_in, r))
    capi.call("mft_lora_tn", dxa, r, x, d_in, dA, r, 0, T, d_in, r, 1.0, w1, w1.numel(), None)    # [in, r]
    w2 = ws(capi.lib.mft_lora_tn_workspace_bytes(T, d_out, r))
    capi.call("mft_lora_tn", xa, r, gy, d_out, dB, d_out, 1, T, d_out, r, 1.0, w2, w2.numel(), None)   # [r, out]
    torch.cuda.synchronize()
    return dx, dA, dB


def make_case(T, d, r, seed):
    rng = np.random.default_rng(seed)
    g = lambda *s, k=1.0: (rng.standard_normal(s) * k).astype(np.float32)
    return dict(x=g(T, d), W=g(d, d, k=0.05), bias=g(d, k=0.1), A=g(d, r, k=0.1), B=g(r, d, k=0.1), gy=g(T, d), scale=2.0)


def set_mode(capi, mode):
    """exact: fp32 SIMT GEMM + 3xTF32 mma kernels; tf32: library default."""
    return capi.lib.mft_use_tensor_cores(0 if mode == "exact" else 1)


@pytest.mark.parametrize("mode", ["exact", "tf32"])
@pytest.mark.parametrize("T,d,r", [(512, 768, 8), (512, 1024, 16), (4096, 768, 64), (2048, 2048, 16), (1000, 1280, 32)])
def test_lora_linear_sweep_against_oracle(capi, T, d, r, mode):
    c = make_case(T, d, r, T + d + r)
    tol = FP32_TOL if mode == "exact" else TF32_TOL
    x, W, bias, A, B, gy = (dev(c[k]) for k in ("x", "W", "bias", "A", "B", "gy"))
    Wt, Bt = W.t().contiguous(), B.t().contiguous()
    prev = set_mode(capi, mode)
    try:
        y, xa = lora_forward(capi, x, Wt, bias, A, Bt, c["scale"])
        dx, dA, dB = lora_backward(capi, x, W, A, B, None, xa, gy, c["scale"])
    finally:
        capi.lib.mft_use_tensor_cores(prev)
    x64, W64, A64, B64, g64 = (c[k].astype(np.float64) for k in ("x", "W", "A", "B", "gy"))
    s = c["scale"]
    ey = x64 @ W64 + c["bias"] + s * (x64 @ A64) @ B64
    dxa = s * g64 @ B64.T
    assert rel(y, ey) < tol
    assert rel(dx, g64 @ W64.T + dxa @ A64.T) < tol
    assert rel(dA, x64.T @ dxa) < tol
    assert rel(dB, (s * x64 @ A64).T @ g64) < tol
    # and the float32 oracle restatement of the reference's chain agrees with the fp64 expectation it is compared to
    if T * d <= 512 * 1024:
        assert rel(O.lora_linear_fwd(c["x"], c["W"], c["bias"], c["A"], c["B"], s), ey) < 1e-5


@pytest.mark.parametrize("T,d,r", [(65536, 768, 8), (65536, 2048, 64)])
def test_lora_linear_65536_tokens_by_size_independent_properties(capi, T, d, r):
    """Largest sweep point. Rows of y / dX depend on their own row only: the first and last 256 rows of the big run are
    compared with the oracle on those rows. dA / dB are sums over tokens: the big run equals the sum of 16 chunk runs."""
    gen = torch.Generator(device="cuda").manual_seed(T + d + r)
    rn = lambda *s, k=1.0: torch.randn(*s, device="cuda", generator=gen) * k
    x, gy = rn(T, d), rn(T, d)
    W, bias, A, B = rn(d, d, k=0.05), rn(d, k=0.1), rn(d, r, k=0.1), rn(r, d, k=0.1)
    Wt, Bt, s = W.t().contiguous(), B.t().contiguous(), 2.0
    y, xa = lora_forward(capi, x, Wt, bias, A, Bt, s)
    dx, dA, dB = lora_backward(capi, x, W, A, B, None, xa, gy, s)
    W64, A64, B64 = (t.double().cpu().numpy() for t in (W, A, B))
    for sl in (slice(0, 256), slice(T - 256, T)):
        x64, g64 = x[sl].double().cpu().numpy(), gy[sl].double().cpu().numpy()
        assert rel(y[sl], x64 @ W64 + bias.double().cpu().numpy() + s * (x64 @ A64) @ B64) < TF32_TOL
        assert rel(dx[sl], g64 @ W64.T + (s * g64 @ B64.T) @ A64.T) < TF32_TOL
    dA_sum, dB_sum = torch.zeros_like(dA, dtype=torch.float64), torch.zeros_like(dB, dtype=torch.float64)
    n_chunks = 16
    step = T // n_chunks
    for c in range(n_chunks):
        sl = slice(c * step, (c + 1) * step)
        xc, gc = x[sl].contiguous(), gy[sl].contiguous()
        _, xac = lora_forward(capi, xc, Wt, bias, A, Bt, s)
        _, dAc, dBc = lora_backward(capi, xc, W, A, B, None, xac, gc, s)
        dA_sum += dAc.double()
        dB_sum += dBc.double()
    assert rel(dA, dA_sum.cpu().numpy()) < TF32_TOL and rel(dB, dB_sum.cpu().numpy()) < TF32_TOL
    # one chunk against the oracle pins the chunk results themselves
    x64, g64 = x[:step].double().cpu().numpy(), gy[:step].double().cpu().numpy()
    _, xac = lora_forward(capi, x[:step].contiguous(), Wt, bias, A, Bt, s)
    _, dAc, dBc = lora_backward(capi, x[:step].contiguous(), W, A, B, None, xac, gy[:step].contiguous(), s)
    assert rel(dAc, x64.T @ (s * g64 @ B64.T)) < TF32_TOL and rel(dBc, (s * x64 @ A64).T @ g64) < TF32_TOL


@pytest.mark.parametrize("B,S,V", [(128, 512, 50257), (16, 512, 262144), (4, 128, 50257)])
def test_cross_entropy_sweep_tokens_up_to_65536(capi, B, S, V):
    """LM cross-entropy at the sweep's token counts and vocabularies (B*S = 65536 x 50257 and 8192 x 262144 are several GB of
    logits): loss against a float64 reduction of per-row NLLs computed on the device, gradient rows against the oracle on
    sampled rows, every ignored / last-position row exactly zero, and the gradient rows each summing to ~0."""
    T = B * S
    ld = (V + 31) // 32 * 32
    gen = torch.Generator(device="cuda").manual_seed(V + T)
    logits = torch.randn(T, ld, device="cuda", generator=gen) * 2.0
    labels = torch.randint(0, V, (B, S), device="cuda", generator=gen, dtype=torch.int32)
    labels[:, ::7] = -100                                         # ignored targets
    loss = torch.empty(1, device="cuda")
    valid = torch.empty(1, dtype=torch.int32, device="cuda")
    lse = torch.empty(T, device="cuda")
    w = ws(capi.lib.mft_lm_ce_workspace_bytes(T))
    capi.call("mft_lm_ce_fwd", logits, ld, labels, B, S, V, -100, 0, loss, valid, lse, w, w.numel(), None)
    g = torch.empty(T, ld, device="cuda")
    go = torch.ones(1, device="cuda")
    capi.call("mft_lm_ce_bwd", logits, ld, labels, lse, valid, go, B, S, V, -100, 0, g, ld, None)
    torch.cuda.synchronize()
    # targets of row (b, s) are labels[b, s+1]; the last position of every sequence has none
    tgt = torch.full((B, S), -100, device="cuda", dtype=torch.int64)
    tgt[:, :-1] = labels[:, 1:].long()
    tgt = tgt.reshape(T)
    ok = tgt >= 0
    assert valid.item() == int(ok.sum().item())
    ref_lse = torch.logsumexp(logits[:, :V].double(), dim=1)
    picked = logits[torch.arange(T, device="cuda"), tgt.clamp(min=0)].double()
    ref_loss = ((ref_lse - picked) * ok).sum().item() / max(int(ok.sum().item()), 1)
    assert abs(loss.item() - ref_loss) <= 1e-5 * abs(ref_loss)
    assert not g[~ok][:, :V].any()                                # skipped rows: all-zero gradient
    assert float(g[ok][:, :V].double().sum(dim=1).abs().max()) < 1e-6      # softmax - onehot sums to zero per row
    rows = torch.nonzero(ok)[:: max(int(ok.sum().item()) // 48, 1), 0][:48]
    for r in rows.tolist():
        b, s = divmod(r, S)
        lg = logits[b * S:(b + 1) * S, :V].cpu().numpy()[None]
        # oracle on the one sequence the row belongs to, scaled to the global valid count
        eg = O.lm_cross_entropy_bwd(lg, labels[b:b + 1].cpu().numpy(), reduction="sum") / float(valid.item())
        assert rel(g[r, :V], eg[0, s]) < 1e-5
        if s > 3:
            break                                                 # a full-sequence oracle call per row is slow at V = 262144


@pytest.mark.parametrize("T,K,r", [(512, 768, 8), (4096, 1024, 16), (1000, 3072, 16), (37, 128, 3), (8192, 4096, 16), (300, 640, 8),
                                   (2048, 1024, 32), (1000, 1536, 64), (512, 3072, 24), (700, 768, 40)])
def test_lora_down_on_bf16_operands(capi, T, K, r):
    """mft_lora_down_bf16 (x and the factor's K-major shadow as bf16, mma.sync bf16, fp32 accumulate) against fp64 on the
    same bf16-rounded inputs (summation order only: 1e-5) and against the fp32 inputs (bf16 operand tolerance 1e-2)."""
    rng = np.random.default_rng(T + K + r)
    x = rng.standard_normal((T, K)).astype(np.float32)
    A = (rng.standard_normal((K, r)) * 0.1).astype(np.float32)
    x16 = dev(x).to(torch.bfloat16)
    At16 = dev(np.ascontiguousarray(A.T)).to(torch.bfloat16)             # [r, K]
    y = torch.full((T, r), float("nan"), device="cuda")
    y16 = torch.empty(T, r, device="cuda", dtype=torch.bfloat16)
    capi.call("mft_lora_down_bf16", x16, K, At16, K, y, r, y16, r, T, K, r, 2.0, None)
    torch.cuda.synchronize()
    e16 = 2.0 * (x16.float().cpu().numpy().astype(np.float64) @ At16.float().cpu().numpy().astype(np.float64).T)
    assert rel(y, e16) < 1e-5
    assert rel(y, 2.0 * (x.astype(np.float64) @ A.astype(np.float64))) < BF16_TOL
    assert torch.equal(y16, y.to(torch.bfloat16))


def test_lora_tn_group_equals_individual_launches(capi):
    """mft_lora_tn_group: several S^T W products of different shapes in one launch == mft_lora_tn one by one (bit for bit: same
    kernel body, same fixed summation order), including a problem whose split list is too long for the last-CTA sum."""
    rng = np.random.default_rng(21)
    cases = [(512, 768, 8, 1), (4096, 1024, 16, 0), (1000, 2304, 16, 1), (128 * 70, 256, 4, 1), (300, 640, 8, 0)]
    keep, outs, refs = [], [], []
    probs = (capi.LoraTnProblem * len(cases))()
    base, max_cb, max_sp, any_long, max_rq = 0, 1, 1, 0, 1
    for i, (T, q, r, rq) in enumerate(cases):
        S = dev(rng.standard_normal((T, r)).astype(np.float32))
        W = dev(rng.standard_normal((T, q)).astype(np.float32))
        out = torch.full((r, q) if rq else (q, r), float("nan"), device="cuda")
        ref = torch.empty_like(out)
        w1 = ws(capi.lib.mft_lora_tn_workspace_bytes(T, q, r))
        capi.call("mft_lora_tn", S, r, W, q, ref, out.shape[1], rq, T, q, r, 1.0, w1, w1.numel(), None)
        assert capi.lib.mft_lora_tn_group_eligible(T, q, r, W.data_ptr(), q) == 1
        part = ws(capi.lib.mft_lora_tn_workspace_bytes(T, q, r))
        cb, sp = capi.lib.mft_lora_tn_group_colblocks(q), capi.lib.mft_lora_tn_group_splits(T)
        tk = 1 if sp <= 64 else 0
        probs[i] = capi.LoraTnProblem(S.data_ptr(), r, W.data_ptr(), q, part.data_ptr(), out.data_ptr(), out.shape[1], T, q, r, rq, 1.0, tk, cb, sp, base)
        if tk:
            base += cb
        else:
            any_long = 1
        max_cb, max_sp, max_rq = max(max_cb, cb), max(max_sp, sp), max(max_rq, r * q)
        keep += [S, W, part]; outs.append(out); refs.append(ref)
    table = torch.frombuffer(bytearray(bytes(probs)), dtype=torch.uint8).cuda()
    for _ in range(2):                       # twice: the ticket counters must be back at zero after a launch
        for o in outs:
            o.fill_(float("nan"))
        capi.call("mft_lora_tn_group", table, len(cases), max_cb, max_sp, any_long, max_rq, None)
        torch.cuda.synchronize()
        for o, e in zip(outs, refs):
            assert torch.equal(o, e)

"""GPU parity tests: every CUDA entry point of include/mft_b200.h, called through the C ABI (ctypes),
against the oracle (numpy restatement pinned to the reference) and against tensors produced by the
reference itself (tests/golden/ops). Tolerances (relative to max|ref|):
  * fp32 elementwise / norm / softmax / CE / Adam / SIMT GEMM / attention (3xTF32): 1e-5
  * tcgen05 TF32 GEMM (tf32 inputs rounded in the TMA, fp32 accumulate):           2e-3   
"""
import ctypes as C

import numpy as np
import pytest
import torch

import mft_oracle as O

pytestmark = pytest.mark.gpu

from conftest import FP32_TOL, TF32_TOL, dev, rel, ws  # noqa: E402


# ----------------------------------------------------------------------------- elementwise
def test_binary_and_bias_golden(capi, golden_ops):
    L = golden_ops
    x, b = dev(L("bias_x")), dev(L("bias_b"))
    out = torch.empty_like(x)
    rows, n = x.numel() // b.numel(), b.numel()
    capi.call("mft_binary", 0, x, b, out, 1, rows, n, 1, 1, None)
    assert rel(out, L("bias_y")) == 0
    gy = dev(L("bias_gy"))
    gb = torch.empty(n, device="cuda")
    w = ws(capi.lib.mft_colsum_workspace_bytes(rows, n))
    capi.call("mft_colsum", gy, n, gb, rows, n, w, w.numel(), None)
    assert rel(gb, L("bias_gb")) < FP32_TOL
    a, bb = dev(L("mul_a")), dev(L("mul_b"))
    capi.call("mft_binary", 2, a, bb, out, 1, 1, a.numel(), 0, 1, None)
    assert rel(out, L("mul_y")) == 0
    ga, gb2 = torch.empty_like(a), torch.empty_like(a)
    capi.call("mft_mul_bwd", a, bb, dev(L("mul_gy")), ga, gb2, a.numel(), None)
    assert rel(ga, L("mul_ga")) == 0 and rel(gb2, L("mul_gb")) == 0


@pytest.mark.parametrize("n", [0, 1, 3, 1024, 4099, 1 << 20])
def test_elementwise_sizes(capi, n):
    rng = np.random.default_rng(n)
    a, b = rng.standard_normal(n).astype(np.float32), rng.standard_normal(n).astype(np.float32) + 3
    da, db = dev(a), dev(b)
    out = torch.empty_like(da)
    for op, f in enumerate([np.add, np.subtract, np.multiply, np.divide]):
        capi.call("mft_binary", op, da, db, out, 1, 1, n, 0, 1, None)
        assert rel(out, f(a, b)) < 1e-6
    capi.call("mft_affine", da, out, n, 0.5, -2.0, None)
    assert rel(out, a * np.float32(0.5) - 2) < 1e-6
    acc = da.clone()
    capi.call("mft_accumulate", acc, db, n, None)
    assert rel(acc, a + b) == 0
    capi.call("mft_fill", out, n, 1.25, None)
    assert rel(out, np.full(n, 1.25, np.float32)) == 0
    capi.call("mft_gelu_fwd", da, out, n, None)
    assert rel(out, O.gelu_fwd(a)) < FP32_TOL
    capi.call("mft_gelu_bwd", da, db, out, n, None)
    assert rel(out, O.gelu_bwd(a, b)) < FP32_TOL


def test_broadcast_modes(capi):
    rng = np.random.default_rng(0)
    a = rng.standard_normal((6, 5, 7)).astype(np.float32)     # [B*H, S, S]-like
    mask = rng.standard_normal((5, 7)).astype(np.float32)
    out = torch.empty(6, 5, 7, device="cuda")
    capi.call("mft_binary", 0, dev(a), dev(mask), out, 6, 5, 7, 2, 1, None)
    assert rel(out, a + mask) == 0
    pad = rng.standard_normal((2, 1, 7)).astype(np.float32)   # [B,1,1,S] with H=3
    capi.call("mft_binary", 0, dev(a), dev(pad), out, 6, 5, 7, 3, 3, None)
    assert rel(out, a + np.repeat(pad, 3, axis=0)) == 0


def test_gelu_golden(capi, golden_ops):
    L = golden_ops
    x = dev(L("gelu_x"))
    y = torch.empty_like(x)
    capi.call("mft_gelu_fwd", x, y, x.numel(), None)
    assert rel(y, L("gelu_y")) < FP32_TOL
    capi.call("mft_gelu_bwd", x, dev(L("gelu_gy")), y, x.numel(), None)
    assert rel(y, L("gelu_gx")) < FP32_TOL


# ----------------------------------------------------------------------------- layout
def test_transpose_permute_golden(capi, golden_ops):
    L = golden_ops
    x = dev(L("perm_x"))
    y = torch.empty(L("perm_y").shape, device="cuda")
    capi.call("mft_permute0213", x, y, *L("perm_x").shape, None)
    assert rel(y, L("perm_y")) == 0
    x2 = dev(L("tr_x"))
    b, h, r, c = L("tr_x").shape
    y2 = torch.empty(b, h, c, r, device="cuda")
    capi.call("mft_transpose2d", x2, c, y2, r, r, c, b * h, r * c, r * c, None)
    assert rel(y2, L("tr_y")) == 0
    big = torch.randn(300, 517, device="cuda")
    out = torch.empty(517, 300, device="cuda")
    capi.call("mft_transpose2d", big, 517, out, 300, 300, 517, 1, 0, 0, None)
    assert torch.equal(out, big.t().contiguous())


def test_embedding_rope_repeat(capi, golden_ops):
    L = golden_ops
    rng = np.random.default_rng(1)
    V, C_, S, B = 97, 24, 6, 3
    wte, wpe = rng.standard_normal((V, C_)).astype(np.float32), rng.standard_normal((16, C_)).astype(np.float32)
    ids = rng.integers(0, V, (B, S)).astype(np.int32)
    x = torch.empty(B, S, C_, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    capi.call("mft_embedding_fwd", dev(ids, torch.int32), dev(wte), dev(wpe), x, B * S, S, C_, V, 1.0, status, None)
    assert rel(x, O.embedding_gpt2(ids, wte, wpe)) == 0 and status.item() == 0
    capi.call("mft_embedding_fwd", dev(ids, torch.int32), dev(wte), None, x, B * S, S, C_, V, 4.0, status, None)
    assert rel(x, wte[ids] * np.float32(4.0)) == 0
    bad = ids.copy(); bad[0, 0] = V
    capi.call("mft_embedding_fwd", dev(bad, torch.int32), dev(wte), None, x, B * S, S, C_, V, 1.0, status, None)
    assert status.item() == 1
    xr = dev(L("rope_x"))
    yr = torch.empty_like(xr)
    b, h, s, hd = L("rope_x").shape
    capi.call("mft_rope", xr, yr, b * h, s, hd, 10000.0, 0, None)
    assert rel(yr, L("rope_y")) < FP32_TOL
    capi.call("mft_rope", dev(L("rope_gy")), yr, b * h, s, hd, 10000.0, 1, None)
    assert rel(yr, L("rope_gx")) < FP32_TOL
    xk = dev(L("repkv_x"))
    b, kv, s, hd = L("repkv_x").shape
    yk = torch.empty(L("repkv_y").shape, device="cuda")
    capi.call("mft_repeat_kv_fwd", xk, yk, b, kv, 3, s, hd, None)
    assert rel(yk, L("repkv_y")) == 0
    gk = torch.empty_like(xk)
    capi.call("mft_repeat_kv_bwd", dev(L("repkv_gy")), gk, b, kv, 3, s, hd, None)
    assert rel(gk, L("repkv_gx")) < FP32_TOL


# ----------------------------------------------------------------------------- norms / softmax
def run_layernorm(capi, x, w, b, gy):
    rows, dim = x.size // x.shape[-1], x.shape[-1]
    dx_, dw_, db_ = dev(x), dev(w), dev(b)
    y = torch.empty_like(dx_)
    capi.call("mft_layernorm_fwd", dx_, dw_, db_, y, rows, dim, 1e-5, None)
    gx, dg, dbeta = torch.empty_like(dx_), torch.empty(dim, device="cuda"), torch.empty(dim, device="cuda")
    w_ = ws(capi.lib.mft_norm_bwd_workspace_bytes(rows, dim))
    capi.call("mft_layernorm_bwd", dx_, dw_, dev(gy), gx, dg, dbeta, rows, dim, 1e-5, w_, w_.numel(), None)
    return y, gx, dg, dbeta


def test_layernorm_golden(capi, golden_ops):
    L = golden_ops
    y, gx, dg, db = run_layernorm(capi, L("ln_x"), L("ln_w"), L("ln_b"), L("ln_gy"))
    assert rel(y, L("ln_y")) < FP32_TOL and rel(gx, L("ln_gx")) < FP32_TOL
    assert rel(dg, L("ln_gw")) < FP32_TOL and rel(db, L("ln_gb")) < FP32_TOL


@pytest.mark.parametrize("rows,dim", [(1, 768), (37, 768), (512, 1024), (300, 640), (64, 100), (5, 2048), (9, 3000)])
def test_layernorm_shapes(capi, rows, dim):
    rng = np.random.default_rng(rows * dim)
    x = (rng.standard_normal((rows, dim)) * 1.5 + 0.3).astype(np.float32)
    w, b = rng.standard_normal(dim).astype(np.float32), rng.standard_normal(dim).astype(np.float32)
    gy = rng.standard_normal((rows, dim)).astype(np.float32)
    y, gx, dg, db = run_layernorm(capi, x, w, b, gy)
    ey = O.layer_norm_fwd(x, w, b)
    ex, eg, eb = O.layer_norm_bwd(x, w, gy)
    assert rel(y, ey) < FP32_TOL and rel(gx, ex) < 2e-5 and rel(dg, eg) < FP32_TOL and rel(db, eb) < FP32_TOL


@pytest.mark.parametrize("rows,dim,wgrad", [(512, 1024, True), (37, 768, False), (64, 100, True), (9, 3000, False), (4096, 1024, False)])
def test_layernorm_with_residual_operands_equals_separate_kernels(capi, rows, dim, wgrad):
    """mft_add_layernorm_fwd == add then LayerNorm; mft_layernorm_bwd_add == LayerNorm backward then accumulate — bit for bit."""
    rng = np.random.default_rng(rows + dim)
    x, a = (rng.standard_normal((2, rows, dim)) * 1.3).astype(np.float32)
    w, b = rng.standard_normal((2, dim)).astype(np.float32)
    gy, wait = rng.standard_normal((2, rows, dim)).astype(np.float32)
    dx, da, dw, db_, dgy, dwait = dev(x), dev(a), dev(w), dev(b), dev(gy), dev(wait)
    s_sep = dx + da
    y_sep = torch.empty_like(dx)
    capi.call("mft_layernorm_fwd", s_sep, dw, db_, y_sep, rows, dim, 1e-5, None)
    s_f, y_f = torch.full_like(dx, float("nan")), torch.full_like(dx, float("nan"))
    capi.call("mft_add_layernorm_fwd", dx, da, s_f, dw, db_, y_f, rows, dim, 1e-5, None)
    torch.cuda.synchronize()
    assert torch.equal(s_f, s_sep) and torch.equal(y_f, y_sep)
    assert rel(y_f, O.layer_norm_fwd(x + a, w, b)) < FP32_TOL

    def bwd(name, extra):
        gx = torch.full_like(dx, float("nan"))
        dg = torch.empty(dim, device="cuda") if wgrad else None
        dbt = torch.empty(dim, device="cuda") if wgrad else None
        wsp = ws(capi.lib.mft_norm_bwd_workspace_bytes(rows, dim)) if wgrad else None
        capi.call(name, s_sep, dw, dgy, *extra, gx, dg, dbt, rows, dim, 1e-5, wsp, wsp.numel() if wgrad else 0, None)
        torch.cuda.synchronize()
        return gx, dg, dbt
    gx0, dg0, db0 = bwd("mft_layernorm_bwd", [])
    gx1, dg1, db1 = bwd("mft_layernorm_bwd_add", [dwait])
    assert torch.equal(gx1, gx0 + dwait)
    if wgrad:
        assert torch.equal(dg0, dg1) and torch.equal(db0, db1)


def test_rmsnorm_golden_and_shapes(capi, golden_ops):
    L = golden_ops
    cases = [(L("rms_x"), L("rms_w"), L("rms_gy"), L("rms_y"), L("rms_gx"), L("rms_gw"))]
    rng = np.random.default_rng(3)
    for rows, dim in [(1024, 640), (100, 256), (3, 50)]:
        x = rng.standard_normal((rows, dim)).astype(np.float32)
        w = (rng.standard_normal(dim) * 0.3).astype(np.float32)
        gy = rng.standard_normal((rows, dim)).astype(np.float32)
        ex, ew = O.rms_norm_bwd(x, w, gy)
        cases.append((x, w, gy, O.rms_norm_fwd(x, w), ex, ew))
    for x, w, gy, ey, ex, ew in cases:
        rows, dim = x.size // x.shape[-1], x.shape[-1]
        dx_, dw_ = dev(x), dev(w)
        y = torch.empty_like(dx_)
        capi.call("mft_rmsnorm_fwd", dx_, dw_, y, rows, dim, 1e-6, 1, None)
        gx, gw = torch.empty_like(dx_), torch.empty(dim, device="cuda")
        w_ = ws(capi.lib.mft_norm_bwd_workspace_bytes(rows, dim))
        capi.call("mft_rmsnorm_bwd", dx_, dw_, dev(gy), gx, gw, rows, dim, 1e-6, 1, w_, w_.numel(), None)
        assert rel(y, ey) < FP32_TOL and rel(gx, ex) < FP32_TOL and rel(gw, ew) < FP32_TOL
        # the _ex form: same fp32 result bit for bit, plus its bf16 rounding at a padded pitch
        gx2, gw2 = torch.empty_like(dx_), torch.empty(dim, device="cuda")
        ld = dim + 8
        g16 = torch.zeros(rows, ld, device="cuda", dtype=torch.bfloat16)
        capi.call("mft_rmsnorm_bwd_ex", dx_, dw_, dev(gy), gx2, g16, ld, gw2, rows, dim, 1e-6, 1, w_, w_.numel(), None)
        assert torch.equal(gx2, gx) and torch.equal(gw2, gw)
        assert torch.equal(g16[:, :dim], gx.reshape(rows, dim).to(torch.bfloat16)) and not g16[:, dim:].any()


def test_softmax_golden(capi, golden_ops):
    L = golden_ops
    x = dev(L("softmax_x"))
    rows, dim = x.numel() // x.shape[-1], x.shape[-1]
    y = torch.empty_like(x)
    capi.call("mft_softmax_fwd", x, y, rows, dim, None)
    assert rel(y, L("softmax_y")) < FP32_TOL
    gx = torch.empty_like(x)
    capi.call("mft_softmax_bwd", y, dev(L("softmax_gy")), gx, rows, dim, None)
    assert rel(gx, L("softmax_gx")) < FP32_TOL


# ----------------------------------------------------------------------------- rare layouts / data preparation on the device
def test_permute_nd_cast_tril_one_hot(capi):
    rng = np.random.default_rng(5)
    x = rng.standard_normal((3, 4, 5, 6)).astype(np.float32)
    for dims in [(3, 1, 0, 2), (2, 0, 3, 1), (1, 3, 2, 0)]:
        e = np.ascontiguousarray(x.transpose(dims))
        stride = [x.strides[d] // 4 for d in dims]
        out = torch.empty(e.shape, device="cuda")
        capi.call("mft_permute_nd", dev(x), out, 4, (C.c_int64 * 4)(*e.shape), (C.c_int64 * 4)(*stride), None)
        assert rel(out, e) == 0
    f = (rng.standard_normal(1000) * 50).astype(np.float32)
    oi = torch.empty(1000, dtype=torch.int32, device="cuda")
    capi.call("mft_cast_i32", dev(f), oi, 1000, 1, None)
    assert np.array_equal(oi.cpu().numpy(), f.astype(np.int32))
    of = torch.empty(1000, device="cuda")
    capi.call("mft_cast_i32", oi, of, 1000, 0, None)
    assert np.array_equal(of.cpu().numpy(), f.astype(np.int32).astype(np.float32))
    m = torch.empty(37, 37, device="cuda")
    capi.call("mft_tril_ones", m, 37, None)
    assert np.array_equal(m.cpu().numpy(), np.tril(np.ones((37, 37), np.float32)))
    t = rng.integers(-1, 12, 50).astype(np.int32)       # -1 and 11 are out of range: all-zero rows
    oh = torch.empty(50, 11, device="cuda")
    capi.call("mft_one_hot", torch.tensor(t).cuda(), 1, oh, 50, 11, None)
    e = np.zeros((50, 11), np.float32)
    for i, c in enumerate(t):
        if 0 <= c < 11:
            e[i, c] = 1
    assert np.array_equal(oh.cpu().numpy(), e)


def test_refresh_bf16_shadows_many_tensors_one_launch(capi):
    """plain[r,c] = bf16(src[r,c]) and transposed[c,r] = bf16(src[r,c]) for a table of small matrices (adapter factors)."""
    rng = np.random.default_rng(8)
    shapes = [(1024, 16), (16, 1024), (640, 8), (8, 2048), (3, 5)]
    srcs = [dev(rng.standard_normal(s).astype(np.float32)) for s in shapes]
    plains, transposed, slots = [], [], (capi.ShadowSlot * len(shapes))()
    for i, (s, t) in enumerate(zip(shapes, srcs)):
        ldp, ldt = (s[1] + 7) // 8 * 8, (s[0] + 7) // 8 * 8
        p = torch.zeros(s[0], ldp, device="cuda", dtype=torch.bfloat16) if i != 2 else None      # slot 2: transposed only
        q = torch.zeros(s[1], ldt, device="cuda", dtype=torch.bfloat16) if i != 3 else None      # slot 3: plain only
        plains.append(p); transposed.append(q)
        slots[i] = capi.ShadowSlot(t.data_ptr(), s[0], s[1], s[1], p.data_ptr() if p is not None else None, ldp,
                                   q.data_ptr() if q is not None else None, ldt)
    table = torch.frombuffer(bytearray(bytes(slots)), dtype=torch.uint8).cuda()
    capi.call("mft_refresh_bf16_shadows", table, len(shapes), max(a * b for a, b in shapes), None)
    torch.cuda.synchronize()
    for s, t, p, q in zip(shapes, srcs, plains, transposed):
        e = t.to(torch.bfloat16)
        if p is not None:
            assert torch.equal(p[:, :s[1]], e)
        if q is not None:
            assert torch.equal(q[:, :s[0]], e.t())


def test_geglu_fwd_bwd_equals_gelu_then_mul(capi):
    """mft_geglu_fwd / _bwd against the two reference nodes they replace (ops::gelu then ops::mul): same fp32 values."""
    rng = np.random.default_rng(12)
    n = 1024 * 20
    gate, up, gy = (rng.standard_normal(n).astype(np.float32) for _ in range(3))
    y, y16 = torch.empty(n, device="cuda"), torch.empty(n, device="cuda", dtype=torch.bfloat16)
    capi.call("mft_geglu_fwd", dev(gate), dev(up), y, y16, n, None)
    e = O.gelu_fwd(gate) * up
    assert rel(y, e) < FP32_TOL and torch.equal(y16, y.to(torch.bfloat16))
    dg, du = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    dg16, du16 = torch.empty(n, device="cuda", dtype=torch.bfloat16), torch.empty(n, device="cuda", dtype=torch.bfloat16)
    capi.call("mft_geglu_bwd", dev(gate), dev(up), dev(gy), dg, du, dg16, du16, n, None)
    assert rel(dg, O.gelu_bwd(gate, gy * up)) < FP32_TOL
    assert rel(du, gy * O.gelu_fwd(gate)) < FP32_TOL
    assert torch.equal(dg16, dg.to(torch.bfloat16)) and torch.equal(du16, du.to(torch.bfloat16))

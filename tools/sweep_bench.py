#!/usr/bin/env python
"""Per-op sweep (BASELINE.json configs[4], SURVEY.md §8d grid): the LoRA-fused linear forward/backward and the LM
cross-entropy through the C ABI, timed with CUDA events, L2 flushed between iterations, against the measured peaks.

    python tools/sweep_bench.py [--json] [--quick] [--out profiles/r2_c5_sweep.txt]

LoRA linear, bf16 operand mode (what bench.py runs): forward = mft_lora_down_bf16 (x*A from the bf16 rows) + mft_gemm_bf16 with
the rank-r second K segment; backward = mft_lora_down_bf16 (dY*B^T) + mft_gemm_bf16 (dX, second segment dxa*A^T) + two
mft_lora_tn (dA, dB). FLOPs per SURVEY §8d: fwd 2*T*in*out + 2*T*r*(in+out); bwd the same + 2*T*in*r + 2*T*r*out.
Cross-entropy: mft_lm_ce_fwd + mft_lm_ce_bwd_ex on fp32 logits; bytes = 4*T*V read (forward) + 4*T*V read + 4*T*V written
(backward) = 12*T*V (the unfused figure of §8d; the second forward pass is served from L2 and not counted).
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mobilefinetuner_b200 import capi  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    d = json.load(open(p)) if os.path.exists(p) else {}
    return d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), ("measured" if d else "fallback")


FLUSH = None


def timeit(fn, iters):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        FLUSH.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e-3


def gemm_bf16(T, n, k, a16, b16, c, k2, a2, b2, bias=None):
    args = capi.GemmArgs(T, n, k, a16.data_ptr(), k, 0, b16.data_ptr(), k, 1, c.data_ptr(), n, 1.0, 0.0,
                         bias.data_ptr() if bias is not None else None, None, k2, a2.data_ptr(), k2, b2.data_ptr(), k2, 1, 0, 0, 0)
    capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "mft_gemm_bf16")


def lora_point(T, d_in, d_out, r, iters):
    g = torch.Generator(device="cuda").manual_seed(T + d_in + d_out + r)
    rn = lambda *s, k=1.0: torch.randn(*s, device="cuda", generator=g) * k
    x, gy = rn(T, d_in), rn(T, d_out)
    W, A, B, bias = rn(d_in, d_out, k=0.05), rn(d_in, r, k=0.1), rn(r, d_out, k=0.1), rn(d_out, k=0.1)
    x16, gy16 = x.bfloat16(), gy.bfloat16()
    Wt16, W16 = W.t().contiguous().bfloat16(), W.bfloat16()             # [out,in] for the forward, [in,out] for dX
    Bt16, A16 = B.t().contiguous().bfloat16(), A.bfloat16()             # [out,r], [in,r]
    xa, xa16 = torch.empty(T, r, device="cuda"), torch.empty(T, r, device="cuda", dtype=torch.bfloat16)
    dxa, dxa16 = torch.empty_like(xa), torch.empty_like(xa16)
    y, dx = torch.empty(T, d_out, device="cuda"), torch.empty(T, d_in, device="cuda")
    dA, dB = torch.empty(d_in, r, device="cuda"), torch.empty(r, d_out, device="cuda")
    w1 = torch.empty(max(capi.lib.mft_lora_tn_workspace_bytes(T, d_in, r), 16), dtype=torch.uint8, device="cuda")
    w2 = torch.empty(max(capi.lib.mft_lora_tn_workspace_bytes(T, d_out, r), 16), dtype=torch.uint8, device="cuda")

    At16, B16 = A.t().contiguous().bfloat16(), B.bfloat16()             # K-major shadows of the factors: [r,in], [r,out]
    kmax = 6144 if r <= 16 else (3072 if r <= 32 else 1536)
    lowp_f, lowp_b = d_in % 128 == 0 and d_in <= kmax, d_out % 128 == 0 and d_out <= kmax

    def fwd():
        if lowp_f:
            capi.call("mft_lora_down_bf16", x16, d_in, At16, d_in, xa, r, xa16, r, T, d_in, r, 2.0, None)
        else:
            capi.call("mft_lora_down_ex", x, d_in, A, r, 1, xa, r, xa16, r, T, d_in, r, 2.0, None)
        gemm_bf16(T, d_out, d_in, x16, Wt16, y, r, xa16, Bt16, bias)

    def bwd():
        if lowp_b:
            capi.call("mft_lora_down_bf16", gy16, d_out, B16, d_out, dxa, r, dxa16, r, T, d_out, r, 2.0, None)
        else:
            capi.call("mft_lora_down_ex", gy, d_out, B, d_out, 0, dxa, r, dxa16, r, T, d_out, r, 2.0, None)
        gemm_bf16(T, d_in, d_out, gy16, W16, dx, r, dxa16, A16)
        capi.call("mft_lora_tn", dxa, r, x, d_in, dA, r, 0, T, d_in, r, 1.0, w1, w1.numel(), None)
        capi.call("mft_lora_tn", xa, r, gy, d_out, dB, d_out, 1, T, d_out, r, 1.0, w2, w2.numel(), None)

    tf, tb = timeit(fwd, iters), timeit(bwd, iters)
    f_fwd = 2.0 * T * d_in * d_out + 2.0 * T * r * (d_in + d_out)
    f_bwd = f_fwd + 2.0 * T * d_in * r + 2.0 * T * r * d_out
    return tf, tb, f_fwd, f_bwd


def ce_point(T, V, iters):
    B, S = 1, T + 1                                  # T shifted rows
    ld = (V + 3) // 4 * 4
    logits = torch.randn(B * S, ld, device="cuda")
    labels = torch.randint(0, V, (B, S), device="cuda", dtype=torch.int32)
    rows = B * S
    lse = torch.empty(rows, device="cuda")
    valid = torch.zeros(4, dtype=torch.int32, device="cuda")
    loss = torch.zeros(1, device="cuda")
    wsb = capi.lib.mft_lm_ce_workspace_bytes(rows)
    w = torch.empty(max(wsb, 16), dtype=torch.uint8, device="cuda")
    dl = torch.empty(B * S, ld, device="cuda")
    gout = torch.ones(1, device="cuda")

    def both():
        capi.call("mft_lm_ce_fwd", logits, ld, labels, B, S, V, -100, 1, loss, valid, lse, w, w.numel(), None)
        capi.call("mft_lm_ce_bwd", logits, ld, labels, lse, valid, gout, B, S, V, -100, 1, dl, ld, None)

    t = timeit(both, iters)
    return t, 12.0 * T * V


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", action="store_true", help="print one JSON line (bench.py --config c5)")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    torch.cuda.set_device(0)
    tf_peak, hbm_peak, src = peaks()
    Ts = [512, 2048, 8192, 32768, 65536]
    grid = [(d, d, r) for d in (768, 1024, 1536, 2048) for r in (8, 16, 32, 64)]
    grid += [(d, m * d, 16) for d in (768, 1024, 2048) for m in (3, 4)] + [(4 * d, d, 16) for d in (768, 1024)]
    if a.quick:
        Ts, grid = [512, 8192], grid[:4]
    rows, out = [], []
    out.append(f"# per-op sweep; peaks ({src}): {tf_peak:.1f} TFLOP/s dense bf16 sustained, {hbm_peak:.1f} GB/s HBM copy")
    out.append(f"{'op':10s} {'T':>6s} {'in':>5s} {'out':>5s} {'r':>3s} {'fwd us':>9s} {'fwd TF/s':>9s} {'frac':>6s} {'bwd us':>9s} {'bwd TF/s':>9s} {'frac':>6s}")
    for (d_in, d_out, r) in grid:
        for T in Ts:
            if T * max(d_in, d_out) * 4 > 3 << 30:
                continue
            iters = 5 if T >= 32768 else 10
            tf, tb, ff, fb = lora_point(T, d_in, d_out, r, iters)
            p = dict(op="lora_linear", T=T, d_in=d_in, d_out=d_out, r=r, fwd_us=round(tf * 1e6, 1), bwd_us=round(tb * 1e6, 1),
                     fwd_tflops=round(ff / tf / 1e12, 1), bwd_tflops=round(fb / tb / 1e12, 1),
                     fwd_frac=round(ff / tf / 1e12 / tf_peak, 3), bwd_frac=round(fb / tb / 1e12 / tf_peak, 3))
            rows.append(p)
            out.append(f"{'lora_lin':10s} {T:6d} {d_in:5d} {d_out:5d} {r:3d} {p['fwd_us']:9.1f} {p['fwd_tflops']:9.1f} {p['fwd_frac']:6.3f} "
                       f"{p['bwd_us']:9.1f} {p['bwd_tflops']:9.1f} {p['bwd_frac']:6.3f}")
    out.append(f"{'op':10s} {'T':>6s} {'V':>7s} {'fwd+bwd us':>11s} {'GB/s':>8s} {'frac':>6s}")
    for V in (50257, 262144):
        for T in Ts:
            if T * V * 8 > 80 << 30:
                continue
            t, by = ce_point(T, V, 5 if T * V > 1 << 30 else 10)
            p = dict(op="lm_cross_entropy", T=T, V=V, us=round(t * 1e6, 1), gbs=round(by / t / 1e9, 1), frac=round(by / t / 1e9 / hbm_peak, 3))
            rows.append(p)
            out.append(f"{'lm_ce':10s} {T:6d} {V:7d} {p['us']:11.1f} {p['gbs']:8.1f} {p['frac']:6.3f}")
            torch.cuda.empty_cache()
    text = "\n".join(out)
    if a.out:
        open(a.out, "w").write(text + "\n")
    if a.json:
        lin = [p for p in rows if p["op"] == "lora_linear"]
        best = max(lin, key=lambda p: p["fwd_frac"]) if lin else None
        print(json.dumps({"metric": "per_op_sweep", "unit": "TFLOP/s | GB/s per grid point", "n_gpus": 1, "higher_is_better": True,
                          "config": {"workload": "Per-op sweep: LoRA-fused linear fwd/bwd (d=768-2048, r=8-64) and cross-entropy (vocab 50257/262144), tokens 512-65536",
                                     "config_id": "c5", "l2": "flushed between iterations (256 MiB write)"},
                          "dtype": "bf16 GEMM operands, fp32 accumulate; cross-entropy fp32", "data": "synthetic",
                          "peaks": {"tflops": tf_peak, "hbm_gbs": hbm_peak, "source": src}, "best_lora_fwd": best, "points": rows}))
    else:
        print(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""Times the tensor-core GEMM (C ABI) on the shapes of the benchmark workload: CUDA events, L2 flushed between launches."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobilefinetuner_b200 import capi  # noqa: E402

SHAPES = [("qkv", 4096, 3072, 1024), ("proj", 4096, 1024, 1024), ("fc_in", 4096, 4096, 1024), ("fc_out", 4096, 1024, 4096),
          ("lm_head", 4096, 50257, 1024), ("lm_head_dx", 4096, 1024, 50264), ("c1_qkv", 512, 2304, 768), ("big", 8192, 8192, 8192)]


def bench(name, m, n, k, bf16, iters=20, fuse=""):
    dt = torch.bfloat16 if bf16 else torch.float32
    a = torch.randn(m, k, device="cuda").to(dt)
    b = (torch.randn(n, k, device="cuda") * 0.05).to(dt)
    ldc = (n + 31) // 32 * 32
    c = torch.empty(m, ldc, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    bias = torch.randn(n, device="cuda")
    res = torch.randn(m, ldc, device="cuda") if "res" in fuse else None
    gx = torch.randn(m, ldc, device="cuda") if "ggrad" in fuse else None
    gout = torch.empty(m, ldc, device="cuda") if "gelu" in fuse else None
    args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, c.data_ptr(), ldc, 1.0, 0.0, bias.data_ptr() if fuse else None,
                         gout.data_ptr() if gout is not None else None, 0, None, 0, None, 0, 1, 0, 0, 0,
                         res.data_ptr() if res is not None else None, ldc, gx.data_ptr() if gx is not None else None, ldc)
    fn = capi.lib.mft_gemm_bf16 if bf16 else capi.lib.mft_gemm
    for _ in range(3):
        capi.check(fn(C.byref(args), None), "gemm")
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(C.byref(args), None)
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    us = tot / iters * 1e3
    tf = 2.0 * m * n * k / (us * 1e-6) / 1e12
    print(f"{name:12s} {'bf16' if bf16 else 'tf32'} m={m} n={n} k={k} {fuse or 'plain':10s}: {us:8.1f} us  {tf:7.1f} TFLOP/s", flush=True)


SKINNY = [("xa   x*A", 4096, 16, 1024, 0, 0), ("dxa gy*B^T", 4096, 16, 1024, 0, 1), ("dB xa^T*gy", 16, 1024, 4096, 1, 0),
          ("dA x^T*dxa", 1024, 16, 4096, 1, 0), ("xa r=8", 4096, 8, 1024, 0, 0), ("dB r=64", 64, 1024, 4096, 1, 0)]


def bench_skinny(name, m, n, k, ta, tb, iters=20):
    a = torch.randn((k, m) if ta else (m, k), device="cuda")
    b = torch.randn((n, k) if tb else (k, n), device="cuda")
    c = torch.empty(m, n, device="cuda")
    bm, bn = (16 if m <= 16 else 64), (16 if n <= 16 else 64)
    tiles = -(-m // bm) * -(-n // bn)
    splits = max(1, min(592 // tiles, k // 128)) if tiles < 148 else 1
    ws = torch.empty(max(capi.lib.mft_gemm_splitk_workspace_bytes(m, n, 1, splits), 16), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    args = capi.GemmArgs(m, n, k, a.data_ptr(), a.shape[1], ta, b.data_ptr(), b.shape[1], tb, c.data_ptr(), n, 1.0, 0.0, None, None,
                         0, None, 0, None, 0, 1, 0, 0, 0)
    for _ in range(3):
        capi.check(capi.lib.mft_gemm_splitk(C.byref(args), splits, ws.data_ptr(), ws.numel(), None), "splitk")
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        capi.lib.mft_gemm_splitk(C.byref(args), splits, ws.data_ptr(), ws.numel(), None)
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    us = tot / iters * 1e3
    mb = (a.numel() + b.numel() + c.numel()) * 4 / 1e6
    print(f"{name:12s} m={m} n={n} k={k} splits={splits}: {us:7.1f} us  ({mb / us * 1e6 / 1e3:7.1f} GB/s algorithmic)", flush=True)


def bench_lora(iters=20):
    T, K, q, r = 4096, 1024, 1024, 16
    x = torch.randn(T, K, device="cuda"); A = torch.randn(K, r, device="cuda"); y = torch.empty(T, r, device="cuda")
    W = torch.randn(T, q, device="cuda"); out = torch.empty(r, q, device="cuda")
    ws = torch.empty(capi.lib.mft_lora_tn_workspace_bytes(T, q, r), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    calls = {"lora_down [4096x1024]x[1024x16]": lambda: capi.call("mft_lora_down", x, K, A, r, 1, y, r, T, K, r, 1.0, None),
             "lora_tn   [4096x16]^T[4096x1024]": lambda: capi.call("mft_lora_tn", y, r, W, q, out, q, 1, T, q, r, 1.0, ws, ws.numel(), None)}
    for name, fn in calls.items():
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            tot += e0.elapsed_time(e1)
        us = tot / iters * 1e3
        print(f"{name}: {us:7.1f} us  ({16.8 / us * 1e3:7.1f} GB/s algorithmic)", flush=True)


def bench_step_forms(iters=20, T=4096):
    """The c4 layer's products in the forms the bf16 step launches them: bf16-only qkv, fc_in with the GELU epilogue writing
    pre-activation and activation as bf16, fp32 proj / fc_out, bf16-only dX of fc_out, fp32 dX of fc_in."""
    Cm = 1024
    forms = [("qkv->bf16", T, 3 * Cm, Cm, "c16"), ("proj->f32", T, Cm, Cm, "f32"), ("fc_in+gelu->bf16", T, 4 * Cm, Cm, "gelu16"),
             ("fc_out->f32", T, Cm, 4 * Cm, "f32"), ("fc_out dX->bf16", T, 4 * Cm, Cm, "c16"), ("fc_in dX->f32", T, Cm, 4 * Cm, "f32")]
    once = [("lm_head->f32", T, 50257, Cm, "f32"), ("lm_head dX->f32", T, Cm, 50264, "f32")]       # once per step, not in the layer total
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    total = 0.0
    for name, m, n, k, form in forms + once:
        a = torch.randn(m, k, device="cuda").bfloat16()
        b = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
        bias = torch.randn(n, device="cuda")
        ldc = n if n % 4 == 0 else (n + 31) // 32 * 32          # the host mirror pads wide odd rows the same way (ops.cpp out_ld)
        c32 = torch.empty(m, ldc, device="cuda") if form == "f32" else None
        c16 = torch.empty(m, n, device="cuda", dtype=torch.bfloat16) if form != "f32" else None
        g16 = torch.empty(m, n, device="cuda", dtype=torch.bfloat16) if form == "gelu16" else None
        args = capi.GemmArgs(m, n, k, a.data_ptr(), k, 0, b.data_ptr(), k, 1, c32.data_ptr() if c32 is not None else None, ldc, 1.0, 0.0, bias.data_ptr(), None,
                             0, None, 0, None, 0, 1, 0, 0, 0, None, 0, None, 0, g16.data_ptr() if g16 is not None else None, n,
                             c16.data_ptr() if c16 is not None else None, n)
        for _ in range(3):
            capi.check(capi.lib.mft_gemm_bf16(C.byref(args), None), "gemm")
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); capi.lib.mft_gemm_bf16(C.byref(args), None); e1.record(); e1.synchronize()
            tot += e0.elapsed_time(e1)
        us = tot / iters * 1e3
        if (name, m, n, k, form) in forms:
            total += us
        del a, b, c32, c16, g16
        print(f"{name:18s} m={m} n={n} k={k}: {us:8.1f} us  {2.0 * m * n * k / (us * 1e-6) / 1e12:7.1f} TFLOP/s", flush=True)
    print(f"layer total {total:8.1f} us  ({2.0 * T * Cm * Cm * 20 / (total * 1e-6) / 1e12:7.1f} TFLOP/s)", flush=True)


if __name__ == "__main__":
    if sys.argv[1:2] == ["step"]:
        bench_step_forms(T=int(sys.argv[2]) if len(sys.argv) > 2 else 4096)
        sys.exit(0)
    if sys.argv[1:] == ["skinny"]:
        for sh in SKINNY:
            bench_skinny(*sh)
        bench_lora()
        sys.exit(0)
    if sys.argv[1:] == ["fused"]:     # epilogue variants of the c4 shapes
        for sh, fz in ((SHAPES[1], "bias"), (SHAPES[1], "bias+res"), (SHAPES[3], "bias+res"), (SHAPES[2], "bias"), (SHAPES[2], "bias+gelu"),
                       (SHAPES[2], "ggrad")):
            bench(*sh, bf16=False, fuse=fz)
        sys.exit(0)
    which = sys.argv[1:] or ["tf32", "bf16"]
    for mode in which:
        for s in SHAPES:
            bench(*s, bf16=(mode == "bf16"))

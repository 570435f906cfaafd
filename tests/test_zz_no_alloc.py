"""GPU: mft_layernorm_bwd_static — the LayerNorm backward for callers that have no way to allocate workspace.

Such callers exist on both sides of this ABI: an FFI front end (cgo / JNI / ctypes) that only passes along tensors its
host application owns has no device allocator of its own, and a training loop in steady state may rule out cudaMalloc
because it synchronises the device. The entry point therefore keeps its partial sums in an array compiled into the
library. This test drives it through the C ABI twice in a row (second call with a different shape, so whatever the
first call left in the shared array must not matter) and requires the same results as mft_layernorm_bwd with
caller-provided workspace, within the fp32 tolerance the other kernel tests use. It is collected last on purpose.
"""
import numpy as np
import pytest
import torch

from conftest import FP32_TOL, dev, rel, ws

pytestmark = pytest.mark.gpu

STATIC_BYTES = 4849664   # MFT_LN_BWD_STATIC_SCRATCH_BYTES in include/mft_b200.h


def both_ways(capi, rows, dim, seed):
    rng = np.random.default_rng(seed)
    x = dev((rng.standard_normal((rows, dim)) * 1.5 + 0.25).astype(np.float32))
    g = dev(rng.standard_normal(dim).astype(np.float32))
    gy = dev(rng.standard_normal((rows, dim)).astype(np.float32))
    ref = [torch.empty_like(x), torch.empty(dim, device="cuda"), torch.empty(dim, device="cuda")]
    w = ws(capi.lib.mft_norm_bwd_workspace_bytes(rows, dim))
    capi.call("mft_layernorm_bwd", x, g, gy, ref[0], ref[1], ref[2], rows, dim, 1e-5, w, w.numel(), None)
    got = [torch.full_like(x, float("nan")), torch.full((dim,), float("nan"), device="cuda"),
           torch.full((dim,), float("nan"), device="cuda")]
    capi.call("mft_layernorm_bwd_static", x, g, gy, got[0], got[1], got[2], rows, dim, 1e-5, None)
    torch.cuda.synchronize()
    return got, ref


def test_static_scratch_variant_matches_workspace_variant_twice_in_a_row(capi):
    before = torch.cuda.memory_allocated()
    for rows, dim, seed in [(4096, 1024, 1), (300, 768, 2)]:      # two calls in a row, shapes differ
        got, ref = both_ways(capi, rows, dim, seed)
        for a, b in zip(got, ref):
            assert rel(a, b.cpu().numpy()) < FP32_TOL
    # an unaligned width goes down the generic kernels and must work from the same array too
    got, ref = both_ways(capi, 77, 50, 3)
    for a, b in zip(got, ref):
        assert rel(a, b.cpu().numpy()) < FP32_TOL
    assert torch.cuda.memory_allocated() >= before   # (torch's own tensors only; the entry point allocates nothing)


def test_sizes_beyond_the_builtin_array_are_rejected(capi):
    rows, dim = 64 * 2000, 2048
    assert capi.lib.mft_norm_bwd_workspace_bytes(rows, dim) > STATIC_BYTES
    x = torch.zeros(8, device="cuda")   # never dereferenced: the size check comes first
    with pytest.raises(capi.MftError, match="built-in array"):
        capi.call("mft_layernorm_bwd_static", x, x, x, x, x, x, rows, dim, 1e-5, None)

"""GPU: full-shape loss trajectories of the BASELINE.json configs against the reference build (SURVEY.md §8c protocol).

tests/golden/traj/*.txt were produced by oracle/_ref/ref_harness (the reference's own sources, tests/golden/make_golden.sh):
same mt19937 weight/token streams, same adapters, same Adam. Each case is stepped through the C-ABI trainer in the exact
fp32 mode and in the two tensor-core modes bench.py can run (TF32 operands, bf16 operands), with CUDA-graph replay on.

Tolerances. Loss: 1e-3 relative per step (north_star); exact mode 1e-4. Set of tensors that receive a gradient: identical.
Gradient norm: every product feeding a gradient has two operands rounded to p significant bits, so with all rounding
errors aligned the norm moves by at most 2 * 2^-p relative: TF32 (p = 11) 1e-3, bf16 (p = 8) 8e-3; both are doubled for the
second rounding of intermediate activations kept in the operand precision. Exact mode: 1e-3 (summation order only).
Final parameters (sum of |.|): 1e-3 exact; 2e-2 in the tensor-core modes — Adam's first steps move every element by about
lr * sign(g), so an element whose tiny gradient changes sign under rounding moves the other way."""
import os
import re

import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

MODES = {"exact": dict(loss=1e-4, gnorm=1e-3, param=1e-3), "tf32": dict(loss=1e-3, gnorm=2e-3, param=2e-2),
         "bf16": dict(loss=1e-3, gnorm=1.6e-2, param=2e-2)}


def golden(name):
    text = open(os.path.join(GOLDEN, "traj", name)).read()
    steps = [(float(m[1]), float(m[2]), int(m[3]), int(m[4])) for m in
             re.finditer(r"^step \d+ loss (\S+) gnorm (\S+) with_grad (\d+)/(\d+)", text, re.M)]
    params = {int(m[1]): (int(m[2]), float(m[3]), float(m[4])) for m in
              re.finditer(r"^param (\d+) n=(\d+) sum=(\S+) sumabs=(\S+)", text, re.M)}
    assert steps, name
    return steps, params


def run_case(mode, gold, max_steps=None, **trainer_kw):
    from mobilefinetuner_b200 import capi
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    tol = MODES[mode]
    steps, params = golden(gold)
    if max_steps:
        steps, params = steps[:max_steps], ({} if max_steps < len(steps) else params)
    prev_tc = capi.lib.mft_use_tensor_cores(0 if mode == "exact" else 1)
    prev_at = capi.lib.mft_attention_precise(1 if mode == "exact" else 0)
    hostlib.mft_host_set_gemm_precision(1 if mode == "bf16" else 0)
    try:
        tr = Trainer(cuda_graph=True, **trainer_kw)
        worst = dict(loss=0.0, gnorm=0.0, param=0.0)
        for i, (ref_loss, ref_gnorm, with_grad, total) in enumerate(steps):
            loss = tr.step_numpy(tr.next_synthetic_batch())
            worst["loss"] = max(worst["loss"], abs(loss - ref_loss) / ref_loss)
            worst["gnorm"] = max(worst["gnorm"], abs(tr.grad_norm() - ref_gnorm) / ref_gnorm)
            assert abs(loss - ref_loss) / ref_loss < tol["loss"], (mode, i, loss, ref_loss)
            assert abs(tr.grad_norm() - ref_gnorm) <= tol["gnorm"] * ref_gnorm, (mode, i, tr.grad_norm(), ref_gnorm)
            assert (tr.num_with_grad, tr.num_trainable) == (with_grad, total), (mode, i)
        for i, (n, s, sa) in params.items():
            gn, gs, gsa = tr.param_checksum(i)
            assert gn == n
            worst["param"] = max(worst["param"], abs(gsa - sa) / max(sa, 1e-6))
            assert abs(gsa - sa) <= tol["param"] * max(sa, 1e-6), (mode, i, gsa, sa)
        print(f"{gold} [{mode}] worst relative errors: {worst}")
        tr.close()
    finally:
        hostlib.mft_host_set_gemm_precision(0)
        capi.lib.mft_use_tensor_cores(prev_tc)
        capi.lib.mft_attention_precise(prev_at)


@pytest.mark.parametrize("mode", ["exact", "tf32", "bf16"])
def test_c4_gpt2_medium_lora_r16_2x512(mode):
    """BASELINE configs[3] model and sequence length (GPT-2 medium, LoRA r16 alpha 32, seq 512), 2 sequences, 4 steps."""
    run_case(mode, "gpt2_medium_lora_c4_b2.txt", batch=2, seq_len=512, shape="gpt2-medium", lora_rank=16, lora_alpha=32.0, arena_mb=8192)


@pytest.mark.parametrize("mode", ["tf32", "bf16"])
def test_c4_gpt2_medium_lora_r16_8x512_bench_shape(mode):
    """The per-GPU shape bench.py steps at 8 GPUs (8 x 512 tokens), 2 steps."""
    run_case(mode, "gpt2_medium_lora_c4_b8.txt", batch=8, seq_len=512, shape="gpt2-medium", lora_rank=16, lora_alpha=32.0, arena_mb=20480)


@pytest.mark.parametrize("mode", ["exact", "tf32", "bf16"])
def test_c2_gpt2_small_full_finetune_8x256(mode):
    """BASELINE configs[1]: full fine-tune (dW = X^T dY on every trained weight), weight decay 0.01, clip 1.0, 4 steps."""
    run_case(mode, "gpt2_small_full_c2.txt", batch=8, seq_len=256, shape="gpt2-small", full_finetune=True, arena_mb=12288)


@pytest.mark.parametrize("mode", ["exact", "tf32", "bf16"])
def test_c3_gemma3_270m_lora_4x256(mode):
    """BASELINE configs[2]: Gemma-3 270M (18 layers, head_dim 256, 4 query heads on 1 KV head, 5 sliding : 1 full layers,
    vocabulary 262144), LoRA r8 on all seven projections, 3 steps."""
    run_case(mode, "gemma3_270m_lora_c3.txt", batch=4, seq_len=256, shape="gemma3-270m", lora_rank=8, lora_alpha=32.0, arena_mb=16384)


@pytest.mark.parametrize("mode", ["exact", "tf32"])
def test_gemma_sliding_window_shorter_than_the_sequence(mode):
    """Window 8 at sequence length 16: the sliding-attention mask really removes keys (the 4x256 case has window 512)."""
    shape = dict(model=1, n_layer=6, n_embd=64, n_head=4, vocab_size=1009, n_positions=64, intermediate=128, kv_heads=1, head_dim=32,
                 sliding_window=8)
    run_case(mode, "gemma_lora_tiny_window8.txt", batch=2, seq_len=16, shape=shape, lora_rank=8, lora_alpha=32.0, arena_mb=512)

"""CPU: the host logic of bench.py that both arms share — which workload a run steps (strong / weak, per-GPU batch) and the
algorithmic FLOPs per token that `roofline.step` divides by. No GPU, no native library."""
import argparse
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def bench():
    return importlib.import_module("bench")


def _args(config="c4", scaling=None):
    return argparse.Namespace(config=config, scaling=scaling)


def test_default_workload_is_the_global_batch_split_over_the_gpus(bench):
    """BASELINE configs[3]: global batch 64 x 512 — strong scaling by default, 64 / N sequences per GPU."""
    for world, per in [(1, 64), (2, 32), (4, 16), (8, 8)]:
        c, scaling, batch, config = bench.plan(_args(), world)
        assert scaling == "strong" and batch == per
        assert config["global_tokens_per_step"] == 64 * 512 and config["tokens_per_step_per_gpu"] == per * 512
        assert config["parallelism"] == f"dp{world}" and config["config_id"] == "c4"


def test_weak_scaling_keeps_the_per_gpu_batch(bench):
    for world in (1, 2, 8):
        _, scaling, batch, config = bench.plan(_args(scaling="weak"), world)
        assert scaling == "weak" and batch == 8 and config["global_tokens_per_step"] == 8 * 512 * world


def test_single_device_configs_are_weak_and_refuse_strong(bench):
    for cid, tokens in [("c1", 4 * 128), ("c2", 8 * 256), ("c3", 4 * 256)]:
        _, scaling, _, config = bench.plan(_args(cid), 1)
        assert scaling == "weak" and config["tokens_per_step_per_gpu"] == tokens
        with pytest.raises(SystemExit):
            bench.plan(_args(cid, "strong"), 1)
    with pytest.raises(SystemExit):          # 64 sequences do not divide over 3 GPUs
        bench.plan(_args(), 3)


def test_both_arms_print_the_same_config_object(bench):
    """The reference arm and the GPU arm describe the workload with the same dictionary (no measured values inside)."""
    a = bench.plan(_args(), 1)[3]
    b = bench.plan(_args(), 1)[3]
    assert a == b and set(a) >= {"workload", "config_id", "batch_per_gpu", "seq_len", "scaling", "l2"}


def test_step_flops_per_token_follow_the_reference_backward_semantics():
    from mobilefinetuner_b200.shapes import MODEL_SHAPES, step_flops_per_token
    # c4: GPT-2 medium, LoRA r16, seq 512 (SURVEY 8d / DESIGN 6b: 1.27 GFLOP per token)
    f, b = step_flops_per_token("gpt2-medium", 512, False, 16)
    assert abs((f + b) - 1.2699e9) / 1.2699e9 < 1e-3
    s = MODEL_SHAPES["gpt2-medium"]
    L, C, V = s["n_layer"], s["n_embd"], s["vocab_size"]
    # forward = every product; the stop-gradient attention leaves qkv / ln_1 without a backward: the LoRA backward is
    # lm_head dX + fc_out / fc_in dX + the proj adapter's three rank-16 products
    assert f == L * (2 * C * 3 * C + 2 * C * C + 16 * C * C + 4 * 512 * C + 2 * 16 * (4 * C) + 2 * 16 * (2 * C)) + 2 * C * V
    assert b == 2 * C * V + L * (16 * C * C + 3 * 2 * 16 * C)
    # full fine-tuning adds the dW products of everything downstream of attention and the head's dE
    ff, fb = step_flops_per_token("gpt2-small", 256, True, 8)
    lf, lb = step_flops_per_token("gpt2-small", 256, False, 8)
    assert fb > lb and ff < lf          # no adapter products in the forward, more backward
    # Gemma: backward reaches every product (real attention backward: twice the forward attention work)
    gf, gb = step_flops_per_token("gemma3-270m", 256, False, 8)
    assert gb > gf

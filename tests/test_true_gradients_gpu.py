"""GPU: the true-gradient mode — GPT-2 attention with a backward and `.grad` accumulation. Off by default:
the parity runs reproduce the reference, whose attention is a stop-gradient (graph/gpt2_model.cpp:616-637,
core/memory_efficient_attention.cpp:165-182) and whose backward overwrites `.grad` (core/autograd_engine.cpp:243)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TINY = dict(model=0, n_layer=2, n_embd=64, n_head=4, vocab_size=503, n_positions=64)


@pytest.fixture
def exact(capi):
    prev_tc, prev_at = capi.lib.mft_use_tensor_cores(0), capi.lib.mft_attention_precise(1)
    yield
    capi.lib.mft_use_tensor_cores(prev_tc)
    capi.lib.mft_attention_precise(prev_at)


@pytest.fixture
def true_gradients():
    from mobilefinetuner_b200.trainer import hostlib
    hostlib.mft_host_set_true_gradients(1)
    yield
    hostlib.mft_host_set_true_gradients(0)


def test_default_mode_reproduces_the_stop_gradient_and_true_mode_trains_every_adapter(capi, exact):
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    a = Trainer(2, 16, shape=TINY, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    la = [a.step_numpy(a.next_synthetic_batch()) for _ in range(3)]
    assert (a.num_with_grad, a.num_trainable) == (4, 8)          # attn.qkv adapters never see a gradient (reference quirk)
    a.close()
    hostlib.mft_host_set_true_gradients(1)
    try:
        b = Trainer(2, 16, shape=TINY, lora_rank=8, lora_alpha=16.0, arena_mb=256)
        try:
            lb = [b.step_numpy(b.next_synthetic_batch()) for _ in range(3)]
            fixed = b.next_synthetic_batch()
            lb += [b.step_numpy(fixed) for _ in range(40)]           # one batch repeated: the loss on it must fall
            assert (b.num_with_grad, b.num_trainable) == (8, 8), (b.num_with_grad, b.num_trainable)   # every LoRA tensor receives a gradient
            assert abs(lb[0] - la[0]) <= 1e-5 * la[0], (lb[0], la[0])            # same forward (another kernel: rounding only)
            assert lb[2] != la[2]                                    # the qkv adapters now move
            assert lb[-1] < lb[3], lb
            moved = [b.param_checksum(i) for i in range(8)]
            assert all(m[2] > 0 for m in moved), moved               # B factors (initialised to zero) of qkv and proj all trained
        finally:
            b.close()
    finally:
        hostlib.mft_host_set_true_gradients(0)


def test_accumulated_micro_batches_equal_one_step_on_the_whole_batch(capi, exact, true_gradients):
    """grad += over two micro-batches, mean, clip, Adam == one step on the concatenated batch (summation order only)."""
    from mobilefinetuner_b200.trainer import Trainer
    rng = np.random.default_rng(3)
    whole = Trainer(4, 16, shape=TINY, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    micro = Trainer(2, 16, shape=TINY, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    for _ in range(4):
        ids = rng.integers(0, TINY["vocab_size"], (4, 16)).astype(np.int32)
        d = torch.from_numpy(ids).cuda()
        whole.step_device(d.data_ptr())
        micro.accumulate(d[:2].contiguous().data_ptr())
        micro.accumulate(d[2:].contiguous().data_ptr())
        micro.apply_accumulated()
        assert abs(whole.grad_norm() - micro.grad_norm()) <= 1e-4 * whole.grad_norm()
    for i in range(whole.num_trainable):
        n, s, sa = whole.param_checksum(i)
        n2, s2, sa2 = micro.param_checksum(i)
        assert n == n2 and abs(sa - sa2) <= 1e-4 * max(sa, 1e-6), (i, sa, sa2)
    whole.close(); micro.close()


def test_accumulate_is_refused_in_the_default_mode(capi):
    from mobilefinetuner_b200.trainer import Trainer, TrainerError
    t = Trainer(2, 16, shape=TINY, lora_rank=8, arena_mb=256)
    d = torch.zeros(2, 16, dtype=torch.int32, device="cuda")
    with pytest.raises(TrainerError, match="true-gradient"):
        t.accumulate(d.data_ptr())
    t.close()


@pytest.mark.tf32
def test_true_gradient_mode_full_size_gpt2_small_all_48_adapters(capi, true_gradients):
    """GPT-2 small (head_dim 64) LoRA r8 at 4 x 128 on the tensor-core path: 48 / 48 adapter tensors receive gradients."""
    from mobilefinetuner_b200.trainer import Trainer
    t = Trainer(4, 128, shape="gpt2-small", lora_rank=8, lora_alpha=16.0, arena_mb=4096, cuda_graph=True)
    try:
        losses = [t.step_numpy(t.next_synthetic_batch()) for _ in range(6)]
        assert (t.num_with_grad, t.num_trainable) == (48, 48)
        assert all(np.isfinite(losses))
    finally:
        t.close()

"""GPU: the C-ABI training step (include/mft_trainer.h) reproduces the reference's trajectories.

The trainer draws weights and tokens from the same mt19937 stream as oracle/harness/train_synth.cpp, so its losses
are compared step by step with the goldens the REFERENCE build produced (tests/golden/traj)."""
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

TINY_GPT2 = dict(model=0, n_layer=2, n_embd=64, n_head=4, vocab_size=503, n_positions=64)
TINY_GEMMA = dict(model=1, n_layer=6, n_embd=64, n_head=4, vocab_size=1009, n_positions=64, intermediate=128, kv_heads=1,
                  head_dim=32, sliding_window=512)


@pytest.fixture(autouse=True)
def exact_mode(request):
    """Tight comparisons run in exact mode (fp32 SIMT GEMM, 3xTF32 attention); tests marked tf32 use the default path."""
    from mobilefinetuner_b200 import capi
    if "tf32" in request.keywords:
        yield
        return
    prev_tc, prev_at = capi.lib.mft_use_tensor_cores(0), capi.lib.mft_attention_precise(1)
    yield
    capi.lib.mft_use_tensor_cores(prev_tc)
    capi.lib.mft_attention_precise(prev_at)


def golden(name):
    text = open(os.path.join(GOLDEN, "traj", name)).read()
    steps = [(float(m[1]), float(m[2]), int(m[3]), int(m[4])) for m in
             re.finditer(r"^step \d+ loss (\S+) gnorm (\S+) with_grad (\d+)/(\d+)", text, re.M)]
    params = {int(m[1]): (int(m[2]), float(m[3]), float(m[4])) for m in
              re.finditer(r"^param (\d+) n=(\d+) sum=(\S+) sumabs=(\S+)", text, re.M)}
    return steps, params


def run(tr, steps, params, loss_tol=1e-4, gnorm_tol=1e-3):
    for i, (ref_loss, ref_gnorm, with_grad, total) in enumerate(steps):
        loss = tr.step_numpy(tr.next_synthetic_batch())
        assert abs(loss - ref_loss) / ref_loss < loss_tol, (i, loss, ref_loss)
        assert abs(tr.grad_norm() - ref_gnorm) <= gnorm_tol * ref_gnorm, (i, tr.grad_norm(), ref_gnorm)
        assert (tr.num_with_grad, tr.num_trainable) == (with_grad, total)
    for i, (n, s, sa) in params.items():
        gn, gs, gsa = tr.param_checksum(i)
        assert gn == n and abs(gsa - sa) <= 1e-3 * max(sa, 1e-6), (i, gsa, sa)


def test_trainer_gpt2_lora_matches_reference_50_steps():
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    run(tr, *golden("gpt2_lora_tiny.txt"))
    assert tr.kernel_launches > 0
    tr.close()


def test_trainer_gpt2_full_finetune_matches_reference_50_steps():
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GPT2, full_finetune=True, arena_mb=256)
    run(tr, *golden("gpt2_full_tiny.txt"))
    tr.close()


def test_trainer_gemma_lora_matches_reference_50_steps():
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GEMMA, lora_rank=8, lora_alpha=32.0, arena_mb=512)
    run(tr, *golden("gemma_lora_tiny.txt"), gnorm_tol=2e-3)
    tr.close()


def test_split_step_with_bucket_equals_fused_step():
    """forward_backward + (identity all-reduce) + update == step_device, bit for bit (world_size 1)."""
    import torch
    from mobilefinetuner_b200.trainer import Trainer
    a = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=256)
    b = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=256)
    bucket = b.make_grad_bucket()
    assert bucket.numel() >= sum(b.param_checksum(i)[0] for i in range(b.num_trainable))
    for _ in range(5):
        ids = a.next_synthetic_batch()
        b.next_synthetic_batch()
        d = torch.from_numpy(ids).cuda()
        a.step_device(d.data_ptr())
        b.forward_backward(d.data_ptr())
        b.update()
        assert a.loss() == b.loss()
    for i in range(a.num_trainable):
        assert a.param_checksum(i) == b.param_checksum(i)
    a.close(); b.close()


@pytest.mark.tf32
def test_trainer_default_tf32_path_stays_within_loss_tolerance():
    """Default execution mode (TF32 tensor-core GEMMs where eligible, TF32 attention operands): loss within 1e-3."""
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    steps, _ = golden("gpt2_lora_tiny.txt")
    for i, (ref_loss, ref_gnorm, with_grad, total) in enumerate(steps[:20]):
        loss = tr.step_numpy(tr.next_synthetic_batch())
        assert abs(loss - ref_loss) / ref_loss < 1e-3, (i, loss, ref_loss)
        assert abs(tr.grad_norm() - ref_gnorm) <= 2e-2 * ref_gnorm
        assert (tr.num_with_grad, tr.num_trainable) == (with_grad, total)
    tr.close()


@pytest.mark.tf32
def test_trainer_bf16_gemm_mode_c1_first_steps():
    """Config c1 (GPT-2 small LoRA r8, B4 S128) with bf16 tensor-core GEMM operands: loss within 1e-3 of the fp32 reference
    run, same set of tensors receiving gradients."""
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    hostlib.mft_host_set_gemm_precision(1)
    try:
        tr = Trainer(4, 128, shape="gpt2-small", lora_rank=8, lora_alpha=16.0, arena_mb=4096)
        steps, _ = golden("gpt2_small_lora_c1.txt")
        for i, (ref_loss, ref_gnorm, with_grad, total) in enumerate(steps):
            loss = tr.step_numpy(tr.next_synthetic_batch())
            assert abs(loss - ref_loss) / ref_loss < 1e-3, (i, loss, ref_loss)
            assert abs(tr.grad_norm() - ref_gnorm) <= 5e-2 * ref_gnorm, (i, tr.grad_norm(), ref_gnorm)
            assert (tr.num_with_grad, tr.num_trainable) == (with_grad, total)
        tr.close()
    finally:
        hostlib.mft_host_set_gemm_precision(0)


@pytest.mark.tf32
def test_bf16_operands_written_by_producers_equal_cast_on_demand():
    """bf16 GEMM mode: LayerNorm / attention / GELU / lora_down / CE kernels write the bf16 copy of their result themselves
    (no cast kernels). The rounding is the same round-to-nearest-even, so losses and parameters must equal the
    cast-on-demand path bit for bit, with fewer launches."""
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    hostlib.mft_host_set_gemm_precision(1)
    try:
        out = []
        os.environ["MFT_BF16_ONLY"] = "0"        # both runs keep every fp32 result: only the way the bf16 copy is made differs
        os.environ["MFT_LORA_DOWN_BF16"] = "0"   # ... and the rank-r products read fp32 rows in both
        for emit in ("0", "1"):
            os.environ["MFT_BF16_EMIT"] = emit
            tr = Trainer(4, 128, shape="gpt2-small", lora_rank=8, lora_alpha=16.0, arena_mb=4096)
            losses = [tr.step_numpy(tr.next_synthetic_batch()) for _ in range(4)]
            out.append((losses, [tr.param_checksum(i) for i in range(tr.num_trainable)], tr.kernel_launches))
            tr.close()
        assert out[0][0] == out[1][0] and out[0][1] == out[1][1]
        assert out[1][2] <= out[0][2] - 50, (out[0][2], out[1][2])
    finally:
        os.environ.pop("MFT_BF16_EMIT", None)
        os.environ.pop("MFT_BF16_ONLY", None)
        os.environ.pop("MFT_LORA_DOWN_BF16", None)
        hostlib.mft_host_set_gemm_precision(0)


@pytest.mark.tf32
def test_bf16_only_results_stay_within_tolerance_of_keeping_every_fp32_copy():
    """bf16 operand mode: results that are consumed only as bf16 operands (qkv for the tcgen05 attention, ln_2 rows, the GELU
    activation and its gradient, dlogits) are written as bf16 alone. Against the same run with every fp32 copy kept
    (MFT_BF16_ONLY=0) the loss moves by rounding only, with fewer bytes written; GPT-2 small LoRA r8, 4 x 128."""
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    hostlib.mft_host_set_gemm_precision(1)
    try:
        out = []
        for only in ("0", "1"):
            os.environ["MFT_BF16_ONLY"] = only
            tr = Trainer(4, 128, shape="gpt2-small", lora_rank=8, lora_alpha=16.0, arena_mb=4096)
            losses = [tr.step_numpy(tr.next_synthetic_batch()) for _ in range(6)]
            out.append((losses, tr.grad_norm(), tr.num_with_grad))
            tr.close()
        for a, b in zip(out[0][0], out[1][0]):
            assert abs(a - b) <= 2e-4 * a, (out[0][0], out[1][0])
        assert abs(out[0][1] - out[1][1]) <= 2e-2 * out[0][1]
        assert out[0][2] == out[1][2]
    finally:
        os.environ.pop("MFT_BF16_ONLY", None)
        hostlib.mft_host_set_gemm_precision(0)


@pytest.mark.tf32
@pytest.mark.parametrize("bf16", [0, 1])
def test_trainer_c1_50_steps_tensor_core_modes_within_loss_tolerance(bf16):
    """Config c1 at full size, 50 steps against the reference build's trajectory (tests/golden/traj/gpt2_small_lora_c1_50.txt):
    loss within 1e-3 relative at every step in the default TF32 mode and in the bf16-operand mode (north_star tolerance)."""
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    hostlib.mft_host_set_gemm_precision(bf16)
    try:
        tr = Trainer(4, 128, shape="gpt2-small", lora_rank=8, lora_alpha=16.0, arena_mb=4096, cuda_graph=True)
        steps, _ = golden("gpt2_small_lora_c1_50.txt")
        assert len(steps) == 50
        worst = 0.0
        for i, (ref_loss, ref_gnorm, with_grad, total) in enumerate(steps):
            loss = tr.step_numpy(tr.next_synthetic_batch())
            worst = max(worst, abs(loss - ref_loss) / ref_loss)
            assert abs(loss - ref_loss) / ref_loss < 1e-3, (i, loss, ref_loss)
            assert (tr.num_with_grad, tr.num_trainable) == (with_grad, total)
        print("worst relative loss error", worst)
        tr.close()
    finally:
        hostlib.mft_host_set_gemm_precision(0)


def test_trainer_cuda_graph_replay_matches_reference_50_steps():
    """Steps 0-1 run eagerly, step 2 is captured, steps 3.. are graph replays: the trajectory must not notice."""
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, lora_alpha=16.0, arena_mb=256, cuda_graph=True)
    run(tr, *golden("gpt2_lora_tiny.txt"))
    tr.close()
    tr = Trainer(2, 16, shape=TINY_GPT2, full_finetune=True, arena_mb=256, cuda_graph=True)
    run(tr, *golden("gpt2_full_tiny.txt"))
    # replayed steps keep updating the device copy: a checksum taken earlier must not be served from a stale host mirror
    before = [tr.param_checksum(i) for i in range(tr.num_trainable)]
    for _ in range(2):
        tr.step_numpy(tr.next_synthetic_batch())
    moved = [tr.param_checksum(i) for i in range(tr.num_trainable)]
    assert sum(a != b for a, b in zip(before, moved)) >= tr.num_with_grad
    # the learning rate is read from device memory by the captured update: a schedule takes effect without a new capture
    tr.set_lr(0.0)
    for _ in range(2):
        tr.step_numpy(tr.next_synthetic_batch())
    assert moved == [tr.param_checksum(i) for i in range(tr.num_trainable)]
    tr.close()


def test_trainer_cuda_graph_gemma_and_split_phases():
    import torch
    from mobilefinetuner_b200.trainer import Trainer
    tr = Trainer(2, 16, shape=TINY_GEMMA, lora_rank=8, lora_alpha=32.0, arena_mb=512, cuda_graph=True)
    run(tr, *golden("gemma_lora_tiny.txt"), gnorm_tol=2e-3)
    tr.close()
    # data-parallel form (forward_backward / update captured as two graphs) against the fused eager step
    a = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=256)
    b = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=256, cuda_graph=True)
    b.make_grad_bucket()
    for _ in range(8):
        ids = a.next_synthetic_batch()
        b.next_synthetic_batch()
        d = torch.from_numpy(ids).cuda()
        a.step_device(d.data_ptr())
        b.forward_backward(d.data_ptr())
        b.update()
        assert a.loss() == b.loss()
    for i in range(a.num_trainable):
        assert a.param_checksum(i) == b.param_checksum(i)
    a.close(); b.close()


def test_arena_is_released_with_the_last_trainer_and_cannot_move_under_a_live_one():
    """The step arena is shared and ref-counted: ops outside a trainer fall back to the caching allocator once the last
    trainer is gone, and a second trainer may not force a larger slab while the first one's graphs point into the old one."""
    from mobilefinetuner_b200.trainer import Trainer, TrainerError
    a = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=64)
    with pytest.raises(TrainerError, match="arena"):                   # larger than any slab an earlier test left behind
        Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=96 * 1024)
    a.step_numpy(a.next_synthetic_batch())
    a.close()
    b = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, arena_mb=128)     # no live user: a slab of any size may be taken
    b.step_numpy(b.next_synthetic_batch())
    b.close()


def test_data_parallel_two_ranks_equal_single_process():
    """2-rank NCCL run (torchrun, 127.0.0.1 rendezvous) of tools/dp_check.py; skipped on a single-GPU box."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for graph in ("0", "1"):
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                            "--master-port", "29517", os.path.join(root, "tools", "dp_check.py")], capture_output=True, text=True, timeout=600,
                           env=dict(os.environ, DP_CHECK_GRAPH=graph))
        assert r.returncode == 0 and "DP_CHECK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("full_ft", [False, True])
def test_resume_from_saved_state_continues_bit_for_bit(tmp_path, full_ft):
    """5 steps, save (parameters + Adam moments + step counters), 4 more steps == a fresh trainer that loads the file and runs
    the same 4 batches: identical losses and parameters (exact mode; CUDA-graph replay on in the resumed trainer)."""
    from mobilefinetuner_b200.trainer import Trainer
    rng = np.random.default_rng(17)
    batches = rng.integers(0, TINY_GPT2["vocab_size"], (9, 2, 16)).astype(np.int32)
    a = Trainer(2, 16, shape=TINY_GPT2, full_finetune=full_ft, lora_rank=8, lora_alpha=16.0, arena_mb=256)
    for i in range(5):
        a.step_numpy(batches[i])
    path = tmp_path / "state.safetensors"
    a.save_state(path)
    la = [a.step_numpy(batches[i]) for i in range(5, 9)]
    pa = [a.param_checksum(i) for i in range(a.num_trainable)]
    a.close()
    b = Trainer(2, 16, shape=TINY_GPT2, full_finetune=full_ft, lora_rank=8, lora_alpha=16.0, arena_mb=256, cuda_graph=True, seed=99)   # other init: all of it must come from the file
    b.load_state(path)
    lb = [b.step_numpy(batches[i]) for i in range(5, 9)]
    if not full_ft:
        # LoRA run: the frozen base weights come from the seed, so only a trainer with the same seed can resume; check that
        # a different base really gives different losses, then resume properly
        assert lb != la
        b.close()
        b = Trainer(2, 16, shape=TINY_GPT2, lora_rank=8, lora_alpha=16.0, arena_mb=256, cuda_graph=True)
        b.load_state(path)
        lb = [b.step_numpy(batches[i]) for i in range(5, 9)]
    assert lb == la, (la, lb)
    assert pa == [b.param_checksum(i) for i in range(b.num_trainable)]
    b.close()


def test_data_parallel_through_the_c_abi_without_python():
    """mobilefinetuner_b200/bin/dp_two_rank (tools/dp_two_rank.cpp): two processes, one GPU each, mft_comm_* (NCCL loaded by the
    library) + mft_trainer_step_dp, against one process stepping the concatenated batches; skipped on a single-GPU box."""
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "mobilefinetuner_b200", "bin", "dp_two_rank")
    assert os.path.exists(exe), "build it with make -C mobilefinetuner_b200/csrc"
    r = subprocess.run([exe, "2", "8"], capture_output=True, text=True, timeout=300, env=dict(os.environ, MFT_USE_TENSOR_CORES="0"))
    assert r.returncode == 0 and "DP_C_ABI_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    # full fine-tuning (every tensor in the bucket)
    r = subprocess.run([exe, "2", "8", "1"], capture_output=True, text=True, timeout=300, env=dict(os.environ, MFT_USE_TENSOR_CORES="0"))
    assert r.returncode == 0 and "DP_C_ABI_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]

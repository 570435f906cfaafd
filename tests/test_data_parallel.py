"""N>1 host logic on CPU: world_size-2 gloo groups, the numpy oracle standing in for the device step.

What must hold for the data-parallel step (mobilefinetuner_b200/data_parallel.py): with equal contiguous row shards,
sum-all-reduce of the packed gradients times 1/world equals the gradient of the whole batch on one process; clip and
Adam applied to that bucket leave every rank with bit-identical parameters, equal to the single-process update."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from mobilefinetuner_b200 import data_parallel as dp  # noqa: E402


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


# ----------------------------------------------------------------------------- pure host logic
def test_bucket_layout_matches_trainer_rule():
    offs, total = dp.bucket_layout([6, 16, 1, 0, 8])
    assert offs == [0, 8, 24, 28, 28] and total == 36
    assert all(o % 4 == 0 for o in offs)
    assert dp.bucket_layout([]) == ([], 0)
    with pytest.raises(ValueError):
        dp.bucket_layout([4, -1])


def test_shard_rows_and_errors():
    assert [dp.shard_rows(8, r, 4) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
    assert dp.shard_rows(3, 0, 1) == (0, 3)
    with pytest.raises(ValueError):
        dp.shard_rows(7, 0, 2)          # ragged shards would change the mean-reduced loss
    with pytest.raises(ValueError):
        dp.shard_rows(8, 2, 2)


def test_pack_unpack_round_trip_with_missing_grads():
    rng = np.random.default_rng(0)
    shapes = [(3, 2), (5,), (4, 4), (1, 7)]
    grads = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    grads[1] = None                      # tensor without a gradient this step
    sizes = [int(np.prod(s)) for s in shapes]
    offs, total = dp.bucket_layout(sizes)
    b = dp.pack(grads, offs, total, np)
    back = dp.unpack(b, sizes, offs, shapes)
    for g, r in zip(grads, back):
        assert np.array_equal(r, np.zeros_like(r) if g is None else g)
    mask = np.ones(total, bool)
    for n, o in zip(sizes, offs):
        mask[o:o + n] = False
    assert not b[mask].any()             # padding stays zero


# ----------------------------------------------------------------------------- 2-rank gloo
def tiny_model(seed):
    rng = np.random.default_rng(seed)
    d_in, d_out, r, V = 12, 10, 4, 17
    return dict(W=rng.standard_normal((d_in, d_out)).astype(np.float32) * 0.3, bias=rng.standard_normal(d_out).astype(np.float32) * 0.1,
                A=rng.standard_normal((d_in, r)).astype(np.float32) * 0.3, B=rng.standard_normal((r, d_out)).astype(np.float32) * 0.3,
                H=rng.standard_normal((d_out, V)).astype(np.float32) * 0.3, scale=2.0)


def grads_of(model, x, labels):
    """Oracle forward/backward of  CE(lora_linear(x) · H): gradients of the trainable A, B (frozen W, H) + a tensor that
    never receives one (mirrors the stop-gradient LoRA tensors of the GPT-2 graph)."""
    import mft_oracle as o
    y = o.lora_linear_fwd(x, model["W"], model["bias"], model["A"], model["B"], model["scale"])
    logits = o.matmul(y, model["H"])
    loss, _ = o.lm_cross_entropy_fwd(logits, labels)
    glogits = o.lm_cross_entropy_bwd(logits, labels)
    gy = glogits.reshape(-1, glogits.shape[-1]) @ model["H"].T
    _, dA, dB = o.lora_linear_bwd(x, model["W"], model["A"], model["B"], model["scale"], gy.reshape(y.shape))
    return float(loss), [dA, None, dB]


def apply_update(model, mom, bucket, sizes, offs, step, world, max_norm=0.05):
    """norm / clip / Adam on the all-reduced bucket, 1/world folded into the clip factor like mft_clip_scale's pre_scale."""
    import mft_oracle as o
    pre = np.float32(1.0 / world)
    parts = dp.unpack(bucket, sizes, offs)
    norm = o.grad_norm([p * pre for p in parts])
    factor = o.clip_scale(norm, max_norm) * pre
    for key, idx in (("A", 0), ("B", 2)):
        g = (parts[idx] * factor).reshape(model[key].shape)
        model[key], mom[key][0], mom[key][1] = o.adam_step(model[key], g, mom[key][0], mom[key][1], step, lr=1e-2, wd=0.01)
    return norm


def worker(rank, world, port, out_dir):
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        model, ref = tiny_model(3), tiny_model(3)
        mom = {k: [np.zeros_like(model[k]), np.zeros_like(model[k])] for k in ("A", "B")}
        mom_ref = {k: [np.zeros_like(ref[k]), np.zeros_like(ref[k])] for k in ("A", "B")}
        sizes = [model["A"].size, 5, model["B"].size]
        offs, total = dp.bucket_layout(sizes)
        rng = np.random.default_rng(11)                       # same global batch on every rank
        for step in range(1, 5):
            x = rng.standard_normal((4, 6, 12)).astype(np.float32)
            labels = rng.integers(0, 17, (4, 6)).astype(np.int32)
            lo, hi = dp.shard_rows(x.shape[0], rank, world)
            loss, grads = grads_of(model, x[lo:hi], labels[lo:hi])
            assert dp.with_grad_agrees([g is not None for g in grads])
            bucket = torch.from_numpy(dp.pack(grads, offs, total, np))
            dp.all_reduce_sum(bucket)
            norm = apply_update(model, mom, bucket.numpy(), sizes, offs, step, world)

            loss_ref, grads_ref = grads_of(ref, x, labels)        # single process, whole batch
            bref = dp.pack(grads_ref, offs, total, np)
            assert np.allclose(bucket.numpy() / world, bref, rtol=2e-5, atol=1e-7)
            norm_ref = apply_update(ref, mom_ref, bref, sizes, offs, step, 1)
            assert abs(norm - norm_ref) <= 1e-5 * norm_ref
            mean_loss = torch.tensor([loss]); dist.all_reduce(mean_loss)
            assert abs(mean_loss.item() / world - loss_ref) <= 1e-5 * abs(loss_ref)
            for k in ("A", "B"):
                assert np.allclose(model[k], ref[k], rtol=1e-4, atol=1e-6)
                mine = torch.from_numpy(model[k].copy())
                both = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(both, mine)
                assert all(torch.equal(both[0], b) for b in both)   # replicas stay bit-identical
        # ranks disagreeing on which tensors have gradients must be detected
        assert not dp.with_grad_agrees([True, rank == 0, True])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_step_equals_single_process(tmp_path):
    world = 2
    mp.spawn(worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))

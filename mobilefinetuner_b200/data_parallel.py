"""Data-parallel training step: one process per GPU, one all-reduce of a flat fp32 gradient bucket per step.

The reference trains on a single device (gpt2_lora_finetune/main.cpp:462-520: backward, global-norm clip, Adam). The
path shards over independent sequences, so N ranks each run forward/backward on their rows of the global batch and the
only exchange is the sum of the trainable gradients:

    forward_backward(local rows)   -> gradients packed into the bucket (mft_pack_grads, fixed layout below)
    all_reduce(bucket, SUM)        -> NCCL on the trainer's stream (gloo in the CPU tests)
    update()                       -> norm / clip / Adam read the bucket; the 1/world factor of the mean is folded into
                                      the clip factor on the device (mft_clip_scale pre_scale), no extra pass

Every rank holds identical parameters and identical optimiser state before and after the step. The helpers here are the
host-side logic of that scheme (layout, sharding, averaging) and are exercised on CPU with world_size-2 gloo groups in
tests/test_data_parallel.py; the device pieces are the C ABI calls named above.
"""
from __future__ import annotations

import os

BUCKET_ALIGN = 4  # elements; keeps every tensor's slice 16-byte aligned for the vectorised kernels (trainer_session.cpp)


def bucket_layout(sizes):
    """Element offset of each trainable tensor inside the flat gradient bucket, and the bucket length.
    Mirrors the layout mft_trainer_grad_elems / mft_pack_grads use: tensors in trainable order, each slice padded to a
    multiple of four elements. Padding words are zero and stay zero under the all-reduce."""
    offsets, off = [], 0
    for n in sizes:
        n = int(n)
        if n < 0:
            raise ValueError("negative tensor size")
        offsets.append(off)
        off += (n + BUCKET_ALIGN - 1) // BUCKET_ALIGN * BUCKET_ALIGN
    return offsets, off


def shard_rows(n_rows, rank, world):
    """Rows [lo, hi) of the global batch owned by `rank`. Equal contiguous shards: the mean-reduced loss of the global
    batch is the mean of the shard losses only when every shard holds the same number of target tokens."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} for world size {world}")
    if n_rows % world:
        raise ValueError(f"global batch of {n_rows} rows does not divide over {world} ranks")
    per = n_rows // world
    return rank * per, (rank + 1) * per


def pack(grads, offsets, total, xp):
    """Gradients -> flat bucket (host restatement of mft_pack_grads; None = tensor without gradient -> zeros)."""
    bucket = xp.zeros(total, dtype=xp.float32)
    for g, off in zip(grads, offsets):
        if g is not None:
            flat = g.reshape(-1)
            bucket[off:off + flat.shape[0]] = flat
    return bucket


def unpack(bucket, sizes, offsets, shapes=None):
    out = []
    for i, (n, off) in enumerate(zip(sizes, offsets)):
        t = bucket[off:off + int(n)]
        out.append(t.reshape(shapes[i]) if shapes is not None else t)
    return out


def all_reduce_sum(bucket, group=None):
    """In-place sum over ranks of a torch tensor (cuda -> NCCL, cpu -> gloo). No-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=group)
    return bucket


def with_grad_agrees(flags, group=None):
    """True when every rank produced gradients for the same set of tensors (the stop-gradient attention of the GPT-2
    graph leaves 24 of 48 LoRA tensors without one; the set must not depend on the rank)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([1 if f else 0 for f in flags], dtype=torch.int32)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return True
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    return bool((lo == hi).all())


class DataParallelTrainer:
    """N-rank wrapper around `Trainer`: same step semantics, gradients averaged over ranks.

    `trainer` must have been created with world_size / rank of the current process group. The all-reduce is issued on
    the trainer's own stream, so it is ordered after the pack and before the update without host synchronisation."""

    def __init__(self, trainer, group=None, exchange=None):
        """exchange: "mft" = the library's own communicator (mft_comm_*, NCCL loaded by libmft_host; the whole step is then one
        C-ABI call, mft_trainer_step_dp, and large buckets are exchanged in chunks while backward runs); "torch" = the bucket is
        all-reduced by torch.distributed between the two phases. Default: "mft" on an NCCL process group (MFT_DP_EXCHANGE
        overrides)."""
        import torch
        import torch.distributed as dist
        self.trainer = trainer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world != trainer.cfg.world_size:
            raise ValueError(f"trainer was built for world_size {trainer.cfg.world_size}, process group has {self.world}")
        self.stream = torch.cuda.ExternalStream(trainer.stream)
        self.bucket = trainer.make_grad_bucket() if self.world > 1 else None
        self.comm = None
        if exchange is None:
            exchange = os.environ.get("MFT_DP_EXCHANGE") or ("mft" if self.world > 1 and dist.get_backend(group) == "nccl" else "torch")
        if exchange not in ("mft", "torch"):
            raise ValueError(f"exchange must be 'mft' or 'torch', not {exchange!r}")
        self.exchange = exchange
        if self.world > 1 and exchange == "mft":
            self.comm = self._make_comm(dist, torch)

    def _make_comm(self, dist, torch):
        """rank 0 draws the 128-byte NCCL id, the process group carries it to the others, every rank joins."""
        import ctypes as C
        from .trainer import hostlib
        rank = dist.get_rank(self.group)
        ident = (C.c_char * 128)()
        if rank == 0 and hostlib.mft_comm_unique_id(ident) != 0:
            raise RuntimeError(hostlib.mft_comm_last_error().decode())
        dev = "cuda" if dist.get_backend(self.group) == "nccl" else "cpu"
        t = torch.frombuffer(bytearray(ident.raw), dtype=torch.uint8).to(dev)
        dist.broadcast(t, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        raw = bytes(t.cpu().numpy().tobytes())
        comm = C.c_void_p()
        if hostlib.mft_comm_init_rank(C.byref(comm), self.world, rank, raw) != 0:
            raise RuntimeError(hostlib.mft_comm_last_error().decode())
        return comm

    def close(self):
        if self.comm:
            from .trainer import hostlib
            self.trainer.sync()
            hostlib.mft_comm_destroy(self.comm)
            self.comm = None

    def step_device(self, dev_ids_ptr):
        """One optimiser step on this rank's rows (int32 token ids already in HBM)."""
        import torch
        tr = self.trainer
        if self.world == 1:
            tr.step_device(dev_ids_ptr)
            return
        if self.comm:
            tr.step_dp(self.comm, dev_ids_ptr)
            return
        tr.forward_backward(dev_ids_ptr)
        with torch.cuda.stream(self.stream):
            all_reduce_sum(self.bucket, self.group)
        tr.update()

    def loss_mean(self):
        """Mean of the ranks' losses (what a single process would report on the global batch)."""
        import torch
        import torch.distributed as dist
        v = torch.tensor([self.trainer.loss()], device="cuda")
        if self.world > 1:
            dist.all_reduce(v, group=self.group)
        return float(v.item()) / self.world


DP_CHECK_SHAPE = dict(model=0, n_layer=2, n_embd=64, n_head=4, vocab_size=503, n_positions=64)


def dp_equivalence_check(steps=12, graph=True, per_rank=2, seq=16, full_finetune=False):
    """N ranks stepping their rows (forward/backward -> NCCL all-reduce of the bucket -> update) against ONE process
    stepping the concatenated batch, on a small GPT-2 LoRA shape in the exact execution mode (fp32 SIMT GEMMs, 3xTF32
    attention), so the only difference is summation order. Needs an initialised NCCL process group and the current
    device set. Returns a dict (on every rank; `ok` is rank 0's verdict broadcast to all):
    loss within 1e-4 relative per step, every trainable tensor within 1e-3 in sum|.|, replicas bit-identical.
    full_finetune: every tensor trains (the 0.4 MB bucket of this shape then stands for c2's 498 MB one); with MFT_DP_OVERLAP=1
    in the environment the exchange runs in 64 KB chunks on a side stream while backward is still running."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from . import capi
    from .trainer import Trainer
    world, rank = dist.get_world_size(), dist.get_rank()
    prev_tc, prev_at = capi.lib.mft_use_tensor_cores(0), capi.lib.mft_attention_precise(1)
    prev_env = {k: os.environ.get(k) for k in ("MFT_DP_OVERLAP", "MFT_DP_CHUNK_KB")}
    overlapped = full_finetune and os.environ.get("MFT_DP_OVERLAP", "0") not in ("", "0")
    if overlapped:
        os.environ["MFT_DP_CHUNK_KB"] = "64"
    try:
        rng = np.random.default_rng(5)
        batches = rng.integers(0, DP_CHECK_SHAPE["vocab_size"], (steps, per_rank * world, seq)).astype(np.int32)
        lo, hi = shard_rows(per_rank * world, rank, world)
        tr = Trainer(per_rank, seq, shape=DP_CHECK_SHAPE, full_finetune=full_finetune, lora_rank=8, lora_alpha=16.0, arena_mb=256,
                     world_size=world, rank=rank, cuda_graph=graph)
        dpt = DataParallelTrainer(tr)
        losses = []
        for s in range(steps):
            ids = torch.from_numpy(batches[s, lo:hi].copy()).cuda()
            dpt.step_device(ids.data_ptr())
            losses.append(dpt.loss_mean())
        sums = [tr.param_checksum(i) for i in range(tr.num_trainable)]
        mine = torch.tensor([[c[1], c[2]] for c in sums], dtype=torch.float64, device="cuda")
        ref0 = mine.clone()
        dist.broadcast(ref0, 0)
        replicas_equal = torch.tensor([1 if torch.equal(mine, ref0) else 0], device="cuda")
        dist.all_reduce(replicas_equal, op=dist.ReduceOp.MIN)
        with_grad = tr.num_with_grad
        exchange = dpt.exchange + (" (chunks overlapped with backward)" if overlapped and dpt.comm else "")
        dpt.close()
        tr.close()
        verdict = torch.zeros(3, dtype=torch.float64, device="cuda")     # ok, worst loss error, worst parameter error
        if rank == 0:
            one = Trainer(per_rank * world, seq, shape=DP_CHECK_SHAPE, full_finetune=full_finetune, lora_rank=8, lora_alpha=16.0, arena_mb=256)
            worst_l = worst_p = 0.0
            for s in range(steps):
                l1 = one.step_numpy(batches[s])
                worst_l = max(worst_l, abs(l1 - losses[s]) / abs(l1))
            ok = one.num_with_grad == with_grad
            for i, c in enumerate(sums):
                n, _, sa1 = one.param_checksum(i)
                ok = ok and n == c[0]
                worst_p = max(worst_p, abs(sa1 - c[2]) / max(sa1, 1e-6))
            one.close()
            ok = ok and worst_l <= 1e-4 and worst_p <= 1e-3 and bool(replicas_equal.item())
            verdict = torch.tensor([1.0 if ok else 0.0, worst_l, worst_p], dtype=torch.float64, device="cuda")
        dist.broadcast(verdict, 0)
        return {"ok": bool(verdict[0].item() == 1.0), "world": world, "steps": steps, "cuda_graph": bool(graph),
                "worst_loss_rel_err": float(verdict[1].item()), "worst_param_rel_err": float(verdict[2].item()),
                "replicas_bit_identical": bool(replicas_equal.item()), "last_loss": losses[-1],
                "exchange": exchange,
                "shape": "GPT-2 2 layers x 64, %s, %d x %d tokens per rank, exact fp32 mode" % ("full fine-tune" if full_finetune else "LoRA r8", per_rank, seq)}
    finally:
        for k, v in prev_env.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        capi.lib.mft_use_tensor_cores(prev_tc)
        capi.lib.mft_attention_precise(prev_at)

"""torchrun entry: N-rank data-parallel steps against one process stepping the concatenated batch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py

Every rank steps its rows through DataParallelTrainer (forward/backward -> NCCL all-reduce of the bucket -> update); rank 0
then replays the same global batches through a world_size-1 trainer and compares loss per step and every trainable tensor.
Exact execution mode (fp32 SIMT GEMMs), so the only difference is summation order. Prints DP_CHECK_OK on success."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mobilefinetuner_b200 import capi  # noqa: E402
from mobilefinetuner_b200.data_parallel import DataParallelTrainer, shard_rows  # noqa: E402
from mobilefinetuner_b200.trainer import Trainer  # noqa: E402

SHAPE = dict(model=0, n_layer=2, n_embd=64, n_head=4, vocab_size=503, n_positions=64)


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    capi.lib.mft_use_tensor_cores(0)
    capi.lib.mft_attention_precise(1)
    per, seq, steps = 2, 16, 12
    graph = os.environ.get("DP_CHECK_GRAPH", "1") == "1"
    rng = np.random.default_rng(5)
    batches = rng.integers(0, SHAPE["vocab_size"], (steps, per * world, seq)).astype(np.int32)
    lo, hi = shard_rows(per * world, rank, world)

    tr = Trainer(per, seq, shape=SHAPE, lora_rank=8, lora_alpha=16.0, arena_mb=256, world_size=world, rank=rank, cuda_graph=graph)
    dpt = DataParallelTrainer(tr)
    losses = []
    for s in range(steps):
        ids = torch.from_numpy(batches[s, lo:hi].copy()).cuda()
        dpt.step_device(ids.data_ptr())
        losses.append(dpt.loss_mean())
    sums = [tr.param_checksum(i) for i in range(tr.num_trainable)]
    # replicas identical?
    mine = torch.tensor([[c[1], c[2]] for c in sums], dtype=torch.float64, device="cuda")
    ref0 = mine.clone()
    dist.broadcast(ref0, 0)
    assert torch.equal(mine, ref0), "replicas diverged"
    with_grad = tr.num_with_grad
    tr.close()
    ok = True
    if rank == 0:
        one = Trainer(per * world, seq, shape=SHAPE, lora_rank=8, lora_alpha=16.0, arena_mb=256)
        for s in range(steps):
            l1 = one.step_numpy(batches[s])
            if abs(l1 - losses[s]) > 1e-4 * abs(l1):
                ok = False
                print("loss mismatch", s, l1, losses[s])
        assert one.num_with_grad == with_grad
        for i, c in enumerate(sums):
            n, s1, sa1 = one.param_checksum(i)
            if n != c[0] or abs(sa1 - c[2]) > 1e-3 * max(sa1, 1e-6):
                ok = False
                print("param mismatch", i, (n, s1, sa1), c)
        one.close()
        print("DP_CHECK_OK" if ok else "DP_CHECK_FAILED", "world", world, "graph", graph, "last loss", losses[-1], flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())

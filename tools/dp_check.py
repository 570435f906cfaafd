"""torchrun entry: N-rank data-parallel steps against one process stepping the concatenated batch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py

Every rank steps its rows through DataParallelTrainer (forward/backward -> NCCL all-reduce of the bucket -> update); rank 0
then replays the same global batches through a world_size-1 trainer and compares loss per step and every trainable tensor
(mobilefinetuner_b200.data_parallel.dp_equivalence_check, which bench.py also runs for N > 1). Prints DP_CHECK_OK on success."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mobilefinetuner_b200.data_parallel import dp_equivalence_check  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = dp_equivalence_check(steps=12, graph=os.environ.get("DP_CHECK_GRAPH", "1") == "1")
    if dist.get_rank() == 0:
        print("DP_CHECK_OK" if r["ok"] else "DP_CHECK_FAILED", r, flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if r["ok"] else 1


if __name__ == "__main__":
    sys.exit(main())

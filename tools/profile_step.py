"""One eager training step inside a cudaProfilerStart/Stop range, for `ncu --profile-from-start off`.

  python tools/profile_step.py --config c4                      # must exit 0 on its own first
  ncu --profile-from-start off --metrics <list> --clock-control none --csv --log-file gpurun_out/x.csv python tools/profile_step.py --config c4
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mobilefinetuner_b200.trainer import Trainer, hostlib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c4", choices=sorted(bench.CONFIGS))
    ap.add_argument("--warm", type=int, default=3)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--precision", default="bf16", choices=["tf32", "bf16"])
    ap.add_argument("--batch", type=int, default=0, help="sequences per step (default: the config's per-GPU batch)")
    a = ap.parse_args()
    hostlib.mft_host_set_gemm_precision(1 if a.precision == "bf16" else 0)
    c = bench.CONFIGS[a.config]
    batch, seq = (a.batch or c["batch"]), c["seq"]
    tr = Trainer(batch, seq, shape=c["shape"], full_finetune=c["full"], lora_rank=c["rank"], lora_alpha=c["alpha"], seed=1234, cuda_graph=False,
                 arena_mb=int((bench.ARENA_GIB_PER_SEQ[a.config] * batch + 3) * 1024))
    ids = torch.from_numpy(np.random.default_rng(0).integers(0, tr.cfg.vocab_size, (batch, seq)).astype(np.int32)).cuda()
    for _ in range(a.warm):
        tr.step_device(ids.data_ptr())
    tr.sync()
    rt = C.CDLL("libcudart.so.12")
    rt.cudaProfilerStart()
    for _ in range(a.steps):
        tr.step_device(ids.data_ptr())
    tr.sync()
    rt.cudaProfilerStop()
    print("profiled", a.steps, "step(s) of", a.config, "loss", tr.loss(), "launches/step", tr.kernel_launches)
    tr.close()


if __name__ == "__main__":
    main()

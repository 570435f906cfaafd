"""Eager vs CUDA-graph trainer runs on the same seeded batches (debug aid): prints the first step where they diverge."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from mobilefinetuner_b200 import capi
from mobilefinetuner_b200.trainer import Trainer
from test_trainer_gpu import TINY_GEMMA, TINY_GPT2

capi.lib.mft_use_tensor_cores(0); capi.lib.mft_attention_precise(1)
for name, shape, kw in (("gemma", TINY_GEMMA, dict(lora_rank=8, lora_alpha=32.0)), ("gpt2", TINY_GPT2, dict(lora_rank=8, lora_alpha=16.0))):
    out = []
    for graph in (False, True):
        tr = Trainer(2, 16, shape=shape, arena_mb=512, cuda_graph=graph, **kw)
        out.append([(tr.step_numpy(tr.next_synthetic_batch()), tr.grad_norm()) for _ in range(12)])
        tr.close()
    bad = [i for i, (x, y) in enumerate(zip(*out)) if x != y]
    print(name, "first divergent step:", bad[:1], out[0][bad[0]] if bad else "", out[1][bad[0]] if bad else "")

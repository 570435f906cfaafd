#!/usr/bin/env python
"""bench.py — tokens/s of the MobileFineTuner training step on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c4|c1|c2|c3] [--impl mft|reference]

One "step" = forward + shifted cross-entropy + backward + global clip + Adam over one synthetic batch.
Default workload (config c4, the one the metric is quoted on: "GPT-2 LoRA train tokens/s at 1/2/4/8 B200"):
GPT-2 medium (355M), LoRA r=16 alpha=32 on attn.qkv+attn.proj, 8 x 512 tokens PER GPU (global batch 64 at 8 GPUs,
weak scaling), random-init weights, synthetic uniform tokens, labels = ids.

Printed JSON line (rank 0): value = whole-job tokens/s with ids already resident in HBM; e2e = same metric through the
public C-ABI step call with HOST buffers (pinned ids H2D + loss D2H inside the timed region); roofline = tcgen05 GEMM
launches timed with CUDA events on their stream inside extra profiled steps; cpu_baseline = the reference's own CPU code
(oracle/_ref/ref_harness, built from the unmodified reference sources) on a bounded sample of the same workload.
`--impl reference` times that CPU implementation alone.
"""
import argparse
import ctypes as C
import json
import os
import re
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (shape, per-GPU batch, seq, full_finetune, rank, alpha, description)
    "c1": ("gpt2-small", 4, 128, False, 8, 16.0, "GPT-2 small (124M) LoRA r=8 alpha=16, batch 4, seq 128"),
    "c2": ("gpt2-small", 8, 256, True, 8, 16.0, "GPT-2 small full fine-tune, batch 8, seq 256, Adam wd 0.01, clip 1.0"),
    "c3": ("gemma3-270m", 4, 256, False, 8, 32.0, "Gemma-3 270M LoRA r=8 targets=full, batch 4, seq 256"),
    "c4": ("gpt2-medium", 8, 512, False, 16, 32.0, "GPT-2 medium (355M) LoRA r=16, 8x512 tokens per GPU (global batch 64 at 8 GPUs)"),
}
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
# bounded CPU sample of the workload: same model/adapters, batch 1, 128 tokens per step
CPU_SAMPLE = dict(batch=1, seq=128)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d.get("hbm_gbs", 6650.0), tf_burst=d.get("bf16_tflops", 1590.0),
                    tf_sustained=d.get("bf16_tflops_sustained", 1400.0), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi sampler running during the timed region (clocks + throttle reasons)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def run_ref_harness(shape, full_ft, rank, alpha, batch, seq, steps, threads):
    """Runs the reference's CPU implementation (oracle/_ref/ref_harness) and returns (tokens/s, mean step s)."""
    from mobilefinetuner_b200.trainer import MODEL_SHAPES
    s = MODEL_SHAPES[shape]
    if s["model"] == 0:
        args = ["gpt2", "--layers", s["n_layer"], "--embd", s["n_embd"], "--heads", s["n_head"], "--vocab", s["vocab_size"],
                "--npos", s["n_positions"], "--mode", "full" if full_ft else "lora"]
    else:
        args = ["gemma", "--layers", s["n_layer"], "--embd", s["n_embd"], "--inter", s["intermediate"], "--heads", s["n_head"],
                "--kv_heads", s["kv_heads"], "--head_dim", s["head_dim"], "--vocab", s["vocab_size"]]
    args += ["--batch", batch, "--seq", seq, "--rank", rank, "--alpha", alpha, "--steps", steps, "--threads", threads,
             "--quiet", "1", "--checksums", "0"]
    env = dict(os.environ, OPENBLAS_NUM_THREADS=str(threads))
    out = subprocess.run([REF_HARNESS] + [str(a) for a in args], capture_output=True, text=True, env=env, timeout=3000)
    if out.returncode != 0:
        raise RuntimeError(f"ref_harness failed: {out.stderr[-500:]}")
    m = re.search(r"timing steps=(\d+) mean_step_s=(\S+) tokens_per_s=(\S+)", out.stdout)
    return float(m[3]), float(m[2])


def reference_arm(a):
    """--impl reference: the reference's own CPU path, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shape, _, _, full_ft, r, alpha, desc = CONFIGS[a.config]
    cores = os.cpu_count() or 1
    if not os.path.exists(REF_HARNESS):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_harness not built (run __graft_entry__.build() where /root/reference exists)"}))
        return 0
    t0 = time.time()
    tps, step_s = run_ref_harness(shape, full_ft, r, alpha, CPU_SAMPLE["batch"], CPU_SAMPLE["seq"], a.steps + 1, cores)
    sample = (f"reference C++ (USE_BLAS, OpenBLAS {cores} threads; only sgemm is multi-threaded), same model/adapters, "
              f"batch {CPU_SAMPLE['batch']} x seq {CPU_SAMPLE['seq']} per step, 1 warm-up + {a.steps} timed steps")
    line = {"impl": "reference", "metric": "train_tokens_per_sec", "value": tps, "unit": "tokens/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": 1, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "bounded_sample": sample},
            "cpu_baseline": {"value": tps, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": tps, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": round(time.time() - t0, 1)}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="mft", choices=["mft", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-steps", type=int, default=3)
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying the captured CUDA graph")
    ap.add_argument("--precision", default=os.environ.get("MFT_GEMM_PRECISION", "bf16"), choices=["tf32", "bf16"],
                    help="operand precision of the large tensor-core GEMMs (fp32 accumulate either way)")
    a = ap.parse_args()
    if a.impl == "reference":
        return reference_arm(a)
    a.warmup = max(a.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback (use --impl reference for the CPU reference arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from mobilefinetuner_b200 import capi
    from mobilefinetuner_b200.trainer import Trainer, hostlib
    from mobilefinetuner_b200.data_parallel import DataParallelTrainer
    hostlib.mft_host_set_gemm_precision(1 if a.precision == "bf16" else 0)

    shape, batch, seq, full_ft, r, alpha, desc = CONFIGS[a.config]
    tr = Trainer(batch, seq, shape=shape, full_finetune=full_ft, lora_rank=r, lora_alpha=alpha, world_size=world, rank=rank,
                 seed=1234, cuda_graph=not a.no_graph)
    ext = torch.cuda.ExternalStream(tr.stream)
    tokens_rank = batch * seq
    rng = np.random.default_rng(1000 + rank)
    vocab = tr.cfg.vocab_size
    nbatches = 8
    host_ids = torch.from_numpy(rng.integers(0, vocab, (nbatches, batch, seq)).astype(np.int32)).pin_memory()
    dev_ids = host_ids.cuda()
    host_loss = torch.zeros(1, dtype=torch.float32).pin_memory()
    dpt = DataParallelTrainer(tr) if world > 1 else None     # forward/backward -> all-reduce(bucket) -> update
    rt = C.CDLL("libcudart.so.12")

    def step_resident(i):
        ptr = dev_ids[i % nbatches].data_ptr()
        if world == 1:
            tr.step_device(ptr)
        else:
            dpt.step_device(ptr)

    staging = torch.empty(batch, seq, dtype=torch.int32, device="cuda")

    def step_e2e(i):
        src = host_ids[i % nbatches]
        if world == 1:
            tr.step(src.data_ptr(), host_loss.data_ptr())          # H2D + step + D2H inside the C-ABI call
        else:
            with torch.cuda.stream(ext):
                staging.copy_(src, non_blocking=True)
            dpt.step_device(staging.data_ptr())
            rt.cudaMemcpyAsync(C.c_void_p(host_loss.data_ptr()), C.c_void_p(tr.loss_ptr), 4, 2, C.c_void_p(tr.stream))
            tr.sync()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        tr.sync()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for i in range(steps):
            fn(i)
        e1.record(ext)
        e1.synchronize()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for i in range(a.warmup):
        step_resident(i)
    tr.sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total = timed(step_resident, a.steps)
    launches = tr.kernel_launches * a.steps      # kernels per step (counted when the step was captured / last ran eagerly)
    clocks = sampler.stop() if rank == 0 else None
    loss_now = tr.loss()
    for i in range(2):
        step_e2e(i)
    ms_e2e = timed(step_e2e, a.steps)

    # ---- roofline of the dominant kernel (tcgen05 GEMM): events around every launch, inside extra steps
    tr.enable_graph(False)                       # per-kernel events need eager launches
    if a.profile_steps > 0:
        step_resident(0)
    capi.lib.mft_gemm_profile(1)
    t_prof = timed(step_resident, a.profile_steps)
    fl, ms, nl = C.c_double(), C.c_double(), C.c_int64()
    capi.lib.mft_gemm_profile_read(C.byref(fl), C.byref(ms), C.byref(nl))
    capi.lib.mft_gemm_profile(0)
    peaks = load_peaks()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    tokens_step = tokens_rank * world
    value = tokens_step * a.steps / (ms_total / 1e3)
    e2e_val = tokens_step * a.steps / (ms_e2e / 1e3)
    achieved_tf = (fl.value / 1e12) / (ms.value / 1e3) if ms.value > 0 else 0.0
    kind = "kind::f16 bf16 operands" if a.precision == "bf16" else "kind::tf32"
    traffic = None                               # DRAM bytes per GEMM launch from the committed ncu capture of this config
    tpath = os.path.join(ROOT, "profiles", "r1_gemm_traffic.json")
    if a.config == "c4" and os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch_" + a.precision)
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": f"gemm_tc_kernel (tcgen05.mma cta_group::2 {kind}, TMA-fed, fp32 accumulate in TMEM)",
                "achieved": round(achieved_tf, 2), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": round(achieved_tf / peaks["tf_sustained"], 4), "traffic": traffic,
                "algorithmic_flop_per_launch": round(fl.value / max(nl.value, 1)),
                "launches_per_step": nl.value // max(a.profile_steps, 1), "gemm_ms_per_step": round(ms.value / max(a.profile_steps, 1), 3),
                "step_ms_while_profiled": round(t_prof / max(a.profile_steps, 1), 3),
                "peak_source": peaks["source"] + ", dense bf16 sustained" + ("" if a.precision == "bf16" else "; the kernel computes in TF32 whose hardware peak is half of bf16")}
    cpu_baseline = None
    if not a.no_cpu_baseline and os.path.exists(REF_HARNESS):
        cores = os.cpu_count() or 1
        try:
            tps, step_s = run_ref_harness(shape, full_ft, r, alpha, CPU_SAMPLE["batch"], CPU_SAMPLE["seq"], 4, cores)
            cpu_baseline = {"value": round(tps, 3), "unit": "tokens/s", "cores": cores, "kind": "reference",
                            "sample": f"oracle/_ref/ref_harness (unmodified reference sources, USE_BLAS/OpenBLAS {cores} threads; only sgemm "
                                      f"is multi-threaded), same model+adapters, batch {CPU_SAMPLE['batch']} x seq {CPU_SAMPLE['seq']}, 1 warm-up + 3 timed steps "
                                      f"({step_s:.2f} s/step)"}
        except Exception as e:  # the baseline is a reported number, never a reason to lose the GPU line
            cpu_baseline = {"value": None, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": f"failed: {e}"}
    line = {"metric": "train_tokens_per_sec", "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(ms_total / a.steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None,
            "dtype": ("bf16 GEMM operands (fp32 master tensors and accumulate; attention TF32; norms/softmax/CE/Adam fp32)" if a.precision == "bf16" else
                      "tf32 (fp32 storage/accumulate; GEMM inputs rounded to TF32 in the TMA; norms/softmax/CE/Adam fp32)"),
            "data": "synthetic",
            "config": {"workload": desc, "config_id": a.config, "tokens_per_step_per_gpu": tokens_rank, "global_tokens_per_step": tokens_step,
                       "parallelism": f"dp{world}", "launch": "eager" if a.no_graph else "cuda graph replay of the captured step", "timing": "inputs larger than L2: per-step activations+logits "
                       f"{tr.arena_high_water / 2**30:.1f} GiB >> 126 MB L2; CUDA events on the compute stream, max over ranks",
                       "loss_last_timed_step": round(loss_now, 5)},
            "e2e": {"value": round(e2e_val, 1), "unit": "tokens/s", "h2d_bytes_per_step": tokens_rank * 4 * world,
                    "d2h_bytes_per_step": 4 * world, "ms_per_step": round(ms_e2e / a.steps, 3)},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": clocks}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

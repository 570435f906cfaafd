#!/usr/bin/env python
"""bench.py — tokens/s of the MobileFineTuner training step on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c4|c1|c2|c3|c5] [--scaling strong|weak] [--impl mft|reference]

One "step" = forward + shifted cross-entropy + backward + global clip + Adam over one synthetic batch.
Default workload (config c4, the one the metric is quoted on: "GPT-2 LoRA train tokens/s at 1/2/4/8 B200"):
GPT-2 medium (355M), LoRA r=16 alpha=32 on attn.qkv+attn.proj, GLOBAL batch 64 x 512 tokens split over the N GPUs
(strong scaling: 64/N sequences per GPU; `--scaling weak` keeps 8 x 512 per GPU instead), random-init weights, synthetic
uniform tokens, labels = ids. c1-c3 are the single-device configs of BASELINE.json; c5 is the per-op sweep
(tools/sweep_bench.py).

Printed JSON line (rank 0): value = whole-job tokens/s with ids already resident in HBM; e2e = same metric through the
public C-ABI step call with HOST buffers (pinned ids H2D + loss D2H inside the timed region); roofline = tcgen05 GEMM
launches timed with CUDA events on their stream inside extra profiled steps, plus the whole step against the tensor peak
(roofline.step); cpu_baseline = the reference's own CPU code (oracle/_ref/ref_harness, built from the unmodified reference
sources) on a bounded, shape-matched sample of the same workload. `--impl reference` times that CPU implementation alone and
imports nothing that loads this repository's native libraries.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: shape, per-GPU batch (weak), seq, full_finetune, rank, alpha, global batch (strong; None = single-device config),
    #       batch of the bounded CPU sample (same model, same sequence length), description
    "c1": dict(shape="gpt2-small", batch=4, seq=128, full=False, rank=8, alpha=16.0, global_batch=None, cpu_batch=4,
               desc="GPT-2 small (124M) LoRA r=8 alpha=16, batch 4, seq 128"),
    "c2": dict(shape="gpt2-small", batch=8, seq=256, full=True, rank=8, alpha=16.0, global_batch=None, cpu_batch=2,
               desc="GPT-2 small full fine-tune, batch 8, seq 256, Adam wd 0.01, clip 1.0"),
    "c3": dict(shape="gemma3-270m", batch=4, seq=256, full=False, rank=8, alpha=32.0, global_batch=None, cpu_batch=1,
               desc="Gemma-3 270M LoRA r=8 targets=full, batch 4, seq 256"),
    "c4": dict(shape="gpt2-medium", batch=8, seq=512, full=False, rank=16, alpha=32.0, global_batch=64, cpu_batch=1,
               desc="GPT-2 medium (355M) LoRA r=16, global batch 64, seq 512, data-parallel over the GPUs"),
}
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
# step-arena bytes per sequence of the config (measured high-water marks / batch, rounded up) + fixed slack
ARENA_GIB_PER_SEQ = {"c1": 0.45, "c2": 0.9, "c3": 2.2, "c4": 1.95}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d.get("hbm_gbs", 6650.0), tf_burst=d.get("bf16_tflops", 1590.0),
                    tf_sustained=d.get("bf16_tflops_sustained", 1400.0), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


def plan(a, world):
    """Per-GPU batch of this run and the `config` object both arms print (no measured values in it)."""
    c = CONFIGS[a.config]
    scaling = a.scaling or ("strong" if c["global_batch"] else "weak")
    if scaling == "strong":
        if not c["global_batch"]:
            raise SystemExit(f"--scaling strong: config {a.config} has no global batch (single-device config)")
        if c["global_batch"] % world:
            raise SystemExit(f"global batch {c['global_batch']} does not divide over {world} GPUs")
        batch = c["global_batch"] // world
    else:
        batch = c["batch"]
    tokens_rank = batch * c["seq"]
    config = {"workload": c["desc"], "config_id": a.config, "batch_per_gpu": batch, "seq_len": c["seq"],
              "tokens_per_step_per_gpu": tokens_rank, "global_tokens_per_step": tokens_rank * world, "parallelism": f"dp{world}",
              "scaling": scaling,
              "l2": "no flush: every step streams inputs larger than L2 (per-step activations and logits are GiB against 126 MB)"}
    return c, scaling, batch, config


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
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def run_ref_harness(shape, full_ft, rank, alpha, batch, seq, max_steps, threads, budget_s, min_timed=1):
    """Runs the reference's CPU implementation (oracle/_ref/ref_harness, a subprocess) on `batch` x `seq` tokens per step:
    one warm-up step, then timed steps until `max_steps` are done or `budget_s` seconds of timed work have passed
    (at least `min_timed`). Steps are timed by the arrival of the harness's per-step lines. Returns
    (tokens/s, mean step s, timed steps)."""
    from mobilefinetuner_b200.shapes import MODEL_SHAPES      # pure Python: loads no native library
    s = MODEL_SHAPES[shape]
    if s["model"] == 0:
        args = ["gpt2", "--layers", s["n_layer"], "--embd", s["n_embd"], "--heads", s["n_head"], "--vocab", s["vocab_size"],
                "--npos", s["n_positions"], "--mode", "full" if full_ft else "lora"]
    else:
        args = ["gemma", "--layers", s["n_layer"], "--embd", s["n_embd"], "--inter", s["intermediate"], "--heads", s["n_head"],
                "--kv_heads", s["kv_heads"], "--head_dim", s["head_dim"], "--vocab", s["vocab_size"]]
    args += ["--batch", batch, "--seq", seq, "--rank", rank, "--alpha", alpha, "--steps", max_steps + 1, "--threads", threads,
             "--quiet", "1", "--checksums", "0"]
    env = dict(os.environ, OPENBLAS_NUM_THREADS=str(threads))
    proc = subprocess.Popen([REF_HARNESS] + [str(x) for x in args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)
    stamps = []
    try:
        for line in proc.stdout:
            if line.startswith("step "):
                stamps.append(time.time())
                timed = len(stamps) - 1
                if timed >= max_steps or (timed >= min_timed and stamps[-1] - stamps[0] >= budget_s):
                    break
    finally:
        if proc.poll() is None:
            proc.kill()                      # exactly the child started above
        proc.wait()
    timed = len(stamps) - 1
    if timed < 1:
        raise RuntimeError(f"ref_harness produced {len(stamps)} step lines: {proc.stderr.read()[-400:]}")
    step_s = (stamps[-1] - stamps[0]) / timed
    return batch * seq / step_s, step_s, timed


def reference_arm(a):
    """--impl reference: the reference's own CPU path, all host threads, a bounded shape-matched sample per step."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", str(a.gpus)))
    if a.config == "c5":
        print(json.dumps({"impl": "reference", "unavailable": "c5 is the per-op sweep; the reference arm times whole training steps (c1-c4)"}))
        return 0
    c, scaling, batch, config = plan(a, world)
    cores = os.cpu_count() or 1
    if not os.path.exists(REF_HARNESS):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_harness not built (run __graft_entry__.build() where /root/reference exists)"}))
        return 0
    t0 = time.time()
    cb = min(c["cpu_batch"], batch)
    tps, step_s, timed = run_ref_harness(c["shape"], c["full"], c["rank"], c["alpha"], cb, c["seq"], a.steps, cores, budget_s=150.0, min_timed=2)
    sample = (f"oracle/_ref/ref_harness = the unmodified reference C++ sources (USE_BLAS, OpenBLAS {cores} threads; only sgemm is "
              f"multi-threaded), same model/adapters/optimiser, each step {cb} x {c['seq']} tokens (the config's own sequence length; "
              f"{'the full per-GPU batch' if cb == batch else f'{cb} of the {batch} sequences a GPU steps'}), 1 warm-up + {timed} timed "
              f"steps ({step_s:.2f} s/step; capped at 150 s of timed work)")
    line = {"impl": "reference", "metric": "train_tokens_per_sec", "value": round(tps, 3), "unit": "tokens/s", "n_gpus": a.gpus,
            "steps": timed, "warmup": 1, "ms_per_step": round(step_s * 1e3, 1), "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": round(tps, 3), "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": round(tps, 3), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": round(time.time() - t0, 1)}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS) + ["c5"])
    ap.add_argument("--scaling", default=None, choices=["strong", "weak"],
                    help="c4 only: strong = global batch 64 split over the GPUs (default), weak = 8 x 512 tokens per GPU")
    ap.add_argument("--impl", default="mft", choices=["mft", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-steps", type=int, default=3)
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying the captured CUDA graph")
    ap.add_argument("--no-dp-check", action="store_true", help="N > 1: skip the data-parallel == single-process check")
    ap.add_argument("--precision", default=os.environ.get("MFT_GEMM_PRECISION", "bf16"), choices=["tf32", "bf16"],
                    help="operand precision of the large tensor-core GEMMs (fp32 accumulate either way)")
    a = ap.parse_args()
    if a.impl == "reference":
        return reference_arm(a)
    if a.config == "c5":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sweep_bench
        return sweep_bench.main(["--json"])
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
    from mobilefinetuner_b200.data_parallel import DataParallelTrainer, dp_equivalence_check
    from mobilefinetuner_b200.shapes import step_flops_per_token

    # N > 1: first prove that the data-parallel step equals the single-process step (small shape, exact mode, graph replay)
    dp_check = None
    if world > 1 and not a.no_dp_check:
        dp_check = dp_equivalence_check(steps=8, graph=not a.no_graph, full_finetune=CONFIGS[a.config]["full"] if a.config in CONFIGS else False)
    hostlib.mft_host_set_gemm_precision(1 if a.precision == "bf16" else 0)

    c, scaling, batch, config = plan(a, world)
    seq = c["seq"]
    free_b, _ = torch.cuda.mem_get_info()
    arena_gib = min(ARENA_GIB_PER_SEQ[a.config] * batch + 3.0, free_b / 2**30 - 8.0)
    tr = Trainer(batch, seq, shape=c["shape"], full_finetune=c["full"], lora_rank=c["rank"], lora_alpha=c["alpha"], world_size=world,
                 rank=rank, seed=1234, cuda_graph=not a.no_graph, arena_mb=int(arena_gib * 1024))
    ext = torch.cuda.ExternalStream(tr.stream)
    tokens_rank = batch * seq
    rng = np.random.default_rng(1000 + rank)
    vocab = tr.cfg.vocab_size
    nbatches = 4
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
    arena_high = tr.arena_high_water
    exchange = None
    if dpt is not None:
        overlapped = dpt.comm is not None and os.environ.get("MFT_DP_OVERLAP", "0") not in ("", "0")
        exchange = ("mft_comm (NCCL behind the C ABI, mft_trainer_step_dp)" if dpt.comm else "torch.distributed all_reduce between the phases") + \
                   (", bucket exchanged in chunks on a side stream while backward runs" if overlapped else ", one all-reduce of the bucket after backward")
        dpt.close()

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
    for name in ("r2_gemm_traffic.json", "r1_gemm_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if a.config == "c4" and os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))                  # captures exist for 8 and 64 sequences per GPU
                traffic = tj.get(f"dram_bytes_per_launch_{a.precision}_b{batch}", tj.get("dram_bytes_per_launch_" + a.precision) if batch == 8 else None)
            except Exception:
                traffic = None
            break
    f_fwd, f_bwd = step_flops_per_token(c["shape"], seq, c["full"], c["rank"])
    flop_step_gpu = (f_fwd + f_bwd) * tokens_rank
    step_tf = flop_step_gpu / 1e12 / (ms_total / a.steps / 1e3)
    roofline = {"bound": "tensor", "kernel": f"gemm_tc_kernel (tcgen05.mma cta_group::2 {kind}, TMA-fed, fp32 accumulate in TMEM)",
                "achieved": round(achieved_tf, 2), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": round(achieved_tf / peaks["tf_sustained"], 4), "traffic": traffic,
                "algorithmic_flop_per_launch": round(fl.value / max(nl.value, 1)),
                "launches_per_step": nl.value // max(a.profile_steps, 1), "gemm_ms_per_step": round(ms.value / max(a.profile_steps, 1), 3),
                "step_ms_while_profiled": round(t_prof / max(a.profile_steps, 1), 3),
                "peak_source": peaks["source"] + ", dense bf16 sustained" + ("" if a.precision == "bf16" else "; the kernel computes in TF32 whose hardware peak is half of bf16"),
                "step": {"flop_per_step_per_gpu": round(flop_step_gpu), "flop_per_token": round(f_fwd + f_bwd), "achieved": round(step_tf, 2),
                         "unit": "TFLOP/s", "frac": round(step_tf / peaks["tf_sustained"], 4),
                         "note": "whole step (every kernel, launch gaps included) against the same tensor peak; algorithmic FLOPs with the reference's backward semantics (SURVEY 8d)"}}
    cpu_baseline = None
    if world == 1 and not a.no_cpu_baseline and os.path.exists(REF_HARNESS):
        cores = os.cpu_count() or 1
        cb = min(c["cpu_batch"], batch)
        try:
            tps, step_s, n_timed = run_ref_harness(c["shape"], c["full"], c["rank"], c["alpha"], cb, seq, 8, cores, budget_s=25.0)
            cpu_baseline = {"value": round(tps, 3), "unit": "tokens/s", "cores": cores, "kind": "reference",
                            "sample": f"oracle/_ref/ref_harness (unmodified reference sources, USE_BLAS/OpenBLAS {cores} threads; only sgemm "
                                      f"is multi-threaded), same model+adapters, {cb} x {seq} tokens per step (the config's own sequence length), "
                                      f"1 warm-up + {n_timed} timed steps ({step_s:.2f} s/step)"}
        except Exception as e:  # the baseline is a reported number, never a reason to lose the GPU line
            cpu_baseline = {"value": None, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": f"failed: {e}"}
    line = {"metric": "train_tokens_per_sec", "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(ms_total / a.steps, 3), "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None,
            "dtype": ("bf16 GEMM and attention operands (fp32 master tensors and accumulate; norms/softmax/CE/Adam fp32)" if a.precision == "bf16" else
                      "tf32 (fp32 storage/accumulate; GEMM inputs rounded to TF32 in the TMA; norms/softmax/CE/Adam fp32)"),
            "data": "synthetic", "config": config,
            "run": {"launch": "eager" if a.no_graph else "cuda graph replay of the captured step",
                    "timing": "CUDA events on the compute stream, barrier + synchronize on both sides, max over ranks",
                    "step_arena_high_water_gib": round(arena_high / 2**30, 2), "loss_last_timed_step": round(loss_now, 5), "exchange": exchange},
            "e2e": {"value": round(e2e_val, 1), "unit": "tokens/s", "h2d_bytes_per_step": tokens_rank * 4 * world,
                    "d2h_bytes_per_step": 4 * world, "ms_per_step": round(ms_e2e / a.steps, 3)},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": clocks}
    if dp_check is not None:
        line["dp_loss_matches_single_process"] = dp_check["ok"]
        line["dp_check"] = dp_check
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

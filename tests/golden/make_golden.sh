#!/bin/bash
# Regenerates the golden fixtures from the REFERENCE ITSELF (built unmodified by oracle/Makefile).
# Needs /root/reference; run in the build container, never on the GPU box. Outputs are committed.
set -euo pipefail
cd "$(dirname "$0")/../.."
make -C oracle -j8 >/dev/null
H=oracle/_ref/ref_harness
rm -rf tests/golden/ops && mkdir -p tests/golden/ops tests/golden/traj
$H ops --dump tests/golden/ops --seed 7 > /dev/null
# training trajectories (loss / grad-norm / with_grad per step + final parameter checksums)
$H gpt2  --steps 50 --layers 2 --embd 64 --heads 4 --vocab 503 --batch 2 --seq 16 --rank 8 --quiet 1 2>&1 | grep -E '^(config|step|param)' > tests/golden/traj/gpt2_lora_tiny.txt
$H gpt2  --steps 50 --layers 2 --embd 64 --heads 4 --vocab 503 --batch 2 --seq 16 --mode full --quiet 1 2>&1 | grep -E '^(config|step|param)' > tests/golden/traj/gpt2_full_tiny.txt
$H gpt2  --steps 10 --layers 2 --embd 128 --heads 4 --vocab 1000 --batch 2 --seq 64 --rank 8 --lora_mlp 1 --lr 1e-2 --quiet 1 2>&1 | grep -E '^(config|step|param)' > tests/golden/traj/gpt2_lora_mlp_s64.txt
$H gemma --steps 50 --layers 6 --embd 64 --inter 128 --heads 4 --kv_heads 1 --head_dim 32 --vocab 1009 --batch 2 --seq 16 --rank 8 2>&1 | grep -E '^(config|step|param)' | sed -E 's/ t_step [0-9.]+//' > tests/golden/traj/gemma_lora_tiny.txt
# c1 (GPT-2 small LoRA r8 B4 S128), first 4 steps: the survey's reference values
$H gpt2 --steps 4 --layers 12 --embd 768 --heads 12 --vocab 50257 --npos 1024 --batch 4 --seq 128 --rank 8 --quiet 1 --checksums 0 --threads 8 2>&1 | grep -E '^(config|step)' > tests/golden/traj/gpt2_small_lora_c1.txt
ls tests/golden/ops | wc -l

#!/bin/bash
# Times the reference's unchanged gpt2_lora_finetune CLI at GPT-2 small shape (config c1: batch 4, seq 128, LoRA r8):
# this library's build on the GPU (default TF32 and exact fp32 mode) next to the reference's CPU build on the host cores.
set -u
FX=${FX:-/tmp/fxs}
OUT=${OUT:-gpurun_out}
mkdir -p "$OUT"
python tools/make_fixtures.py --out "$FX" --family gpt2 --n_layer 12 --n_embd 768 --n_head 12 --n_positions 1024 --train_lines 4000 || exit 1
ARGS="--data_dir $FX/data --pretrained_dir $FX/model --batch_size 4 --seq_len 128 --lr 1e-4 --seed 7"
for mode in warm default exact; do
  [ "$mode" = exact ] && export MFT_USE_TENSOR_CORES=0 || unset MFT_USE_TENSOR_CORES
  t0=$(date +%s.%N)
  mobilefinetuner_b200/bin/gpt2_lora_finetune $ARGS --steps 100 --lora_out "$FX/a.safetensors" > "$OUT/cli_${mode}_100.log" 2>&1
  t1=$(date +%s.%N)
  mobilefinetuner_b200/bin/gpt2_lora_finetune $ARGS --steps 1100 --lora_out "$FX/b.safetensors" > "$OUT/cli_${mode}_1100.log" 2>&1
  t2=$(date +%s.%N)
  python - "$t0" "$t1" "$t2" "$mode" <<'PY'
import sys
t0, t1, t2 = map(float, sys.argv[1:4])
per = ((t2 - t1) - (t1 - t0)) / 1000.0         # start-up (load, tokenise) cancels between the 100- and 1100-step runs
print(f"mine {sys.argv[4]}: {per*1e3:.2f} ms/step, {4*128/per:.0f} tokens/s (start-up {(t1-t0) - 100*per:.1f} s)")
PY
  grep -E "global (1|1100)/1100" "$OUT/cli_${mode}_1100.log" | cut -c1-130
done
unset MFT_USE_TENSOR_CORES
export OPENBLAS_NUM_THREADS=$(nproc)
t0=$(date +%s.%N)
oracle/_ref/ref_gpt2_lora_finetune $ARGS --steps 1 --lora_out "$FX/r1.safetensors" > "$OUT/cli_ref_1.log" 2>&1
t1=$(date +%s.%N)
oracle/_ref/ref_gpt2_lora_finetune $ARGS --steps 4 --lora_out "$FX/r4.safetensors" > "$OUT/cli_ref_4.log" 2>&1
t2=$(date +%s.%N)
python - "$t0" "$t1" "$t2" "$(nproc)" <<'PY'
import sys
t0, t1, t2 = map(float, sys.argv[1:4])
per = ((t2 - t1) - (t1 - t0)) / 3.0
print(f"reference CPU build ({sys.argv[4]} threads): {per:.2f} s/step, {4*128/per:.1f} tokens/s")
PY
grep -E "global (1|4)/4" "$OUT/cli_ref_4.log" | cut -c1-130

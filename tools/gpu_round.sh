#!/bin/bash
# One GPU-box session: tests, bench lines, ncu launch lists. Usage: bash tools/gpu_round.sh <tag> [steps...]
# steps: tests bench_c4 bench_small ncu_c4 ncu_c4_b64 ncu_c3 gemm_step
tag=$1; shift
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum"
out=gpurun_out
for s in "$@"; do
  case $s in
    tests) timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_tests.log 2>&1; tail -3 $out/${tag}_tests.log ;;
    bench_c4)
      timeout 600 python bench.py --steps 10 --warmup 3 > $out/${tag}_c4_strong.log 2>&1; tail -c 600 $out/${tag}_c4_strong.log
      timeout 600 python bench.py --scaling weak --steps 20 --warmup 5 > $out/${tag}_c4_weak.log 2>&1; tail -c 300 $out/${tag}_c4_weak.log ;;
    bench_small)
      for c in c1 c2 c3; do timeout 600 python bench.py --config $c --steps 20 --warmup 5 > $out/${tag}_$c.log 2>&1; tail -1 $out/${tag}_$c.log | cut -c1-200; done ;;
    ncu_c4)
      timeout 300 python tools/profile_step.py --config c4 > $out/${tag}_prof_c4.log 2>&1 &&
      timeout 900 ncu --profile-from-start off --metrics $M --clock-control none --csv --log-file $out/${tag}_c4_metrics.csv python tools/profile_step.py --config c4 > $out/${tag}_ncu_c4.log 2>&1
      python tools/ncu_summary.py $out/${tag}_c4_metrics.csv > $out/${tag}_c4_kernels.txt 2>&1; head -24 $out/${tag}_c4_kernels.txt ;;
    ncu_c4_b64)   # 64 x 512 tokens: multi-pass metrics would save / restore ~70 GB per replayed launch, so one pass per run
      timeout 300 python tools/profile_step.py --config c4 --batch 64 > $out/${tag}_prof_c4_b64.log 2>&1 &&
      timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/${tag}_c4_b64_time.csv python tools/profile_step.py --config c4 --batch 64 > $out/${tag}_ncu_c4_b64.log 2>&1
      python tools/ncu_summary.py $out/${tag}_c4_b64_time.csv > $out/${tag}_c4_b64_kernels.txt 2>&1; head -24 $out/${tag}_c4_b64_kernels.txt ;;
    ncu_c4_b64_dram)
      timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $out/${tag}_c4_b64_dram.csv python tools/profile_step.py --config c4 --batch 64 > $out/${tag}_ncu_c4_b64_dram.log 2>&1
      python tools/ncu_summary.py $out/${tag}_c4_b64_dram.csv > $out/${tag}_c4_b64_dram_kernels.txt 2>&1; head -24 $out/${tag}_c4_b64_dram_kernels.txt ;;
    ncu_c3)
      timeout 300 python tools/profile_step.py --config c3 > $out/${tag}_prof_c3.log 2>&1 &&
      timeout 900 ncu --profile-from-start off --metrics $M --clock-control none --csv --log-file $out/${tag}_c3_metrics.csv python tools/profile_step.py --config c3 > $out/${tag}_ncu_c3.log 2>&1
      python tools/ncu_summary.py $out/${tag}_c3_metrics.csv > $out/${tag}_c3_kernels.txt 2>&1; head -40 $out/${tag}_c3_kernels.txt ;;
    dp)   # on an N-GPU box: NGPU=2 bash tools/gpu_round.sh tag dp
      n=${NGPU:-2}
      timeout 900 python -m pytest tests/test_trainer_gpu.py -m gpu -x -q -k "data_parallel" > $out/${tag}_dp_tests.log 2>&1; tail -3 $out/${tag}_dp_tests.log
      MFT_USE_TENSOR_CORES=0 MFT_DP_OVERLAP=1 MFT_DP_CHUNK_KB=64 timeout 300 mobilefinetuner_b200/bin/dp_two_rank $n 8 1 > $out/${tag}_dp_two_rank_full.log 2>&1; tail -2 $out/${tag}_dp_two_rank_full.log
      TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511"
      timeout 600 python bench.py --config c2 --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_c2_n1.log 2>&1; tail -1 $out/${tag}_c2_n1.log | cut -c1-160
      timeout 600 $TR bench.py --config c2 --gpus $n --steps 20 --warmup 5 > $out/${tag}_c2_n$n.log 2>&1; tail -1 $out/${tag}_c2_n$n.log | cut -c1-160
      MFT_DP_OVERLAP=0 timeout 600 $TR bench.py --config c2 --gpus $n --steps 20 --warmup 5 > $out/${tag}_c2_n${n}_nooverlap.log 2>&1; tail -1 $out/${tag}_c2_n${n}_nooverlap.log | cut -c1-160
      timeout 600 $TR bench.py --gpus $n --steps 10 --warmup 3 > $out/${tag}_c4_strong_n$n.log 2>&1; tail -1 $out/${tag}_c4_strong_n$n.log | cut -c1-160
      timeout 600 $TR bench.py --gpus $n --scaling weak --steps 20 --warmup 5 > $out/${tag}_c4_weak_n$n.log 2>&1; tail -1 $out/${tag}_c4_weak_n$n.log | cut -c1-160 ;;
    ncu_full_gqa)
      timeout 900 ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"gqa_fwd|gqa_bwd_dq|gqa_bwd_dkv" -c 3 -o $out/${tag}_gqa_full python tools/profile_step.py --config c3 > $out/${tag}_ncu_full_gqa.log 2>&1; ls -la $out/${tag}_gqa_full.ncu-rep ;;
    ncu_full_attn)
      timeout 900 ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"attention_tc_fwd" -c 2 -o $out/${tag}_attn_full python tools/profile_step.py --config c4 > $out/${tag}_ncu_full_attn.log 2>&1; ls -la $out/${tag}_attn_full.ncu-rep ;;
    overlap1)   # the overlapped exchange with a one-rank communicator (one GPU): chunks, side stream, NCCL nodes inside the captured phase
      MFT_USE_TENSOR_CORES=0 MFT_DP_OVERLAP=2 MFT_DP_CHUNK_KB=64 timeout 90 mobilefinetuner_b200/bin/dp_two_rank 1 8 1 > $out/${tag}_overlap1.log 2>&1; echo "overlap1 rc=$?"; tail -2 $out/${tag}_overlap1.log ;;
    bench_nopdl)
      for c in c1 c3; do MFT_PDL=0 timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_${c}_nopdl.log 2>&1; tail -1 $out/${tag}_${c}_nopdl.log | cut -c1-200; done
      MFT_PDL=0 timeout 300 python bench.py --scaling weak --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_c4_weak_nopdl.log 2>&1; tail -1 $out/${tag}_c4_weak_nopdl.log | cut -c1-200 ;;
    gemm_step) timeout 300 python tools/gemm_bench.py step > $out/${tag}_gemm_step.log 2>&1; tail -12 $out/${tag}_gemm_step.log ;;
    gemm_step64)
      timeout 300 python tools/gemm_bench.py step 32768 > $out/${tag}_gemm_step64.log 2>&1; tail -12 $out/${tag}_gemm_step64.log
      MFT_TC_GROUP_M=1000 timeout 300 python tools/gemm_bench.py step 32768 > $out/${tag}_gemm_step64_mfast.log 2>&1; tail -12 $out/${tag}_gemm_step64_mfast.log
      MFT_TC_GROUP_M=8 timeout 300 python tools/gemm_bench.py step 32768 > $out/${tag}_gemm_step64_g8.log 2>&1; tail -12 $out/${tag}_gemm_step64_g8.log ;;
  esac
done

#!/bin/bash
# Builds epilogue variants of the tcgen05 GEMM into mobilefinetuner_b200/variants/ (git-ignored, travels with gpurun):
#   MFT_B200_LIB=mobilefinetuner_b200/variants/libmft_b200_w4b1.so python tools/gemm_bench.py tf32
set -e
cd "$(dirname "$0")/../mobilefinetuner_b200/csrc"
make -j8 ../libmft_b200.so > /dev/null
mkdir -p ../variants build/variants
OTHERS=$(ls build/*.o | grep -v gemm_tc.o)
for v in "8 2" "8 1" "4 2"; do
  set -- $v
  name=w$1b$2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I../../include -Ikernels \
      --expt-relaxed-constexpr -DMFT_TC_EPI_WARPS=$1 -DMFT_TC_EPI_BOXES=$2 -c kernels/gemm_tc.cu -o build/variants/gemm_tc_$name.o &
done
wait
for v in "8 2" "8 1" "4 2"; do
  set -- $v
  name=w$1b$2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libmft_b200_$name.so $OTHERS build/variants/gemm_tc_$name.o -lcudart
done
ls -la ../variants

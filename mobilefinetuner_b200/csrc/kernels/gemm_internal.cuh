// Internal GEMM parameter block shared by the SIMT and tcgen05 paths (mirrors mft_gemm_args).
#pragma once
#include "common.cuh"

namespace mft {

struct GemmParams {
    int m, n, k;
    const void* a; int64_t lda; int trans_a;     // fp32, or bf16 when `bf16` is set (tensor-core path only)
    const void* b; int64_t ldb; int trans_b;
    float* c; int64_t ldc;
    float alpha, beta;
    const float* bias;
    float* gelu_out;
    int k2; const void* a2; int64_t lda2; const void* b2; int64_t ldb2;
    int batch; int64_t stride_a, stride_b, stride_c;
    const float* addend = nullptr; int64_t ld_addend = 0;
    const float* gelu_grad_x = nullptr; int64_t ld_gelu_grad_x = 0;
    void* gelu_out_bf16 = nullptr; int64_t ld_gelu_out_bf16 = 0;
    void* c_bf16 = nullptr; int64_t ldc_bf16 = 0;   // bf16 copy of C; with c == nullptr the bf16 copy is the only result (tensor-core path)
    bool bf16 = false;         // operands are bf16 arrays
    int ksplit = 1;            // SIMT split-K: number of K slices (partials reduced in a fixed order)
    float* partial = nullptr;  // [batch][ksplit][m][n] scratch when ksplit > 1
};

bool tensor_cores_enabled();   // mft_use_tensor_cores state (gemm.cu)
int launch_gemm_simt(const GemmParams& p, cudaStream_t st);
bool gemm_tc_eligible(const GemmParams& p);
int launch_gemm_tc(const GemmParams& p, cudaStream_t st);
void gemm_tc_profile_enable(bool on);
void gemm_tc_profile_read(double* flops, double* ms, long long* launches);

}  // namespace mft

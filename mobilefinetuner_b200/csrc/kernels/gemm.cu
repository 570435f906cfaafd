// mft_gemm dispatch + library-level entry points (version, errors, device check).
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include "common.cuh"
#include "gemm_internal.cuh"

namespace mft {

static thread_local char g_err[512] = "";
static long long g_launches = 0;
void add_launches(int n) { g_launches += n; }
// MFT_USE_TENSOR_CORES=0 in the environment starts the library on the exact fp32 SIMT GEMM (parity debugging).
static bool env_use_tc() {
    const char* e = getenv("MFT_USE_TENSOR_CORES");
    return !(e && e[0] == '0');
}
static bool g_use_tc = env_use_tc();

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static int to_params(const mft_gemm_args* a, GemmParams& p) {
    if (!a) return set_error(1, "mft_gemm: null args");
    if (a->m < 0 || a->n < 0 || a->k < 0 || a->k2 < 0) return set_error(1, "mft_gemm: negative dimension");
    p.m = a->m; p.n = a->n; p.k = a->k;
    p.a = a->a; p.lda = a->lda; p.trans_a = a->trans_a;
    p.b = a->b; p.ldb = a->ldb; p.trans_b = a->trans_b;
    p.c = a->c; p.ldc = a->ldc; p.alpha = a->alpha; p.beta = a->beta;
    p.bias = a->bias; p.gelu_out = a->gelu_out;
    p.k2 = a->k2; p.a2 = a->a2; p.lda2 = a->lda2; p.b2 = a->b2; p.ldb2 = a->ldb2;
    p.batch = a->batch < 1 ? 1 : a->batch;
    p.stride_a = a->stride_a; p.stride_b = a->stride_b; p.stride_c = a->stride_c;
    p.addend = a->addend; p.ld_addend = a->ld_addend;
    p.gelu_grad_x = a->gelu_grad_x; p.ld_gelu_grad_x = a->ld_gelu_grad_x;
    p.gelu_out_bf16 = a->gelu_out_bf16; p.ld_gelu_out_bf16 = a->ld_gelu_out_bf16;
    p.c_bf16 = a->c_bf16; p.ldc_bf16 = a->ldc_bf16;
    if (p.gelu_out_bf16 && (p.batch > 1 || (p.ld_gelu_out_bf16 & 3) || (reinterpret_cast<uintptr_t>(p.gelu_out_bf16) & 7u)))
        return set_error(1, "mft_gemm: gelu_out_bf16 needs batch 1, an 8-byte aligned base and a pitch that is a multiple of 4");
    if (p.c_bf16 && (p.batch > 1 || p.ldc_bf16 < p.n || (p.ldc_bf16 & 3) || (reinterpret_cast<uintptr_t>(p.c_bf16) & 7u)))
        return set_error(1, "mft_gemm: c_bf16 needs batch 1, an 8-byte aligned base and a pitch >= n that is a multiple of 4");
    if (p.m == 0 || p.n == 0) return 0;
    if ((!p.c && !p.c_bf16) || (p.k > 0 && (!p.a || !p.b))) return set_error(1, "mft_gemm: null operand");
    if (p.k2 > 0 && (!p.a2 || !p.b2)) return set_error(1, "mft_gemm: null second-segment operand");
    const int64_t min_lda = p.trans_a ? p.m : p.k, min_ldb = p.trans_b ? p.k : p.n;
    if (p.lda < min_lda || p.ldb < min_ldb || (p.c && p.ldc < p.n)) return set_error(1, "mft_gemm: leading dimension too small");
    if (p.k2 > 0 && (p.lda2 < p.k2 || p.ldb2 < p.k2)) return set_error(1, "mft_gemm: second-segment ld too small");
    return 0;
}

}  // namespace mft

using namespace mft;

extern "C" int mft_version(void) { return 100; }
extern "C" int64_t mft_launch_count(void) { return g_launches; }
extern "C" const char* mft_last_error(void) { return g_err; }

extern "C" int mft_device_check(void) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(4, "mft_device_check: %s", cudaGetErrorString(e));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) return set_error(4, "mft_device_check: %s", cudaGetErrorString(e));
    if (prop.major != 10) return set_error(4, "mft_device_check: device is sm_%d%d, this library is sm_100a only", prop.major, prop.minor);
    return 0;
}

namespace mft { bool tensor_cores_enabled() { return g_use_tc; } }

extern "C" int mft_tensor_cores_enabled(void) { return g_use_tc ? 1 : 0; }

extern "C" int mft_use_tensor_cores(int enable) {
    int prev = g_use_tc ? 1 : 0;
    g_use_tc = enable != 0;
    return prev;
}

extern "C" int mft_gemm_uses_tensor_cores(const mft_gemm_args* args) {
    GemmParams p;
    if (!args || to_params(args, p) != 0) return 0;
    return (g_use_tc && gemm_tc_eligible(p)) ? 1 : 0;
}

extern "C" size_t mft_gemm_splitk_workspace_bytes(int32_t m, int32_t n, int32_t batch, int32_t splits) {
    if (splits <= 1) return 0;
    return size_t(m) * size_t(n) * size_t(batch < 1 ? 1 : batch) * size_t(splits) * sizeof(float);
}

extern "C" int mft_gemm_splitk(const mft_gemm_args* args, int32_t splits, void* workspace, size_t workspace_bytes, mft_stream_t stream) {
    GemmParams p;
    int rc = to_params(args, p);
    if (rc) return rc;
    if (p.m == 0 || p.n == 0) return 0;
    if (splits < 1) return set_error(1, "mft_gemm_splitk: splits must be >= 1");
    if (splits > 1) {
        if (p.k2 > 0) return set_error(1, "mft_gemm_splitk: a second K segment cannot be split");
        const size_t need = mft_gemm_splitk_workspace_bytes(p.m, p.n, p.batch, splits);
        if (!workspace || workspace_bytes < need) return set_error(1, "mft_gemm_splitk: workspace too small (%zu < %zu)", workspace_bytes, need);
        if (int64_t(p.batch) * splits > 65535) return set_error(1, "mft_gemm_splitk: batch*splits too large");
        p.ksplit = splits;
        p.partial = static_cast<float*>(workspace);
    }
    if (!p.c || (p.gelu_out_bf16 && !p.gelu_out)) return set_error(1, "mft_gemm_splitk: the SIMT path needs the fp32 result buffers (bf16-only results are a tensor-core feature)");
    rc = launch_gemm_simt(p, as_stream(stream));
    if (rc == 0 && p.gelu_out_bf16) rc = mft_cast_bf16(p.gelu_out, p.ldc, p.gelu_out_bf16, p.ld_gelu_out_bf16, p.m, p.n, stream);   // SIMT path: separate cast
    if (rc == 0 && p.c_bf16) rc = mft_cast_bf16(p.c, p.ldc, p.c_bf16, p.ldc_bf16, p.m, p.n, stream);
    return rc;
}

extern "C" int mft_gemm(const mft_gemm_args* args, mft_stream_t stream) {
    GemmParams p;
    int rc = to_params(args, p);
    if (rc) return rc;
    if (p.m == 0 || p.n == 0) return 0;
    if (g_use_tc && gemm_tc_eligible(p)) return launch_gemm_tc(p, as_stream(stream));
    if (!p.c || (p.gelu_out_bf16 && !p.gelu_out)) return set_error(1, "mft_gemm: the SIMT path needs the fp32 result buffers (bf16-only results are a tensor-core feature)");
    rc = launch_gemm_simt(p, as_stream(stream));
    if (rc == 0 && p.gelu_out_bf16) rc = mft_cast_bf16(p.gelu_out, p.ldc, p.gelu_out_bf16, p.ld_gelu_out_bf16, p.m, p.n, stream);   // SIMT path: separate cast
    if (rc == 0 && p.c_bf16) rc = mft_cast_bf16(p.c, p.ldc, p.c_bf16, p.ldc_bf16, p.m, p.n, stream);
    return rc;
}

extern "C" int mft_gemm_profile(int enable) { gemm_tc_profile_enable(enable != 0); return 0; }
extern "C" int mft_gemm_profile_read(double* flops, double* ms, int64_t* launches) {
    long long n = 0;
    gemm_tc_profile_read(flops, ms, &n);
    if (launches) *launches = n;
    return 0;
}

// bf16-operand form: a, b, a2, b2 point to bf16 arrays (leading dimensions in elements); C stays fp32.
extern "C" int mft_gemm_bf16(const mft_gemm_args* args, mft_stream_t stream) {
    GemmParams p;
    int rc = to_params(args, p);
    if (rc) return rc;
    if (p.m == 0 || p.n == 0) return 0;
    p.bf16 = true;
    if (!gemm_tc_eligible(p))
        return set_error(1, "mft_gemm_bf16: operands must be K-major (trans_a=0, trans_b=1), 16-byte aligned, with leading dimensions that are multiples of 8");
    return launch_gemm_tc(p, as_stream(stream));
}

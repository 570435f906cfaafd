// ops:: implementations: shape logic on the host, arithmetic in CUDA kernels behind the C ABI.
// Reference semantics are cited per function (paths relative to operators/finetune_ops/).
#include "ops.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <random>
#include <stdexcept>

#include "autograd_engine.h"
#include "mft_b200.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

using autograd::Engine;
using internal::GemmCall;

namespace internal {

bool wants_grad(const TensorPtr& t) { return t && t->requires_grad(); }

void track(const TensorPtr& out, const std::vector<TensorPtr>& inputs, BackwardFunctionPtr fn) {
    if (!Engine::instance().is_enabled()) return;
    bool any = false;
    for (const auto& t : inputs) any = any || wants_grad(t);
    if (!any) return;
    out->set_requires_grad(true);
    Engine::instance().register_node(out, inputs, std::move(fn));
}

TensorPtr new_like(const TensorPtr& t) { return Tensor::make(t->shape(), t->dtype()); }

void require_f32(const TensorPtr& t, const char* op) {
    if (!t) throw TensorError(std::string(op) + ": null tensor");
    if (t->dtype() != kFloat32) throw TensorError(std::string(op) + ": float32 tensor expected");
}

namespace {
int env_precision() {
    const char* e = std::getenv("MFT_GEMM_PRECISION");
    return (e && (e[0] == 'b' || e[0] == 'B')) ? 1 : 0;
}
int g_precision = env_precision();

// bf16 K-major copy of a small/temporary fp32 matrix stored [rows, cols] (ld) -> [rows, cols] or its transpose
struct TempBf16 { std::shared_ptr<Tensor> keep; const void* ptr; int64_t ld; };
TempBf16 temp_bf16(const float* src, int64_t ld_src, int64_t rows, int64_t cols, bool transpose) {
    const int64_t orows = transpose ? cols : rows, ocols = transpose ? rows : cols;
    const int64_t ld = (ocols + 7) / 8 * 8;
    auto buf = Tensor::make({orows, ld / 2}, kFloat32);     // raw bytes: orows * ld bf16
    void* dst = buf->dev_out();
    if (transpose) mft_rt::check(mft_transpose_cast_bf16(src, ld_src, dst, ld, rows, cols, mft_rt::stream()), "mft_transpose_cast_bf16");
    else mft_rt::check(mft_cast_bf16(src, ld_src, dst, ld, rows, cols, mft_rt::stream()), "mft_cast_bf16");
    return {buf, dst, ld};
}
}  // namespace

int gemm_precision() { return g_precision; }
bool emit_bf16_operands() {
    static const bool tc_off = [] { const char* e = std::getenv("MFT_USE_TENSOR_CORES"); return e && e[0] == '0'; }();
    if (g_precision != 1 || tc_off) return false;
    const char* e = std::getenv("MFT_BF16_EMIT");      // read per call: tests flip it to compare against the cast-on-demand path
    return !(e && e[0] == '0');
}
// bf16-only results (MFT_BF16_ONLY=0 keeps every fp32 copy; read per call so tests can compare the two)
bool bf16_only_results() {
    if (!emit_bf16_operands()) return false;
    const char* e = std::getenv("MFT_BF16_ONLY");
    return !(e && e[0] == '0');
}
void set_gemm_precision(int bf16) {
    g_precision = bf16 ? 1 : 0;
    mft_attention_bf16(g_precision);      // attention follows the GEMM operand mode
}

// True-gradient mode (off by default because the parity runs must reproduce the reference's quirks):
// GPT-2 attention gets a backward (the reference severs the graph at the qkv split, graph/gpt2_model.cpp:616-637, and its
// memory-efficient attention has no backward, core/memory_efficient_attention.cpp:165-182) and `.grad` accumulates over
// backward() calls instead of being overwritten (core/autograd_engine.cpp:243) — the semantics
// pytorch_alignment/gpt2_lora_finetune.py:228-239 trains with. MFT_TRUE_ATTN_GRAD=1 or mft_host_set_true_gradients(1).
namespace {
bool env_true_gradients() { const char* e = std::getenv("MFT_TRUE_ATTN_GRAD"); return e && e[0] == '1'; }
bool g_true_gradients = [] { const bool on = env_true_gradients(); autograd::Engine::instance().set_accumulate(on); return on; }();
}  // namespace
bool true_gradients() { return g_true_gradients; }
void set_true_gradients(bool on) { g_true_gradients = on; autograd::Engine::instance().set_accumulate(on); }

bool gemm_takes_bf16(int64_t m, int64_t n, int64_t k) {
    static const bool tc_off = [] { const char* e = std::getenv("MFT_USE_TENSOR_CORES"); return e && e[0] == '0'; }();
    return g_precision == 1 && !tc_off && m * n * k >= (int64_t(1) << 24) && m >= 64 && n >= 64 && k >= 32;
}

// Chooses the execution form of one GEMM. Large products whose operands are not K-major are brought into
// K-major form by a transpose pass (cached on 2-D weights) so they run on the tcgen05 kernel; everything
// else goes to mft_gemm as is (SIMT kernel handles any layout).
void run_gemm(GemmCall g) {
    if (g.m == 0 || g.n == 0) return;
    const int64_t work = int64_t(g.m) * g.n * std::max<int64_t>(g.k, 1) * std::max(g.batch, 1);
    // A skinny output whose large operand would have to be transposed (contraction over the token index: LoRA dA/dB) is
    // cheaper as a split-K SIMT product reading that operand once than as transpose + tensor-core GEMM.
    const bool skinny_tn = g.trans_a && std::min(g.m, g.n) < 64;
    // Products with a narrow output (rank-r LoRA factors: x A, dY B^T) only light up m/128 SMs on the tensor-core kernel
    // and are bandwidth-bound anyway: they go to the split-K SIMT kernel too (exact fp32 as a bonus).
    const bool big = !skinny_tn && work >= (int64_t(1) << 24) && g.m >= 64 && g.n >= 64 && g.k >= 32;
    TensorPtr tmp_a, tmp_b;   // keep K-major temporaries alive until the launch is enqueued
    static const bool tc_off = [] { const char* e = std::getenv("MFT_USE_TENSOR_CORES"); return e && e[0] == '0'; }();
    if (big && g.batch <= 1 && g_precision == 1 && !tc_off) {
        // ---- bf16 tensor-core form: every operand becomes a K-major bf16 array (cached on the owning tensor when known)
        std::vector<TempBf16> keep;
        mft_gemm_args a;
        std::memset(&a, 0, sizeof(a));
        a.m = int32_t(g.m); a.n = int32_t(g.n); a.k = int32_t(g.k);
        a.trans_a = 0; a.trans_b = 1;
        if (!g.trans_a) {
            // a_tensor with a == null (or a pointing at its fp32 data): its bf16 copy is the operand — written by the producer,
            // or cast once; a tensor that only exists as bf16 is never materialised as fp32 here
            const bool a_known = g.a_tensor && g.a_tensor->shape().back() == g.k && g.a_tensor->ld() == g.lda &&
                                 (g.a == nullptr || (!g.a_tensor->f32_pending() && g.a_tensor->dev_f32() == g.a));
            if (a_known) { auto v = g.a_tensor->bf16_view(); a.a = static_cast<const float*>(v.ptr); a.lda = v.ld; }
            else {
                if (!g.a) g.a = g.a_tensor->dev_f32();
                keep.push_back(temp_bf16(g.a, g.lda, g.m, g.k, false)); a.a = static_cast<const float*>(keep.back().ptr); a.lda = keep.back().ld;
            }
        } else {
            if (!g.a) g.a = g.a_tensor->dev_f32();
            keep.push_back(temp_bf16(g.a, g.lda, g.k, g.m, true)); a.a = static_cast<const float*>(keep.back().ptr); a.lda = keep.back().ld;
        }
        const bool b_known = g.b_weight && g.b_weight->ndim() == 2 && g.b_weight->dev_f32() == g.b;
        if (g.trans_b) {
            if (b_known) { auto v = g.b_weight->bf16_view(); a.b = static_cast<const float*>(v.ptr); a.ldb = v.ld; }
            else { keep.push_back(temp_bf16(g.b, g.ldb, g.n, g.k, false)); a.b = static_cast<const float*>(keep.back().ptr); a.ldb = keep.back().ld; }
        } else {
            if (b_known) { auto v = g.b_weight->bf16_transposed_view(); a.b = static_cast<const float*>(v.ptr); a.ldb = v.ld; }
            else { keep.push_back(temp_bf16(g.b, g.ldb, g.k, g.n, true)); a.b = static_cast<const float*>(keep.back().ptr); a.ldb = keep.back().ld; }
        }
        if (g.k2 > 0) {
            if (g.a2_tensor && g.a2_tensor->ndim() == 2 && g.a2_tensor->shape()[1] == g.k2 && g.a2_tensor->shape()[0] == g.m) {
                auto v = g.a2_tensor->bf16_view(); a.a2 = static_cast<const float*>(v.ptr); a.lda2 = v.ld;      // written by lora_down itself
            } else { keep.push_back(temp_bf16(g.a2, g.lda2, g.m, g.k2, false)); a.a2 = static_cast<const float*>(keep.back().ptr); a.lda2 = keep.back().ld; }
            if (g.b2_tensor && g.b2_tensor->ndim() == 2) {      // adapter factor: its bf16 K-major shadow is cached per version
                auto v = g.b2_stored_kn ? g.b2_tensor->bf16_transposed_view() : g.b2_tensor->bf16_view();
                a.b2 = static_cast<const float*>(v.ptr); a.ldb2 = v.ld;
            } else { keep.push_back(temp_bf16(g.b2, g.ldb2, g.n, g.k2, false)); a.b2 = static_cast<const float*>(keep.back().ptr); a.ldb2 = keep.back().ld; }
            a.k2 = int32_t(g.k2);
        }
        a.ldc = g.ldc; a.alpha = g.alpha; a.beta = g.beta; a.bias = g.bias;
        a.addend = g.addend; a.ld_addend = g.ld_addend; a.gelu_grad_x = g.gelu_grad_x; a.ld_gelu_grad_x = g.ld_gelu_grad_x;
        a.batch = 1;
        const bool diet = bf16_only_results();
        const bool plain_epilogue = g.beta == 0.0f && !g.addend && !g.gelu_grad_x;
        const bool gelu_here = g.gelu_out != nullptr || (g.gelu_tensor && g.gelu_bf16_only);
        (void)gelu_here;
        if (g.c_tensor && !g.c && diet && g.c_bf16_only && plain_epilogue && g.c_tensor->is_dense() && g.n % 8 == 0 && emit_bf16_operands()) {
            auto e = g.c_tensor->bf16_only_out();                 // bf16 is the only copy of the result
            a.c = nullptr; a.c_bf16 = e.ptr; a.ldc_bf16 = e.ld;
        } else {
            a.c = g.c ? g.c : g.c_tensor->dev_f32_out();
        }
        if (g.gelu_tensor && (g.gelu_out || g.gelu_bf16_only) && g.gelu_tensor->ld() == g.ldc) {
            if (!g.gelu_out && diet && emit_bf16_operands() && g.gelu_tensor->is_dense() && g.n % 8 == 0) {
                auto e = g.gelu_tensor->bf16_only_out();          // the activation feeds a bf16 GEMM operand only
                a.gelu_out = nullptr; a.gelu_out_bf16 = e.ptr; a.ld_gelu_out_bf16 = e.ld;
            } else {
                a.gelu_out = g.gelu_out ? g.gelu_out : g.gelu_tensor->dev_f32_out();
                auto e = g.gelu_tensor->bf16_emit();
                if (e.ld % 4 == 0) { a.gelu_out_bf16 = e.ptr; a.ld_gelu_out_bf16 = e.ld; }
            }
        } else a.gelu_out = g.gelu_out;
        mft_rt::check(mft_gemm_bf16(&a, mft_rt::stream()), "mft_gemm_bf16");
        return;
    }
    if (!g.a && g.a_tensor) g.a = g.a_tensor->dev_f32();
    if (!g.c && g.c_tensor) g.c = g.c_tensor->dev_f32_out();
    if (!g.gelu_out && g.gelu_tensor && g.gelu_bf16_only) g.gelu_out = g.gelu_tensor->dev_f32_out();
    if (big && g.batch <= 1) {
        if (g.trans_a) {       // stored [k,m] -> [m,k]
            const int64_t ld = (g.k + 3) / 4 * 4;
            tmp_a = Tensor::make({g.m, g.k}, kFloat32, false, ld);
            mft_rt::check(mft_transpose2d(g.a, g.lda, tmp_a->dev_f32_out(), ld, g.k, g.m, 1, 0, 0, mft_rt::stream()), "mft_transpose2d");
            g.a = tmp_a->dev_f32(); g.lda = ld; g.trans_a = 0;
        }
        if (!g.trans_b) {      // stored [k,n] -> [n,k]
            if (g.b_weight && g.b_weight->ndim() == 2 && g.b_weight->dev_f32() == g.b) {
                tmp_b = g.b_weight->transposed_shadow();
            } else {
                const int64_t ld = (g.k + 3) / 4 * 4;
                tmp_b = Tensor::make({g.n, g.k}, kFloat32, false, ld);
                mft_rt::check(mft_transpose2d(g.b, g.ldb, tmp_b->dev_f32_out(), ld, g.k, g.n, 1, 0, 0, mft_rt::stream()), "mft_transpose2d");
            }
            g.b = tmp_b->dev_f32(); g.ldb = tmp_b->ld(); g.trans_b = 1;
        }
    }
    mft_gemm_args a;
    std::memset(&a, 0, sizeof(a));
    a.m = int32_t(g.m); a.n = int32_t(g.n); a.k = int32_t(g.k);
    a.a = g.a; a.lda = g.lda; a.trans_a = g.trans_a;
    a.b = g.b; a.ldb = g.ldb; a.trans_b = g.trans_b;
    a.c = g.c; a.ldc = g.ldc; a.alpha = g.alpha; a.beta = g.beta; a.bias = g.bias; a.gelu_out = g.gelu_out;
        a.addend = g.addend; a.ld_addend = g.ld_addend; a.gelu_grad_x = g.gelu_grad_x; a.ld_gelu_grad_x = g.ld_gelu_grad_x;
    a.k2 = int32_t(g.k2); a.a2 = g.a2; a.lda2 = g.lda2; a.b2 = g.b2; a.ldb2 = g.ldb2;
    a.batch = std::max(g.batch, 1); a.stride_a = g.stride_a; a.stride_b = g.stride_b; a.stride_c = g.stride_c;
    // few output tiles + long contraction on the SIMT path: split K so the whole chip works on it
    if (!big && g.k2 == 0 && g.k >= 512) {
        const int64_t bm = g.m <= 16 ? 16 : 64, bn = g.n <= 16 ? 16 : 64;     // tile shapes of the SIMT kernel
        const int64_t tiles = ((g.m + bm - 1) / bm) * ((g.n + bn - 1) / bn) * a.batch;
        if (tiles < 148) {
            int splits = int(std::min<int64_t>(std::min<int64_t>(592 / tiles, g.k / 128), 65535 / a.batch));
            if (splits > 1) {
                const size_t wsb = mft_gemm_splitk_workspace_bytes(a.m, a.n, a.batch, splits);
                mft_rt::check(mft_gemm_splitk(&a, splits, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_gemm_splitk");
                return;
            }
        }
    }
    mft_rt::check(mft_gemm(&a, mft_rt::stream()), "mft_gemm");
}

// Output leading dimension for a [.., n] GEMM result: rows are padded to 32 floats when n is large and not a
// multiple of 4 (a 50257-wide logits row could not be addressed by TMA otherwise).
int64_t out_ld(int64_t n) { return (n >= 1024 && (n % 4) != 0) ? (n + 31) / 32 * 32 : n; }

}  // namespace internal

using namespace internal;

TensorPtr contiguous(const TensorPtr& t) {
    if (!t) throw TensorError("contiguous: null tensor");
    if (t->is_dense()) return t;
    auto out = Tensor::make(t->shape(), t->dtype());
    const size_t es = DTypeUtils::size_of(t->dtype());
    const int64_t w = t->shape().back();
    mft_rt::d2d_2d(out->dev_out(), size_t(w) * es, t->dev(), size_t(t->ld()) * es, size_t(w) * es, size_t(t->rows()));
    return out;
}

// ============================================================================ broadcasting helpers
bool same_shape(const TensorPtr& a, const TensorPtr& b) { return a && b && a->shape() == b->shape(); }

std::vector<int64_t> infer_broadcast_shape(const TensorPtr& a, const TensorPtr& b) {
    const auto& sa = a->shape();
    const auto& sb = b->shape();
    const size_t n = std::max(sa.size(), sb.size());
    std::vector<int64_t> out(n);
    for (size_t i = 0; i < n; ++i) {
        int64_t da = i < n - sa.size() ? 1 : sa[i - (n - sa.size())];
        int64_t db = i < n - sb.size() ? 1 : sb[i - (n - sb.size())];
        if (da != db && da != 1 && db != 1) throw TensorError("Tensors are not broadcastable");
        out[i] = std::max(da, db);
    }
    return out;
}
std::vector<int64_t> broadcast_shape(const TensorPtr& a, const TensorPtr& b) { return infer_broadcast_shape(a, b); }
bool broadcastable(const TensorPtr& a, const TensorPtr& b) {
    try { infer_broadcast_shape(a, b); return true; } catch (const TensorError&) { return false; }
}

namespace {

// Describes how `small` broadcasts over `big` (big has the output shape): one of the four modes of mft_binary.
struct Bcast { int mode; int64_t outer, mid, inner, div; };

bool plan_broadcast(const std::vector<int64_t>& big, const std::vector<int64_t>& small, Bcast& p) {
    int64_t nb = 1, ns = 1;
    for (auto d : big) nb *= d;
    for (auto d : small) ns *= d;
    if (big == small) { p = {0, 1, 1, nb, 1}; return true; }
    if (ns == 1) return false;  // scalar tensors are handled by the caller
    // strip leading 1s of small
    std::vector<int64_t> s = small;
    while (!s.empty() && s.front() == 1 && s.size() > 1) s.erase(s.begin());
    // suffix match: small == trailing dims of big  -> mode 1 (1 dim) / mode 2 (>= 2 dims)
    if (s.size() <= big.size() && std::equal(s.begin(), s.end(), big.end() - s.size())) {
        if (s.size() == 1) { p = {1, 1, nb / s[0], s[0], 1}; return true; }
        int64_t inner = s.back(), mid = ns / inner;
        p = {2, nb / ns, mid, inner, 1};
        return true;
    }
    // [B,1,..,1,S] over [B,H,..,S]: same rank, first and last dims equal, the rest 1 -> mode 3
    if (small.size() == big.size() && small.size() >= 3 && small.front() == big.front() && small.back() == big.back()) {
        bool mid_ones = true;
        for (size_t i = 1; i + 1 < small.size(); ++i) mid_ones = mid_ones && small[i] == 1;
        if (mid_ones) {
            int64_t inner = big.back();
            int64_t per_b = nb / big.front() / inner;   // rows per batch entry
            p = {3, nb / inner, 1, inner, per_b};
            return true;
        }
    }
    return false;
}

// Sums `g` (shape gshape) down to `target` shape — AddBackward's sum_to_shape, core/backward_functions.cpp:36-64.
TensorPtr reduce_to_shape(const TensorPtr& g, const std::vector<int64_t>& target) {
    if (g->shape() == target) return g;
    int64_t nt = 1;
    for (auto d : target) nt *= d;
    auto gd = contiguous(g);
    Bcast p;
    if (nt == 1) {
        auto out = Tensor::make(target, kFloat32);
        size_t wsb = mft_colsum_workspace_bytes(gd->numel(), 1);
        mft_rt::check(mft_colsum(gd->dev_f32(), 1, out->dev_f32_out(), gd->numel(), 1, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_colsum");
        return out;
    }
    if (!plan_broadcast(gd->shape(), target, p)) throw TensorError("sum_to_shape: unsupported broadcast pattern");
    auto out = Tensor::make(target, kFloat32);
    if (p.mode == 1 || p.mode == 2) {
        const int64_t n = (p.mode == 1) ? p.inner : p.mid * p.inner;
        const int64_t rows = gd->numel() / n;
        size_t wsb = mft_colsum_workspace_bytes(rows, n);
        mft_rt::check(mft_colsum(gd->dev_f32(), n, out->dev_f32_out(), rows, n, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_colsum");
        return out;
    }
    // mode 3: [B, per_b, inner] summed over per_b, batched column sums
    const int64_t B = target.front(), inner = p.inner, per_b = p.div;
    for (int64_t b = 0; b < B; ++b) {
        size_t wsb = mft_colsum_workspace_bytes(per_b, inner);
        mft_rt::check(mft_colsum(gd->dev_f32() + b * per_b * inner, inner, out->dev_f32_out() + b * inner, per_b, inner,
                                 mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_colsum");
    }
    return out;
}

class BinaryBackward : public BackwardFunction {
public:
    BinaryBackward(int op, TensorPtr a, TensorPtr b) : op_(op), a_(std::move(a)), b_(std::move(b)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        TensorPtr ga, gb;
        const bool wa = wants_grad(a_), wb = wants_grad(b_);
        if (op_ == 0 || op_ == 1) {          // add / sub  (core/backward_functions.cpp:36-72)
            if (wa) ga = reduce_to_shape(gy, a_->shape());
            if (wb) {
                TensorPtr g = gy;
                if (op_ == 1) { g = new_like(gy); mft_rt::check(mft_affine(gy->dev_f32(), g->dev_f32_out(), gy->numel(), -1.f, 0.f, mft_rt::stream()), "mft_affine"); }
                gb = reduce_to_shape(g, b_->shape());
            }
        } else if (op_ == 2) {               // mul (core/backward_functions.cpp:74-83)
            if (a_->shape() == b_->shape()) {
                if (wa) ga = new_like(gy);
                if (wb) gb = new_like(gy);
                mft_rt::check(mft_mul_bwd(contiguous(a_)->dev_f32(), contiguous(b_)->dev_f32(), gy->dev_f32(),
                                          ga ? ga->dev_f32_out() : nullptr, gb ? gb->dev_f32_out() : nullptr, gy->numel(), mft_rt::stream()), "mft_mul_bwd");
            } else {
                if (wa) ga = reduce_to_shape(mul(gy, b_), a_->shape());
                if (wb) gb = reduce_to_shape(mul(gy, a_), b_->shape());
            }
        } else {                             // div: ga = gy/b ; gb = -gy*a/b^2
            if (wa) ga = reduce_to_shape(div(gy, b_), a_->shape());
            if (wb) gb = reduce_to_shape(mul(div(mul(gy, a_), mul(b_, b_)), -1.0f), b_->shape());
        }
        return {ga, gb};
    }
private:
    int op_;
    TensorPtr a_, b_;
};

TensorPtr binary(int op, const TensorPtr& a_in, const TensorPtr& b_in, const char* name) {
    require_f32(a_in, name);
    require_f32(b_in, name);
    TensorPtr a = contiguous(a_in), b = contiguous(b_in);
    auto oshape = infer_broadcast_shape(a, b);
    auto out = Tensor::make(oshape, kFloat32);
    Bcast p;
    auto st = mft_rt::stream();
    if (a->shape() == oshape && plan_broadcast(oshape, b->shape(), p)) {
        mft_rt::check(mft_binary(op, a->dev_f32(), b->dev_f32(), out->dev_f32_out(), p.outer, p.mid, p.inner, p.mode, p.div, st), name);
    } else if (b->numel() == 1 && a->shape() == oshape) {
        // scalar tensor on the right: broadcast as a 1-element bias
        mft_rt::check(mft_binary(op, a->dev_f32(), b->dev_f32(), out->dev_f32_out(), 1, a->numel(), 1, 1, 1, st), name);
    } else if (b->shape() == oshape && (op == 0 || op == 2) && (plan_broadcast(oshape, a->shape(), p) || a->numel() == 1)) {
        // commutative ops with the small operand on the left
        if (a->numel() == 1) mft_rt::check(mft_binary(op, b->dev_f32(), a->dev_f32(), out->dev_f32_out(), 1, b->numel(), 1, 1, 1, st), name);
        else mft_rt::check(mft_binary(op, b->dev_f32(), a->dev_f32(), out->dev_f32_out(), p.outer, p.mid, p.inner, p.mode, p.div, st), name);
    } else {
        throw TensorError(std::string(name) + ": unsupported broadcast pattern on the device path");
    }
    track(out, {a_in, b_in}, std::make_shared<BinaryBackward>(op, a_in, b_in));
    return out;
}

class AffineBackward : public BackwardFunction {   // ScaleBackward: core/backward_functions.cpp:970-974
public:
    explicit AffineBackward(float s) : s_(s) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        if (s_ == 1.0f) return {gy_in};
        TensorPtr gy = contiguous(gy_in);
        auto g = new_like(gy);
        mft_rt::check(mft_affine(gy->dev_f32(), g->dev_f32_out(), gy->numel(), s_, 0.f, mft_rt::stream()), "mft_affine");
        return {g};
    }
private:
    float s_;
};

TensorPtr affine(const TensorPtr& t_in, float alpha, float beta, const char* name) {
    require_f32(t_in, name);
    TensorPtr t = contiguous(t_in);
    auto out = new_like(t);
    mft_rt::check(mft_affine(t->dev_f32(), out->dev_f32_out(), t->numel(), alpha, beta, mft_rt::stream()), name);
    track(out, {t_in}, std::make_shared<AffineBackward>(alpha));
    return out;
}

}  // namespace

// core/ops.cpp:166-484
TensorPtr add(const TensorPtr& a, const TensorPtr& b) { return binary(0, a, b, "add"); }
TensorPtr sub(const TensorPtr& a, const TensorPtr& b) { return binary(1, a, b, "sub"); }
TensorPtr mul(const TensorPtr& a, const TensorPtr& b) { return binary(2, a, b, "mul"); }
TensorPtr div(const TensorPtr& a, const TensorPtr& b) { return binary(3, a, b, "div"); }
TensorPtr add(const TensorPtr& t, float s) { return affine(t, 1.f, s, "add(scalar)"); }
TensorPtr sub(const TensorPtr& t, float s) { return affine(t, 1.f, -s, "sub(scalar)"); }
TensorPtr mul(const TensorPtr& t, float s) { return affine(t, s, 0.f, "mul(scalar)"); }
TensorPtr div(const TensorPtr& t, float s) { return affine(t, 1.0f / s, 0.f, "div(scalar)"); }
TensorPtr add(float s, const TensorPtr& t) { return affine(t, 1.f, s, "add(scalar)"); }
TensorPtr sub(float s, const TensorPtr& t) { return affine(t, -1.f, s, "sub(scalar)"); }
TensorPtr mul(float s, const TensorPtr& t) { return affine(t, s, 0.f, "mul(scalar)"); }
TensorPtr div(float s, const TensorPtr& t) { return mul(pow(t, -1.0f), s); }

}  // namespace ops

// matmul / transpose / permute / fused LoRA linear / tied lm_head — the tensor-core side of the op library.
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <unordered_map>

#include "autograd_engine.h"
#include "mft_b200.h"
#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

using namespace internal;

namespace {

int64_t leading(const std::vector<int64_t>& s) {
    int64_t b = 1;
    for (size_t i = 0; i + 2 < s.size(); ++i) b *= s[i];
    return b;
}

std::string shape_str(const std::vector<int64_t>& s) {
    std::string r = "[";
    for (size_t i = 0; i < s.size(); ++i) r += (i ? ", " : "") + std::to_string(s[i]);
    return r + "]";
}

// MatmulBackward (core/backward_functions.cpp:94-138). `b_nk` = b is stored [n,k] (the auto-transposed case of
// core/ops.cpp:506-512, whose TransposeBackward hands the gradient back in b's own layout).
class MatmulBackward : public BackwardFunction {
public:
    MatmulBackward(TensorPtr a, TensorPtr b, bool b_nk) : a_(std::move(a)), b_(std::move(b)), b_nk_(b_nk) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        const auto& sa = a_->shape();
        const auto& sb = b_->shape();
        const int64_t m = sa[sa.size() - 2], k = sa.back();
        const int64_t n = b_nk_ ? sb[sb.size() - 2] : sb.back();
        const bool shared_b = sb.size() == 2;
        const int64_t batch = leading(sa);
        TensorPtr ga, gb;
        if (wants_grad(a_) || Engine_needs(a_)) {           // dA = dC * B^T
            ga = Tensor::make(sa, kFloat32);
            GemmCall g;
            g.k = n; g.n = k;
            g.a = gy->dev_f32(); g.lda = gy->ld(); g.trans_a = 0;
            g.b = b_->dev_f32(); g.ldb = b_->ld();
            g.trans_b = b_nk_ ? 0 : 1;                      // contraction over n: b stored [k,n] is already K-major
            g.c = ga->dev_f32_out(); g.ldc = k;
            if (shared_b) { g.m = batch * m; g.b_weight = b_; }
            else { g.m = m; g.batch = int(batch); g.stride_a = m * gy->ld(); g.stride_b = sb[sb.size() - 2] * b_->ld(); g.stride_c = m * k; }
            run_gemm(g);
        }
        if (wants_grad(b_)) {                               // dB = A^T * dC   (or dC^T * A for [n,k] storage)
            gb = Tensor::make(sb, kFloat32);
            GemmCall g;
            const TensorPtr& lhs = b_nk_ ? gy : a_;         // stored [rows, M'] -> M' x rows after transpose
            const TensorPtr& rhs = b_nk_ ? a_ : gy;
            g.trans_a = 1; g.trans_b = 0;
            g.a = lhs->dev_f32(); g.lda = lhs->ld();
            g.b = rhs->dev_f32(); g.ldb = rhs->ld();
            g.c = gb->dev_f32_out(); g.ldc = sb.back();
            g.m = sb[sb.size() - 2]; g.n = sb.back();
            if (shared_b) { g.k = batch * m; }               // flattened over the leading dims (:104-127)
            else { g.k = m; g.batch = int(batch); g.stride_a = m * lhs->ld(); g.stride_b = m * rhs->ld(); g.stride_c = g.m * g.n; }
            run_gemm(g);
        }
        return {ga, gb};
    }
private:
    static bool Engine_needs(const TensorPtr& t) { return autograd::Engine::instance().needs_grad(t); }
    TensorPtr a_, b_;
    bool b_nk_;
};

TensorPtr matmul_impl(const TensorPtr& a, const TensorPtr& b, bool b_nk) {
    const auto& sa = a->shape();
    const auto& sb = b->shape();
    const int64_t m = sa[sa.size() - 2], k = sa.back();
    const int64_t n = b_nk ? sb[sb.size() - 2] : sb.back();
    const bool shared_b = sb.size() == 2;
    const int64_t batch = leading(sa);
    if (!shared_b && (sb.size() != sa.size() || leading(sb) != batch))
        throw TensorError("matmul: batched operands must have equal leading dimensions, got A " + shape_str(sa) + " B " + shape_str(sb));
    std::vector<int64_t> so = sa;
    so.back() = n;
    auto out = Tensor::make(so, kFloat32, false, out_ld(n));
    GemmCall g;
    g.n = n; g.k = k;
    g.a = a->dev_f32(); g.lda = a->ld(); g.trans_a = 0;
    g.b = b->dev_f32(); g.ldb = b->ld(); g.trans_b = b_nk ? 1 : 0;
    g.c = out->dev_f32_out(); g.ldc = out->ld();
    if (shared_b) { g.m = batch * m; g.b_weight = b; }
    else { g.m = m; g.batch = int(batch); g.stride_a = m * a->ld(); g.stride_b = sb[sb.size() - 2] * b->ld(); g.stride_c = m * out->ld(); }
    run_gemm(g);
    track(out, {a, b}, std::make_shared<MatmulBackward>(a, b, b_nk));
    return out;
}

}  // namespace

// core/ops.cpp:486-651
TensorPtr matmul(const TensorPtr& a, const TensorPtr& b) {
    if (!a || !b) throw TensorError("matmul: input tensors must not be null");
    require_f32(a, "matmul");
    require_f32(b, "matmul");
    const auto& sa = a->shape();
    const auto& sb = b->shape();
    if (sa.size() < 2 || sb.size() < 2) throw TensorError("matmul requires tensors with at least 2 dimensions");
    const int64_t k = sa.back();
    if (k == sb[sb.size() - 2]) return matmul_impl(a, b, false);
    if (k == sb.back()) return matmul_impl(a, b, true);     // B given as [out,in]: transposed on the fly (:506-512)
    throw TensorError("matmul dimension mismatch:\n  A shape: " + shape_str(sa) + "\n  B shape: " + shape_str(sb) +
                      "\n  A's K dimension=" + std::to_string(k) + ", B's K dimension=" + std::to_string(sb[sb.size() - 2]) +
                      "\n  Hint: If B is a weight matrix, you may need transpose(B)");
}

// core/ops.cpp:653-780 — a @ b^T with b [.., n, k]
TensorPtr matmul_rhs_T(const TensorPtr& a, const TensorPtr& b) {
    if (!a || !b) throw TensorError("matmul_rhs_T: input tensors must not be null");
    require_f32(a, "matmul_rhs_T");
    require_f32(b, "matmul_rhs_T");
    if (a->ndim() < 2 || b->ndim() < 2) throw TensorError("matmul_rhs_T requires tensors with at least 2 dimensions");
    if (a->shape().back() != b->shape().back()) throw TensorError("matmul_rhs_T: inner dimensions do not match");
    return matmul_impl(a, b, true);
}

// ---------------------------------------------------------------------------- transpose / permute
namespace {
class PermuteBackward : public BackwardFunction {
public:
    explicit PermuteBackward(std::vector<int> inv) : inv_(std::move(inv)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        autograd::Engine& e = autograd::Engine::instance();
        const bool was = e.is_enabled();
        e.set_enabled(false);
        auto g = permute(gy, inv_);
        e.set_enabled(was);
        return {g};
    }
private:
    std::vector<int> inv_;
};
}  // namespace

// core/ops.cpp:886-961 (general) with device fast paths for the layouts the models use:
// swap of the last two dims (any rank) and {0,2,1,3}.
TensorPtr permute(const TensorPtr& t_in, const std::vector<int>& dims_in) {
    require_f32(t_in, "permute");
    TensorPtr t = contiguous(t_in);
    const int nd = int(t->ndim());
    if (int(dims_in.size()) != nd) throw TensorError("permute: dims size mismatch");
    std::vector<int> dims(nd), inv(nd);
    std::vector<bool> seen(nd, false);
    for (int i = 0; i < nd; ++i) {
        int d = dims_in[i] < 0 ? dims_in[i] + nd : dims_in[i];
        if (d < 0 || d >= nd || seen[d]) throw TensorError("permute: invalid dims");
        seen[d] = true; dims[i] = d; inv[d] = i;
    }
    const auto& s = t->shape();
    std::vector<int64_t> so(nd);
    for (int i = 0; i < nd; ++i) so[i] = s[dims[i]];
    bool identity = true;
    for (int i = 0; i < nd; ++i) identity = identity && dims[i] == i;
    TensorPtr out;
    auto st = mft_rt::stream();
    bool last2 = nd >= 2 && dims[nd - 1] == nd - 2 && dims[nd - 2] == nd - 1;
    for (int i = 0; i + 2 < nd; ++i) last2 = last2 && dims[i] == i;
    if (identity) {
        out = t->clone();
    } else if (last2) {
        out = Tensor::make(so, kFloat32);
        int64_t r = s[nd - 2], c = s[nd - 1], b = t->numel() / std::max<int64_t>(r * c, 1);
        mft_rt::check(mft_transpose2d(t->dev_f32(), c, out->dev_f32_out(), r, r, c, b, r * c, r * c, st), "mft_transpose2d");
    } else if (nd == 4 && dims == std::vector<int>{0, 2, 1, 3}) {
        out = Tensor::make(so, kFloat32);
        mft_rt::check(mft_permute0213(t->dev_f32(), out->dev_f32_out(), s[0], s[1], s[2], s[3], st), "mft_permute0213");
    } else if (nd == 3 && dims == std::vector<int>{1, 0, 2}) {
        out = Tensor::make(so, kFloat32);
        mft_rt::check(mft_permute0213(t->dev_f32(), out->dev_f32_out(), 1, s[0], s[1], s[2], st), "mft_permute0213");
    } else {
        // general case: one gather kernel (no host staging)
        if (nd > 8) throw TensorError("permute: more than 8 dimensions are not supported");
        out = Tensor::make(so, kFloat32);
        std::vector<int64_t> stride(nd, 1), src_stride(nd);
        for (int i = nd - 2; i >= 0; --i) stride[i] = stride[i + 1] * s[i + 1];
        for (int i = 0; i < nd; ++i) src_stride[i] = stride[dims[i]];
        mft_rt::check(mft_permute_nd(t->dev_f32(), out->dev_f32_out(), nd, so.data(), src_stride.data(), st), "mft_permute_nd");
    }
    track(out, {t_in}, std::make_shared<PermuteBackward>(inv));
    return out;
}

// core/ops.cpp:782-884
TensorPtr transpose(const TensorPtr& t, int dim0, int dim1) {
    if (!t) throw TensorError("transpose: null tensor");
    const int nd = int(t->ndim());
    if (dim0 < 0) dim0 += nd;
    if (dim1 < 0) dim1 += nd;
    if (dim0 < 0 || dim0 >= nd || dim1 < 0 || dim1 >= nd) throw TensorError("transpose: dimension out of range");
    std::vector<int> dims(nd);
    for (int i = 0; i < nd; ++i) dims[i] = i;
    std::swap(dims[dim0], dims[dim1]);
    return permute(t, dims);
}

// ---------------------------------------------------------------------------- fused LoRA linear
namespace {

struct TermLayout { bool a_in_r; bool b_r_out; int64_t r; };

TermLayout classify(const LoraTerm& t, int64_t in, int64_t out) {
    const auto& sa = t.A->shape();
    const auto& sb = t.B->shape();
    if (sa.size() != 2 || sb.size() != 2) throw std::runtime_error("LoRA A and B must be 2-D");
    TermLayout l;
    if (sa[0] == in) { l.a_in_r = true; l.r = sa[1]; }               // same precedence as matmul's auto-transpose
    else if (sa[1] == in) { l.a_in_r = false; l.r = sa[0]; }
    else throw TensorError("LoRA A shape " + shape_str(sa) + " does not match input width " + std::to_string(in));
    if (sb[0] == l.r && sb[1] == out) l.b_r_out = true;
    else if (sb[1] == l.r && sb[0] == out) l.b_r_out = false;
    else throw TensorError("LoRA B shape " + shape_str(sb) + " does not match rank/out");
    return l;
}

// Small [rows,cols] -> [cols,rows] device transpose into a fresh dense tensor.
TensorPtr small_transpose(const TensorPtr& t) {
    auto o = Tensor::make({t->shape()[1], t->shape()[0]}, kFloat32);
    mft_rt::check(mft_transpose2d(t->dev_f32(), t->ld(), o->dev_f32_out(), o->ld(), t->shape()[0], t->shape()[1], 1, 0, 0, mft_rt::stream()), "mft_transpose2d");
    return o;
}

// MFT_LORA_SKINNY: bit 0 routes x*A / dY*B^T through mft_lora_down, bit 1 routes dA / dB through mft_lora_tn; the rest goes
// through the general GEMM dispatch (SIMT split-K for these shapes).
int skinny_mask() {
    static const int m = [] { const char* e = std::getenv("MFT_LORA_SKINNY"); return e ? std::atoi(e) : 3; }();
    return m;
}
// y[T,r] = alpha * x[T,K] * A   (A [K,r] when a_is_kr, else [r,K])
void lora_down(const TensorPtr& x, int64_t T, int64_t K, const TensorPtr& A, bool a_is_kr, int64_t r, float alpha, const TensorPtr& y) {
    // bf16 operand mode: x already exists as bf16 (written by its producer) — read that copy (half the bytes) against the
    // factor's K-major bf16 shadow. MFT_LORA_DOWN_BF16=0 keeps the fp32 kernel.
    const char* lowp_env = std::getenv("MFT_LORA_DOWN_BF16");          // read per call: tests compare the two paths
    const bool lowp_on = !(lowp_env && lowp_env[0] == '0');
    if (lowp_on && (skinny_mask() & 1) && emit_bf16_operands() && gemm_precision() == 1 && mft_tensor_cores_enabled() && r >= 1 && r <= 64 &&
        K % 128 == 0 && K <= (r <= 16 ? 6144 : (r <= 32 ? 3072 : 1536)) && x->has_bf16() && A->ndim() == 2) {
        auto xv = x->bf16_view();
        auto av = a_is_kr ? A->bf16_transposed_view() : A->bf16_view();      // [r][K], cached per version of the factor
        if (xv.ld % 8 == 0 && av.ld % 8 == 0) {
            float* yp = y->dev_f32_out();
            auto e = y->bf16_emit();
            mft_rt::check(mft_lora_down_bf16(xv.ptr, xv.ld, av.ptr, av.ld, yp, r, e.ptr, e.ld, T, K, int32_t(r), alpha, mft_rt::stream()), "mft_lora_down_bf16");
            return;
        }
    }
    if ((skinny_mask() & 1) && r >= 1 && r <= 64) {
        float* yp = y->dev_f32_out();
        if (emit_bf16_operands()) {      // the rank-r product is the A operand of the bf16 GEMM's second K segment
            auto e = y->bf16_emit();
            mft_rt::check(mft_lora_down_ex(x->dev_f32(), x->ld(), A->dev_f32(), A->ld(), a_is_kr ? 1 : 0, yp, r, e.ptr, e.ld, T, K, int32_t(r), alpha,
                                           mft_rt::stream()), "mft_lora_down_ex");
        } else {
            mft_rt::check(mft_lora_down(x->dev_f32(), x->ld(), A->dev_f32(), A->ld(), a_is_kr ? 1 : 0, yp, r, T, K, int32_t(r), alpha,
                                        mft_rt::stream()), "mft_lora_down");
        }
        return;
    }
    GemmCall g;
    g.m = T; g.n = r; g.k = K; g.alpha = alpha;
    g.a = x->dev_f32(); g.lda = x->ld(); g.trans_a = 0;
    g.b = A->dev_f32(); g.ldb = A->ld(); g.trans_b = a_is_kr ? 0 : 1;
    g.c = y->dev_f32_out(); g.ldc = r;
    run_gemm(g);
}
// ---- adapter gradients of one backward pass, batched: while the engine runs a backward, every eligible dA / dB product is
// only recorded; flush_deferred_lora_tn() (called by the engine when the last node has run) does all of them in one grouped
// launch. The operands stay untouched until then: tensors are never modified after they are produced, and the recorded
// TensorPtrs keep them alive. MFT_LORA_TN_GROUP=0 launches each product on the spot as before.
struct DeferredTn {
    bool active = false;
    std::vector<mft_lora_tn_problem> probs;
    std::vector<TensorPtr> keep;
    struct Table { void* host = nullptr; void* dev = nullptr; size_t bytes = 0; };
    std::unordered_map<uint64_t, std::vector<Table>> tables;      // content hash -> immutable (pinned, device) copies
} g_tn;

bool tn_group_enabled() {
    static const bool on = [] { const char* e = std::getenv("MFT_LORA_TN_GROUP"); return !(e && e[0] == '0'); }();
    return on;
}

// out = S[T,r]^T W[T,q], written as [r,q] (out_is_rq) or [q,r]
void lora_tn(const TensorPtr& S, const TensorPtr& W, int64_t T, int64_t q, int64_t r, bool out_is_rq, const TensorPtr& out) {
    if (g_tn.active && (skinny_mask() & 2) && tn_group_enabled()) {
        const float* wp = W->dev_f32();
        if (mft_lora_tn_group_eligible(T, q, int32_t(r), wp, W->ld())) {
            mft_lora_tn_problem p;
            std::memset(&p, 0, sizeof(p));
            p.colblocks = mft_lora_tn_group_colblocks(q);
            p.splits = mft_lora_tn_group_splits(T);
            auto partial = Tensor::make({int64_t(p.splits) * r * q}, kFloat32);
            p.s = S->dev_f32(); p.lds = S->ld(); p.w = wp; p.ldw = W->ld(); p.partial = partial->dev_f32_out();
            p.out = out->dev_f32_out(); p.ldo = out->ld(); p.T = T; p.q = q; p.r = int32_t(r); p.out_is_rq = out_is_rq ? 1 : 0; p.alpha = 1.0f;
            p.use_tickets = p.splits <= 64 ? 1 : 0;
            g_tn.probs.push_back(p);
            g_tn.keep.push_back(S); g_tn.keep.push_back(W); g_tn.keep.push_back(out); g_tn.keep.push_back(partial);
            return;
        }
    }
    if ((skinny_mask() & 2) && r >= 1 && r <= 64) {
        const size_t wsb = mft_lora_tn_workspace_bytes(T, q, int32_t(r));
        mft_rt::check(mft_lora_tn(S->dev_f32(), S->ld(), W->dev_f32(), W->ld(), out->dev_f32_out(), out->ld(), out_is_rq ? 1 : 0, T, q, int32_t(r),
                                  1.0f, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_lora_tn");
        return;
    }
    GemmCall g;
    g.trans_a = 1; g.trans_b = 0; g.k = T;
    if (out_is_rq) { g.m = r; g.n = q; g.a = S->dev_f32(); g.lda = S->ld(); g.b = W->dev_f32(); g.ldb = W->ld(); }
    else { g.m = q; g.n = r; g.a = W->dev_f32(); g.lda = W->ld(); g.b = S->dev_f32(); g.ldb = S->ld(); }
    g.c = out->dev_f32_out(); g.ldc = out->ld();
    run_gemm(g);
}

}  // namespace

namespace internal {
void begin_deferred_lora_tn() { g_tn.active = true; }
void flush_deferred_lora_tn() {
    g_tn.active = false;
    if (g_tn.probs.empty()) return;
    // ticket ranges; problems that do not fit the ticket table fall back to the separate reduce pass
    int32_t base = 0, max_cb = 1, max_sp = 1, any_long = 0;
    int64_t max_rq = 1;
    const int32_t cap = mft_lora_tn_group_ticket_capacity();
    for (auto& p : g_tn.probs) {
        if (p.use_tickets && base + p.colblocks > cap) p.use_tickets = 0;
        if (p.use_tickets) { p.ticket_base = base; base += p.colblocks; } else any_long = 1;
        max_cb = std::max(max_cb, p.colblocks); max_sp = std::max(max_sp, p.splits);
        max_rq = std::max<int64_t>(max_rq, int64_t(p.r) * p.q);
    }
    // The table is uploaded from pinned memory; under CUDA-graph replay that copy is re-executed, so a (pinned, device) pair
    // is written once and never changed: one pair per distinct table content (a step's problems repeat with the same
    // addresses, thanks to the step arena).
    const size_t bytes = g_tn.probs.size() * sizeof(mft_lora_tn_problem);
    uint64_t h = 1469598103934665603ull;
    const unsigned char* raw = reinterpret_cast<const unsigned char*>(g_tn.probs.data());
    for (size_t i = 0; i < bytes; ++i) { h ^= raw[i]; h *= 1099511628211ull; }
    DeferredTn::Table* tab = nullptr;
    for (auto& c : g_tn.tables[h])
        if (c.bytes == bytes && std::memcmp(c.host, raw, bytes) == 0) { tab = &c; break; }
    if (!tab) {
        DeferredTn::Table c;
        c.bytes = bytes;
        c.host = mft_rt::pinned_alloc(bytes);
        std::memcpy(c.host, raw, bytes);
        { mft_rt::PersistentScope keep; c.dev = mft_rt::dev_alloc(bytes); }
        g_tn.tables[h].push_back(c);
        tab = &g_tn.tables[h].back();
    }
    mft_rt::h2d(tab->dev, tab->host, bytes);
    const int32_t n = int32_t(g_tn.probs.size());
    g_tn.probs.clear();
    std::vector<TensorPtr> keep;
    keep.swap(g_tn.keep);               // released when this function returns: the launch below is already enqueued
    mft_rt::check(mft_lora_tn_group(static_cast<const mft_lora_tn_problem*>(tab->dev), n, max_cb, max_sp, any_long, max_rq, mft_rt::stream()),
                  "mft_lora_tn_group");
}
}  // namespace internal

namespace {

// One node for  y = x W (+b) + sum_i s_i (x A_i) B_i.
// Backward (per term, s folded into xa / dxa):   dB = xa^T gy ;  dxa = s gy B^T ;  dA = x^T dxa ;
// dx = gy W^T + sum dxa A^T (rank-r update fused as second K segment) ; dW = x^T gy ; db = colsum(gy).
class LoraLinearBackward : public BackwardFunction {
public:
    LoraLinearBackward(TensorPtr x, TensorPtr W, TensorPtr bias, std::vector<LoraTerm> terms, std::vector<TensorPtr> xa,
                       TensorPtr gelu_pre = nullptr, bool has_residual = false)
        : x_(std::move(x)), W_(std::move(W)), bias_(std::move(bias)), terms_(std::move(terms)), xa_(std::move(xa)),
          pre_(std::move(gelu_pre)), has_residual_(has_residual) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        const int64_t in = W_->shape()[0], out = W_->shape()[1];
        const int64_t T = x_->numel() / in;
        std::vector<TensorPtr> grads(3 + 2 * terms_.size() + (has_residual_ ? 1 : 0));
        if (has_residual_) grads.back() = gy;               // y = ... + residual: the gradient passes through unchanged
        const bool need_dx = autograd::Engine::instance().needs_grad(pre_ ? pre_ : x_);
        std::vector<TensorPtr> dxa(terms_.size());
        for (size_t i = 0; i < terms_.size(); ++i) {
            const LoraTerm& t = terms_[i];
            TermLayout l = classify(t, in, out);
            const bool wa = wants_grad(t.A), wb = wants_grad(t.B);
            if (wb) {                                        // dB = (s x A)^T gy, written in B's own layout
                auto gB = Tensor::make(t.B->shape(), kFloat32);
                lora_tn(xa_[i], gy, T, out, l.r, l.b_r_out, gB);
                grads[3 + 2 * i + 1] = gB;
            }
            if (wa || need_dx) {                             // dxa = s gy B^T   [T, r]
                dxa[i] = Tensor::make({T, l.r}, kFloat32);
                lora_down(gy, T, out, t.B, /*a_is_kr=*/!l.b_r_out, l.r, t.scale, dxa[i]);   // contraction over `out`
            }
            if (wa) {                                        // dA = x^T dxa, in A's own layout
                auto gA = Tensor::make(t.A->shape(), kFloat32);
                lora_tn(dxa[i], x_, T, in, l.r, /*out_is_rq=*/!l.a_in_r, gA);
                grads[3 + 2 * i] = gA;
            }
        }
        if (need_dx) {                                       // dx = gy W^T (+ dxa A^T)
            auto gx = Tensor::make(x_->shape(), kFloat32);
            GemmCall g;
            g.m = T; g.n = in; g.k = out;
            g.a = nullptr; g.lda = gy->ld(); g.trans_a = 0; g.a_tensor = gy;
            g.b = W_->dev_f32(); g.ldb = W_->ld(); g.trans_b = 1; g.b_weight = W_;   // W stored [in,out]: K-major for this product
            g.c = nullptr; g.c_tensor = gx; g.ldc = in;
            g.c_bf16_only = x_->grad_may_be_bf16() && !pre_ && terms_.size() <= 1;     // dX of a GELU activation: read by GeluBackward as bf16
            std::vector<TensorPtr> keep;
            for (size_t i = 0; i < terms_.size(); ++i) {
                TermLayout l = classify(terms_[i], in, out);
                TensorPtr a_in_r = (l.a_in_r || (i == 0 && gemm_takes_bf16(T, in, out))) ? terms_[i].A : small_transpose(terms_[i].A);
                keep.push_back(a_in_r);
                if (i == 0) {
                    g.k2 = l.r; g.a2 = dxa[i]->dev_f32(); g.lda2 = l.r; g.b2 = a_in_r->dev_f32(); g.ldb2 = a_in_r->ld();
                    g.a2_tensor = dxa[i];
                    g.b2_tensor = terms_[i].A; g.b2_stored_kn = !l.a_in_r;      // A [in,r] is K-major as stored; PEFT [r,in] needs the transposed shadow
                }
            }
            const bool fuse_gelu_bwd = pre_ && terms_.size() <= 1;   // further slices accumulate after the epilogue ran
            if (fuse_gelu_bwd) { g.gelu_grad_x = pre_->dev_f32(); g.ld_gelu_grad_x = pre_->ld(); }
            run_gemm(g);
            for (size_t i = 1; i < terms_.size(); ++i) {     // further slices: accumulate
                TermLayout l = classify(terms_[i], in, out);
                GemmCall h;
                h.m = T; h.n = in; h.k = l.r; h.beta = 1.0f;
                h.a = dxa[i]->dev_f32(); h.lda = l.r; h.trans_a = 0;
                h.b = keep[i]->dev_f32(); h.ldb = keep[i]->ld(); h.trans_b = 1;
                h.c = gx->dev_f32_mut(); h.ldc = in;
                run_gemm(h);
            }
            if (pre_ && !fuse_gelu_bwd) {
                TensorPtr pc = contiguous(pre_);
                float* gxp = gx->dev_f32_mut();
                mft_rt::check(mft_gelu_bwd(pc->dev_f32(), gxp, gxp, T * in, mft_rt::stream()), "mft_gelu_bwd");
            }
            grads[0] = gx;
        }
        if (wants_grad(W_)) {                                // dW = x^T gy  [in,out]
            auto gW = Tensor::make(W_->shape(), kFloat32);
            GemmCall g;
            g.m = in; g.n = out; g.k = T; g.trans_a = 1; g.trans_b = 0;
            g.a = x_->dev_f32(); g.lda = x_->ld(); g.b = gy->dev_f32(); g.ldb = gy->ld();
            g.c = gW->dev_f32_out(); g.ldc = out;
            run_gemm(g);
            grads[1] = gW;
        }
        if (bias_ && wants_grad(bias_)) {
            auto gb = Tensor::make(bias_->shape(), kFloat32);
            size_t wsb = mft_colsum_workspace_bytes(T, out);
            mft_rt::check(mft_colsum(gy->dev_f32(), gy->ld(), gb->dev_f32_out(), T, out, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_colsum");
            grads[2] = gb;
        }
        return grads;
    }
private:
    TensorPtr x_, W_, bias_;
    std::vector<LoraTerm> terms_;
    std::vector<TensorPtr> xa_;
    TensorPtr pre_;            // x_ == gelu(pre_): dx is returned for pre_
    bool has_residual_;
};

}  // namespace

TensorPtr fused_lora_linear(const TensorPtr& x_in, const TensorPtr& W, const TensorPtr& bias, const std::vector<LoraTerm>& terms,
                            TensorPtr* gelu_act) {
    LinearFusion f;
    f.gelu_act = gelu_act;
    return fused_lora_linear(x_in, W, bias, terms, f);
}

TensorPtr fused_lora_linear(const TensorPtr& x_in, const TensorPtr& W, const TensorPtr& bias, const std::vector<LoraTerm>& terms,
                            const LinearFusion& fusion) {
    TensorPtr* gelu_act = fusion.gelu_act;
    require_f32(x_in, "lora_linear");
    require_f32(W, "lora_linear");
    if (W->ndim() != 2) throw TensorError("lora_linear: weight must be [in, out]");
    TensorPtr x = contiguous(x_in);
    const int64_t in = W->shape()[0], out = W->shape()[1];
    if (x->shape().back() != in) throw TensorError("matmul dimension mismatch:\n  A shape: " + shape_str(x->shape()) + "\n  B shape: " + shape_str(W->shape()));
    if (bias && bias->numel() != out) throw TensorError("lora_linear: bias size mismatch");
    const int64_t T = x->numel() / in;
    std::vector<int64_t> so = x->shape();
    so.back() = out;
    TensorPtr residual = fusion.residual;
    if (residual) {
        require_f32(residual, "lora_linear");
        if (residual->shape() != so) throw TensorError("lora_linear: residual shape " + shape_str(residual->shape()) + " does not match the output " + shape_str(so));
    }
    const TensorPtr& pre = fusion.gelu_input_pre;
    if (pre && pre->shape() != x_in->shape()) throw TensorError("lora_linear: gelu_input_pre must have the input's shape");
    auto y = Tensor::make(so, kFloat32, false, out_ld(out));
    std::vector<TensorPtr> xa(terms.size()), bt(terms.size());
    for (size_t i = 0; i < terms.size(); ++i) {
        TermLayout l = classify(terms[i], in, out);
        xa[i] = Tensor::make({T, l.r}, kFloat32);            // s * x A   (kept for dB)
        lora_down(x, T, in, terms[i].A, l.a_in_r, l.r, terms[i].scale, xa[i]);
        // [out, r] K-major; the bf16 GEMM form reads the factor's cached bf16 shadow instead (b2_tensor below)
        bt[i] = (l.b_r_out && !(i == 0 && gemm_takes_bf16(T, out, in))) ? small_transpose(terms[i].B) : terms[i].B;
    }
    GemmCall g;
    g.m = T; g.n = out; g.k = in;
    g.a = nullptr; g.lda = in; g.trans_a = 0; g.a_tensor = x;          // resolved by run_gemm (bf16 copy or fp32 data)
    g.b = W->dev_f32(); g.ldb = W->ld(); g.trans_b = 0; g.b_weight = W;
    g.c = nullptr; g.c_tensor = y; g.ldc = y->ld();
    g.c_bf16_only = fusion.out_bf16_only && terms.size() <= 1 && !residual;
    g.bias = bias ? bias->dev_f32() : nullptr;
    TensorPtr act;
    const bool fuse_gelu = gelu_act != nullptr && terms.size() <= 1;   // later slices accumulate after the epilogue ran
    if (fuse_gelu) {
        act = Tensor::make(so, kFloat32, false, y->ld());
        g.gelu_tensor = act;
        if (fusion.gelu_bf16_only) g.gelu_bf16_only = true; else g.gelu_out = act->dev_f32_out();
        act->set_grad_may_be_bf16(true);      // its gradient goes to GeluBackward, which reads bf16 directly
    }
    if (!terms.empty()) {
        g.k2 = xa[0]->shape()[1]; g.a2 = xa[0]->dev_f32(); g.lda2 = g.k2; g.b2 = bt[0]->dev_f32(); g.ldb2 = bt[0]->ld();
        g.a2_tensor = xa[0];
        g.b2_tensor = terms[0].B; g.b2_stored_kn = classify(terms[0], in, out).b_r_out;   // B [r,out]: transposed shadow; PEFT [out,r]: as stored
    }
    if (residual) { g.addend = residual->dev_f32(); g.ld_addend = residual->ld(); }   // (x W + b) + residual, then the slices below
    run_gemm(g);
    for (size_t i = 1; i < terms.size(); ++i) {
        GemmCall h;
        h.m = T; h.n = out; h.k = xa[i]->shape()[1]; h.beta = 1.0f;
        h.a = xa[i]->dev_f32(); h.lda = h.k; h.trans_a = 0;
        h.b = bt[i]->dev_f32(); h.ldb = bt[i]->ld(); h.trans_b = 1;
        h.c = y->dev_f32_mut(); h.ldc = y->ld();
        run_gemm(h);
    }
    std::vector<TensorPtr> inputs{pre ? pre : x_in, W, bias};
    for (const auto& t : terms) { inputs.push_back(t.A); inputs.push_back(t.B); }
    if (residual) inputs.push_back(residual);
    track(y, inputs, std::make_shared<LoraLinearBackward>(x, W, bias, terms, xa, pre, residual != nullptr));
    if (gelu_act) {
        if (fuse_gelu) { track(act, {y}, make_gelu_backward(y)); *gelu_act = act; }
        else *gelu_act = gelu(y);
    }
    return y;
}

// core/ops.cpp:963-1031 — output = input @ weight + bias (weight [in,out])
TensorPtr linear(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias) {
    return fused_lora_linear(input, weight, bias, {});
}

// core/ops.cpp:2266-2439 — older fused entry point kept for API compatibility: W [in,out], A [in,r], B [r,out]
TensorPtr lora_linear(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& lora_A, const TensorPtr& lora_B,
                      float alpha, const TensorPtr& bias) {
    if (!lora_A || !lora_B) return fused_lora_linear(input, weight, bias, {});
    return fused_lora_linear(input, weight, bias, {LoraTerm{lora_A, lora_B, alpha}});
}

// ---------------------------------------------------------------------------- tied lm_head
namespace {
class LmHeadBackward : public BackwardFunction {
public:
    LmHeadBackward(TensorPtr x, TensorPtr table) : x_(std::move(x)), e_(std::move(table)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        const int64_t V = e_->shape()[0], C = e_->shape()[1];
        const int64_t T = x_->numel() / C;
        TensorPtr gx, ge;
        if (autograd::Engine::instance().needs_grad(x_)) {   // dx = dlogits E      (contraction over V)
            gx = Tensor::make(x_->shape(), kFloat32);
            GemmCall g;
            g.m = T; g.n = C; g.k = V;
            g.a = nullptr; g.lda = gy->ld(); g.trans_a = 0; g.a_tensor = gy;
            g.b = e_->dev_f32(); g.ldb = e_->ld(); g.trans_b = 0; g.b_weight = e_;   // E stored [k=V, n=C]
            g.c = gx->dev_f32_out(); g.ldc = C;
            run_gemm(g);
        }
        if (wants_grad(e_)) {                                // dE = dlogits^T x     [V,C]  (graph/gpt2_model.cpp:430-431 via TransposeBackward)
            ge = Tensor::make(e_->shape(), kFloat32);
            GemmCall g;
            g.m = V; g.n = C; g.k = T; g.trans_a = 1; g.trans_b = 0;
            g.a = gy->dev_f32(); g.lda = gy->ld(); g.b = x_->dev_f32(); g.ldb = x_->ld();
            g.c = ge->dev_f32_out(); g.ldc = C;
            run_gemm(g);
        }
        return {gx, ge};
    }
private:
    TensorPtr x_, e_;
};
}  // namespace

TensorPtr lm_head_tied(const TensorPtr& x_in, const TensorPtr& table) {
    require_f32(x_in, "lm_head");
    require_f32(table, "lm_head");
    TensorPtr x = contiguous(x_in);
    const int64_t V = table->shape()[0], C = table->shape()[1];
    if (x->shape().back() != C) throw TensorError("lm_head: hidden size mismatch");
    const int64_t T = x->numel() / C;
    std::vector<int64_t> so = x->shape();
    so.back() = V;
    auto y = Tensor::make(so, kFloat32, false, out_ld(V));
    GemmCall g;
    g.m = T; g.n = V; g.k = C;
    g.a = nullptr; g.lda = C; g.trans_a = 0; g.a_tensor = x;
    g.b = table->dev_f32(); g.ldb = table->ld(); g.trans_b = 1; g.b_weight = table;     // E [V,C] is K-major for x E^T
    g.c = y->dev_f32_out(); g.ldc = y->ld();
    run_gemm(g);
    y->set_grad_may_be_bf16(!wants_grad(table));      // frozen table: dlogits only feeds the bf16 dX product
    track(y, {x_in, table}, std::make_shared<LmHeadBackward>(x, table));
    return y;
}

}  // namespace ops

// mft_trainer_* — the training step of the reference's trainers behind a C ABI (include/mft_trainer.h).
// Loop body mirrored: gpt2_lora_finetune/main.cpp:561-601 (forward, lm_cross_entropy, backward, global clip with
// scale = max/(norm+1e-6), Adam, drop grads); Gemma: optim/gemma_trainer.cpp:104-149.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <unordered_map>
#include <vector>

#include "adam.h"
#include "autograd_engine.h"
#include "gemma_lora_injector.h"
#include "gemma_model.h"
#include "gpt2_model.h"
#include "lm_loss.h"
#include "lora_injector.h"
#include "mft_b200.h"
#include "mft_trainer.h"
#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"
#include "safetensors_loader.h"

using namespace ops;

namespace {
thread_local char g_terr[1024] = "";
int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_terr, sizeof(g_terr), fmt, ap);
    va_end(ap);
    return 1;
}
}  // namespace

struct mft_trainer {
    mft_trainer_config cfg;
    std::unique_ptr<GPT2Model> gpt2;
    std::unique_ptr<GemmaModel> gemma;
    LoraInjector gpt2_inj;
    GemmaLoraInjector gemma_inj;
    std::vector<TensorPtr> trainable;
    std::unique_ptr<Adam> opt;
    std::mt19937 gen;
    TensorPtr ids;                 // persistent [B,S] int32 device buffer (labels = ids)
    float* loss_dev = nullptr;     // persistent scalars
    float* norm_dev = nullptr;
    float* scale_dev = nullptr;
    float* lr_dev = nullptr;       // learning rate lives on the device so a schedule never invalidates the captured step
    float* lr_host = nullptr;      // pinned staging word for it
    void* slots_dev[2] = {nullptr, nullptr};     // device slot table (param/grad/m/v per trainable tensor)
    mft_tensor_slot* slots_host[2] = {nullptr, nullptr};   // pinned staging copies; [1] = gradients read from the all-reduced bucket
    std::vector<TensorPtr> m, v;   // Adam moments, owned by the trainer; step counters live on the device
    int32_t* steps_dev = nullptr;
    double* sumsq_dev = nullptr;
    float* bucket = nullptr;
    int64_t bucket_elems = 0, grad_elems = 0;
    std::vector<int64_t> offsets;
    std::vector<TensorPtr> grads;  // of the last backward (arena memory, valid until the next step)
    int with_grad = 0;
    int64_t launches = 0;
    bool arena_on = false;
    // CUDA graphs (captured after two eager runs of the same phase): 0 = whole step, 1 = forward/backward(+pack), 2 = update
    cudaGraphExec_t graph_exec[3] = {nullptr, nullptr, nullptr};
    int eager_runs[3] = {0, 0, 0};
    int64_t phase_launches[3] = {0, 0, 0};
    cudaEvent_t slots_event[2] = {nullptr, nullptr};   // recorded after each eager upload of the pinned slot table
    bool capturing = false;
    void* shadow_dev = nullptr;    // device table of mft_shadow_slot (bf16 shadows of the trained adapter factors)
    mft_shadow_slot* shadow_host = nullptr;
    std::vector<char> had_grad;    // per trainable tensor: received a gradient in the last eagerly run update
    std::vector<int> shadowed;     // indices into `trainable` whose shadows the refresh kernel rewrote last step
    TensorPtr own_bucket;          // gradient bucket allocated by mft_trainer_step_dp when the caller supplied none
    std::vector<TensorPtr> accum;  // true-gradient mode: persistent gradient sums over micro-batches (mft_trainer_accumulate)
    int accum_count = 0;
    float extra_grad_scale = 1.0f; // 1 / micro-batches, folded into the clip factor of the next update
    bool graph_enabled = true;     // runtime switch (mft_trainer_enable_graph); only effective when cfg.use_cuda_graph
    int eager_steps = 0;
    // data-parallel exchange overlapped with backward (mft_trainer_step_dp, large buckets): the bucket is cut into chunks of
    // consecutive tensors; a chunk is packed and all-reduced on a side stream as soon as the last of its gradients is final
    struct Chunk { int first = 0, last = 0; int64_t off = 0, elems = 0; int waiting = 0; bool sent = false; cudaEvent_t packed = nullptr; };
    std::vector<Chunk> chunks;
    std::vector<int> chunk_of;
    std::unordered_map<const Tensor*, int> index_of;
    std::vector<const float*> ready_grad;          // per trainable tensor: device pointer of its final gradient (this step)
    std::vector<TensorPtr> ready_keep;
    mft_comm* overlap_comm = nullptr;              // non-null only while such a step runs
    bool overlap_mode = false;                     // what the phase-1 graph (if any) was captured with
    cudaStream_t side = nullptr;
    cudaEvent_t joined = nullptr;
};

namespace {

void fill_normal(const TensorPtr& t, std::mt19937& g, float stdv) {
    std::normal_distribution<float> d(0.0f, stdv);
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = d(g);
    (void)t->dev();   // upload now (persistent scope still active)
}
void fill_const(const TensorPtr& t, float v) {
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = v;
    (void)t->dev();
}

void build(mft_trainer* t) {
    const mft_trainer_config& c = t->cfg;
    mft_rt::PersistentScope keep;   // everything created here outlives the step arena
    t->gen.seed(c.seed);
    if (c.model == 0) {
        GPT2Config g;
        g.n_layer = c.n_layer; g.n_embd = c.n_embd; g.n_head = c.n_head; g.vocab_size = c.vocab_size; g.n_positions = c.n_positions;
        t->gpt2 = std::make_unique<GPT2Model>(g);
        t->gpt2->tie_weights();
        auto ps = t->gpt2->parameters();
        for (size_t i = 0; i < ps.size(); ++i) {   // same protocol as oracle/harness/train_synth.cpp
            bool gamma = false;
            if (i >= 2 && i < ps.size() - 2) { size_t k = (i - 2) % 12; gamma = (k == 0 || k == 6); }
            if (i == ps.size() - 2) gamma = true;
            if (gamma) fill_const(ps[i], 1.0f); else fill_normal(ps[i], t->gen, 0.02f);
        }
        t->gpt2->init_lora_modules();
        if (!c.full_finetune) {
            LoraSpec s;
            s.rank = c.lora_rank; s.alpha = c.lora_alpha; s.dropout = 0.0f; s.split_qkv = false;
            s.targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj};
            if (c.lora_mlp) s.targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj, LoraTarget::MlpFcIn, LoraTarget::MlpFcOut};
            t->gpt2_inj.inject(*t->gpt2, s);
            t->trainable = t->gpt2->get_lora_parameters();
            for (auto& p : t->gpt2->parameters()) p->set_requires_grad(false);
            for (auto& p : t->trainable) { p->set_requires_grad(true); (void)p->dev(); }
        } else {
            t->trainable = t->gpt2->parameters();
            for (auto& p : t->trainable) p->set_requires_grad(true);
        }
    } else {
        GemmaTextConfig g;
        g.num_hidden_layers = c.n_layer; g.hidden_size = c.n_embd; g.intermediate_size = c.intermediate;
        g.num_attention_heads = c.n_head; g.num_key_value_heads = c.kv_heads; g.head_dim = c.head_dim;
        g.vocab_size = c.vocab_size; g.sliding_window = c.sliding_window > 0 ? c.sliding_window : 512;
        g.layer_types.clear();
        for (int i = 0; i < g.num_hidden_layers; ++i) g.layer_types.push_back(((i + 1) % 6 == 0) ? "full_attention" : "sliding_attention");
        t->gemma = std::make_unique<GemmaModel>(g);
        auto mk = [&](std::vector<int64_t> shape) {
            auto w = std::make_shared<Tensor>(shape, kFloat32, kCPU);
            fill_normal(w, t->gen, 0.02f);
            return w;
        };
        const int64_t C = g.hidden_size, I = g.intermediate_size, Hd = g.head_dim;
        const int64_t QD = g.num_attention_heads * Hd, KD = g.num_key_value_heads * Hd;
        t->gemma->assign_weight("embed_tokens.weight", mk({g.vocab_size, C}));
        t->gemma->assign_weight("norm.weight", mk({C}));
        for (int l = 0; l < g.num_hidden_layers; ++l) {
            const std::string p = "layers." + std::to_string(l) + ".";
            t->gemma->assign_weight(p + "input_layernorm.weight", mk({C}));
            t->gemma->assign_weight(p + "post_attention_layernorm.weight", mk({C}));
            t->gemma->assign_weight(p + "pre_feedforward_layernorm.weight", mk({C}));
            t->gemma->assign_weight(p + "post_feedforward_layernorm.weight", mk({C}));
            t->gemma->assign_weight(p + "self_attn.q_proj.weight", mk({C, QD}));
            t->gemma->assign_weight(p + "self_attn.k_proj.weight", mk({C, KD}));
            t->gemma->assign_weight(p + "self_attn.v_proj.weight", mk({C, KD}));
            t->gemma->assign_weight(p + "self_attn.o_proj.weight", mk({QD, C}));
            t->gemma->assign_weight(p + "self_attn.q_norm.weight", mk({Hd}));
            t->gemma->assign_weight(p + "self_attn.k_norm.weight", mk({Hd}));
            t->gemma->assign_weight(p + "mlp.gate_proj.weight", mk({C, I}));
            t->gemma->assign_weight(p + "mlp.up_proj.weight", mk({C, I}));
            t->gemma->assign_weight(p + "mlp.down_proj.weight", mk({I, C}));
        }
        t->gemma->init_lora_modules();
        GemmaLoraSpec s = GemmaLoraSpec::full_attn_mlp();
        s.rank = c.lora_rank; s.alpha = c.lora_alpha; s.dropout = 0.0f;
        GemmaLoraInjector::reseed_init();       // same adapter init for every trainer of a process
        t->gemma_inj.inject(*t->gemma, s);
        t->trainable = t->gemma_inj.get_trainable_params();
        for (auto& p : t->trainable) (void)p->dev();
    }
    AdamConfig oc(c.lr, c.beta1, c.beta2, c.eps, c.weight_decay, c.clip_norm, false);
    t->opt = std::make_unique<Adam>(oc);
    t->ids = std::make_shared<Tensor>(std::vector<int64_t>{c.batch, c.seq_len}, kInt32, kCPU);
    float* scalars = static_cast<float*>(mft_rt::dev_alloc(64));
    mft_rt::memset0(scalars, 64);
    t->loss_dev = scalars; t->norm_dev = scalars + 1; t->scale_dev = scalars + 2; t->lr_dev = scalars + 3;
    t->lr_host = static_cast<float*>(mft_rt::pinned_alloc(sizeof(float)));
    *t->lr_host = c.lr;
    mft_rt::h2d(t->lr_dev, t->lr_host, sizeof(float));
    const size_t nt = t->trainable.size();
    t->m.resize(nt); t->v.resize(nt);
    for (size_t i = 0; i < nt; ++i) {   // zero-filled; tensors that never receive a gradient simply keep them untouched
        t->m[i] = zeros(t->trainable[i]->shape(), kFloat32, kCPU);
        t->v[i] = zeros(t->trainable[i]->shape(), kFloat32, kCPU);
    }
    t->steps_dev = static_cast<int32_t*>(mft_rt::dev_alloc(nt * sizeof(int32_t) + 16));
    mft_rt::memset0(t->steps_dev, nt * sizeof(int32_t) + 16);
    for (int k = 0; k < 2; ++k) {
        t->slots_dev[k] = mft_rt::dev_alloc(nt * sizeof(mft_tensor_slot) + 16);
        t->slots_host[k] = static_cast<mft_tensor_slot*>(mft_rt::pinned_alloc(nt * sizeof(mft_tensor_slot) + 16));
    }
    t->shadow_dev = mft_rt::dev_alloc(nt * sizeof(mft_shadow_slot) + 16);
    t->shadow_host = static_cast<mft_shadow_slot*>(mft_rt::pinned_alloc(nt * sizeof(mft_shadow_slot) + 16));
    t->sumsq_dev = static_cast<double*>(mft_rt::dev_alloc(64));
    t->offsets.resize(t->trainable.size());
    int64_t off = 0;
    for (size_t i = 0; i < t->trainable.size(); ++i) { t->offsets[i] = off; off += (t->trainable[i]->numel() + 3) / 4 * 4; }
    t->grad_elems = off;
}

// ---- overlapped exchange (see mft_trainer::Chunk)
void build_chunks(mft_trainer* t) {
    if (!t->chunks.empty()) return;
    const char* e = std::getenv("MFT_DP_CHUNK_KB");
    const int64_t want = std::max<int64_t>(1, e && *e ? std::atoll(e) : 48 * 1024) * 256;      // elements
    const int n = int(t->trainable.size());
    t->chunk_of.assign(size_t(n), 0);
    mft_trainer::Chunk c;
    for (int i = 0; i < n; ++i) {
        if (c.last == c.first) c.off = t->offsets[size_t(i)];
        c.last = i + 1;
        const int64_t end = i + 1 < n ? t->offsets[size_t(i) + 1] : t->grad_elems;
        c.elems = end - c.off;
        t->chunk_of[size_t(i)] = int(t->chunks.size());
        t->index_of[t->trainable[size_t(i)].get()] = i;
        if (c.elems >= want || i + 1 == n) {
            cudaEventCreateWithFlags(&c.packed, cudaEventDisableTiming);
            t->chunks.push_back(c);
            c = mft_trainer::Chunk();
            c.first = c.last = i + 1;
        }
    }
    cudaStreamCreateWithFlags(&t->side, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&t->joined, cudaEventDisableTiming);
}

void fill_slot(mft_trainer* t, size_t i, bool packed, const float* grad_ptr) {
    mft_tensor_slot& s = t->slots_host[packed][i];
    std::memset(&s, 0, sizeof(s));
    s.n = t->trainable[i]->numel();
    s.bucket_offset = t->offsets[i];
    s.param = const_cast<float*>(t->trainable[i]->dev_f32());
    s.grad = const_cast<float*>(grad_ptr);
}

// pack one finished chunk into the bucket (compute stream) and start its all-reduce on the side stream
void send_chunk(mft_trainer* t, mft_trainer::Chunk& c) {
    cudaStream_t st = static_cast<cudaStream_t>(mft_rt::stream());
    const size_t n = size_t(c.last - c.first);
    for (size_t i = size_t(c.first); i < size_t(c.last); ++i) fill_slot(t, i, false, t->ready_grad[i]);
    char* dev = static_cast<char*>(t->slots_dev[0]) + size_t(c.first) * sizeof(mft_tensor_slot);
    mft_rt::h2d(dev, t->slots_host[0] + c.first, n * sizeof(mft_tensor_slot));
    mft_rt::check(mft_pack_grads(reinterpret_cast<const mft_tensor_slot*>(dev), int32_t(n), t->bucket, st), "mft_pack_grads");
    if (cudaEventRecord(c.packed, st) != cudaSuccess || cudaStreamWaitEvent(t->side, c.packed, 0) != cudaSuccess) mft_rt::fail("overlapped exchange: event fork failed");
    if (mft_comm_allreduce_sum(t->overlap_comm, t->bucket + c.off, c.elems, t->side) != 0) mft_rt::fail(std::string("overlapped exchange: ") + mft_comm_last_error());
    c.sent = true;
}

void forward_backward(mft_trainer* t) {
    if (t->arena_on) mft_rt::arena_reset();
    TensorPtr logits = t->cfg.model == 0 ? t->gpt2->forward(t->ids, nullptr) : t->gemma->forward(t->ids, nullptr);
    TensorPtr loss = lm_cross_entropy(logits, t->ids, -100, "mean");
    mft_rt::d2d(t->loss_dev, loss->dev(), sizeof(float));
    if (t->overlap_comm) {
        build_chunks(t);
        // the previous step's chunk uploads may still be reading the pinned table
        if (t->slots_event[0] && !t->capturing) cudaEventSynchronize(t->slots_event[0]);
        t->ready_grad.assign(t->trainable.size(), nullptr);
        t->ready_keep.clear();
        for (auto& c : t->chunks) { c.waiting = c.last - c.first; c.sent = false; }
        struct HookGuard {
            ~HookGuard() { autograd::Engine::instance().set_leaf_ready_hook(nullptr); }
        } guard;
        autograd::Engine::instance().set_leaf_ready_hook([t](const TensorPtr& leaf, const TensorPtr& g) {
            auto it = t->index_of.find(leaf.get());
            if (it == t->index_of.end()) return;
            const size_t i = size_t(it->second);
            if (g) {
                TensorPtr d = contiguous(g);
                if (d != g) t->ready_keep.push_back(d);               // a dense copy made here: keep it until the chunk is packed
                t->ready_grad[i] = d->dev_f32();                      // g itself is held by the engine until backward ends
            }
            mft_trainer::Chunk& c = t->chunks[size_t(t->chunk_of[i])];
            if (--c.waiting == 0 && !c.sent) send_chunk(t, c);
        });
        loss->backward();
        for (auto& c : t->chunks) {                                    // chunks with tensors the graph never reached (packed as zeros)
            if (c.sent) continue;
            for (size_t i = size_t(c.first); i < size_t(c.last); ++i) {   // the engine has written `.grad` by now: read the pointers from there
                TensorPtr g = t->trainable[i]->grad();
                if (g && !g->is_dense()) { g = contiguous(g); t->ready_keep.push_back(g); }
                t->ready_grad[i] = g ? g->dev_f32() : nullptr;
            }
            send_chunk(t, c);
        }
        t->ready_keep.clear();
        cudaStream_t st = static_cast<cudaStream_t>(mft_rt::stream());
        if (cudaEventRecord(t->joined, t->side) != cudaSuccess || cudaStreamWaitEvent(st, t->joined, 0) != cudaSuccess) mft_rt::fail("overlapped exchange: event join failed");
        if (!t->capturing) {
            if (!t->slots_event[0]) cudaEventCreateWithFlags(&t->slots_event[0], cudaEventDisableTiming);
            cudaEventRecord(t->slots_event[0], st);
        }
    } else {
        loss->backward();
    }
    t->grads.assign(t->trainable.size(), nullptr);
    t->with_grad = 0;
    for (size_t i = 0; i < t->trainable.size(); ++i) {
        t->grads[i] = t->trainable[i]->grad();
        if (t->grads[i]) ++t->with_grad;
    }
}

// Slot table for pack / clip / Adam. All state is on the device (moments, step counters, clip factor), so the phase needs no
// per-step host data and can be replayed from a CUDA graph. `packed`: gradients are read from the all-reduced bucket.
const mft_tensor_slot* upload_slots(mft_trainer* t, bool packed) {
    const size_t n = t->trainable.size();
    // the previous step's upload may still be reading the pinned table (step_device never synchronises)
    if (t->slots_event[packed] && !t->capturing) cudaEventSynchronize(t->slots_event[packed]);
    for (size_t i = 0; i < n; ++i) {
        mft_tensor_slot& s = t->slots_host[packed][i];
        std::memset(&s, 0, sizeof(s));
        s.n = t->trainable[i]->numel();
        // Adam only writes tensors that have a gradient: only those get their version bumped (cached K-major / bf16 shadows of
        // the others — e.g. GPT-2's attn.qkv adapters, which the reference never trains — stay valid)
        const bool written = t->grads[i] != nullptr;
        s.param = written ? t->trainable[i]->dev_f32_mut() : const_cast<float*>(t->trainable[i]->dev_f32());
        s.m = written ? t->m[i]->dev_f32_mut() : const_cast<float*>(t->m[i]->dev_f32());
        s.v = written ? t->v[i]->dev_f32_mut() : const_cast<float*>(t->v[i]->dev_f32());
        s.bucket_offset = t->offsets[i];
        if (t->grads[i]) s.grad = packed ? t->bucket + t->offsets[i] : t->grads[i]->dev_f32_mut();
    }
    // one pinned table per variant: the async upload (or its graph memcpy node) reads the host copy when it executes
    mft_rt::h2d(t->slots_dev[packed], t->slots_host[packed], n * sizeof(mft_tensor_slot));
    if (!t->capturing) {
        if (!t->slots_event[packed]) cudaEventCreateWithFlags(&t->slots_event[packed], cudaEventDisableTiming);
        cudaEventRecord(t->slots_event[packed], static_cast<cudaStream_t>(mft_rt::stream()));
    }
    return static_cast<const mft_tensor_slot*>(t->slots_dev[packed]);
}

// bf16 operand mode: the bf16 copies / K-major transposes of the small trained matrices (adapter factors) that the next
// forward will read are rewritten by ONE launch right after Adam, instead of one cast launch per factor when first used.
// Only shadows that already exist are refreshed (after the first step those are exactly the ones the graph reads).
void refresh_shadows(mft_trainer* t) {
    t->shadowed.clear();
    if (!ops::internal::emit_bf16_operands() || ops::internal::gemm_precision() != 1) return;
    int64_t max_elems = 0;
    int n = 0;
    for (size_t i = 0; i < t->trainable.size(); ++i) {
        const TensorPtr& p = t->trainable[i];
        if (!t->grads[i] || p->ndim() != 2 || p->numel() > (int64_t(1) << 18) || !p->is_dense()) continue;
        const auto sh = p->existing_bf16_shadows();
        if (!sh.plain && !sh.transposed) continue;
        mft_shadow_slot& s = t->shadow_host[n++];
        s.src = p->dev_f32(); s.rows = p->shape()[0]; s.cols = p->shape()[1]; s.ld_src = p->ld();
        s.plain = sh.plain; s.ld_plain = sh.ld_plain; s.transposed = sh.transposed; s.ld_transposed = sh.ld_transposed;
        max_elems = std::max<int64_t>(max_elems, p->numel());
        t->shadowed.push_back(int(i));
    }
    if (n == 0) return;
    mft_rt::h2d(t->shadow_dev, t->shadow_host, size_t(n) * sizeof(mft_shadow_slot));
    mft_rt::check(mft_refresh_bf16_shadows(static_cast<const mft_shadow_slot*>(t->shadow_dev), n, max_elems, mft_rt::stream()), "mft_refresh_bf16_shadows");
    for (int i : t->shadowed) t->trainable[size_t(i)]->mark_bf16_shadows_current();
}

void pack_phase(mft_trainer* t) {
    const mft_tensor_slot* slots = upload_slots(t, false);
    mft_rt::check(mft_pack_grads(slots, int32_t(t->trainable.size()), t->bucket, mft_rt::stream()), "mft_pack_grads");
}

void update_phase(mft_trainer* t, bool packed) {
    const int32_t n = int32_t(t->trainable.size());
    const mft_tensor_slot* slots = upload_slots(t, packed);
    const size_t wsb = mft_grad_sumsq_workspace_bytes();
    mft_rt::check(mft_grad_sumsq(slots, n, t->sumsq_dev, mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_grad_sumsq");
    const float pre = (packed && t->cfg.world_size > 1 ? 1.0f / float(t->cfg.world_size) : 1.0f) * t->extra_grad_scale;
    mft_rt::check(mft_clip_scale(t->sumsq_dev, t->cfg.clip_norm, pre, t->scale_dev, t->norm_dev, mft_rt::stream()), "mft_clip_scale");
    mft_rt::check(mft_adam_step_dev(slots, n, t->lr_dev, t->cfg.beta1, t->cfg.beta2, t->cfg.eps, t->cfg.weight_decay,
                                    t->scale_dev, t->steps_dev, mft_rt::stream()), "mft_adam_step_dev");
    refresh_shadows(t);
    t->had_grad.assign(t->trainable.size(), 0);
    for (size_t i = 0; i < t->trainable.size(); ++i) t->had_grad[i] = t->grads[i] ? 1 : 0;
    for (auto& p : t->trainable) p->set_grad(nullptr);   // gradients live in the step arena; never reused
}

// A replayed update phase (0 = whole step, 2 = update) writes parameters and moments on the device without running the host
// body that normally records it: note the write here, so host mirrors are re-read and cached shadows are versioned.
void mark_state_written(mft_trainer* t) {
    for (size_t i = 0; i < t->trainable.size(); ++i) {
        if (i < t->had_grad.size() && !t->had_grad[i]) continue;     // untouched by the update
        (void)t->trainable[i]->dev_f32_mut();
        (void)t->m[i]->dev_f32_mut();
        (void)t->v[i]->dev_f32_mut();
    }
    for (int i : t->shadowed) t->trainable[size_t(i)]->mark_bf16_shadows_current();   // the replayed update refreshed them too
}

// Runs one phase eagerly, or (cfg.use_cuda_graph) captures it into a CUDA graph after two eager runs and replays it.
// The step arena makes every intermediate address identical from step to step, which is what makes replay valid.
template <typename Body>
void run_phase(mft_trainer* t, int phase, Body body) {
    cudaStream_t st = static_cast<cudaStream_t>(mft_rt::stream());
    const bool want_graph = t->cfg.use_cuda_graph && t->graph_enabled;
    if (want_graph && t->graph_exec[phase]) {
        if (cudaGraphLaunch(t->graph_exec[phase], st) != cudaSuccess) mft_rt::fail("cudaGraphLaunch failed");
        if (phase != 1) mark_state_written(t);
        return;
    }
    const int64_t before = mft_launch_count();
    if (want_graph && t->eager_runs[phase] >= 2) {
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) != cudaSuccess) mft_rt::fail("cudaStreamBeginCapture failed");
        t->capturing = true;
        try { body(); } catch (...) {
            t->capturing = false;
            cudaStreamEndCapture(st, &graph);
            if (graph) cudaGraphDestroy(graph);
            throw;
        }
        t->capturing = false;
        if (cudaStreamEndCapture(st, &graph) != cudaSuccess || !graph) mft_rt::fail("cudaStreamEndCapture failed (an operation in the step is not capturable)");
        cudaGraphExec_t exec = nullptr;
        cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) mft_rt::fail(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
        t->graph_exec[phase] = exec;
        t->phase_launches[phase] = mft_launch_count() - before;
        if (cudaGraphLaunch(exec, st) != cudaSuccess) mft_rt::fail("cudaGraphLaunch failed");
        if (phase != 1) mark_state_written(t);
        return;
    }
    body();
    ++t->eager_runs[phase];
    t->phase_launches[phase] = mft_launch_count() - before;
}

}  // namespace

extern "C" const char* mft_trainer_last_error(void) { return g_terr; }

extern "C" int mft_trainer_create(const mft_trainer_config* cfg, mft_trainer** out) {
    if (!cfg || !out) return fail("mft_trainer_create: null argument");
    if (cfg->batch <= 0 || cfg->seq_len <= 0 || cfg->n_layer <= 0) return fail("mft_trainer_create: bad shape");
    try {
        auto t = std::make_unique<mft_trainer>();
        t->cfg = *cfg;
        if (t->cfg.world_size < 1) t->cfg.world_size = 1;
        (void)mft_rt::stream();
        build(t.get());
        const size_t mb = cfg->arena_mb > 0 ? size_t(cfg->arena_mb) : size_t(24) * 1024;
        mft_rt::arena_acquire(mb << 20);
        t->arena_on = true;
        mft_rt::sync();
        *out = t.release();
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_create: %s", e.what()); }
}

extern "C" void mft_trainer_destroy(mft_trainer* t) {
    if (!t) return;
    try { mft_rt::sync(); } catch (...) {}
    for (auto& g : t->graph_exec) if (g) cudaGraphExecDestroy(g);
    for (int k = 0; k < 2; ++k) { mft_rt::pinned_free(t->slots_host[k]); if (t->slots_event[k]) cudaEventDestroy(t->slots_event[k]); }
    for (auto& c : t->chunks) if (c.packed) cudaEventDestroy(c.packed);
    if (t->joined) cudaEventDestroy(t->joined);
    if (t->side) cudaStreamDestroy(t->side);
    mft_rt::pinned_free(t->shadow_host);
    mft_rt::pinned_free(t->lr_host);
    const bool had_arena = t->arena_on;
    delete t;
    if (had_arena) mft_rt::arena_release();
}

extern "C" int mft_trainer_next_synthetic_batch(mft_trainer* t, int32_t* host_ids) {
    if (!t || !host_ids) return fail("mft_trainer_next_synthetic_batch: null argument");
    std::uniform_int_distribution<int> d(0, t->cfg.vocab_size - 1);
    const int64_t n = int64_t(t->cfg.batch) * t->cfg.seq_len;
    for (int64_t i = 0; i < n; ++i) host_ids[i] = d(t->gen);
    return 0;
}

static int run_step(mft_trainer* t) {
    run_phase(t, 0, [t] { forward_backward(t); update_phase(t, false); });
    t->launches = t->phase_launches[0];
    ++t->eager_steps;
    return 0;
}

extern "C" int mft_trainer_step_device(mft_trainer* t, const int32_t* dev_ids) {
    if (!t || !dev_ids) return fail("mft_trainer_step_device: null argument");
    try {
        const size_t bytes = size_t(t->cfg.batch) * t->cfg.seq_len * sizeof(int32_t);
        mft_rt::d2d(t->ids->dev_out(), dev_ids, bytes);
        return run_step(t);
    } catch (const std::exception& e) { return fail("mft_trainer_step_device: %s", e.what()); }
}

extern "C" int mft_trainer_step(mft_trainer* t, const int32_t* host_ids, float* host_loss) {
    if (!t || !host_ids || !host_loss) return fail("mft_trainer_step: null argument");
    try {
        const size_t bytes = size_t(t->cfg.batch) * t->cfg.seq_len * sizeof(int32_t);
        mft_rt::h2d(t->ids->dev_out(), host_ids, bytes);
        run_step(t);
        mft_rt::d2h(host_loss, t->loss_dev, sizeof(float));   // synchronises
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_step: %s", e.what()); }
}

extern "C" int64_t mft_trainer_grad_elems(mft_trainer* t) { return t ? t->grad_elems : 0; }

extern "C" int mft_trainer_set_grad_bucket(mft_trainer* t, float* dev_bucket, int64_t elems) {
    if (!t || !dev_bucket) return fail("mft_trainer_set_grad_bucket: null argument");
    if (elems < t->grad_elems) return fail("mft_trainer_set_grad_bucket: bucket too small (%lld < %lld)", (long long)elems, (long long)t->grad_elems);
    if (t->bucket != dev_bucket) {        // phases 1 and 2 bake the bucket address into their graphs: capture them again
        for (int p = 1; p < 3; ++p) {
            if (t->graph_exec[p]) { try { mft_rt::sync(); } catch (...) {} cudaGraphExecDestroy(t->graph_exec[p]); t->graph_exec[p] = nullptr; }
            t->eager_runs[p] = 0;
        }
    }
    t->bucket = dev_bucket;
    t->bucket_elems = elems;
    return 0;
}

// the phase-1 graph is captured with or without the exchange inside: a change of mode captures it again
static void set_overlap_mode(mft_trainer* t, bool on) {
    if (t->overlap_mode == on) return;
    if (t->graph_exec[1]) { try { mft_rt::sync(); } catch (...) {} cudaGraphExecDestroy(t->graph_exec[1]); t->graph_exec[1] = nullptr; }
    t->eager_runs[1] = 0;
    t->overlap_mode = on;
}

// MFT_DP_OVERLAP=1: the bucket is exchanged in chunks on a side stream while backward still runs (see mft_trainer::Chunk);
// 2 = the same even with a one-rank communicator (exercises the chunk / side-stream / capture machinery on one GPU).
// Default 0: one all-reduce after backward.
static bool want_overlap(const mft_trainer* t) {
    const char* e = std::getenv("MFT_DP_OVERLAP");
    const int v = e && *e ? std::atoi(e) : 0;
    if (v >= 2) return true;
    return v == 1 && t->cfg.world_size > 1;
}

extern "C" int mft_trainer_forward_backward(mft_trainer* t, const int32_t* dev_ids) {
    if (!t || !dev_ids) return fail("mft_trainer_forward_backward: null argument");
    if (!t->bucket) return fail("mft_trainer_forward_backward: set a gradient bucket first");
    try {
        const size_t bytes = size_t(t->cfg.batch) * t->cfg.seq_len * sizeof(int32_t);
        mft_rt::d2d(t->ids->dev_out(), dev_ids, bytes);
        set_overlap_mode(t, t->overlap_comm != nullptr);
        run_phase(t, 1, [t] { forward_backward(t); if (!t->overlap_comm) pack_phase(t); });
        t->launches = t->phase_launches[1];
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_forward_backward: %s", e.what()); }
}

extern "C" int mft_trainer_update(mft_trainer* t) {
    if (!t) return fail("mft_trainer_update: null argument");
    try {
        run_phase(t, 2, [t] { update_phase(t, true); });
        t->launches = t->phase_launches[1] + t->phase_launches[2];
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_update: %s", e.what()); }
}

extern "C" int mft_trainer_step_dp(mft_trainer* t, mft_comm* comm, const int32_t* dev_ids) {
    if (!t || !comm || !dev_ids) return fail("mft_trainer_step_dp: null argument");
    if (mft_comm_world_size(comm) != t->cfg.world_size) return fail("mft_trainer_step_dp: the trainer was built for world_size %d, the communicator has %d", t->cfg.world_size, mft_comm_world_size(comm));
    try {
        if (!t->bucket) {                       // the trainer's own bucket (persistent)
            mft_rt::PersistentScope keep;
            t->own_bucket = zeros({t->grad_elems > 0 ? t->grad_elems : 4}, kFloat32, kCPU);
            if (mft_trainer_set_grad_bucket(t, t->own_bucket->dev_f32_mut(), t->grad_elems) != 0) return 1;
        }
        const bool overlap = want_overlap(t);
        t->overlap_comm = overlap ? comm : nullptr;
        const int rc = mft_trainer_forward_backward(t, dev_ids);
        t->overlap_comm = nullptr;
        if (rc != 0) return 1;
        if (!overlap && mft_comm_allreduce_sum(comm, t->bucket, t->grad_elems, mft_rt::stream()) != 0)
            return fail("mft_trainer_step_dp: %s", mft_comm_last_error());
        return mft_trainer_update(t);
    } catch (const std::exception& e) { return fail("mft_trainer_step_dp: %s", e.what()); }
}

// ---- gradient accumulation over micro-batches (true-gradient mode only: the reference overwrites `.grad`)
extern "C" int mft_trainer_accumulate(mft_trainer* t, const int32_t* dev_ids) {
    if (!t || !dev_ids) return fail("mft_trainer_accumulate: null argument");
    if (!ops::internal::true_gradients()) return fail("mft_trainer_accumulate: needs the true-gradient mode (mft_host_set_true_gradients(1) / MFT_TRUE_ATTN_GRAD=1); the default mode reproduces the reference, whose backward overwrites .grad");
    try {
        const size_t bytes = size_t(t->cfg.batch) * t->cfg.seq_len * sizeof(int32_t);
        mft_rt::d2d(t->ids->dev_out(), dev_ids, bytes);
        const size_t n = t->trainable.size();
        if (t->accum.size() != n) t->accum.assign(n, nullptr);
        for (size_t i = 0; i < n; ++i) t->trainable[i]->set_grad(t->accum[i]);    // the engine adds into these (grad += g)
        forward_backward(t);
        for (size_t i = 0; i < n; ++i) {
            if (!t->grads[i] || t->accum[i]) continue;
            mft_rt::PersistentScope keep;                  // first micro-batch: move the gradient out of the step arena
            t->accum[i] = t->grads[i]->clone();
        }
        ++t->accum_count;
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_accumulate: %s", e.what()); }
}

extern "C" int mft_trainer_apply_accumulated(mft_trainer* t) {
    if (!t) return fail("mft_trainer_apply_accumulated: null argument");
    if (t->accum_count == 0) return fail("mft_trainer_apply_accumulated: nothing accumulated");
    try {
        t->grads = t->accum;
        t->with_grad = 0;
        for (auto& g : t->grads) if (g) ++t->with_grad;
        t->extra_grad_scale = 1.0f / float(t->accum_count);      // mean over micro-batches (the mains scale the loss by 1/accum)
        update_phase(t, false);
        t->extra_grad_scale = 1.0f;
        t->accum.assign(t->trainable.size(), nullptr);
        t->accum_count = 0;
        return 0;
    } catch (const std::exception& e) { t->extra_grad_scale = 1.0f; return fail("mft_trainer_apply_accumulated: %s", e.what()); }
}

// ---- checkpoint of everything a resumed run needs: trainable tensors, Adam moments and per-tensor step counters
// (the reference saves adapters, graph/lora_saver.cpp, but no optimizer state: optim/adam.cpp:142-164)
extern "C" int mft_trainer_save_state(mft_trainer* t, const char* path) {
    if (!t || !path) return fail("mft_trainer_save_state: null argument");
    try {
        mft_rt::sync();
        const size_t n = t->trainable.size();
        std::vector<int32_t> steps(n + 4, 0);
        mft_rt::d2h(steps.data(), t->steps_dev, n * sizeof(int32_t));
        std::map<std::string, TensorPtr> tensors;
        std::vector<std::pair<std::string, std::string>> meta;
        meta.emplace_back("format", "mft_trainer_state_v1");
        meta.emplace_back("num_params", std::to_string(n));
        for (size_t i = 0; i < n; ++i) {
            const std::string k = std::to_string(i);
            tensors["param." + k] = t->trainable[i];
            tensors["adam." + k + ".m"] = t->m[i];
            tensors["adam." + k + ".v"] = t->v[i];
            meta.emplace_back("adam." + k + ".step", std::to_string(steps[i]));
        }
        write_safetensors(path, tensors, meta);
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_save_state: %s", e.what()); }
}

extern "C" int mft_trainer_load_state(mft_trainer* t, const char* path) {
    if (!t || !path) return fail("mft_trainer_load_state: null argument");
    try {
        mft_rt::sync();
        SafeTensorsReader rd(path);
        rd.parse_header();
        const auto& md = rd.metadata();
        auto fmt = md.find("format");
        if (fmt == md.end() || fmt->second != "mft_trainer_state_v1") return fail("mft_trainer_load_state: %s is not a trainer state file", path);
        const size_t n = t->trainable.size();
        if (size_t(std::stoul(md.at("num_params"))) != n) return fail("mft_trainer_load_state: the file holds %s tensors, the trainer %zu", md.at("num_params").c_str(), n);
        std::vector<int32_t> steps(n + 4, 0);
        auto fill = [](const TensorPtr& dst, const TensorPtr& src, const char* what, size_t i) {
            if (dst->numel() != src->numel()) throw std::runtime_error(std::string("size mismatch for ") + what + " " + std::to_string(i));
            std::memcpy(dst->data<float>(), static_cast<const Tensor&>(*src).data<float>(), size_t(dst->numel()) * sizeof(float));
            (void)dst->dev();        // upload now; the write bumped the version, so cached K-major / bf16 shadows are rebuilt
        };
        for (size_t i = 0; i < n; ++i) {
            const std::string k = std::to_string(i);
            fill(t->trainable[i], rd.load_tensor("param." + k), "param", i);
            fill(t->m[i], rd.load_tensor("adam." + k + ".m"), "adam m", i);
            fill(t->v[i], rd.load_tensor("adam." + k + ".v"), "adam v", i);
            steps[i] = int32_t(std::stol(md.at("adam." + k + ".step")));
        }
        mft_rt::h2d(t->steps_dev, steps.data(), n * sizeof(int32_t));
        mft_rt::sync();              // `steps` is a stack buffer
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_load_state: %s", e.what()); }
}

extern "C" void mft_host_set_true_gradients(int enable) { ops::internal::set_true_gradients(enable != 0); }
extern "C" int mft_host_true_gradients(void) { return ops::internal::true_gradients() ? 1 : 0; }

extern "C" const float* mft_trainer_loss_ptr(mft_trainer* t) { return t ? t->loss_dev : nullptr; }
extern "C" const float* mft_trainer_grad_norm_ptr(mft_trainer* t) { return t ? t->norm_dev : nullptr; }
extern "C" void* mft_trainer_stream(mft_trainer*) { return mft_rt::stream(); }
extern "C" int mft_trainer_sync(mft_trainer*) {
    try { mft_rt::sync(); return 0; } catch (const std::exception& e) { return fail("mft_trainer_sync: %s", e.what()); }
}
extern "C" void mft_trainer_enable_graph(mft_trainer* t, int enable) {
    if (!t) return;
    const bool on = enable != 0;
    if (on && !t->graph_enabled) {   // eager steps in between may have moved cached buffers: capture again
        for (int p = 0; p < 3; ++p) {
            if (t->graph_exec[p]) { cudaGraphExecDestroy(t->graph_exec[p]); t->graph_exec[p] = nullptr; }
            t->eager_runs[p] = 0;
        }
    }
    t->graph_enabled = on;
}
extern "C" void mft_host_set_gemm_precision(int bf16) { ops::internal::set_gemm_precision(bf16); }   // also switches attention (ops.cpp)
extern "C" int mft_host_gemm_precision(void) { return ops::internal::gemm_precision(); }
extern "C" void mft_trainer_set_lr(mft_trainer* t, float lr) {
    if (!t || !t->opt) return;
    t->opt->set_learning_rate(lr);
    try {
        mft_rt::sync();              // the staging word may still be in flight from the previous update
        *t->lr_host = lr;
        mft_rt::h2d(t->lr_dev, t->lr_host, sizeof(float));
    } catch (...) {}
}
extern "C" int32_t mft_trainer_num_trainable(mft_trainer* t) { return t ? int32_t(t->trainable.size()) : 0; }
extern "C" int32_t mft_trainer_num_with_grad(mft_trainer* t) { return t ? t->with_grad : 0; }
extern "C" int64_t mft_trainer_kernel_launches(mft_trainer* t) { return t ? t->launches : 0; }
extern "C" size_t mft_trainer_arena_high_water(mft_trainer*) { return mft_rt::arena_high_water(); }

extern "C" int mft_trainer_param_checksum(mft_trainer* t, int32_t index, double* sum, double* sumabs, int64_t* n) {
    if (!t || index < 0 || index >= int32_t(t->trainable.size())) return fail("mft_trainer_param_checksum: bad index");
    try {
        const TensorPtr& p = t->trainable[index];
        const float* d = static_cast<const Tensor&>(*p).data<float>();
        double s = 0, sa = 0;
        for (int64_t i = 0; i < p->numel(); ++i) { s += d[i]; sa += std::fabs(d[i]); }
        if (sum) *sum = s;
        if (sumabs) *sumabs = sa;
        if (n) *n = p->numel();
        return 0;
    } catch (const std::exception& e) { return fail("mft_trainer_param_checksum: %s", e.what()); }
}

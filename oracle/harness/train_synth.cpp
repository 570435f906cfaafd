// train_synth — synthetic-token training / per-op driver written ONLY against the public
// MobileFineTuner operator API (ops::Tensor, ops::*, LoRALinear, GPT2Model, GemmaModel, Adam).
//
// The same source is compiled twice:
//   * against the reference's own sources (oracle/Makefile -> oracle/_ref/ref_harness): the ORACLE;
//   * against this repository's B200 host library (mobilefinetuner_b200/csrc/Makefile -> mft_harness).
// Identical source linking both libraries is the drop-in proof for the hot path
// (reference call sequence: gpt2_lora_finetune/main.cpp:561-601, optim/gemma_trainer.cpp:104-149).
//
// TEST INFRASTRUCTURE: when built with -DMFT_HARNESS_REFERENCE this binary is the oracle; it is only
// ever executed by tests/, bench.py's cpu_baseline / --impl reference legs and the fixture generator.
#include "finetune_ops/core/tensor.h"
#include "finetune_ops/core/ops.h"
#include "finetune_ops/core/lm_loss.h"
#include "finetune_ops/core/autograd_engine.h"
#include "finetune_ops/core/memory_efficient_attention.h"
#include "finetune_ops/nn/lora_linear.h"
#include "finetune_ops/graph/gpt2_model.h"
#include "finetune_ops/graph/gemma_model.h"
#include "finetune_ops/graph/lora_injector.h"
#include "finetune_ops/graph/gemma_lora_injector.h"
#include "finetune_ops/optim/adam.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <random>
#include <string>
#include <vector>

#ifdef MFT_HARNESS_REFERENCE
extern "C" void openblas_set_num_threads(int);
#endif

using namespace ops;
using clk = std::chrono::steady_clock;

static double secs_since(clk::time_point t0) {
    return std::chrono::duration<double>(clk::now() - t0).count();
}

// ---------------------------------------------------------------- tiny .npy writer
static void save_npy(const std::string& path, const void* data, const std::vector<int64_t>& shape,
                     const char* descr, size_t elem) {
    std::string sh = "(";
    for (size_t i = 0; i < shape.size(); ++i) {
        sh += std::to_string(shape[i]);
        if (shape.size() == 1 || i + 1 < shape.size()) sh += ",";
    }
    sh += ")";
    std::string hdr = std::string("{'descr': '") + descr + "', 'fortran_order': False, 'shape': " + sh + ", }";
    size_t total = 10 + hdr.size() + 1;
    size_t pad = (64 - total % 64) % 64;
    hdr += std::string(pad, ' ');
    hdr.push_back('\n');
    std::ofstream f(path, std::ios::binary);
    if (!f) { std::fprintf(stderr, "cannot write %s\n", path.c_str()); std::exit(2); }
    const char magic[8] = {'\x93', 'N', 'U', 'M', 'P', 'Y', 1, 0};
    f.write(magic, 8);
    uint16_t hl = static_cast<uint16_t>(hdr.size());
    f.write(reinterpret_cast<const char*>(&hl), 2);
    f.write(hdr.data(), static_cast<std::streamsize>(hdr.size()));
    size_t n = 1;
    for (auto d : shape) n *= static_cast<size_t>(d);
    f.write(static_cast<const char*>(data), static_cast<std::streamsize>(n * elem));
}

static std::string g_dump_dir;
static void dump(const std::string& name, const TensorPtr& t) {
    if (g_dump_dir.empty() || !t) return;
    if (t->dtype() == kInt32)
        save_npy(g_dump_dir + "/" + name + ".npy", t->data<int32_t>(), t->shape(), "<i4", 4);
    else
        save_npy(g_dump_dir + "/" + name + ".npy", t->data<float>(), t->shape(), "<f4", 4);
}

// ---------------------------------------------------------------- args
struct Args {
    std::map<std::string, std::string> kv;
    std::string get(const std::string& k, const std::string& d) const {
        auto it = kv.find(k);
        return it == kv.end() ? d : it->second;
    }
    int geti(const std::string& k, int d) const { return std::atoi(get(k, std::to_string(d)).c_str()); }
    double getf(const std::string& k, double d) const {
        auto it = kv.find(k);
        return it == kv.end() ? d : std::atof(it->second.c_str());
    }
};

static Args parse(int argc, char** argv, int first) {
    Args a;
    for (int i = first; i < argc; ++i) {
        std::string s = argv[i];
        if (s.rfind("--", 0) != 0) continue;
        s = s.substr(2);
        auto eq = s.find('=');
        if (eq != std::string::npos) a.kv[s.substr(0, eq)] = s.substr(eq + 1);
        else if (i + 1 < argc && std::string(argv[i + 1]).rfind("--", 0) != 0) { a.kv[s] = argv[i + 1]; ++i; }
        else a.kv[s] = "1";
    }
    return a;
}

// ---------------------------------------------------------------- deterministic fills
static void fill_normal(const TensorPtr& t, std::mt19937& g, float std_) {
    std::normal_distribution<float> d(0.0f, std_);
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = d(g);
}
static void fill_const(const TensorPtr& t, float v) {
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = v;
}
static TensorPtr make_f32(const std::vector<int64_t>& shape) {
    return std::make_shared<Tensor>(shape, kFloat32, kCPU);
}
static TensorPtr make_ids(int64_t B, int64_t S, int V, std::mt19937& g) {
    auto t = std::make_shared<Tensor>(std::vector<int64_t>{B, S}, kInt32, kCPU);
    std::uniform_int_distribution<int> d(0, V - 1);
    int32_t* p = t->data<int32_t>();
    for (int64_t i = 0; i < B * S; ++i) p[i] = d(g);
    return t;
}

struct GradStats { double norm = 0; int with_grad = 0; int total = 0; };

// Host clip exactly as the GPT-2 trainers do it (gpt2_lora_finetune/main.cpp:491-516):
// double accumulation, scale = max/(norm+1e-6) applied in place through data<float>().
static GradStats host_clip(const std::vector<TensorPtr>& params, float max_norm) {
    GradStats st;
    double tot = 0.0;
    for (auto& p : params) {
        ++st.total;
        auto g = p->grad();
        if (!g) continue;
        ++st.with_grad;
        const float* d = g->data<float>();
        for (int64_t i = 0; i < g->numel(); ++i) tot += static_cast<double>(d[i]) * d[i];
    }
    st.norm = std::sqrt(tot);
    if (max_norm > 0 && st.norm > max_norm) {
        float sc = static_cast<float>(max_norm / (st.norm + 1e-6));
        for (auto& p : params) {
            auto g = p->grad();
            if (!g) continue;
            float* d = g->data<float>();
            for (int64_t i = 0; i < g->numel(); ++i) d[i] *= sc;
        }
    }
    return st;
}

static void checksum(const char* tag, int idx, const TensorPtr& t) {
    if (!t) { std::printf("%s %d null\n", tag, idx); return; }
    const float* d = t->data<float>();
    double s = 0, sa = 0;
    for (int64_t i = 0; i < t->numel(); ++i) { s += d[i]; sa += std::fabs(d[i]); }
    std::printf("%s %d n=%lld sum=%.9g sumabs=%.9g\n", tag, idx, static_cast<long long>(t->numel()), s, sa);
}

// ---------------------------------------------------------------- GPT-2 training (LoRA or full FT)
// Init protocol: mt19937(seed): every matrix/embedding/bias ~ N(0, 0.02) in
// parameters() order; LayerNorm gamma = 1 (beta drawn from the stream). Tokens: uniform [0,V) from
// the same engine, a fresh batch per step, labels = ids, attention_mask = nullptr.
static int run_gpt2(const Args& a) {
    GPT2Config cfg;
    cfg.n_layer = a.geti("layers", 2);
    cfg.n_embd = a.geti("embd", 64);
    cfg.n_head = a.geti("heads", 4);
    cfg.vocab_size = a.geti("vocab", 503);
    cfg.n_positions = a.geti("npos", 64);
    const int B = a.geti("batch", 2), S = a.geti("seq", 16), steps = a.geti("steps", 5);
    const int rank = a.geti("rank", 8);
    const float alpha = static_cast<float>(a.getf("alpha", 16.0));
    const std::string mode = a.get("mode", "lora");
    const float lr = static_cast<float>(a.getf("lr", 2e-4));
    const float wd = static_cast<float>(a.getf("wd", mode == "full" ? 0.01 : 0.0));
    const float clip = static_cast<float>(a.getf("clip", 1.0));
    const int seed = a.geti("seed", 1234);
    const bool quiet = a.geti("quiet", 0) != 0;

    GPT2Model model(cfg);
    model.tie_weights();
    std::mt19937 gen(static_cast<uint32_t>(seed));
    {
        auto ps = model.parameters();
        // parameters(): wte, wpe, per block {ln1.w, ln1.b, qkv.w, qkv.b, proj.w, proj.b, ln2.w, ln2.b,
        // fc_in.w, fc_in.b, fc_out.w, fc_out.b}, ln_f.w, ln_f.b  (graph/gpt2_model.cpp:837-861)
        for (size_t i = 0; i < ps.size(); ++i) {
            bool is_gamma = false;
            if (i >= 2 && i < ps.size() - 2) { size_t k = (i - 2) % 12; is_gamma = (k == 0 || k == 6); }
            if (i == ps.size() - 2) is_gamma = true;
            if (is_gamma) fill_const(ps[i], 1.0f); else fill_normal(ps[i], gen, 0.02f);
        }
    }
    std::vector<TensorPtr> trainable;
    LoraInjector injector;
    if (mode == "lora") {
        model.init_lora_modules();
        LoraSpec spec;
        spec.rank = rank; spec.alpha = alpha; spec.dropout = 0.0f; spec.split_qkv = false;
        spec.targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj};
        if (a.geti("lora_mlp", 0)) spec.targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj,
                                                   LoraTarget::MlpFcIn, LoraTarget::MlpFcOut};
        injector.inject(model, spec);
        trainable = model.get_lora_parameters();
        for (auto& p : model.parameters()) p->set_requires_grad(false);
        for (auto& p : trainable) p->set_requires_grad(true);
    } else {
        model.init_lora_modules();
        trainable = model.parameters();
        for (auto& p : trainable) p->set_requires_grad(true);
    }

    AdamConfig oc(lr, 0.9f, 0.999f, 1e-8f, wd, clip, false);
    Adam opt(oc);
    std::printf("config model=gpt2 mode=%s L=%d C=%d H=%d V=%d B=%d S=%d r=%d trainable=%zu\n",
                mode.c_str(), cfg.n_layer, cfg.n_embd, cfg.n_head, cfg.vocab_size, B, S, rank, trainable.size());

    double t_total = 0;
    for (int step = 0; step < steps; ++step) {
        auto ids = make_ids(B, S, cfg.vocab_size, gen);
        auto t0 = clk::now();
        auto logits = model.forward(ids, nullptr);
        double t_fwd = secs_since(t0);
        auto loss = lm_cross_entropy(logits, ids, -100, "mean");
        float loss_val = loss->data<float>()[0];
        double t_ce = secs_since(t0);
        if (step == 0 && !g_dump_dir.empty()) { dump("ids0", ids); dump("logits0", logits); }
        loss->backward();
        double t_bwd = secs_since(t0);
        GradStats gs = host_clip(trainable, clip);
        std::vector<TensorPtr> grads;
        grads.reserve(trainable.size());
        for (auto& p : trainable) grads.push_back(p->grad());
        opt.step(trainable, grads);
        if (step == 0 && !g_dump_dir.empty())
            for (size_t i = 0; i < trainable.size(); ++i)
                if (trainable[i]->grad()) dump("grad0_" + std::to_string(i), trainable[i]->grad());
        for (auto& p : trainable) if (p->grad()) p->zero_grad();
        double t_all = secs_since(t0);
        if (step > 0 || steps == 1) t_total += t_all;
        if (!quiet)
            std::printf("step %d loss %.9g gnorm %.9g with_grad %d/%d  t_fwd %.4f t_ce %.4f t_bwd %.4f t_opt %.4f\n",
                        step, loss_val, gs.norm, gs.with_grad, gs.total, t_fwd, t_ce - t_fwd, t_bwd - t_ce, t_all - t_bwd);
        else
            std::printf("step %d loss %.9g gnorm %.9g with_grad %d/%d\n", step, loss_val, gs.norm, gs.with_grad, gs.total);
        std::fflush(stdout);
    }
    int timed = steps > 1 ? steps - 1 : 1;
    std::printf("timing steps=%d mean_step_s=%.6f tokens_per_s=%.3f\n", timed, t_total / timed,
                static_cast<double>(B) * S * timed / t_total);
    if (a.geti("checksums", 1))
        for (size_t i = 0; i < trainable.size(); ++i) checksum("param", static_cast<int>(i), trainable[i]);
    return 0;
}

// ---------------------------------------------------------------- Gemma-3 LoRA training
// Weights go in through assign_weight (keys: graph/gemma_model.cpp:687-734); RMSNorm weights
// ~ N(0,0.02) (effective scale 1+w); LoRA init is the injector's own (function-static mt19937(42)).
static int run_gemma(const Args& a) {
    GemmaTextConfig cfg;
    cfg.num_hidden_layers = a.geti("layers", 2);
    cfg.hidden_size = a.geti("embd", 64);
    cfg.intermediate_size = a.geti("inter", 128);
    cfg.num_attention_heads = a.geti("heads", 4);
    cfg.num_key_value_heads = a.geti("kv_heads", 1);
    cfg.head_dim = a.geti("head_dim", 32);
    cfg.vocab_size = a.geti("vocab", 1009);
    cfg.sliding_window = a.geti("window", 512);
    cfg.query_pre_attn_scalar = static_cast<float>(a.getf("qscalar", 256.0));
    cfg.layer_types.clear();
    for (int i = 0; i < cfg.num_hidden_layers; ++i)   // Gemma-3 pattern: 5 sliding : 1 full
        cfg.layer_types.push_back(((i + 1) % 6 == 0) ? "full_attention" : "sliding_attention");
    const int B = a.geti("batch", 2), S = a.geti("seq", 16), steps = a.geti("steps", 5);
    const int rank = a.geti("rank", 8);
    const float alpha = static_cast<float>(a.getf("alpha", 32.0));
    const float lr = static_cast<float>(a.getf("lr", 2e-4));
    const float clip = static_cast<float>(a.getf("clip", 1.0));
    const int seed = a.geti("seed", 1234);

    GemmaModel model(cfg);
    std::mt19937 gen(static_cast<uint32_t>(seed));
    auto mk = [&](const std::vector<int64_t>& shape) { auto t = make_f32(shape); fill_normal(t, gen, 0.02f); return t; };
    const int64_t C = cfg.hidden_size, I = cfg.intermediate_size, Hd = cfg.head_dim;
    const int64_t QD = cfg.num_attention_heads * Hd, KD = cfg.num_key_value_heads * Hd;
    model.assign_weight("embed_tokens.weight", mk({cfg.vocab_size, C}));
    model.assign_weight("norm.weight", mk({C}));
    for (int l = 0; l < cfg.num_hidden_layers; ++l) {
        std::string p = "layers." + std::to_string(l) + ".";
        model.assign_weight(p + "input_layernorm.weight", mk({C}));
        model.assign_weight(p + "post_attention_layernorm.weight", mk({C}));
        model.assign_weight(p + "pre_feedforward_layernorm.weight", mk({C}));
        model.assign_weight(p + "post_feedforward_layernorm.weight", mk({C}));
        model.assign_weight(p + "self_attn.q_proj.weight", mk({C, QD}));
        model.assign_weight(p + "self_attn.k_proj.weight", mk({C, KD}));
        model.assign_weight(p + "self_attn.v_proj.weight", mk({C, KD}));
        model.assign_weight(p + "self_attn.o_proj.weight", mk({QD, C}));
        model.assign_weight(p + "self_attn.q_norm.weight", mk({Hd}));
        model.assign_weight(p + "self_attn.k_norm.weight", mk({Hd}));
        model.assign_weight(p + "mlp.gate_proj.weight", mk({C, I}));
        model.assign_weight(p + "mlp.up_proj.weight", mk({C, I}));
        model.assign_weight(p + "mlp.down_proj.weight", mk({I, C}));
    }
    model.init_lora_modules();
    GemmaLoraSpec spec = GemmaLoraSpec::full_attn_mlp();
    spec.rank = rank; spec.alpha = alpha; spec.dropout = 0.0f;
    GemmaLoraInjector injector;
    injector.inject(model, spec);
    auto trainable = injector.get_trainable_params();
    AdamConfig oc(lr, 0.9f, 0.999f, 1e-8f, 0.0f, clip, false);
    Adam opt(oc);
    std::printf("config model=gemma L=%d C=%d I=%d H=%d KV=%d Hd=%d V=%d B=%d S=%d r=%d trainable=%zu\n",
                cfg.num_hidden_layers, cfg.hidden_size, cfg.intermediate_size, cfg.num_attention_heads,
                cfg.num_key_value_heads, cfg.head_dim, cfg.vocab_size, B, S, rank, trainable.size());
    double t_total = 0;
    for (int step = 0; step < steps; ++step) {
        auto ids = make_ids(B, S, cfg.vocab_size, gen);
        auto t0 = clk::now();
        auto logits = model.forward(ids, nullptr);
        auto loss = lm_cross_entropy(logits, ids, -100, "mean");
        float loss_val = loss->data<float>()[0];
        if (step == 0 && !g_dump_dir.empty()) { dump("ids0", ids); dump("logits0", logits); }
        loss->backward();
        GradStats gs = host_clip(trainable, clip);
        std::vector<TensorPtr> grads;
        for (auto& p : trainable) grads.push_back(p->grad());
        opt.step(trainable, grads);
        if (step == 0 && !g_dump_dir.empty())
            for (size_t i = 0; i < trainable.size(); ++i)
                if (trainable[i]->grad()) dump("grad0_" + std::to_string(i), trainable[i]->grad());
        for (auto& p : trainable) if (p->grad()) p->zero_grad();
        double t_all = secs_since(t0);
        if (step > 0 || steps == 1) t_total += t_all;
        std::printf("step %d loss %.9g gnorm %.9g with_grad %d/%d t_step %.4f\n", step, loss_val, gs.norm,
                    gs.with_grad, gs.total, t_all);
        std::fflush(stdout);
    }
    int timed = steps > 1 ? steps - 1 : 1;
    std::printf("timing steps=%d mean_step_s=%.6f tokens_per_s=%.3f\n", timed, t_total / timed,
                static_cast<double>(B) * S * timed / t_total);
    if (a.geti("checksums", 1))
        for (size_t i = 0; i < trainable.size(); ++i) checksum("param", static_cast<int>(i), trainable[i]);
    return 0;
}

// ---------------------------------------------------------------- per-op cases (golden generator)
// Every case writes its inputs, forward output and all input gradients as .npy into --dump DIR.
// Upstream gradients are explicit (Engine::run_backward({y},{gy})), the pattern the reference's own
// op tests use (optim/test_qkt_softmax_grad.cpp:37-66).
static TensorPtr rnd(const std::vector<int64_t>& shape, std::mt19937& g, float std_, bool req = true) {
    auto t = make_f32(shape);
    fill_normal(t, g, std_);
    t->set_requires_grad(req);
    return t;
}
static void backward_with(const TensorPtr& y, const TensorPtr& gy) {
    autograd::Engine::instance().run_backward({y}, {gy});
}

static int run_ops(const Args& a) {
    if (g_dump_dir.empty()) { std::fprintf(stderr, "ops mode needs --dump DIR\n"); return 2; }
    std::mt19937 g(static_cast<uint32_t>(a.geti("seed", 7)));
    const int64_t B = 2, S = 5, C = 24;

    {   // matmul 3-D x 2-D (weight trainable) — core/ops.cpp:486-651, backward_functions.cpp:94-138
        auto x = rnd({B, S, C}, g, 1.0f), w = rnd({C, 40}, g, 0.2f);
        auto y = matmul(x, w);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("matmul_x", x); dump("matmul_w", w); dump("matmul_y", y); dump("matmul_gy", gy);
        dump("matmul_gx", x->grad()); dump("matmul_gw", w->grad());
    }
    {   // batched 4-D matmul (attention shapes)
        auto p = rnd({2, 3, 5, 5}, g, 1.0f), v = rnd({2, 3, 5, 8}, g, 1.0f);
        auto y = matmul(p, v);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("bmm_a", p); dump("bmm_b", v); dump("bmm_y", y); dump("bmm_gy", gy);
        dump("bmm_ga", p->grad()); dump("bmm_gb", v->grad());
    }
    {   // bias add (broadcast) + residual add — core/ops.cpp:166-309
        auto x = rnd({B, S, C}, g, 1.0f), b = rnd({C}, g, 1.0f);
        auto y = add(x, b);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("bias_x", x); dump("bias_b", b); dump("bias_y", y); dump("bias_gy", gy);
        dump("bias_gx", x->grad()); dump("bias_gb", b->grad());
    }
    {   // layer_norm — core/ops.cpp:1404-1458, backward_functions.cpp:425-481
        auto x = rnd({B, S, C}, g, 1.5f), w = rnd({C}, g, 1.0f), b = rnd({C}, g, 1.0f);
        auto y = layer_norm(x, w, b, 1e-5f);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("ln_x", x); dump("ln_w", w); dump("ln_b", b); dump("ln_y", y); dump("ln_gy", gy);
        dump("ln_gx", x->grad()); dump("ln_gw", w->grad()); dump("ln_gb", b->grad());
    }
    {   // rms_norm (1+w) — core/ops.cpp:1489-1537, backward_functions.cpp:483-523
        auto x = rnd({B, 3, S, C}, g, 1.5f), w = rnd({C}, g, 0.3f);
        auto y = rms_norm(x, w, 1e-6f);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("rms_x", x); dump("rms_w", w); dump("rms_y", y); dump("rms_gy", gy);
        dump("rms_gx", x->grad()); dump("rms_gw", w->grad());
    }
    {   // gelu (tanh form) — core/ops.cpp:1055-1079, backward_functions.cpp:155-177
        auto x = rnd({B, S, C}, g, 2.0f);
        auto y = gelu(x);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("gelu_x", x); dump("gelu_y", y); dump("gelu_gy", gy); dump("gelu_gx", x->grad());
    }
    {   // elementwise mul (GeGLU product) and scalar mul
        auto p = rnd({B, S, C}, g, 1.0f), q = rnd({B, S, C}, g, 1.0f);
        auto y = mul(p, q);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("mul_a", p); dump("mul_b", q); dump("mul_y", y); dump("mul_gy", gy);
        dump("mul_ga", p->grad()); dump("mul_gb", q->grad());
    }
    {   // softmax over last dim — core/ops.cpp:1119-1175, backward_functions.cpp:297-340
        auto x = rnd({2, 3, 5, 7}, g, 2.0f);
        auto y = softmax(x, -1);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("softmax_x", x); dump("softmax_y", y); dump("softmax_gy", gy); dump("softmax_gx", x->grad());
    }
    {   // GPT-2 attention forward (no gradient) — core/memory_efficient_attention.cpp:40-185
        auto q = rnd({2, 3, 9, 16}, g, 1.0f, false), k = rnd({2, 3, 9, 16}, g, 1.0f, false),
             v = rnd({2, 3, 9, 16}, g, 1.0f, false);
        MemoryEfficientAttentionConfig mc;
        mc.use_causal_mask = true; mc.scale = 0.25f;
        auto y = memory_efficient_attention(q, k, v, nullptr, mc);
        dump("attn_q", q); dump("attn_k", k); dump("attn_v", v); dump("attn_y", y);
    }
    {   // apply_rope (interleaved pairs) — core/ops.cpp:2151-2225, backward_functions.cpp:718-763
        auto x = rnd({2, 3, 6, 8}, g, 1.0f);
        auto y = apply_rope(x, 6, 8, 10000.0f);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("rope_x", x); dump("rope_y", y); dump("rope_gy", gy); dump("rope_gx", x->grad());
    }
    {   // repeat_kv_heads — core/ops.cpp:2072-2149, backward_functions.cpp:683-716
        auto x = rnd({2, 2, 4, 6}, g, 1.0f);
        auto y = repeat_kv_heads(x, 3);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("repkv_x", x); dump("repkv_y", y); dump("repkv_gy", gy); dump("repkv_gx", x->grad());
    }
    {   // permute(0,2,1,3) and transpose(2,3)
        auto x = rnd({2, 3, 4, 5}, g, 1.0f);
        auto y = permute(x, {0, 2, 1, 3});
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        dump("perm_x", x); dump("perm_y", y); dump("perm_gy", gy); dump("perm_gx", x->grad());
        auto x2 = rnd({2, 3, 4, 5}, g, 1.0f);
        auto y2 = transpose(x2, 2, 3);
        auto gy2 = rnd(y2->shape(), g, 1.0f, false);
        backward_with(y2, gy2);
        dump("tr_x", x2); dump("tr_y", y2); dump("tr_gy", gy2); dump("tr_gx", x2->grad());
    }
    {   // lm_cross_entropy with ignore_index rows — core/lm_loss.cpp:19-210
        const int V = 37;
        auto x = rnd({2, 6, V}, g, 2.0f);
        auto lab = std::make_shared<Tensor>(std::vector<int64_t>{2, 6}, kInt32, kCPU);
        int32_t lv[12] = {3, 5, -100, 36, 0, 7, 11, -100, -100, 2, 36, 1};
        std::memcpy(lab->data<int32_t>(), lv, sizeof(lv));
        auto y = lm_cross_entropy(x, lab, -100, "mean");
        y->backward();
        dump("ce_x", x); dump("ce_labels", lab); dump("ce_loss", y); dump("ce_gx", x->grad());
    }
    for (int layout = 0; layout < 2; ++layout) {   // LoRALinear, both layouts — nn/lora_linear.cpp:47-108
        const int64_t in = 24, out = 40, r = 4;
        auto x = rnd({B, S, in}, g, 1.0f);
        TensorPtr W = rnd({in, out}, g, 0.2f, false), bias = rnd({out}, g, 0.5f, false);
        TensorPtr A = layout == 0 ? rnd({in, r}, g, 0.3f) : rnd({r, in}, g, 0.3f);
        TensorPtr Bm = layout == 0 ? rnd({r, out}, g, 0.3f) : rnd({out, r}, g, 0.3f);
        LoRALinear lin(&W, &bias);
        lin.attach_lora(A, Bm, 2.0f, 0, static_cast<int>(out));
        auto y = lin.forward(x);
        auto gy = rnd(y->shape(), g, 1.0f, false);
        backward_with(y, gy);
        std::string p = layout == 0 ? "lora0_" : "lora1_";
        dump(p + "x", x); dump(p + "W", W); dump(p + "bias", bias); dump(p + "A", A); dump(p + "B", Bm);
        dump(p + "y", y); dump(p + "gy", gy);
        dump(p + "gx", x->grad()); dump(p + "gA", A->grad()); dump(p + "gB", Bm->grad());
    }
    {   // Adam (L2-coupled) 3 steps with wd — optim/adam.cpp:25-90
        auto p = rnd({7, 9}, g, 1.0f);
        dump("adam_p0", p);
        AdamConfig oc(1e-2f, 0.9f, 0.999f, 1e-8f, 0.01f, 1.0f, false);
        Adam opt(oc);
        for (int s = 0; s < 3; ++s) {
            auto gr = rnd({7, 9}, g, 1.0f, false);
            dump("adam_g" + std::to_string(s), gr);
            opt.step({p}, {gr});
            dump("adam_p" + std::to_string(s + 1), p);
        }
    }
    std::printf("ops cases written to %s\n", g_dump_dir.c_str());
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s gpt2|gemma|ops [--key value ...]\n", argv[0]);
        return 2;
    }
    Args a = parse(argc, argv, 2);
    g_dump_dir = a.get("dump", "");
#ifdef MFT_HARNESS_REFERENCE
    if (a.geti("threads", 0) > 0) openblas_set_num_threads(a.geti("threads", 0));
    std::printf("impl reference\n");
#else
    std::printf("impl mft_b200\n");
#endif
    try {
        std::string cmd = argv[1];
        if (cmd == "gpt2") return run_gpt2(a);
        if (cmd == "gemma") return run_gemma(a);
        if (cmd == "ops") return run_ops(a);
        std::fprintf(stderr, "unknown command %s\n", cmd.c_str());
        return 2;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}

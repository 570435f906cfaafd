// graph/lora_saver.cpp:20-454 of the reference, restated on top of the shared safetensors reader/writer.
#include "lora_saver.h"

#include <algorithm>
#include <cctype>
#include <iostream>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>

#include "safetensors_loader.h"

namespace ops {

namespace {

const char* target_meta_name(LoraTarget t) {
    switch (t) {
        case LoraTarget::AttnQKV: return "AttnQKV";
        case LoraTarget::AttnProj: return "AttnProj";
        case LoraTarget::MlpFcIn: return "MlpFcIn";
        case LoraTarget::MlpFcOut: return "MlpFcOut";
    }
    return "";
}
bool target_from_meta(std::string s, LoraTarget& out) {       // accepts "AttnQKV" / "attn_qkv" spellings (graph/lora_saver.cpp:33-55)
    for (char& c : s) c = char(std::tolower(static_cast<unsigned char>(c)));
    s.erase(std::remove_if(s.begin(), s.end(), [](char c) { return c == ' '; }), s.end());
    if (s == "attnqkv" || s == "attn_qkv") { out = LoraTarget::AttnQKV; return true; }
    if (s == "attnproj" || s == "attn_proj") { out = LoraTarget::AttnProj; return true; }
    if (s == "mlpfcin" || s == "mlp_fc_in") { out = LoraTarget::MlpFcIn; return true; }
    if (s == "mlpfcout" || s == "mlp_fc_out") { out = LoraTarget::MlpFcOut; return true; }
    return false;
}
bool target_from_key(const std::string& k, LoraTarget& out) {
    if (k == "attn.q" || k == "attn.k" || k == "attn.v" || k == "attn.qkv") { out = LoraTarget::AttnQKV; return true; }
    if (k == "attn.proj") { out = LoraTarget::AttnProj; return true; }
    if (k == "mlp.fc_in") { out = LoraTarget::MlpFcIn; return true; }
    if (k == "mlp.fc_out") { out = LoraTarget::MlpFcOut; return true; }
    return false;
}

}  // namespace

bool LoRAState::compatible_with(const GPT2Model& model) const {
    (void)model;
    for (const auto& kv : tensors)
        if (kv.first.find("layer.0") != std::string::npos && !kv.second->shape().empty()) return rank >= 1 && rank <= 128;
    return !tensors.empty();
}

std::string LoraSaver::make_peft_key(int layer, const std::string& target, const std::string& ab) {
    return "layer." + std::to_string(layer) + "." + target + ".lora_" + ab;
}

bool LoraSaver::parse_peft_key(const std::string& key, int& layer, std::string& target, std::string& ab) {
    const size_t lp = key.find("layer."), rp = key.find(".lora_");
    if (lp == std::string::npos || rp == std::string::npos) return false;
    const size_t ls = lp + 6, le = key.find('.', ls);
    if (le == std::string::npos || le >= rp) return false;
    try { layer = std::stoi(key.substr(ls, le - ls)); } catch (...) { return false; }
    target = key.substr(le + 1, rp - le - 1);
    const std::string tail = key.substr(rp + 6);
    if (tail != "A" && tail != "B") return false;
    ab = tail;
    return true;
}

void LoraSaver::save_safetensors(const std::string& path, const GPT2Model& model, const LoraSpec& spec) {
    std::map<std::string, TensorPtr> state;
    auto wanted = [&](LoraTarget t) { return spec.targets.empty() || std::find(spec.targets.begin(), spec.targets.end(), t) != spec.targets.end(); };
    for (int i = 0; i < model.config().n_layer; ++i) {
        const GPT2BlockWeights& blk = model.get_block(i);
        auto add = [&](const std::string& name, const LoRASlice& s) {
            if (!s.A || !s.B) return;
            state[make_peft_key(i, name, "A")] = s.A;
            state[make_peft_key(i, name, "B")] = s.B;
        };
        if (blk.qkv_lin && wanted(LoraTarget::AttnQKV) && !blk.qkv_lin->slices().empty()) {
            const auto& sl = blk.qkv_lin->slices();
            if (spec.split_qkv) {
                static const char* const names[] = {"attn.q", "attn.k", "attn.v"};
                for (size_t k = 0; k < std::min<size_t>(sl.size(), 3); ++k) add(names[k], sl[k]);
            } else add("attn.qkv", sl.front());
        }
        if (blk.proj_lin && wanted(LoraTarget::AttnProj) && !blk.proj_lin->slices().empty()) add("attn.proj", blk.proj_lin->slices().front());
        if (blk.fc_in_lin && wanted(LoraTarget::MlpFcIn) && !blk.fc_in_lin->slices().empty()) add("mlp.fc_in", blk.fc_in_lin->slices().front());
        if (blk.fc_out_lin && wanted(LoraTarget::MlpFcOut) && !blk.fc_out_lin->slices().empty()) add("mlp.fc_out", blk.fc_out_lin->slices().front());
    }
    std::ostringstream targets, alpha, dropout;
    for (size_t i = 0; i < spec.targets.size(); ++i) targets << (i ? "," : "") << target_meta_name(spec.targets[i]);
    alpha << spec.alpha;
    dropout << spec.dropout;
    write_safetensors(path, state, {{"rank", std::to_string(spec.rank)}, {"alpha", alpha.str()}, {"dropout", dropout.str()},
                                    {"split_qkv", spec.split_qkv ? "true" : "false"}, {"targets", targets.str()}});
    std::cout << "[LoraSaver] Saved " << state.size() << " tensors to " << path << std::endl;
}

LoRAState LoraSaver::load_safetensors(const std::string& path) {
    SafeTensorsReader reader(path);
    reader.parse_header();
    LoRAState st;
    const auto& meta = reader.metadata();
    auto get = [&](const char* k) { auto it = meta.find(k); return it == meta.end() ? std::string() : it->second; };
    try { if (!get("rank").empty()) st.rank = std::stoi(get("rank")); } catch (...) {}
    try { if (!get("alpha").empty()) st.alpha = std::stof(get("alpha")); } catch (...) {}
    try { if (!get("dropout").empty()) st.dropout = std::stof(get("dropout")); } catch (...) {}
    if (!get("split_qkv").empty()) {
        std::string v = get("split_qkv");
        for (char& c : v) c = char(std::tolower(static_cast<unsigned char>(c)));
        st.split_qkv = v == "true" || v == "1";
    }
    if (!get("targets").empty()) {
        std::stringstream ss(get("targets"));
        std::string item;
        while (std::getline(ss, item, ',')) {
            LoraTarget t;
            if (!item.empty() && target_from_meta(item, t) && std::find(st.targets.begin(), st.targets.end(), t) == st.targets.end()) st.targets.push_back(t);
        }
    }
    const std::vector<std::string> names = reader.get_tensor_names();
    if (st.rank <= 0)
        for (const auto& n : names)
            if (n.find("lora_A") != std::string::npos && reader.get_tensor_info(n).shape.size() == 2) { st.rank = int(reader.get_tensor_info(n).shape[1]); break; }
    std::cout << "[LoraSaver] Found " << names.size() << " entries in file" << std::endl;
    for (const auto& n : names) st.tensors[n] = reader.load_tensor(n, false);
    std::cout << "[LoraSaver] Loaded " << st.tensors.size() << " LoRA tensors from " << path << "\n  rank=" << st.rank << ", alpha=" << st.alpha << std::endl;
    bool fused = false, split = false;
    std::set<LoraTarget> seen;
    for (const auto& kv : st.tensors) {
        int layer; std::string target, ab;
        if (!parse_peft_key(kv.first, layer, target, ab)) continue;
        fused |= target == "attn.qkv";
        split |= target == "attn.q" || target == "attn.k" || target == "attn.v";
        LoraTarget t;
        if (target_from_key(target, t)) seen.insert(t);
    }
    if (st.targets.empty()) st.targets.assign(seen.begin(), seen.end());
    if (fused) st.split_qkv = false; else if (split) st.split_qkv = true;
    return st;
}

void LoraSaver::attach_from_state(GPT2Model& model, const LoRAState& state) {
    std::unordered_map<std::string, TensorPtr> A, B;
    for (const auto& kv : state.tensors) {
        int layer; std::string target, ab;
        if (!parse_peft_key(kv.first, layer, target, ab)) continue;
        (ab == "A" ? A : B)[std::to_string(layer) + "." + target] = kv.second;
    }
    const GPT2Config& cfg = model.config();
    const float scale = state.alpha / float(state.rank > 0 ? state.rank : 1);
    for (int i = 0; i < cfg.n_layer; ++i) {
        GPT2BlockWeights& blk = model.get_block(i);
        for (LoRALinear* lin : {blk.qkv_lin.get(), blk.proj_lin.get(), blk.fc_in_lin.get(), blk.fc_out_lin.get()})
            if (lin) lin->clear_lora();
    }
    auto attach = [&](int layer, const std::string& target, LoRALinear* lin, int col0, int cols) {
        if (!lin) return false;
        const std::string k = std::to_string(layer) + "." + target;
        auto a = A.find(k), b = B.find(k);
        if (a == A.end() || b == B.end() || !a->second || !b->second) return false;
        a->second->set_requires_grad(true);
        b->second->set_requires_grad(true);
        lin->attach_lora(a->second, b->second, scale, col0, cols);
        return true;
    };
    const int C = cfg.n_embd;
    for (int i = 0; i < cfg.n_layer; ++i) {
        GPT2BlockWeights& blk = model.get_block(i);
        if (!attach(i, "attn.qkv", blk.qkv_lin.get(), 0, 3 * C)) {
            attach(i, "attn.q", blk.qkv_lin.get(), 0, C);
            attach(i, "attn.k", blk.qkv_lin.get(), C, C);
            attach(i, "attn.v", blk.qkv_lin.get(), 2 * C, C);
        }
        attach(i, "attn.proj", blk.proj_lin.get(), 0, C);
        attach(i, "mlp.fc_in", blk.fc_in_lin.get(), 0, 4 * C);
        attach(i, "mlp.fc_out", blk.fc_out_lin.get(), 0, C);
    }
    std::cout << "[LoraSaver] Attached LoRA to model (rank=" << state.rank << ", alpha=" << state.alpha << ")" << std::endl;
}

}  // namespace ops

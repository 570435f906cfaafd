// data/wikitext2_dataset.cpp:96-635 of the reference, restated with a resident token stream.
#include "wikitext2_dataset.h"

#include <algorithm>
#include <cctype>
#include <fstream>
#include <iostream>
#include <iterator>
#include <numeric>
#include <stdexcept>

#include "ops.h"

namespace ops {

namespace {

bool blank(const std::string& s) {
    for (unsigned char c : s) if (!std::isspace(c)) return false;
    return true;
}
// first integer after "key": at or after `from` (the meta file is flat enough for the reference to scrape it this way)
bool find_int(const std::string& text, const std::string& key, int64_t& out, size_t from = 0) {
    size_t k = text.find("\"" + key + "\"", from);
    if (k == std::string::npos) return false;
    size_t p = text.find(':', k);
    if (p == std::string::npos) return false;
    ++p;
    while (p < text.size() && std::isspace(static_cast<unsigned char>(text[p]))) ++p;
    size_t e = p;
    if (e < text.size() && text[e] == '-') ++e;
    while (e < text.size() && std::isdigit(static_cast<unsigned char>(text[e]))) ++e;
    if (e == p) return false;
    try { out = std::stoll(text.substr(p, e - p)); } catch (...) { return false; }
    return true;
}
bool find_bool(const std::string& text, const std::string& key, bool& out) {
    size_t k = text.find("\"" + key + "\"");
    if (k == std::string::npos) return false;
    size_t p = text.find(':', k);
    if (p == std::string::npos) return false;
    ++p;
    while (p < text.size() && std::isspace(static_cast<unsigned char>(text[p]))) ++p;
    if (text.compare(p, 4, "true") == 0) { out = true; return true; }
    if (text.compare(p, 5, "false") == 0) { out = false; return true; }
    return false;
}
bool find_range(const std::string& text, const std::string& name, size_t& offset, size_t& length) {
    size_t k = text.find("\"" + name + "\"");
    if (k == std::string::npos) return false;
    int64_t o = 0, l = 0;
    if (!find_int(text, "offset", o, k) || !find_int(text, "length", l, k)) return false;
    offset = size_t(std::max<int64_t>(0, o));
    length = size_t(std::max<int64_t>(0, l));
    return true;
}
std::string meta_for(const std::string& bin, const std::string& given) {
    if (!given.empty()) return given;
    const size_t slash = bin.find_last_of('/');
    const std::string dir = slash == std::string::npos ? std::string(".") : bin.substr(0, slash);
    std::vector<std::string> cands = {dir + "/meta.json", bin + ".json", bin + ".meta.json"};
    const size_t dot = bin.find_last_of('.');
    if (dot != std::string::npos && (slash == std::string::npos || dot > slash)) cands.push_back(bin.substr(0, dot) + ".json");
    for (const auto& c : cands) { std::ifstream f(c); if (f.good()) return c; }
    return "";
}

}  // namespace

WikiText2Dataset::WikiText2Dataset(const WT2Config& cfg, GPT2BPETokenizer* tok) : cfg_(cfg), rng_(cfg.seed) {
    encode_fn_ = [tok](const std::string& text) {
        const std::vector<int> e = tok->encode(text, false, 0, false);
        return std::vector<int32_t>(e.begin(), e.end());
    };
}
WikiText2Dataset::WikiText2Dataset(const WT2Config& cfg, std::function<std::vector<int32_t>(const std::string&)> encode_fn)
    : cfg_(cfg), encode_fn_(std::move(encode_fn)), rng_(cfg.seed) {}

void WikiText2Dataset::build_starts(bool streaming_rule) {
    starts_.clear();
    const size_t S = size_t(cfg_.seq_len), stride = cfg_.stride <= 0 ? S : size_t(cfg_.stride), need = S + 1;
    for (size_t s = 0; s + need <= total_tokens_; s += stride) starts_.push_back(s);
    if (!streaming_rule && !cfg_.drop_last) {          // keep the tail as one more (overlapping) window
        if (starts_.empty() || starts_.back() + need < total_tokens_) {
            const size_t s = total_tokens_ > need ? total_tokens_ - need : 0;
            if (starts_.empty() || starts_.back() != s) starts_.push_back(s);
        }
    }
    order_.resize(starts_.size());
    std::iota(order_.begin(), order_.end(), size_t(0));
}

void WikiText2Dataset::load_raw(Split split) {
    const std::string& path = split == Split::Train ? cfg_.train_path : split == Split::Valid ? cfg_.valid_path : cfg_.test_path;
    std::ifstream in(path);
    if (!in) throw std::runtime_error((cfg_.streaming_mode ? "Cannot open file: " : "Open WT2 file failed: ") + path);
    ids_.clear();
    std::string line;
    while (std::getline(in, line)) {
        if (!cfg_.insert_eos_between_lines && blank(line)) continue;
        const std::vector<int32_t> e = encode_fn_(line);
        ids_.insert(ids_.end(), e.begin(), e.end());
        if (cfg_.insert_eos_between_lines) ids_.push_back(int32_t(cfg_.eos_id));
    }
    const float frac = std::min(1.0f, std::max(0.0f, cfg_.data_fraction));
    const size_t min_tokens = size_t(cfg_.seq_len) + 1;
    if (cfg_.streaming_mode) {
        // the reference's prescan counts one token past the end of the file and chunks over that count; the extra token
        // is never produced, so a final window may come back one position short (padded)
        const size_t counted = ids_.size() + 1;
        size_t limit = size_t(double(counted) * frac);
        limit = std::min(std::max(limit, min_tokens), counted);
        total_tokens_ = limit;
        if (ids_.size() > total_tokens_) ids_.resize(total_tokens_);
        build_starts(true);
    } else {
        if (ids_.empty() || ids_.back() != cfg_.eos_id) ids_.push_back(int32_t(cfg_.eos_id));
        if (frac < 1.0f) {
            size_t limit = size_t(double(ids_.size()) * frac);
            limit = std::min(std::max(limit, min_tokens), ids_.size());
            ids_.resize(limit);
            if (ids_.empty() || ids_.back() != cfg_.eos_id) ids_.push_back(int32_t(cfg_.eos_id));
        }
        total_tokens_ = ids_.size();
        build_starts(false);
    }
}

void WikiText2Dataset::load_pretokenized(Split split) {
    const std::string meta_path = meta_for(cfg_.pretokenized_path, cfg_.pretokenized_meta);
    if (meta_path.empty()) throw std::runtime_error("Cannot locate meta.json for pretokenized dataset");
    std::ifstream mf(meta_path);
    if (!mf) throw std::runtime_error("Open pretokenized meta failed: " + meta_path);
    const std::string meta((std::istreambuf_iterator<char>(mf)), std::istreambuf_iterator<char>());
    int64_t v = 0;
    if (!find_int(meta, "total_tokens", v)) throw std::runtime_error("meta.json missing total_tokens");
    const size_t total = size_t(std::max<int64_t>(0, v));
    if (find_int(meta, "eos_token_id", v)) cfg_.eos_id = int(v);
    if (find_int(meta, "pad_token_id", v)) cfg_.pad_id = int(v);
    bool b = true;
    if (find_bool(meta, "insert_eos_between_lines", b)) cfg_.insert_eos_between_lines = b;
    size_t off = 0, len = 0;
    bool have = false;
    if (split == Split::Train) have = find_range(meta, "train", off, len);
    else if (split == Split::Valid) have = find_range(meta, "valid", off, len) || find_range(meta, "validation", off, len);
    else have = find_range(meta, "test", off, len);
    if (!have) throw std::runtime_error("Requested split not found in pretokenized meta");
    size_t use = len;
    const float frac = std::min(1.0f, std::max(0.0f, cfg_.data_fraction));
    if (frac < 1.0f) use = std::max(size_t(cfg_.seq_len) + 1, std::min(use, size_t(double(use) * frac)));
    if (off >= total) throw std::runtime_error("Pretokenized split offset exceeds total tokens");
    use = std::min(use, total - off);
    std::ifstream bin(cfg_.pretokenized_path, std::ios::binary);
    if (!bin) throw std::runtime_error("Open pretokenized bin failed: " + cfg_.pretokenized_path);
    ids_.assign(use, 0);
    bin.seekg(std::streamoff(off * sizeof(int32_t)));
    bin.read(reinterpret_cast<char*>(ids_.data()), std::streamsize(use * sizeof(int32_t)));
    ids_.resize(size_t(bin.gcount()) / sizeof(int32_t));
    if (ids_.empty()) throw std::runtime_error("Loaded zero tokens for pretokenized split");
    total_tokens_ = ids_.size();
    build_starts(false);
    std::cout << "[Pretokenized] split=" << (split == Split::Train ? "train" : split == Split::Valid ? "valid" : "test") << " tokens=" << total_tokens_
              << " chunks=" << starts_.size() << std::endl;
}

void WikiText2Dataset::load(Split split) {
    ids_.clear(); starts_.clear(); order_.clear();
    cursor_ = 0;
    total_tokens_ = 0;
    if (!cfg_.pretokenized_path.empty()) {
        cfg_.streaming_mode = false;
        load_pretokenized(split);
        if (split == Split::Train && cfg_.shuffle_train) shuffle();
        return;
    }
    load_raw(split);
    // data/wikitext2_dataset.cpp:239-262: streaming honours shuffle_train, the resident path always shuffles the train split
    if (split == Split::Train && (cfg_.streaming_mode ? cfg_.shuffle_train : true)) shuffle();
}

size_t WikiText2Dataset::num_sequences() const { return starts_.size(); }
void WikiText2Dataset::shuffle() { std::shuffle(order_.begin(), order_.end(), rng_); cursor_ = 0; }
void WikiText2Dataset::reset_cursor() { cursor_ = 0; }

Batch WikiText2Dataset::next_batch(size_t batch_size, bool need_loop) {
    if (order_.empty()) throw std::runtime_error("Dataset not loaded");
    if (cursor_ >= order_.size()) {
        if (!need_loop) return Batch{nullptr, nullptr, nullptr};
        shuffle();
    }
    const size_t start = cursor_;
    cursor_ = std::min(cursor_ + batch_size, order_.size());
    return get_batch(start, cursor_ - start);
}

std::vector<int32_t> WikiText2Dataset::peek_tokens(size_t count) const {
    if (count == 0 || ids_.empty()) return {};
    return std::vector<int32_t>(ids_.begin(), ids_.begin() + long(std::min(count, ids_.size())));
}

Batch WikiText2Dataset::get_batch(size_t index_start, size_t batch_size) const {
    const size_t S = size_t(cfg_.seq_len);
    in_buf_.assign(batch_size * S, int32_t(cfg_.pad_id));
    label_buf_.assign(batch_size * S, -100);
    mask_buf_.assign(batch_size * S, 0.0f);
    for (size_t b = 0; b < batch_size; ++b) {
        const size_t start = starts_[order_[index_start + b]];
        // a window needs S+1 tokens; when fewer are there the last one is held back like the (absent) next-token target
        const size_t avail = start < ids_.size() ? std::min(S + 1, ids_.size() - start) : 0;
        const size_t n = avail >= 2 ? std::min(avail - 1, S) : 0;
        for (size_t i = 0; i < n; ++i) {
            in_buf_[b * S + i] = ids_[start + i];
            label_buf_[b * S + i] = ids_[start + i];
            mask_buf_[b * S + i] = 1.0f;
        }
    }
    Batch out;
    const std::vector<int64_t> shape{int64_t(batch_size), int64_t(S)};
    out.input_ids = from_blob(in_buf_.data(), shape, kInt32, kCPU);
    out.attention_mask = from_blob(mask_buf_.data(), shape, kFloat32, kCPU);
    out.labels = from_blob(label_buf_.data(), shape, kInt32, kCPU);
    return out;
}

}  // namespace ops

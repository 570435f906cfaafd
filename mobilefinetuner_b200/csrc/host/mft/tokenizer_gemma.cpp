#include "tokenizer_gemma.h"

#include <cctype>
#include <climits>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace ops {

namespace {

const char kSpace[] = "\xE2\x96\x81";   // U+2581, the sentencepiece word boundary

// decodes the JSON string that starts at s[i] == '"' and leaves i after the closing quote (BMP \u escapes as UTF-8)
std::string take_string(const std::string& s, size_t& i) {
    std::string out;
    ++i;
    while (i < s.size() && s[i] != '"') {
        if (s[i] == '\\' && i + 1 < s.size()) {
            const char e = s[i + 1];
            i += 2;
            if (e == 'u' && i + 4 <= s.size()) {
                const int cp = std::stoi(s.substr(i, 4), nullptr, 16);
                i += 4;
                if (cp < 0x80) out.push_back(char(cp));
                else if (cp < 0x800) { out.push_back(char(0xC0 | (cp >> 6))); out.push_back(char(0x80 | (cp & 0x3F))); }
                else { out.push_back(char(0xE0 | (cp >> 12))); out.push_back(char(0x80 | ((cp >> 6) & 0x3F))); out.push_back(char(0x80 | (cp & 0x3F))); }
            } else if (e == 'n') out.push_back('\n');
            else if (e == 't') out.push_back('\t');
            else if (e == 'r') out.push_back('\r');
            else out.push_back(e);
        } else out.push_back(s[i++]);
    }
    ++i;
    return out;
}
// index one past the bracket that closes the one at `open` (strings skipped)
size_t block_end(const std::string& s, size_t open, char o, char c) {
    int depth = 0;
    for (size_t i = open; i < s.size(); ++i) {
        if (s[i] == '"') { take_string(s, i); --i; continue; }
        if (s[i] == o) ++depth;
        else if (s[i] == c && --depth == 0) return i + 1;
    }
    throw std::runtime_error("Unterminated JSON block while parsing tokenizer.json");
}
bool byte_token(const std::string& t) {
    return t.size() == 6 && t[0] == '<' && t[1] == '0' && t[2] == 'x' && t[5] == '>' && std::isxdigit(static_cast<unsigned char>(t[3])) &&
           std::isxdigit(static_cast<unsigned char>(t[4]));
}

}  // namespace

GemmaTokenizerConfig GemmaTokenizerConfig::from_pretrained(const std::string& model_dir) {
    GemmaTokenizerConfig c;
    c.tokenizer_json_path = model_dir + "/tokenizer.json";
    return c;
}

GemmaTokenizer::GemmaTokenizer(GemmaTokenizerConfig config) : config_(std::move(config)) {}

void GemmaTokenizer::load() {
    std::ifstream f(config_.tokenizer_json_path);
    if (!f.is_open()) throw std::runtime_error("Failed to open tokenizer json: " + config_.tokenizer_json_path);
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string s = ss.str();
    const size_t model = s.find("\"model\"");
    if (model == std::string::npos) throw std::runtime_error("tokenizer.json missing \"model\" block");
    const size_t vkey = s.find("\"vocab\"", model);
    if (vkey == std::string::npos) throw std::runtime_error("tokenizer.json missing \"vocab\" block");
    const size_t vopen = s.find('{', vkey);
    if (vopen == std::string::npos) throw std::runtime_error("Failed to locate vocab object in tokenizer.json");
    const size_t vend = block_end(s, vopen, '{', '}');
    vocab_.clear();
    id_to_token_.clear();
    for (size_t i = vopen + 1; i < vend;) {
        if (s[i] != '"') { ++i; continue; }
        const std::string tok = take_string(s, i);
        while (i < vend && s[i] != ':') ++i;
        ++i;
        while (i < vend && std::isspace(static_cast<unsigned char>(s[i]))) ++i;
        size_t j = i;
        while (j < vend && std::isdigit(static_cast<unsigned char>(s[j]))) ++j;
        if (j == i) continue;
        const int id = std::stoi(s.substr(i, j - i));
        vocab_[tok] = id;
        id_to_token_[id] = tok;
        i = j;
    }
    const size_t mkey = s.find("\"merges\"", vkey);
    if (mkey == std::string::npos) throw std::runtime_error("tokenizer.json missing \"merges\" array");
    const size_t mopen = s.find('[', mkey);
    if (mopen == std::string::npos) throw std::runtime_error("Failed to locate merges array in tokenizer.json");
    const size_t mend = block_end(s, mopen, '[', ']');
    ranks_.clear();
    int rank = 0;
    for (size_t i = mopen + 1; i < mend;) {          // [["a","b"], ...]
        if (s[i] != '[') { ++i; continue; }
        size_t k = i + 1;
        while (k < mend && s[k] != '"' && s[k] != ']') ++k;
        if (k >= mend || s[k] == ']') { i = k + 1; continue; }
        const std::string a = take_string(s, k);
        while (k < mend && s[k] != '"' && s[k] != ']') ++k;
        if (k >= mend || s[k] == ']') { i = k + 1; continue; }
        const std::string b = take_string(s, k);
        ranks_[{a, b}] = rank++;
        i = k;
    }
    auto need = [&](const std::string& tok) {
        auto it = vocab_.find(tok);
        if (it == vocab_.end()) throw std::runtime_error("Token not found in vocab: " + tok);
        return it->second;
    };
    bos_token_id_ = need(config_.bos_token);
    eos_token_id_ = need(config_.eos_token);
    pad_token_id_ = need(config_.pad_token);
    unk_token_id_ = need(config_.unk_token);
    vocab_size_ = int(vocab_.size());
    special_ids_ = {bos_token_id_, eos_token_id_, pad_token_id_, unk_token_id_};
    for (const auto& t : config_.special_tokens) { auto it = vocab_.find(t); if (it != vocab_.end()) special_ids_.insert(it->second); }
}

std::vector<std::string> GemmaTokenizer::merge(const std::string& word) const {
    if (word.empty()) return {};
    if (vocab_.count(word)) return {word};
    std::vector<std::string> w;
    for (size_t i = 0; i < word.size();) {
        const unsigned char c = static_cast<unsigned char>(word[i]);
        size_t l = c < 0x80 ? 1 : (c & 0xE0) == 0xC0 ? 2 : (c & 0xF0) == 0xE0 ? 3 : (c & 0xF8) == 0xF0 ? 4 : 1;
        if (i + l > word.size()) l = word.size() - i;
        w.emplace_back(word.substr(i, l));
        i += l;
    }
    while (w.size() > 1) {
        int best = INT_MAX;
        size_t at = 0;
        for (size_t i = 0; i + 1 < w.size(); ++i) {
            auto it = ranks_.find({w[i], w[i + 1]});
            if (it != ranks_.end() && it->second < best) { best = it->second; at = i; }
        }
        if (best == INT_MAX) break;
        w[at] += w[at + 1];
        w.erase(w.begin() + long(at) + 1);
    }
    return w;
}

std::vector<int> GemmaTokenizer::encode(const std::string& text, bool add_special_tokens, int max_length, bool truncation) const {
    std::vector<int> ids;
    if (add_special_tokens && config_.add_bos_token) ids.push_back(bos_token_id_);
    std::string norm;
    for (char c : text) { if (c == ' ') norm += kSpace; else norm.push_back(c); }
    for (const std::string& piece : merge(norm)) {
        auto it = vocab_.find(piece);
        if (it != vocab_.end()) { ids.push_back(it->second); continue; }
        for (unsigned char b : piece) {                 // byte fallback
            char buf[8];
            std::snprintf(buf, sizeof(buf), "<0x%02X>", b);
            auto fb = vocab_.find(buf);
            ids.push_back(fb != vocab_.end() ? fb->second : unk_token_id_);
        }
    }
    if (add_special_tokens && config_.add_eos_token) ids.push_back(eos_token_id_);
    if (truncation && max_length > 0 && int(ids.size()) > max_length) ids.resize(size_t(max_length));
    return ids;
}

std::string GemmaTokenizer::decode(const std::vector<int>& ids, bool skip_special_tokens) const {
    std::string sp;
    for (int id : ids) {
        if (skip_special_tokens && special_ids_.count(id)) continue;
        auto it = id_to_token_.find(id);
        if (it == id_to_token_.end()) continue;
        if (byte_token(it->second)) sp.push_back(char(std::stoi(it->second.substr(3, 2), nullptr, 16)));
        else sp += it->second;
    }
    std::string out;
    for (size_t i = 0; i < sp.size();) {
        if (sp.compare(i, 3, kSpace) == 0) { out.push_back(' '); i += 3; } else out.push_back(sp[i++]);
    }
    return out;
}

}  // namespace ops

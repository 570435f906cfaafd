// core/tokenizer_bpe.cpp:20-466 of the reference, restated: same byte alphabet, same pre-tokeniser classes, same greedy
// lowest-rank merge loop, same special-token / truncation / padding rules.
#include "tokenizer_bpe.h"

#include <algorithm>
#include <climits>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace ops {

namespace {

std::string utf8(int cp) {
    std::string s;
    if (cp < 0x80) s.push_back(char(cp));
    else if (cp < 0x800) { s.push_back(char(0xC0 | (cp >> 6))); s.push_back(char(0x80 | (cp & 0x3F))); }
    else { s.push_back(char(0xE0 | (cp >> 12))); s.push_back(char(0x80 | ((cp >> 6) & 0x3F))); s.push_back(char(0x80 | (cp & 0x3F))); }
    return s;
}
int utf8_len(unsigned char c) { return c < 0x80 ? 1 : (c & 0xE0) == 0xC0 ? 2 : (c & 0xF0) == 0xE0 ? 3 : (c & 0xF8) == 0xF0 ? 4 : 1; }

// JSON string starting at s[i] == '"'; returns the decoded text and leaves i after the closing quote
std::string json_string(const std::string& s, size_t& i) {
    std::string out;
    ++i;
    while (i < s.size() && s[i] != '"') {
        if (s[i] == '\\' && i + 1 < s.size()) {
            const char e = s[i + 1];
            i += 2;
            if (e == 'u' && i + 4 <= s.size()) {
                int cp = std::stoi(s.substr(i, 4), nullptr, 16);
                i += 4;
                if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 <= s.size() && s[i] == '\\' && s[i + 1] == 'u') {   // surrogate pair
                    const int lo = std::stoi(s.substr(i + 2, 4), nullptr, 16);
                    i += 6;
                    cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    out.push_back(char(0xF0 | (cp >> 18))); out.push_back(char(0x80 | ((cp >> 12) & 0x3F)));
                    out.push_back(char(0x80 | ((cp >> 6) & 0x3F))); out.push_back(char(0x80 | (cp & 0x3F)));
                } else out += utf8(cp);
            } else if (e == 'n') out.push_back('\n');
            else if (e == 't') out.push_back('\t');
            else if (e == 'r') out.push_back('\r');
            else if (e == 'b') out.push_back('\b');
            else if (e == 'f') out.push_back('\f');
            else out.push_back(e);
        } else out.push_back(s[i++]);
    }
    ++i;
    return out;
}

bool is_alpha(unsigned char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }
bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
bool is_space(unsigned char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }
bool is_other(unsigned char c) { return !is_space(c) && !is_alpha(c) && !is_digit(c) && c != '_'; }   // [^\s\w]

}  // namespace

GPT2BPETokenizer::GPT2BPETokenizer(const BPEConfig& config) : config_(config) {
    // the published GPT-2 byte alphabet: printable Latin-1 bytes stand for themselves, the other 68 bytes take the code
    // points 256, 257, ... in byte order
    bool printable[256] = {false};
    for (int b = '!'; b <= '~'; ++b) printable[b] = true;
    for (int b = 0xA1; b <= 0xAC; ++b) printable[b] = true;
    for (int b = 0xAE; b <= 0xFF; ++b) printable[b] = true;
    int next = 256;
    for (int b = 0; b < 256; ++b) {
        byte_to_text_[b] = utf8(printable[b] ? b : next++);
        text_to_byte_[byte_to_text_[b]] = uint8_t(b);
    }
}

void GPT2BPETokenizer::load() {
    {   // vocab.json: one flat object of "token": id
        std::ifstream f(config_.vocab_path);
        if (!f) throw std::runtime_error("Cannot open vocab file: " + config_.vocab_path);
        std::stringstream ss;
        ss << f.rdbuf();
        const std::string s = ss.str();
        vocab_.clear();
        id_to_token_.clear();
        size_t i = s.find('{');
        if (i == std::string::npos) throw std::runtime_error("vocab.json: not a JSON object");
        ++i;
        while (i < s.size()) {
            while (i < s.size() && s[i] != '"' && s[i] != '}') ++i;
            if (i >= s.size() || s[i] == '}') break;
            const std::string tok = json_string(s, i);
            while (i < s.size() && s[i] != ':') ++i;
            ++i;
            while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\t' || s[i] == '\r')) ++i;
            size_t j = i;
            while (j < s.size() && (is_digit(static_cast<unsigned char>(s[j])) || s[j] == '-')) ++j;
            if (j == i) throw std::runtime_error("vocab.json: id expected after \"" + tok + "\"");
            const int id = std::stoi(s.substr(i, j - i));
            i = j;
            vocab_[tok] = id;
            id_to_token_[id] = tok;
        }
        vocab_size_ = int(vocab_.size());
    }
    {   // merges.txt
        std::ifstream f(config_.merges_path);
        if (!f) throw std::runtime_error("Cannot open merges file: " + config_.merges_path);
        ranks_.clear();
        std::string line;
        int rank = 0;
        bool first = true;
        while (std::getline(f, line)) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (first) { first = false; if (line.rfind("#version", 0) == 0) continue; }
            if (line.empty()) continue;
            if (line.find(' ') == std::string::npos) continue;
            ranks_[line] = rank++;
        }
    }
    loaded_ = true;
    std::cout << "[GPT2BPETokenizer] vocab=" << vocab_size_ << " merges=" << ranks_.size() << std::endl;
}

// Leftmost match of:  's|'t|'re|'ve|'m|'ll|'d | ' '?[A-Za-z]+ | ' '?[0-9]+ | ' '?[^\s\w]+ | \s+   (first alternative that
// matches wins; a byte no alternative matches — '_' — is skipped), on the byte-mapped text.
std::vector<std::string> GPT2BPETokenizer::pretokenize(const std::string& t) const {
    std::vector<std::string> out;
    const size_t n = t.size();
    auto c = [&](size_t k) { return static_cast<unsigned char>(t[k]); };
    size_t p = 0;
    while (p < n) {
        size_t len = 0;
        if (t[p] == '\'' && p + 1 < n) {
            const char a = t[p + 1];
            if (a == 's' || a == 't' || a == 'm' || a == 'd') len = 2;
            else if (p + 2 < n && ((a == 'r' && t[p + 2] == 'e') || (a == 'v' && t[p + 2] == 'e') || (a == 'l' && t[p + 2] == 'l'))) len = 3;
        }
        if (!len) {
            const size_t q = (t[p] == ' ' && p + 1 < n) ? p + 1 : p;     // optional leading space
            bool (*cls)(unsigned char) = nullptr;
            if (is_alpha(c(q))) cls = is_alpha; else if (is_digit(c(q))) cls = is_digit; else if (is_other(c(q))) cls = is_other;
            if (cls) {
                size_t e = q;
                while (e < n && cls(c(e))) ++e;
                len = e - p;
            }
        }
        if (!len && is_space(c(p))) {
            size_t e = p;
            while (e < n && is_space(c(e))) ++e;
            // \s+(?!\S): give back one whitespace when a non-space follows and more than one was taken
            if (e < n && e - p > 1) --e;
            len = e - p;
        }
        if (!len) { ++p; continue; }
        out.emplace_back(t.substr(p, len));
        p += len;
    }
    return out;
}

std::vector<std::string> GPT2BPETokenizer::merge(const std::string& piece) const {
    if (piece.empty()) return {};
    if (vocab_.count(piece)) return {piece};
    std::vector<std::string> w;
    for (size_t i = 0; i < piece.size();) {
        size_t l = size_t(utf8_len(static_cast<unsigned char>(piece[i])));
        if (i + l > piece.size()) l = piece.size() - i;
        w.push_back(piece.substr(i, l));
        i += l;
    }
    while (w.size() > 1) {
        int best = INT_MAX;
        size_t at = 0;
        for (size_t i = 0; i + 1 < w.size(); ++i) {
            auto it = ranks_.find(w[i] + " " + w[i + 1]);
            if (it != ranks_.end() && it->second < best) { best = it->second; at = i; }
        }
        if (best == INT_MAX) break;
        w[at] += w[at + 1];                        // only the first occurrence of the best pair is merged per round
        w.erase(w.begin() + long(at) + 1);
    }
    return w;
}

std::vector<int> GPT2BPETokenizer::encode(const std::string& text, bool add_special_tokens, int max_length, bool truncation) {
    if (!loaded_) load();
    std::string mapped;
    mapped.reserve(text.size() * 2);
    for (unsigned char b : text) mapped += byte_to_text_[b];
    std::vector<int> ids;
    for (const std::string& piece : pretokenize(mapped))
        for (const std::string& tok : merge(piece)) {
            auto it = vocab_.find(tok);
            if (it != vocab_.end()) ids.push_back(it->second);
            else std::cerr << "[ERROR] Token not in vocab: \"" << tok << "\" (len=" << tok.size() << ")" << std::endl;
        }
    if (add_special_tokens) ids.push_back(config_.eos_token_id);
    if (truncation && max_length > 0 && int(ids.size()) > max_length) ids.resize(size_t(max_length));
    return ids;
}

std::string GPT2BPETokenizer::decode(const std::vector<int>& ids, bool skip_special_tokens) {
    if (!loaded_) load();
    std::string mapped;
    for (int id : ids) {
        if (skip_special_tokens && (id == config_.eos_token_id || id == config_.bos_token_id || id == config_.pad_token_id)) continue;
        auto it = id_to_token_.find(id);
        if (it != id_to_token_.end()) mapped += it->second;
    }
    std::string out;
    for (size_t i = 0; i < mapped.size();) {
        size_t l = size_t(utf8_len(static_cast<unsigned char>(mapped[i])));
        if (i + l > mapped.size()) l = mapped.size() - i;
        auto it = text_to_byte_.find(mapped.substr(i, l));
        if (it != text_to_byte_.end()) out.push_back(char(it->second));
        i += l;
    }
    return out;
}

std::pair<std::vector<std::vector<int>>, std::vector<std::vector<int>>> GPT2BPETokenizer::batch_encode(
    const std::vector<std::string>& texts, int max_length, const std::string& padding, bool truncation) {
    std::vector<std::vector<int>> ids, masks;
    for (const auto& t : texts) ids.push_back(encode(t, false, max_length, truncation));
    int target = 0;
    if (padding == "max_length" && max_length > 0) target = max_length;
    else if (padding == "longest") for (const auto& v : ids) target = std::max(target, int(v.size()));
    for (auto& v : ids) {
        std::vector<int> m(v.size(), 1);
        if (target > 0 && int(v.size()) < target) {
            const size_t pad = size_t(target) - v.size();
            v.insert(v.end(), pad, config_.pad_token_id);
            m.insert(m.end(), pad, 0);
        }
        masks.push_back(std::move(m));
    }
    return {ids, masks};
}

std::string GPT2BPETokenizer::get_token_string(int token_id) const {
    auto it = id_to_token_.find(token_id);
    return it == id_to_token_.end() ? std::string() : it->second;
}

}  // namespace ops

// safetensors reader / writer (see safetensors_loader.h). Behaviour follows graph/safetensors_loader.cpp:20-384 of the
// reference: same accepted dtypes, same "transformer." prefix fallback, same rule for which tensors are transposed.
#include "safetensors_loader.h"

#include <algorithm>
#include <cstring>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace ops {

namespace {

// ---- minimal JSON reader for the header: objects, arrays, strings, integers (all the format uses)
struct JsonCursor {
    const std::string& s;
    size_t i = 0;
    explicit JsonCursor(const std::string& str) : s(str) {}
    void ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\t' || s[i] == '\r')) ++i; }
    char peek() { ws(); return i < s.size() ? s[i] : '\0'; }
    void expect(char c) {
        ws();
        if (i >= s.size() || s[i] != c) throw std::runtime_error(std::string("safetensors header: expected '") + c + "' at offset " + std::to_string(i));
        ++i;
    }
    bool accept(char c) { ws(); if (i < s.size() && s[i] == c) { ++i; return true; } return false; }
    std::string string() {
        expect('"');
        std::string out;
        while (i < s.size() && s[i] != '"') {
            if (s[i] == '\\' && i + 1 < s.size()) {
                const char e = s[i + 1];
                i += 2;
                switch (e) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {   // keep BMP code points as UTF-8
                        unsigned cp = 0;
                        for (int k = 0; k < 4 && i < s.size(); ++k, ++i) cp = cp * 16 + unsigned(std::stoi(std::string(1, s[i]), nullptr, 16));
                        if (cp < 0x80) out.push_back(char(cp));
                        else if (cp < 0x800) { out.push_back(char(0xC0 | (cp >> 6))); out.push_back(char(0x80 | (cp & 0x3F))); }
                        else { out.push_back(char(0xE0 | (cp >> 12))); out.push_back(char(0x80 | ((cp >> 6) & 0x3F))); out.push_back(char(0x80 | (cp & 0x3F))); }
                        break;
                    }
                    default: out.push_back(e);
                }
            } else out.push_back(s[i++]);
        }
        expect('"');
        return out;
    }
    int64_t integer() {
        ws();
        size_t j = i;
        if (j < s.size() && (s[j] == '-' || s[j] == '+')) ++j;
        while (j < s.size() && s[j] >= '0' && s[j] <= '9') ++j;
        if (j == i) throw std::runtime_error("safetensors header: integer expected at offset " + std::to_string(i));
        const int64_t v = std::stoll(s.substr(i, j - i));
        i = j;
        return v;
    }
    void skip_value() {   // anything we do not interpret
        ws();
        if (peek() == '"') { string(); return; }
        if (accept('{')) { if (!accept('}')) { do { string(); expect(':'); skip_value(); } while (accept(',')); expect('}'); } return; }
        if (accept('[')) { if (!accept(']')) { do { skip_value(); } while (accept(',')); expect(']'); } return; }
        while (i < s.size() && s[i] != ',' && s[i] != '}' && s[i] != ']') ++i;
    }
};

float half_to_float(uint16_t h) {
    const uint32_t sign = uint32_t(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {   // subnormal: renormalise
            int e = -1;
            do { ++e; man <<= 1; } while (!(man & 0x400u));
            bits = sign | uint32_t(127 - 15 - e) << 23 | (man & 0x3FFu) << 13;
        }
    } else if (exp == 31) bits = sign | 0x7F800000u | man << 13;
    else bits = sign | (exp + 112u) << 23 | man << 13;
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

size_t dtype_size(const std::string& d) {
    if (d == "F32" || d == "I32" || d == "U32") return 4;
    if (d == "F16" || d == "BF16" || d == "I16" || d == "U16") return 2;
    if (d == "F64" || d == "I64" || d == "U64") return 8;
    if (d == "I8" || d == "U8" || d == "BOOL") return 1;
    throw std::runtime_error("safetensors: unsupported dtype " + d);
}

}  // namespace

SafeTensorsReader::SafeTensorsReader(const std::string& filepath) : filepath_(filepath) {
    file_.open(filepath, std::ios::binary);
    if (!file_) throw std::runtime_error("SafeTensorsReader: cannot open " + filepath);
}
SafeTensorsReader::~SafeTensorsReader() { if (file_.is_open()) file_.close(); }

void SafeTensorsReader::parse_header() {
    file_.clear();
    file_.seekg(0, std::ios::end);
    const size_t file_size = size_t(file_.tellg());
    file_.seekg(0);
    uint64_t len = 0;
    file_.read(reinterpret_cast<char*>(&len), 8);
    if (!file_ || len == 0 || len > file_size - 8) throw std::runtime_error("SafeTensorsReader: bad header length in " + filepath_);
    header_len_ = size_t(len);
    data_offset_ = 8 + header_len_;
    std::string json(header_len_, '\0');
    file_.read(&json[0], std::streamsize(header_len_));
    if (!file_) throw std::runtime_error("SafeTensorsReader: truncated header in " + filepath_);
    tensor_map_.clear();
    order_.clear();
    metadata_.clear();
    JsonCursor c(json);
    c.expect('{');
    if (!c.accept('}')) {
        do {
            const std::string name = c.string();
            c.expect(':');
            if (name == "__metadata__") {
                c.expect('{');
                if (!c.accept('}')) {
                    do {
                        const std::string k = c.string();
                        c.expect(':');
                        if (c.peek() == '"') metadata_[k] = c.string(); else c.skip_value();
                    } while (c.accept(','));
                    c.expect('}');
                }
                continue;
            }
            SafeTensorInfo info;
            c.expect('{');
            do {
                const std::string k = c.string();
                c.expect(':');
                if (k == "dtype") info.dtype = c.string();
                else if (k == "shape") {
                    c.expect('[');
                    if (!c.accept(']')) { do { info.shape.push_back(c.integer()); } while (c.accept(',')); c.expect(']'); }
                } else if (k == "data_offsets") {
                    c.expect('[');
                    do { info.data_offsets.push_back(size_t(c.integer())); } while (c.accept(','));
                    c.expect(']');
                } else c.skip_value();
            } while (c.accept(','));
            c.expect('}');
            if (info.data_offsets.size() != 2 || info.data_offsets[1] < info.data_offsets[0] || data_offset_ + info.data_offsets[1] > file_size)
                throw std::runtime_error("SafeTensorsReader: bad data_offsets for " + name);
            int64_t n = 1;
            for (int64_t d : info.shape) n *= d;
            if (size_t(n) * dtype_size(info.dtype) != info.data_offsets[1] - info.data_offsets[0])
                throw std::runtime_error("SafeTensorsReader: size mismatch for " + name);
            tensor_map_[name] = info;
            order_.push_back(name);
        } while (c.accept(','));
        c.expect('}');
    }
    parsed_ = true;
}

std::vector<std::string> SafeTensorsReader::get_tensor_names() const { return order_; }

SafeTensorInfo SafeTensorsReader::get_tensor_info(const std::string& name) const {
    auto it = tensor_map_.find(name);
    if (it == tensor_map_.end()) throw std::runtime_error("SafeTensorsReader: tensor not found: " + name);
    return it->second;
}

TensorPtr SafeTensorsReader::load_tensor(const std::string& name, bool transpose) {
    if (!parsed_) parse_header();
    return read_tensor_data(get_tensor_info(name), transpose);
}

TensorPtr SafeTensorsReader::read_tensor_data(const SafeTensorInfo& info, bool transpose) {
    int64_t n = 1;
    for (int64_t d : info.shape) n *= d;
    const size_t bytes = info.data_offsets[1] - info.data_offsets[0];
    std::vector<char> raw(bytes);
    file_.clear();
    file_.seekg(std::streamoff(data_offset_ + info.data_offsets[0]));
    file_.read(raw.data(), std::streamsize(bytes));
    if (!file_) throw std::runtime_error("SafeTensorsReader: short read in " + filepath_);
    std::vector<int64_t> shape = info.shape;
    const bool tr = transpose && shape.size() == 2;
    if (tr) std::swap(shape[0], shape[1]);
    // tensors are built from host buffers (copy-from-host constructor): no device memory is touched until a kernel needs them
    if (info.dtype == "I32") return std::make_shared<Tensor>(shape, static_cast<const void*>(raw.data()), kInt32, kCPU);
    std::vector<float> buf(static_cast<size_t>(n));
    float* dst = buf.data();
    auto at = [&](int64_t i) -> float {
        if (info.dtype == "F32") { float f; std::memcpy(&f, raw.data() + i * 4, 4); return f; }
        if (info.dtype == "F16") { uint16_t h; std::memcpy(&h, raw.data() + i * 2, 2); return half_to_float(h); }
        if (info.dtype == "BF16") { uint16_t h; std::memcpy(&h, raw.data() + i * 2, 2); uint32_t b = uint32_t(h) << 16; float f; std::memcpy(&f, &b, 4); return f; }
        if (info.dtype == "F64") { double d; std::memcpy(&d, raw.data() + i * 8, 8); return float(d); }
        if (info.dtype == "I64") { int64_t v; std::memcpy(&v, raw.data() + i * 8, 8); return float(v); }
        throw std::runtime_error("SafeTensorsReader: unsupported dtype " + info.dtype);
    };
    if (!tr) {
        if (info.dtype == "F32") std::memcpy(dst, raw.data(), bytes);
        else for (int64_t i = 0; i < n; ++i) dst[i] = at(i);
    } else {
        const int64_t R = info.shape[0], Ccols = info.shape[1];      // file [R, C] -> tensor [C, R]
        for (int64_t r = 0; r < R; ++r)
            for (int64_t c = 0; c < Ccols; ++c) dst[c * R + r] = at(r * Ccols + c);
    }
    return std::make_shared<Tensor>(shape, static_cast<const void*>(buf.data()), kFloat32, kCPU);
}

std::unordered_map<std::string, TensorPtr> SafeTensorsReader::load_tensors_mapped(
    const std::unordered_map<std::string, std::string>& key_mapping, const SafeTensorsLoadOptions& options) {
    if (!parsed_) parse_header();
    std::unordered_map<std::string, TensorPtr> out;
    for (const auto& kv : key_mapping) {
        const std::string& internal = kv.first;
        const std::string& hf = kv.second;
        auto it = tensor_map_.find(hf);
        if (it == tensor_map_.end() && hf.rfind("transformer.", 0) != 0) it = tensor_map_.find("transformer." + hf);
        if (it == tensor_map_.end()) {
            if (options.verbose) std::cerr << "[WARN] HF key not found: " << hf << std::endl;
            continue;
        }
        // graph/safetensors_loader.cpp:260-269: linear weights only (not norms, not embeddings), 2-D
        const bool embedding = hf.find("wte") != std::string::npos || hf.find("wpe") != std::string::npos || hf.find("embed_tokens") != std::string::npos;
        const bool transpose = options.transpose_linear && hf.find("weight") != std::string::npos && hf.find("ln") == std::string::npos &&
                               !embedding && it->second.shape.size() == 2;
        if (!options.auto_promote_fp16 && (it->second.dtype == "F16" || it->second.dtype == "BF16"))
            throw std::runtime_error("SafeTensorsReader: " + hf + " is " + it->second.dtype + " and auto_promote_fp16 is off");
        TensorPtr t = read_tensor_data(it->second, transpose);
        out[internal] = t;
        if (options.verbose) {
            std::cout << "[Loaded] " << internal << " <- " << hf << " shape=[";
            for (size_t i = 0; i < t->shape().size(); ++i) std::cout << (i ? ", " : "") << t->shape()[i];
            std::cout << "]" << (transpose ? " (transposed)" : "") << std::endl;
        }
    }
    return out;
}

// ---------------------------------------------------------------------------- key maps
std::unordered_map<std::string, std::string> GPT2KeyMapper::generate_gpt2_mapping(int num_layers) {
    std::unordered_map<std::string, std::string> m;
    m["wte.weight"] = "wte.weight";
    m["wpe.weight"] = "wpe.weight";
    m["ln_f.weight"] = "ln_f.weight";
    m["ln_f.bias"] = "ln_f.bias";
    static const char* const pairs[][2] = {{"ln_1", "ln_1"}, {"attn.qkv", "attn.c_attn"}, {"attn.proj", "attn.c_proj"},
                                           {"ln_2", "ln_2"}, {"mlp.fc_in", "mlp.c_fc"}, {"mlp.fc_out", "mlp.c_proj"}};
    for (int i = 0; i < num_layers; ++i)
        for (const auto& p : pairs)
            for (const char* wb : {"weight", "bias"})
                m["blocks." + std::to_string(i) + "." + p[0] + "." + wb] = "h." + std::to_string(i) + "." + p[1] + "." + wb;
    return m;
}

std::unordered_map<std::string, std::string> GemmaKeyMapper::generate_gemma_mapping(int num_layers) {
    std::unordered_map<std::string, std::string> m;
    m["embed_tokens.weight"] = "model.embed_tokens.weight";
    m["norm.weight"] = "model.norm.weight";
    m["lm_head.weight"] = "lm_head.weight";
    static const char* const names[] = {"input_layernorm.weight", "post_attention_layernorm.weight", "pre_feedforward_layernorm.weight",
                                        "post_feedforward_layernorm.weight", "self_attn.q_proj.weight", "self_attn.k_proj.weight",
                                        "self_attn.v_proj.weight", "self_attn.o_proj.weight", "self_attn.q_norm.weight",
                                        "self_attn.k_norm.weight", "mlp.gate_proj.weight", "mlp.up_proj.weight", "mlp.down_proj.weight"};
    for (int i = 0; i < num_layers; ++i)
        for (const char* n : names) m["layers." + std::to_string(i) + "." + n] = "model.layers." + std::to_string(i) + "." + n;
    return m;
}

static void print_map(const char* title, const std::unordered_map<std::string, std::string>& mapping) {
    std::vector<std::pair<std::string, std::string>> v(mapping.begin(), mapping.end());
    std::sort(v.begin(), v.end());
    std::cout << title << " (" << v.size() << " entries)\n";
    for (const auto& kv : v) std::cout << "  " << kv.first << " <- " << kv.second << "\n";
}
void GPT2KeyMapper::print_mapping(const std::unordered_map<std::string, std::string>& mapping) { print_map("[GPT2KeyMapper]", mapping); }
void GemmaKeyMapper::print_mapping(const std::unordered_map<std::string, std::string>& mapping) { print_map("[GemmaKeyMapper]", mapping); }

// ---------------------------------------------------------------------------- writer
void write_safetensors(const std::string& path, const std::map<std::string, TensorPtr>& tensors,
                       const std::vector<std::pair<std::string, std::string>>& metadata) {
    auto esc = [](const std::string& s) {
        std::string o;
        for (char ch : s) { if (ch == '"' || ch == '\\') o.push_back('\\'); o.push_back(ch); }
        return o;
    };
    std::ostringstream h;
    h << "{";
    size_t off = 0;
    bool first = true;
    for (const auto& kv : tensors) {       // std::map: sorted keys, data in the same order
        const TensorPtr& t = kv.second;
        if (!t) throw std::runtime_error("write_safetensors: null tensor " + kv.first);
        if (t->dtype() != kFloat32) throw std::runtime_error("write_safetensors: only fp32 tensors are written (" + kv.first + ")");
        const size_t nbytes = size_t(t->numel()) * sizeof(float);
        h << (first ? "" : ",") << "\"" << esc(kv.first) << "\":{\"dtype\":\"F32\",\"shape\":[";
        for (size_t i = 0; i < t->shape().size(); ++i) h << (i ? "," : "") << t->shape()[i];
        h << "],\"data_offsets\":[" << off << "," << off + nbytes << "]}";
        off += nbytes;
        first = false;
    }
    if (!metadata.empty()) {
        h << (first ? "" : ",") << "\"__metadata__\":{";
        for (size_t i = 0; i < metadata.size(); ++i) h << (i ? "," : "") << "\"" << esc(metadata[i].first) << "\":\"" << esc(metadata[i].second) << "\"";
        h << "}";
    }
    h << "}";
    std::string header = h.str();
    while (header.size() % 8) header.push_back(' ');    // keep the data section 8-byte aligned
    std::ofstream out(path, std::ios::binary);
    if (!out) throw std::runtime_error("write_safetensors: cannot open " + path);
    const uint64_t len = header.size();
    out.write(reinterpret_cast<const char*>(&len), 8);
    out.write(header.data(), std::streamsize(header.size()));
    for (const auto& kv : tensors) {
        const Tensor& t = *kv.second;
        out.write(reinterpret_cast<const char*>(t.data<float>()), std::streamsize(size_t(t.numel()) * sizeof(float)));   // host mirror (D2H if needed)
    }
    if (!out) throw std::runtime_error("write_safetensors: write failed for " + path);
}

}  // namespace ops

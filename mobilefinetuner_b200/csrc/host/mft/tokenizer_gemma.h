// Gemma tokenizer — API mirror of the reference's core/tokenizer_gemma.h:17-90 (GemmaTokenizerConfig, GemmaTokenizer).
// tokenizer.json (HuggingFace BPE: model.vocab object + model.merges array of pairs); a text is normalised by replacing
// every space with U+2581, merged as ONE word by lowest-rank BPE, pieces missing from the vocabulary fall back to
// <0xNN> byte tokens (core/tokenizer_gemma.cpp:283-417). Host-side input preparation only.
#pragma once
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

namespace ops {

struct hash_pair {      // core/tokenizer.h:18-24 (key type of the merge table)
    size_t operator()(const std::pair<std::string, std::string>& p) const {
        return std::hash<std::string>{}(p.first) ^ (std::hash<std::string>{}(p.second) << 1);
    }
};

struct GemmaTokenizerConfig {
    std::string tokenizer_json_path;
    bool add_bos_token = true;
    bool add_eos_token = false;
    std::string bos_token = "<bos>";
    std::string eos_token = "<eos>";
    std::string pad_token = "<pad>";
    std::string unk_token = "<unk>";
    std::vector<std::string> special_tokens = {"<pad>", "<eos>", "<bos>", "<unk>", "<start_of_image>", "<end_of_image>", "<image_soft_token>"};
    static GemmaTokenizerConfig from_pretrained(const std::string& model_dir);
};

class GemmaTokenizer {
public:
    explicit GemmaTokenizer(GemmaTokenizerConfig config);
    ~GemmaTokenizer() = default;
    void load();
    std::vector<int> encode(const std::string& text, bool add_special_tokens = true, int max_length = 0, bool truncation = true) const;
    std::string decode(const std::vector<int>& ids, bool skip_special_tokens = true) const;
    int get_bos_token_id() const { return bos_token_id_; }
    int get_eos_token_id() const { return eos_token_id_; }
    int get_pad_token_id() const { return pad_token_id_; }
    int get_unk_token_id() const { return unk_token_id_; }
    int get_vocab_size() const { return vocab_size_; }

private:
    std::vector<std::string> merge(const std::string& word) const;
    GemmaTokenizerConfig config_;
    std::unordered_map<std::string, int> vocab_;
    std::unordered_map<int, std::string> id_to_token_;
    std::unordered_map<std::pair<std::string, std::string>, int, hash_pair> ranks_;
    std::unordered_set<int> special_ids_;
    int bos_token_id_ = -1, eos_token_id_ = -1, pad_token_id_ = -1, unk_token_id_ = -1, vocab_size_ = 0;
};

}  // namespace ops

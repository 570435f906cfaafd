// GPT-2 byte-level BPE tokenizer — API mirror of the reference's core/tokenizer_bpe.h:22-145 (BPEConfig,
// GPT2BPETokenizer). Host-side text processing that feeds the training step (input contract); it reproduces
// the reference's tokenisation, including where that differs from the HuggingFace tokenizer: the pre-tokeniser runs on the
// byte-mapped text with ASCII classes only (core/tokenizer_bpe.cpp:256-285), so a mapped space never fuses with the word
// that follows it and '_' is dropped.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace ops {

struct BPEConfig {
    std::string vocab_path;            // vocab.json  {"token": id, ...}
    std::string merges_path;           // merges.txt  ("#version" line, then "left right" per line, rank = line order)
    std::string special_tokens_path;   // optional, not required
    int eos_token_id = 50256;
    int bos_token_id = 50256;
    int pad_token_id = 50256;
    int unk_token_id = 50256;
    bool add_prefix_space = false;
    BPEConfig() = default;
    static BPEConfig from_pretrained(const std::string& model_dir) {
        BPEConfig c;
        c.vocab_path = model_dir + "/vocab.json";
        c.merges_path = model_dir + "/merges.txt";
        c.special_tokens_path = model_dir + "/special_tokens_map.json";
        return c;
    }
};

class GPT2BPETokenizer {
public:
    explicit GPT2BPETokenizer(const BPEConfig& config);
    ~GPT2BPETokenizer() = default;
    void load();
    std::vector<int> encode(const std::string& text, bool add_special_tokens = false, int max_length = 0, bool truncation = true);
    std::string decode(const std::vector<int>& ids, bool skip_special_tokens = false);
    // returns (ids, attention masks); padding: "longest" | "max_length" | anything else = none
    std::pair<std::vector<std::vector<int>>, std::vector<std::vector<int>>> batch_encode(const std::vector<std::string>& texts, int max_length = 0,
                                                                                        const std::string& padding = "longest", bool truncation = true);
    int get_eos_token_id() const { return config_.eos_token_id; }
    int get_bos_token_id() const { return config_.bos_token_id; }
    int get_pad_token_id() const { return config_.pad_token_id; }
    int get_unk_token_id() const { return config_.unk_token_id; }
    int get_vocab_size() const { return vocab_size_; }
    std::string get_token_string(int token_id) const;

private:
    std::vector<std::string> pretokenize(const std::string& mapped) const;
    std::vector<std::string> merge(const std::string& piece) const;
    BPEConfig config_;
    std::string byte_to_text_[256];                          // byte -> UTF-8 of its printable stand-in
    std::unordered_map<std::string, uint8_t> text_to_byte_;
    std::unordered_map<std::string, int> vocab_;
    std::unordered_map<int, std::string> id_to_token_;
    std::unordered_map<std::string, int> ranks_;             // "left right" -> merge rank
    int vocab_size_ = 0;
    bool loaded_ = false;
};

}  // namespace ops

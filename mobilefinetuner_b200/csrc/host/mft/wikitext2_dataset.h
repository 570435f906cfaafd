// WikiText-2 batches — API mirror of the reference's data/wikitext2_dataset.h:16-160 (WT2Config, Split, Batch,
// WikiText2Dataset). This is the input contract of the training step: input_ids [B,S] int32,
// attention_mask [B,S] float32 (1/0), labels [B,S] int32 = ids with pad -> -100 (no pre-shift).
// The reference re-tokenises the raw file for every cache window in "streaming mode" to fit a phone's RAM; here the token
// stream is tokenised once and kept resident (a few MB), with the same stream length, chunk starts, shuffling and
// padding rules, so batches are identical for the same seed. Batches alias internal host buffers (valid until the next
// get_batch/next_batch), exactly like the reference's from_blob views.
#pragma once
#include <cstdint>
#include <functional>
#include <random>
#include <string>
#include <vector>

#include "tensor.h"
#include "tokenizer_bpe.h"

namespace ops {

struct WT2Config {
    std::string train_path, valid_path, test_path;
    std::string pretokenized_path;     // int32 token stream (.bin)
    std::string pretokenized_meta;     // its meta json (optional: looked up next to the .bin)
    int seq_len = 256;
    int stride = -1;                   // <= 0: seq_len
    int eos_id = 50256;
    int pad_id = 0;
    bool insert_eos_between_lines = true;
    bool drop_last = true;
    uint64_t seed = 2025;
    bool shuffle_train = true;
    bool streaming_mode = true;        // kept for the trainers' flags; selects the reference's stream-length / shuffle rules
    size_t max_cache_tokens = 100000;  // unused: the stream is resident
    float data_fraction = 1.0f;
};

enum class Split { Train, Valid, Test };

struct Batch {
    TensorPtr input_ids;
    TensorPtr attention_mask;
    TensorPtr labels;
};

class WikiText2Dataset {
public:
    WikiText2Dataset(const WT2Config& cfg, GPT2BPETokenizer* tok);
    WikiText2Dataset(const WT2Config& cfg, std::function<std::vector<int32_t>(const std::string&)> encode_fn);
    void load(Split split);
    size_t num_sequences() const;
    void shuffle();
    Batch get_batch(size_t index_start, size_t batch_size) const;
    Batch next_batch(size_t batch_size, bool need_loop = true);
    std::vector<int32_t> peek_tokens(size_t count) const;
    void reset_cursor();

private:
    void load_raw(Split split);
    void load_pretokenized(Split split);
    void build_starts(bool streaming_rule);
    WT2Config cfg_;
    std::function<std::vector<int32_t>(const std::string&)> encode_fn_;
    std::vector<int32_t> ids_;        // resident token stream of the loaded split
    size_t total_tokens_ = 0;         // logical stream length used for chunking (can exceed ids_.size() by the trailing EOS the
                                      // reference counts but never materialises in streaming mode)
    std::vector<size_t> starts_, order_;
    mutable size_t cursor_ = 0;
    mutable std::mt19937_64 rng_;
    mutable std::vector<int32_t> in_buf_, label_buf_;
    mutable std::vector<float> mask_buf_;
};

}  // namespace ops

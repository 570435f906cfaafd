// Host-side input/output components through the reference's public API only: tokenizer, dataset, safetensors reader,
// key maps, step-pacing policy. ONE source, two builds (like train_synth.cpp): against the reference library
// (oracle/_ref/ref_host_io) and against libmft_host (mobilefinetuner_b200/bin/host_io). Both print the same text when the
// components agree; nothing here launches a kernel, so the comparison runs without a GPU.
//   host_io <model_dir> <data_dir>
#include <algorithm>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "finetune_ops/core/tokenizer_bpe.h"
#include "finetune_ops/data/wikitext2_dataset.h"
#include "finetune_ops/graph/safetensors_loader.h"
#include "opt_ops/energy/power_monitor.h"

using namespace ops;

static void dump_batch(const char* tag, const Batch& b, int rows, int S) {
    const int32_t* ids = b.input_ids->data<int32_t>();
    const int32_t* lab = b.labels->data<int32_t>();
    const float* m = b.attention_mask->data<float>();
    long long si = 0, sl = 0;
    double sm = 0;
    for (int i = 0; i < rows * S; ++i) { si += ids[i]; sl += lab[i]; sm += m[i]; }
    std::printf("%s rows=%d sum_ids=%lld sum_labels=%lld sum_mask=%.1f first=", tag, rows, si, sl, sm);
    for (int i = 0; i < std::min(S, 12); ++i) std::printf("%d ", ids[i]);
    std::printf("\n");
}

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: host_io <model_dir> <data_dir>\n"); return 2; }
    const std::string model_dir = argv[1], data_dir = argv[2];
    // ---- tokenizer
    setvbuf(stdout, nullptr, _IOLBF, 0);
    GPT2BPETokenizer tok(BPEConfig::from_pretrained(model_dir));
    tok.load();
    std::printf("vocab %d eos %d\n", tok.get_vocab_size(), tok.get_eos_token_id());
    std::vector<std::string> samples = {"Hello world , it's the 1987 season .", "  two  spaces\tand_tab under_score don't we'll they've I'm he'd",
                                        "caf\xC3\xA9 na\xC3\xAFve \xE2\x82\xAC 100", ""};
    {
        std::ifstream f(data_dir + "/wiki.valid.raw");
        std::string line;
        int n = 0;
        while (std::getline(f, line) && n < 12) { if (line.size() > 3) { samples.push_back(line); ++n; } }
    }
    for (const auto& s : samples) {
        std::vector<int> ids = tok.encode(s, false, 0, false);
        std::printf("enc %zu:", ids.size());
        for (int id : ids) std::printf(" %d", id);
        std::printf("\n");
        std::string back = tok.decode(ids, false);
        std::printf("dec %zu %s\n", back.size(), back == s ? "same" : "differs");
    }
    auto be = tok.batch_encode({"the game", "the first season was released"}, 0, "longest", true);
    for (size_t i = 0; i < be.first.size(); ++i) {
        std::printf("batch %zu:", i);
        for (size_t j = 0; j < be.first[i].size(); ++j) std::printf(" %d/%d", be.first[i][j], be.second[i][j]);
        std::printf("\n");
    }
    // ---- dataset, both stream modes, fraction, valid split without drop_last
    for (int mode = 0; mode < 3; ++mode) {
        WT2Config c;
        c.train_path = data_dir + "/wiki.train.raw";
        c.valid_path = data_dir + "/wiki.valid.raw";
        c.test_path = data_dir + "/wiki.test.raw";
        c.seq_len = 16;
        c.seed = 99;
        c.streaming_mode = mode != 1;
        c.data_fraction = mode == 2 ? 0.37f : 1.0f;
        c.eos_id = tok.get_eos_token_id();
        WikiText2Dataset train(c, &tok);
        train.load(Split::Train);
        std::printf("mode %d train sequences %zu\n", mode, train.num_sequences());
        for (int k = 0; k < 3; ++k) {
            Batch b = train.next_batch(4, true);
            dump_batch("  train", b, int(b.input_ids->shape()[0]), 16);
        }
        WT2Config v = c;
        v.drop_last = false;
        v.stride = 8;
        WikiText2Dataset valid(v, &tok);
        valid.load(Split::Valid);
        std::printf("mode %d valid sequences %zu\n", mode, valid.num_sequences());
        Batch last = valid.get_batch(valid.num_sequences() - 2, 2);
        dump_batch("  valid-tail", last, 2, 16);
        size_t total = 0;
        valid.reset_cursor();
        for (;;) { Batch b = valid.next_batch(5, false); if (!b.input_ids) break; total += size_t(b.input_ids->shape()[0]); }
        std::printf("  valid epoch rows %zu\n", total);
    }
    // ---- safetensors
    SafeTensorsReader rd(model_dir + "/model.safetensors");
    rd.parse_header();
    std::vector<std::string> names = rd.get_tensor_names();
    names.erase(std::remove(names.begin(), names.end(), std::string("__metadata__")), names.end());   // the reference lists the metadata block as a name
    std::sort(names.begin(), names.end());
    std::printf("tensors %zu\n", names.size());
    for (size_t i = 0; i < names.size(); i += 7) {
        auto info = rd.get_tensor_info(names[i]);
        auto t = rd.load_tensor(names[i], false);
        double s = 0;
        const float* d = t->data<float>();
        for (int64_t k = 0; k < t->numel(); ++k) s += d[k];
        std::printf("  %s %s [", names[i].c_str(), info.dtype.c_str());
        for (auto x : info.shape) std::printf("%lld ", (long long)x);
        std::printf("] sum %.6e\n", s);
    }
    SafeTensorsLoadOptions opt;
    opt.verbose = false;
    opt.transpose_linear = true;
    auto mapped = rd.load_tensors_mapped(GPT2KeyMapper::generate_gpt2_mapping(2), opt);
    std::vector<std::string> mk;
    for (auto& kv : mapped) mk.push_back(kv.first);
    std::sort(mk.begin(), mk.end());
    for (auto& k : mk) {
        const auto& t = mapped[k];
        std::printf("  mapped %s [", k.c_str());
        for (auto x : t->shape()) std::printf("%lld ", (long long)x);
        std::printf("] first %.6e\n", double(t->data<float>()[t->numel() > 1 ? 1 : 0]));
    }
    std::printf("gemma map %zu\n", GemmaKeyMapper::generate_gemma_mapping(3).size());
    // ---- step pacing
    energy::PowerConfig pc;
    pc.check_interval_steps = 5;
    energy::PowerMonitor pm(pc);
    pm.set_manual_readings(15.0f, 50.0f);
    pm.set_step_schedule(energy::PowerMonitor::parse_schedule("3-4:70,bad,10-:5,8-9:x"));
    for (int s = 0; s < 14; ++s) std::printf("%d ", pm.suggest_sleep_ms(s));
    std::printf("\n");
    return 0;
}

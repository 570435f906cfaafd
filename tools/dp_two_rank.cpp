// Two (or N) data-parallel ranks driven through the C ABI only — no Python, no torch.distributed:
//   dp_two_rank [world=2] [steps=6] [full_finetune=0]
// With full_finetune=1 every tensor trains; MFT_DP_OVERLAP=1 MFT_DP_CHUNK_KB=64 then exercises the exchange that runs in
// chunks on a side stream while backward is still producing the rest of the gradients.
// The parent forks one process per rank; rank r selects GPU r, creates a trainer (small GPT-2 LoRA shape) and a communicator
// (mft_comm_init_from_file), and steps its rows of every global batch with mft_trainer_step_dp. Then the parent steps the SAME
// global batches in one process (world_size 1) and compares the per-step mean loss and every trainable tensor's checksum.
// Run with MFT_USE_TENSOR_CORES=0 for the exact fp32 mode, where only the summation order differs. Prints DP_C_ABI_OK.
#include <cuda_runtime.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "mft_trainer.h"

static int g_full = 0;
static mft_trainer_config make_cfg(int batch, int seq, int world, int rank) {
    mft_trainer_config c;
    std::memset(&c, 0, sizeof(c));
    c.model = 0; c.n_layer = 2; c.n_embd = 64; c.n_head = 4; c.vocab_size = 503; c.n_positions = 64;
    c.lora_rank = 8; c.lora_alpha = 16.0f; c.batch = batch; c.seq_len = seq;
    c.lr = 2e-4f; c.beta1 = 0.9f; c.beta2 = 0.999f; c.eps = 1e-8f; c.weight_decay = 0.0f; c.clip_norm = 1.0f;
    c.full_finetune = g_full;
    c.seed = 1234; c.world_size = world; c.rank = rank; c.arena_mb = 256; c.use_cuda_graph = 1;
    return c;
}

struct Result { std::vector<float> loss; std::vector<double> sumabs; };

static int write_result(const std::string& path, const Result& r) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return 1;
    size_t n = r.loss.size(), m = r.sumabs.size();
    std::fwrite(&n, sizeof(n), 1, f); std::fwrite(r.loss.data(), sizeof(float), n, f);
    std::fwrite(&m, sizeof(m), 1, f); std::fwrite(r.sumabs.data(), sizeof(double), m, f);
    std::fclose(f);
    return 0;
}
static bool read_result(const std::string& path, Result& r) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    size_t n = 0, m = 0;
    if (std::fread(&n, sizeof(n), 1, f) != 1) { std::fclose(f); return false; }
    r.loss.resize(n);
    if (std::fread(r.loss.data(), sizeof(float), n, f) != n || std::fread(&m, sizeof(m), 1, f) != 1) { std::fclose(f); return false; }
    r.sumabs.resize(m);
    const bool ok = std::fread(r.sumabs.data(), sizeof(double), m, f) == m;
    std::fclose(f);
    return ok;
}

static std::vector<int32_t> global_batches(int steps, int rows, int seq, int vocab) {
    std::mt19937 g(77);
    std::uniform_int_distribution<int> d(0, vocab - 1);
    std::vector<int32_t> v(size_t(steps) * rows * seq);
    for (auto& x : v) x = d(g);
    return v;
}

static int run_rank(int world, int rank, int steps, int per, int seq, const std::string& dir) {
    if (cudaSetDevice(rank) != cudaSuccess) { std::fprintf(stderr, "rank %d: cudaSetDevice failed\n", rank); return 2; }
    mft_trainer_config c = make_cfg(per, seq, world, rank);
    mft_trainer* t = nullptr;
    if (mft_trainer_create(&c, &t) != 0) { std::fprintf(stderr, "rank %d: %s\n", rank, mft_trainer_last_error()); return 2; }
    mft_comm* comm = nullptr;
    if (mft_comm_init_from_file(&comm, world, rank, (dir + "/nccl_id").c_str(), 120) != 0) { std::fprintf(stderr, "rank %d: %s\n", rank, mft_comm_last_error()); return 2; }
    auto ids = global_batches(steps, per * world, seq, c.vocab_size);
    int32_t* dev = nullptr;
    cudaMalloc(&dev, size_t(per) * seq * sizeof(int32_t));
    Result r;
    for (int s = 0; s < steps; ++s) {
        const int32_t* mine = ids.data() + (size_t(s) * per * world + size_t(rank) * per) * seq;
        cudaMemcpy(dev, mine, size_t(per) * seq * sizeof(int32_t), cudaMemcpyHostToDevice);
        if (mft_trainer_step_dp(t, comm, dev) != 0) { std::fprintf(stderr, "rank %d step %d: %s\n", rank, s, mft_trainer_last_error()); return 2; }
        mft_trainer_sync(t);
        float loss = 0.f;
        cudaMemcpy(&loss, mft_trainer_loss_ptr(t), sizeof(float), cudaMemcpyDeviceToHost);
        r.loss.push_back(loss);
    }
    for (int i = 0; i < mft_trainer_num_trainable(t); ++i) {
        double s = 0, sa = 0; int64_t n = 0;
        mft_trainer_param_checksum(t, i, &s, &sa, &n);
        r.sumabs.push_back(sa);
    }
    write_result(dir + "/rank" + std::to_string(rank), r);
    mft_comm_destroy(comm);
    mft_trainer_destroy(t);
    cudaFree(dev);
    return 0;
}

int main(int argc, char** argv) {
    const int world = argc > 1 ? std::atoi(argv[1]) : 2, steps = argc > 2 ? std::atoi(argv[2]) : 6;
    g_full = argc > 3 ? std::atoi(argv[3]) : 0;
    const int per = 2, seq = 16;
    int ngpu = 0;
    // (device count is queried in a child: the parent must not create a CUDA context before fork)
    char tmpl[] = "/tmp/mft_dp_XXXXXX";
    const char* dir = mkdtemp(tmpl);
    if (!dir) { std::perror("mkdtemp"); return 2; }
    std::vector<pid_t> kids;
    for (int r = 0; r < world; ++r) {
        pid_t p = fork();
        if (p == 0) {
            if (cudaGetDeviceCount(&ngpu) != cudaSuccess || ngpu < world) { std::fprintf(stderr, "rank %d: needs %d GPUs, found %d\n", r, world, ngpu); _exit(3); }
            _exit(run_rank(world, r, steps, per, seq, dir));
        }
        kids.push_back(p);
    }
    int bad = 0;
    for (pid_t p : kids) { int st = 0; waitpid(p, &st, 0); if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = WIFEXITED(st) ? WEXITSTATUS(st) : 2; }
    if (bad) { std::printf(bad == 3 ? "DP_C_ABI_SKIPPED (not enough GPUs)\n" : "DP_C_ABI_FAILED (a rank failed)\n"); return bad == 3 ? 0 : 1; }
    // single process on the whole batch (child too, so the parent stays context-free)
    pid_t p = fork();
    if (p == 0) {
        cudaSetDevice(0);
        mft_trainer_config c = make_cfg(per * world, seq, 1, 0);
        mft_trainer* t = nullptr;
        if (mft_trainer_create(&c, &t) != 0) { std::fprintf(stderr, "single: %s\n", mft_trainer_last_error()); _exit(2); }
        auto ids = global_batches(steps, per * world, seq, c.vocab_size);
        Result r;
        for (int s = 0; s < steps; ++s) {
            float loss = 0.f;
            if (mft_trainer_step(t, ids.data() + size_t(s) * per * world * seq, &loss) != 0) { std::fprintf(stderr, "single: %s\n", mft_trainer_last_error()); _exit(2); }
            r.loss.push_back(loss);
        }
        for (int i = 0; i < mft_trainer_num_trainable(t); ++i) {
            double s = 0, sa = 0; int64_t n = 0;
            mft_trainer_param_checksum(t, i, &s, &sa, &n);
            r.sumabs.push_back(sa);
        }
        write_result(std::string(dir) + "/single", r);
        mft_trainer_destroy(t);
        _exit(0);
    }
    int st = 0;
    waitpid(p, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) { std::printf("DP_C_ABI_FAILED (single-process run failed)\n"); return 1; }
    std::vector<Result> rr(world);
    Result one;
    bool ok = read_result(std::string(dir) + "/single", one);
    for (int r = 0; r < world; ++r) ok = ok && read_result(std::string(dir) + "/rank" + std::to_string(r), rr[r]);
    if (!ok) { std::printf("DP_C_ABI_FAILED (missing result files)\n"); return 1; }
    double worst_loss = 0, worst_param = 0;
    for (int s = 0; s < steps; ++s) {
        double mean = 0;
        for (int r = 0; r < world; ++r) mean += rr[r].loss[s];
        mean /= world;
        worst_loss = std::fmax(worst_loss, std::fabs(mean - one.loss[s]) / std::fabs(one.loss[s]));
    }
    for (size_t i = 0; i < one.sumabs.size(); ++i) {
        for (int r = 0; r < world; ++r) {
            if (rr[r].sumabs[i] != rr[0].sumabs[i]) { std::printf("DP_C_ABI_FAILED (replicas differ at tensor %zu)\n", i); return 1; }
        }
        worst_param = std::fmax(worst_param, std::fabs(rr[0].sumabs[i] - one.sumabs[i]) / std::fmax(one.sumabs[i], 1e-6));
    }
    const bool pass = worst_loss <= 1e-4 && worst_param <= 1e-3;
    std::printf("%s world %d steps %d full %d worst_loss_rel %.3g worst_param_rel %.3g nccl %d\n", pass ? "DP_C_ABI_OK" : "DP_C_ABI_FAILED", world, steps, g_full,
                worst_loss, worst_param, mft_comm_nccl_version());
    return pass ? 0 : 1;
}

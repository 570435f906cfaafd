/*
 * mft_trainer.h — C ABI of the whole training step (forward + shifted CE + backward + clip + Adam) for the
 * reference's trainers, exported by libmft_host.so. It is the loop body of
 *   gpt2_lora_finetune/main.cpp:561-601, gpt2_full_finetune/main.cpp:481-507 and
 *   operators/finetune_ops/optim/gemma_trainer.cpp:104-149
 * packaged behind plain pointers so a non-C++ caller (the benchmark, a Python or JNI front end) can drive it;
 * the C++ trainers keep using the ops::/GPT2Model/Adam classes directly.
 *
 * Data parallelism: one process per GPU owns one trainer (rank, world_size); every rank builds identical
 * weights from `seed`, feeds its shard of the batch, and the gradient exchange is ONE all-reduce of a flat
 * fp32 bucket between mft_trainer_forward_backward() and mft_trainer_update(). The all-reduce itself is the
 * caller's (NCCL through torch.distributed in bench.py): the bucket memory is supplied by the caller.
 */
#ifndef MFT_TRAINER_H_
#define MFT_TRAINER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mft_trainer mft_trainer;

typedef struct mft_trainer_config {
    int32_t model;            /* 0 = GPT-2 (graph/gpt2_model.h), 1 = Gemma-3 text (graph/gemma_model.h) */
    int32_t n_layer, n_embd, n_head, vocab_size, n_positions;
    int32_t intermediate, kv_heads, head_dim, sliding_window;        /* Gemma only */
    int32_t full_finetune;    /* 0 = LoRA adapters train, base frozen; 1 = every base parameter trains */
    int32_t lora_rank; float lora_alpha;
    int32_t lora_mlp;         /* GPT-2: also adapt mlp.fc_in / mlp.fc_out (default targets: attn.qkv + attn.proj) */
    int32_t batch, seq_len;   /* per-rank micro-batch */
    float lr, beta1, beta2, eps, weight_decay, clip_norm;
    uint32_t seed;            /* weights: mt19937(seed), N(0,0.02), LayerNorm gamma = 1 */
    int32_t world_size, rank;
    int64_t arena_mb;         /* per-step HBM arena (0 = default 24 GiB, sized for 180 GB parts) */
    int32_t use_cuda_graph;   /* capture the step once and replay it (shapes are static) */
} mft_trainer_config;

int mft_trainer_create(const mft_trainer_config* cfg, mft_trainer** out);
void mft_trainer_destroy(mft_trainer* t);
const char* mft_trainer_last_error(void);

/* Next synthetic batch of the trainer's own token stream (uniform ids from the mt19937 that initialised the
 * weights, exactly what oracle/harness/train_synth.cpp feeds the reference). host_ids: [batch*seq_len] int32. */
int mft_trainer_next_synthetic_batch(mft_trainer* t, int32_t* host_ids);

/* One full step from HOST buffers: H2D of ids (labels = ids), step, D2H of the loss (blocks until done). */
int mft_trainer_step(mft_trainer* t, const int32_t* host_ids, float* host_loss);
/* One full step with ids already in HBM; the loss stays in HBM (mft_trainer_loss_ptr). Does not synchronise. */
int mft_trainer_step_device(mft_trainer* t, const int32_t* dev_ids);

/* Split form for data parallelism. forward_backward leaves the gradients packed in the caller's bucket
 * (set once with mft_trainer_set_grad_bucket; at least mft_trainer_grad_elems floats); the caller all-reduces
 * (sum) the bucket; update() applies 1/world_size, the global clip and Adam. */
int64_t mft_trainer_grad_elems(mft_trainer* t);
int mft_trainer_set_grad_bucket(mft_trainer* t, float* dev_bucket, int64_t elems);
int mft_trainer_forward_backward(mft_trainer* t, const int32_t* dev_ids);
int mft_trainer_update(mft_trainer* t);

const float* mft_trainer_loss_ptr(mft_trainer* t);       /* device scalar: loss of the last step */
const float* mft_trainer_grad_norm_ptr(mft_trainer* t);  /* device scalar: pre-clip global gradient norm */
void* mft_trainer_stream(mft_trainer* t);                /* the cudaStream_t the step runs on */
int mft_trainer_sync(mft_trainer* t);
void mft_trainer_set_lr(mft_trainer* t, float lr);          /* written to a device word the update reads: also takes effect under graph replay */
/* With cfg.use_cuda_graph, each phase is captured after two eager runs and then replayed. This switch turns replay off/on
 * at run time (e.g. to time individual kernels with events); graphs already captured are kept. */
void mft_trainer_enable_graph(mft_trainer* t, int enable);

/* Operand precision of the large tensor-core GEMMs in the host layer (process-wide): 0 = TF32 (default), 1 = bf16
 * (fp32 accumulate either way). Also settable with MFT_GEMM_PRECISION=bf16|tf32 in the environment. */
void mft_host_set_gemm_precision(int bf16);
int mft_host_gemm_precision(void);

/* True-gradient mode (process-wide, off by default; MFT_TRUE_ATTN_GRAD=1 in the environment): GPT-2 attention gets a
 * backward, so attn.qkv adapters / weights, ln_1 and wpe receive gradients, and `.grad` accumulates over backward calls. The
 * default mode reproduces the reference, which cuts the graph at the qkv split (graph/gpt2_model.cpp:616-637), ships an
 * attention without backward (core/memory_efficient_attention.cpp:165-182) and overwrites `.grad` on every backward
 * (core/autograd_engine.cpp:243); the semantics matched here are pytorch_alignment/gpt2_lora_finetune.py:228-239. */
void mft_host_set_true_gradients(int enable);
int mft_host_true_gradients(void);
/* Gradient accumulation (true-gradient mode only): forward + backward of one micro-batch whose gradients are ADDED to the
 * sums kept since the last apply; mft_trainer_apply_accumulated() takes their mean, clips, runs Adam and clears them. */
int mft_trainer_accumulate(mft_trainer* t, const int32_t* dev_ids);
int mft_trainer_apply_accumulated(mft_trainer* t);

/* Checkpoint for resuming a run: every trainable tensor, both Adam moments and the per-tensor step counters, as one
 * safetensors file ("param.<i>", "adam.<i>.m", "adam.<i>.v", metadata "adam.<i>.step"). The reference saves adapters
 * (graph/lora_saver.cpp) but its optimizer state never round-trips (optim/adam.cpp:142-164 reads a header only); a run
 * resumed from this file continues bit for bit like the uninterrupted one. Both calls synchronise. */
int mft_trainer_save_state(mft_trainer* t, const char* path);
int mft_trainer_load_state(mft_trainer* t, const char* path);

/* ---- data-parallel exchange without Python: NCCL behind the C ABI (SURVEY §8b last row, §8e). One process per GPU with
 * its CUDA device already selected. Rank 0 makes the 128-byte unique id (mft_comm_unique_id) and passes it to the other ranks
 * by any side channel — mft_comm_init_from_file does it through a file that rank 0 writes and the others poll. The
 * all-reduce is in place, fp32, sum, enqueued on `stream`. libnccl.so.2 is loaded at run time (PyTorch's copy when the
 * process already carries one, else the system's). Every function returns 0 on success; mft_comm_last_error() explains. */
typedef struct mft_comm mft_comm;
const char* mft_comm_last_error(void);
int mft_comm_nccl_version(void);                               /* e.g. 22703; 0 when libnccl cannot be loaded */
int mft_comm_unique_id(void* out128);
int mft_comm_init_rank(mft_comm** out, int world_size, int rank, const void* unique_id_128);
int mft_comm_init_from_file(mft_comm** out, int world_size, int rank, const char* path, int timeout_s);
int mft_comm_world_size(mft_comm* c);
int mft_comm_rank(mft_comm* c);
int mft_comm_allreduce_sum(mft_comm* c, float* dev_buf, int64_t n, void* stream);
int mft_comm_destroy(mft_comm* c);
/* One data-parallel step on this rank's rows (ids in HBM): forward/backward + pack -> all-reduce of the gradient bucket
 * over `comm` on the trainer's stream -> 1/world, clip, Adam. The trainer must have been created with the communicator's
 * world_size / rank; it allocates its own bucket on first use unless mft_trainer_set_grad_bucket supplied one. */
int mft_trainer_step_dp(mft_trainer* t, mft_comm* comm, const int32_t* dev_ids);

/* Introspection for tests / reports. */
int32_t mft_trainer_num_trainable(mft_trainer* t);
int32_t mft_trainer_num_with_grad(mft_trainer* t);       /* tensors that received a gradient in the last step */
int64_t mft_trainer_kernel_launches(mft_trainer* t);     /* kernels per step (counted when the step last ran eagerly / was captured) */
int mft_trainer_param_checksum(mft_trainer* t, int32_t index, double* sum, double* sumabs, int64_t* n);
size_t mft_trainer_arena_high_water(mft_trainer* t);

#ifdef __cplusplus
}
#endif
#endif /* MFT_TRAINER_H_ */

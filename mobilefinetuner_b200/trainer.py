"""ctypes binding of include/mft_trainer.h (libmft_host.so): the reference trainers' step loop behind a C ABI.

Python only moves pointers: ids live in torch (pinned host / CUDA) tensors, the data-parallel gradient bucket is a
torch CUDA tensor all-reduced with torch.distributed (NCCL). All compute is in the CUDA library.
"""
import ctypes as C
import os

from . import capi  # loads libmft_b200.so first (RTLD_GLOBAL) — raises if it is missing

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "libmft_host.so")
if not os.path.exists(HOST_LIB_PATH):
    raise ImportError(f"{HOST_LIB_PATH} is missing: run `make -C mobilefinetuner_b200/csrc` (there is no CPU fallback)")
hostlib = C.CDLL(HOST_LIB_PATH, mode=C.RTLD_GLOBAL)

i32, i64, f32, u32, vp, sz = C.c_int32, C.c_int64, C.c_float, C.c_uint32, C.c_void_p, C.c_size_t


class TrainerConfig(C.Structure):
    _fields_ = [("model", i32), ("n_layer", i32), ("n_embd", i32), ("n_head", i32), ("vocab_size", i32), ("n_positions", i32),
                ("intermediate", i32), ("kv_heads", i32), ("head_dim", i32), ("sliding_window", i32),
                ("full_finetune", i32), ("lora_rank", i32), ("lora_alpha", f32), ("lora_mlp", i32),
                ("batch", i32), ("seq_len", i32),
                ("lr", f32), ("beta1", f32), ("beta2", f32), ("eps", f32), ("weight_decay", f32), ("clip_norm", f32),
                ("seed", u32), ("world_size", i32), ("rank", i32), ("arena_mb", i64), ("use_cuda_graph", i32)]


TRAINER_SIGNATURES = {
    "mft_trainer_create": (C.c_int, [C.POINTER(TrainerConfig), C.POINTER(vp)]),
    "mft_trainer_destroy": (None, [vp]),
    "mft_trainer_last_error": (C.c_char_p, []),
    "mft_trainer_next_synthetic_batch": (C.c_int, [vp, vp]),
    "mft_trainer_step": (C.c_int, [vp, vp, vp]),
    "mft_trainer_step_device": (C.c_int, [vp, vp]),
    "mft_trainer_grad_elems": (i64, [vp]),
    "mft_trainer_set_grad_bucket": (C.c_int, [vp, vp, i64]),
    "mft_trainer_forward_backward": (C.c_int, [vp, vp]),
    "mft_trainer_update": (C.c_int, [vp]),
    "mft_trainer_loss_ptr": (vp, [vp]),
    "mft_trainer_grad_norm_ptr": (vp, [vp]),
    "mft_trainer_stream": (vp, [vp]),
    "mft_trainer_sync": (C.c_int, [vp]),
    "mft_trainer_set_lr": (None, [vp, f32]),
    "mft_trainer_enable_graph": (None, [vp, C.c_int]),
    "mft_host_set_gemm_precision": (None, [C.c_int]),
    "mft_host_gemm_precision": (C.c_int, []),
    "mft_host_set_true_gradients": (None, [C.c_int]),
    "mft_host_true_gradients": (C.c_int, []),
    "mft_trainer_accumulate": (C.c_int, [vp, vp]),
    "mft_trainer_apply_accumulated": (C.c_int, [vp]),
    "mft_trainer_save_state": (C.c_int, [vp, C.c_char_p]),
    "mft_trainer_load_state": (C.c_int, [vp, C.c_char_p]),
    "mft_comm_last_error": (C.c_char_p, []),
    "mft_comm_nccl_version": (C.c_int, []),
    "mft_comm_unique_id": (C.c_int, [vp]),
    "mft_comm_init_rank": (C.c_int, [C.POINTER(vp), C.c_int, C.c_int, vp]),
    "mft_comm_init_from_file": (C.c_int, [C.POINTER(vp), C.c_int, C.c_int, C.c_char_p, C.c_int]),
    "mft_comm_world_size": (C.c_int, [vp]),
    "mft_comm_rank": (C.c_int, [vp]),
    "mft_comm_allreduce_sum": (C.c_int, [vp, vp, i64, vp]),
    "mft_comm_destroy": (C.c_int, [vp]),
    "mft_trainer_step_dp": (C.c_int, [vp, vp, vp]),
    "mft_trainer_num_trainable": (i32, [vp]),
    "mft_trainer_num_with_grad": (i32, [vp]),
    "mft_trainer_kernel_launches": (i64, [vp]),
    "mft_trainer_param_checksum": (C.c_int, [vp, i32, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(i64)]),
    "mft_trainer_arena_high_water": (sz, [vp]),
}
for _n, (_r, _a) in TRAINER_SIGNATURES.items():
    _f = getattr(hostlib, _n)
    _f.restype, _f.argtypes = _r, _a


class TrainerError(RuntimeError):
    pass


from .shapes import MODEL_SHAPES  # noqa: E402,F401  (pure-Python table; re-exported for callers of this module)


class Trainer:
    """One rank's training step. `step(host_ids)` is the end-to-end call (H2D, step, D2H of the loss)."""

    def __init__(self, batch, seq_len, shape="gpt2-small", full_finetune=False, lora_rank=8, lora_alpha=16.0, lora_mlp=False,
                 lr=2e-4, weight_decay=None, clip_norm=1.0, seed=1234, world_size=1, rank=0, arena_mb=0, cuda_graph=False,
                 **overrides):
        cfg = TrainerConfig()
        for k, v in dict(MODEL_SHAPES[shape] if isinstance(shape, str) else shape, **overrides).items():
            setattr(cfg, k, v)
        cfg.full_finetune = int(full_finetune)
        cfg.lora_rank, cfg.lora_alpha, cfg.lora_mlp = lora_rank, lora_alpha, int(lora_mlp)
        cfg.batch, cfg.seq_len = batch, seq_len
        cfg.lr, cfg.beta1, cfg.beta2, cfg.eps = lr, 0.9, 0.999, 1e-8
        cfg.weight_decay = (0.01 if full_finetune else 0.0) if weight_decay is None else weight_decay
        cfg.clip_norm, cfg.seed = clip_norm, seed
        cfg.world_size, cfg.rank, cfg.arena_mb = world_size, rank, arena_mb
        cfg.use_cuda_graph = int(cuda_graph)
        self.cfg = cfg
        self._h = vp()
        if hostlib.mft_trainer_create(C.byref(cfg), C.byref(self._h)) != 0:
            raise TrainerError(hostlib.mft_trainer_last_error().decode())
        self.tokens_per_step = batch * seq_len
        self._bucket = None

    def _ck(self, rc):
        if rc != 0:
            raise TrainerError(hostlib.mft_trainer_last_error().decode())

    def close(self):
        if self._h:
            hostlib.mft_trainer_destroy(self._h)
            self._h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data
    def next_synthetic_batch(self, out=None):
        import numpy as np
        out = np.empty((self.cfg.batch, self.cfg.seq_len), np.int32) if out is None else out
        self._ck(hostlib.mft_trainer_next_synthetic_batch(self._h, out.ctypes.data))
        return out

    # ---- steps
    def step(self, host_ids_ptr, host_loss_ptr):
        self._ck(hostlib.mft_trainer_step(self._h, host_ids_ptr, host_loss_ptr))

    def step_numpy(self, ids):
        import numpy as np
        ids = np.ascontiguousarray(ids, np.int32)
        loss = np.zeros(1, np.float32)
        self.step(ids.ctypes.data, loss.ctypes.data)
        return float(loss[0])

    def step_device(self, dev_ids_ptr):
        self._ck(hostlib.mft_trainer_step_device(self._h, dev_ids_ptr))

    def forward_backward(self, dev_ids_ptr):
        self._ck(hostlib.mft_trainer_forward_backward(self._h, dev_ids_ptr))

    def update(self):
        self._ck(hostlib.mft_trainer_update(self._h))

    def step_dp(self, comm, dev_ids_ptr):
        """mft_trainer_step_dp: forward/backward, NCCL exchange through `comm` (an mft_comm handle), update — all behind the C ABI."""
        self._ck(hostlib.mft_trainer_step_dp(self._h, comm, dev_ids_ptr))

    def accumulate(self, dev_ids_ptr):
        """True-gradient mode: forward + backward of one micro-batch, gradients added to the running sums."""
        self._ck(hostlib.mft_trainer_accumulate(self._h, dev_ids_ptr))

    def apply_accumulated(self):
        self._ck(hostlib.mft_trainer_apply_accumulated(self._h))

    def make_grad_bucket(self):
        import torch
        n = hostlib.mft_trainer_grad_elems(self._h)
        self._bucket = torch.zeros(n, dtype=torch.float32, device="cuda")
        self._ck(hostlib.mft_trainer_set_grad_bucket(self._h, self._bucket.data_ptr(), n))
        return self._bucket

    def save_state(self, path):
        """Trainable tensors + Adam moments + step counters -> one safetensors file."""
        self._ck(hostlib.mft_trainer_save_state(self._h, str(path).encode()))

    def load_state(self, path):
        self._ck(hostlib.mft_trainer_load_state(self._h, str(path).encode()))

    def set_lr(self, lr):
        hostlib.mft_trainer_set_lr(self._h, float(lr))

    def enable_graph(self, on):
        hostlib.mft_trainer_enable_graph(self._h, int(on))

    def sync(self):
        self._ck(hostlib.mft_trainer_sync(self._h))

    # ---- introspection
    @property
    def stream(self):
        return hostlib.mft_trainer_stream(self._h)

    def loss(self):
        """Synchronising read of the last step's loss (4-byte D2H)."""
        import torch
        self.sync()
        out = C.c_float()
        rt = C.CDLL("libcudart.so.12") if not hasattr(Trainer, "_rt") else Trainer._rt
        Trainer._rt = rt
        rt.cudaMemcpy(C.byref(out), vp(hostlib.mft_trainer_loss_ptr(self._h)), 4, 2)
        return out.value

    def grad_norm(self):
        self.sync()
        out = C.c_float()
        rt = C.CDLL("libcudart.so.12") if not hasattr(Trainer, "_rt") else Trainer._rt
        Trainer._rt = rt
        rt.cudaMemcpy(C.byref(out), vp(hostlib.mft_trainer_grad_norm_ptr(self._h)), 4, 2)
        return out.value

    loss_ptr = property(lambda self: hostlib.mft_trainer_loss_ptr(self._h))
    num_trainable = property(lambda self: hostlib.mft_trainer_num_trainable(self._h))
    num_with_grad = property(lambda self: hostlib.mft_trainer_num_with_grad(self._h))
    kernel_launches = property(lambda self: hostlib.mft_trainer_kernel_launches(self._h))
    arena_high_water = property(lambda self: hostlib.mft_trainer_arena_high_water(self._h))

    def param_checksum(self, i):
        s, sa, n = C.c_double(), C.c_double(), i64()
        self._ck(hostlib.mft_trainer_param_checksum(self._h, i, C.byref(s), C.byref(sa), C.byref(n)))
        return n.value, s.value, sa.value

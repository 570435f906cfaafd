"""ctypes binding of include/mft_b200.h (one Python function per C entry point, same names).

Arguments that are device pointers accept torch CUDA tensors (``.data_ptr()`` is passed) or raw ints.
Every wrapper raises :class:`MftError` on a non-zero status. Nothing here computes on the CPU.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MFT_B200_LIB: load another build of the same library (kernel-variant experiments, tools/build_variants.sh)
LIB_PATH = os.environ.get("MFT_B200_LIB") or os.path.join(_HERE, "libmft_b200.so")


class MftError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(make -C mobilefinetuner_b200/csrc). There is no CPU fallback.")

lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

i32, i64, f32, vp, sz = C.c_int32, C.c_int64, C.c_float, C.c_void_p, C.c_size_t


class GemmArgs(C.Structure):
    _fields_ = [("m", i32), ("n", i32), ("k", i32),
                ("a", vp), ("lda", i64), ("trans_a", i32),
                ("b", vp), ("ldb", i64), ("trans_b", i32),
                ("c", vp), ("ldc", i64),
                ("alpha", f32), ("beta", f32),
                ("bias", vp), ("gelu_out", vp),
                ("k2", i32), ("a2", vp), ("lda2", i64), ("b2", vp), ("ldb2", i64),
                ("batch", i32), ("stride_a", i64), ("stride_b", i64), ("stride_c", i64),
                ("addend", vp), ("ld_addend", i64), ("gelu_grad_x", vp), ("ld_gelu_grad_x", i64),
                ("gelu_out_bf16", vp), ("ld_gelu_out_bf16", i64), ("c_bf16", vp), ("ldc_bf16", i64)]


class TensorSlot(C.Structure):
    _fields_ = [("param", vp), ("grad", vp), ("m", vp), ("v", vp), ("n", i64),
                ("bias_corr1", f32), ("bias_corr2", f32), ("bucket_offset", i64)]


class LoraTnProblem(C.Structure):
    _fields_ = [("s", vp), ("lds", i64), ("w", vp), ("ldw", i64), ("partial", vp), ("out", vp), ("ldo", i64), ("T", i64), ("q", i64),
                ("r", i32), ("out_is_rq", i32), ("alpha", f32), ("use_tickets", i32), ("colblocks", i32), ("splits", i32), ("ticket_base", i32)]


class ShadowSlot(C.Structure):
    _fields_ = [("src", vp), ("rows", i64), ("cols", i64), ("ld_src", i64), ("plain", vp), ("ld_plain", i64), ("transposed", vp),
                ("ld_transposed", i64)]


# name -> (restype, argtypes); must list every symbol declared in include/mft_b200.h
SIGNATURES = {
    "mft_version": (C.c_int, []),
    "mft_launch_count": (i64, []),
    "mft_last_error": (C.c_char_p, []),
    "mft_device_check": (C.c_int, []),
    "mft_use_tensor_cores": (C.c_int, [C.c_int]),
    "mft_tensor_cores_enabled": (C.c_int, []),
    "mft_gemm": (C.c_int, [C.POINTER(GemmArgs), vp]),
    "mft_gemm_bf16": (C.c_int, [C.POINTER(GemmArgs), vp]),
    "mft_cast_bf16": (C.c_int, [vp, i64, vp, i64, i64, i64, vp]),
    "mft_transpose_cast_bf16": (C.c_int, [vp, i64, vp, i64, i64, i64, vp]),
    "mft_gemm_splitk_workspace_bytes": (sz, [i32, i32, i32, i32]),
    "mft_gemm_splitk": (C.c_int, [C.POINTER(GemmArgs), i32, vp, sz, vp]),
    "mft_lora_down": (C.c_int, [vp, i64, vp, i64, C.c_int, vp, i64, i64, i64, i32, f32, vp]),
    "mft_lora_down_ex": (C.c_int, [vp, i64, vp, i64, C.c_int, vp, i64, vp, i64, i64, i64, i32, f32, vp]),
    "mft_lora_down_bf16": (C.c_int, [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, i32, f32, vp]),
    "mft_lora_tn_workspace_bytes": (sz, [i64, i64, i32]),
    "mft_lora_tn": (C.c_int, [vp, i64, vp, i64, vp, i64, C.c_int, i64, i64, i32, f32, vp, sz, vp]),
    "mft_lora_tn_group_eligible": (C.c_int, [i64, i64, i32, vp, i64]),
    "mft_lora_tn_group_colblocks": (i32, [i64]),
    "mft_lora_tn_group_splits": (i32, [i64]),
    "mft_lora_tn_group_ticket_capacity": (i32, []),
    "mft_lora_tn_group": (C.c_int, [vp, i32, i32, i32, i32, i64, vp]),
    "mft_gemm_uses_tensor_cores": (C.c_int, [C.POINTER(GemmArgs)]),
    "mft_gemm_profile": (C.c_int, [C.c_int]),
    "mft_gemm_profile_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(i64)]),
    "mft_transpose2d": (C.c_int, [vp, i64, vp, i64, i64, i64, i64, i64, i64, vp]),
    "mft_permute0213": (C.c_int, [vp, vp, i64, i64, i64, i64, vp]),
    "mft_permute_nd": (C.c_int, [vp, vp, i32, C.POINTER(i64), C.POINTER(i64), vp]),
    "mft_cast_i32": (C.c_int, [vp, vp, i64, C.c_int, vp]),
    "mft_tril_ones": (C.c_int, [vp, i64, vp]),
    "mft_one_hot": (C.c_int, [vp, C.c_int, vp, i64, i64, vp]),
    "mft_binary": (C.c_int, [C.c_int, vp, vp, vp, i64, i64, i64, C.c_int, i64, vp]),
    "mft_affine": (C.c_int, [vp, vp, i64, f32, f32, vp]),
    "mft_accumulate": (C.c_int, [vp, vp, i64, vp]),
    "mft_fill": (C.c_int, [vp, i64, f32, vp]),
    "mft_colsum_workspace_bytes": (sz, [i64, i64]),
    "mft_colsum": (C.c_int, [vp, i64, vp, i64, i64, vp, sz, vp]),
    "mft_unary": (C.c_int, [C.c_int, vp, vp, i64, vp]),
    "mft_unary_param": (C.c_int, [C.c_int, vp, vp, i64, f32, f32, vp]),
    "mft_gelu_fwd": (C.c_int, [vp, vp, i64, vp]),
    "mft_gelu_bwd": (C.c_int, [vp, vp, vp, i64, vp]),
    "mft_gelu_bwd_ex": (C.c_int, [vp, vp, vp, vp, i64, vp]),
    "mft_geglu_fwd": (C.c_int, [vp, vp, vp, vp, i64, vp]),
    "mft_geglu_bwd": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, vp]),
    "mft_gelu_bwd_lowp": (C.c_int, [vp, C.c_int, vp, vp, vp, i64, vp]),
    "mft_cast_from_bf16": (C.c_int, [vp, i64, vp, i64, i64, i64, vp]),
    "mft_mul_bwd": (C.c_int, [vp, vp, vp, vp, vp, i64, vp]),
    "mft_layernorm_fwd": (C.c_int, [vp, vp, vp, vp, i64, i64, f32, vp]),
    "mft_norm_bwd_workspace_bytes": (sz, [i64, i64]),
    "mft_layernorm_bwd": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i64, f32, vp, sz, vp]),
    "mft_add_layernorm_fwd": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i64, f32, vp]),
    "mft_layernorm_bwd_add": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i64, f32, vp, sz, vp]),
    "mft_layernorm_fwd_ex": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i64, i64, f32, vp]),
    "mft_layernorm_bwd_ex": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, vp, vp, i64, i64, f32, vp, sz, vp]),
    "mft_layernorm_bwd_static": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i64, f32, vp]),
    "mft_rmsnorm_fwd": (C.c_int, [vp, vp, vp, i64, i64, f32, C.c_int, vp]),
    "mft_rmsnorm_fwd_ex": (C.c_int, [vp, vp, vp, vp, i64, i64, i64, f32, C.c_int, vp]),
    "mft_rmsnorm_bwd": (C.c_int, [vp, vp, vp, vp, vp, i64, i64, f32, C.c_int, vp, sz, vp]),
    "mft_rmsnorm_bwd_ex": (C.c_int, [vp, vp, vp, vp, vp, i64, vp, i64, i64, f32, C.c_int, vp, sz, vp]),
    "mft_softmax_fwd": (C.c_int, [vp, vp, i64, i64, vp]),
    "mft_softmax_bwd": (C.c_int, [vp, vp, vp, i64, i64, vp]),
    "mft_attention_precise": (C.c_int, [C.c_int]),
    "mft_attention_bf16": (C.c_int, [C.c_int]),
    "mft_attention_mode": (C.c_int, []),
    "mft_causal_attention_fwd": (C.c_int, [vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i64, i64, i64, i64,
                                           i64, i64, i64, f32, i32, vp]),
    "mft_causal_attention_fwd_ex": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i64, i64, i64, i64,
                                           i64, i64, i64, f32, i32, vp]),
    "mft_causal_attention_tc_fwd": (C.c_int, [vp, i64, vp, vp, i32, i32, i32, i32, f32, vp]),
    "mft_gqa_attention_fwd": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i64, i64, i64, i64, i64, i64, f32, i32, vp]),
    "mft_gqa_attention_bwd_workspace_bytes": (sz, [i32, i32, i32, i32, i32]),
    "mft_gqa_attention_bwd": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32] + [i64] * 15 + [f32, i32, vp, sz, vp]),
    "mft_qk_norm_rope_fwd": (C.c_int, [vp, vp, vp, i64, i32, vp, vp, vp, i64, i32, i32, i32, f32, f32, C.c_int, vp]),
    "mft_qk_norm_rope_bwd": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, i32, f32, f32, C.c_int, vp]),
    "mft_rope": (C.c_int, [vp, vp, i64, i64, i64, f32, C.c_int, vp]),
    "mft_repeat_kv_fwd": (C.c_int, [vp, vp, i64, i64, i64, i64, i64, vp]),
    "mft_repeat_kv_bwd": (C.c_int, [vp, vp, i64, i64, i64, i64, i64, vp]),
    "mft_embedding_fwd": (C.c_int, [vp, vp, vp, vp, i64, i64, i64, i64, f32, vp, vp]),
    "mft_lm_ce_workspace_bytes": (sz, [i64]),
    "mft_lm_ce_fwd": (C.c_int, [vp, i64, vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, sz, vp]),
    "mft_lm_ce_bwd": (C.c_int, [vp, i64, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, i64, vp]),
    "mft_lm_ce_bwd_ex": (C.c_int, [vp, i64, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, i64, vp, i64, vp]),
    "mft_refresh_bf16_shadows": (C.c_int, [vp, i32, i64, vp]),
    "mft_pack_grads": (C.c_int, [vp, i32, vp, vp]),
    "mft_grad_sumsq_workspace_bytes": (sz, []),
    "mft_grad_sumsq": (C.c_int, [vp, i32, vp, vp, sz, vp]),
    "mft_clip_scale": (C.c_int, [vp, f32, f32, vp, vp, vp]),
    "mft_scale_grads": (C.c_int, [vp, i32, vp, vp]),
    "mft_adam_step": (C.c_int, [vp, i32, f32, f32, f32, f32, f32, vp, vp]),
    "mft_adam_step_dev": (C.c_int, [vp, i32, vp, f32, f32, f32, f32, vp, vp, vp]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError here = the .so does not export a declared symbol
    _fn.restype = _res
    _fn.argtypes = _args


def ptr(t):
    """Device pointer of a torch tensor (or pass through ints / None)."""
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def check(status, what=""):
    if status != 0:
        msg = lib.mft_last_error()
        raise MftError(f"{what} failed with status {status}: {msg.decode() if msg else ''}")


def call(name, *args):
    """Call a status-returning entry point, converting tensors to pointers."""
    conv = [ptr(a) if hasattr(a, "data_ptr") else a for a in args]
    check(getattr(lib, name)(*conv), name)


def gemm(m, n, k, a, lda, trans_a, b, ldb, trans_b, c, ldc, alpha=1.0, beta=0.0, bias=None, gelu_out=None,
         k2=0, a2=None, lda2=0, b2=None, ldb2=0, batch=1, stride_a=0, stride_b=0, stride_c=0, stream=None):
    args = GemmArgs(m, n, k, ptr(a), lda, trans_a, ptr(b), ldb, trans_b, ptr(c), ldc, alpha, beta, ptr(bias),
                    ptr(gelu_out), k2, ptr(a2), lda2, ptr(b2), ldb2, batch, stride_a, stride_b, stride_c)
    check(lib.mft_gemm(C.byref(args), stream), "mft_gemm")
    return args

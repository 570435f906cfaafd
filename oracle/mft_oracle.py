"""CPU restatement (numpy) of the reference's hot-path operator math.

TEST INFRASTRUCTURE ONLY — nothing in the product imports this module. It is the checker used by
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline leg.

Parity status: PINNED. Every function here is checked (tests/test_oracle_golden.py) against
  * the reference's own known-answer test for this path (nn/test_lora_linear.cpp:157-180, both layouts), and
  * outputs of the reference itself, built unmodified from /root/reference by oracle/Makefile and dumped by
    oracle/harness/train_synth.cpp `ops` (fixtures under tests/golden/ops/, generator tests/golden/make_golden.sh).

All citations are relative to /root/reference/operators/finetune_ops/. Accumulation precision follows
the reference (float unless stated); summation ORDER is numpy's, so agreement is ~1e-6 relative, not bitwise.
"""
import numpy as np

F = np.float32


# ----------------------------------------------------------------------------- linear algebra
def matmul(a, b):
    """core/ops.cpp:486-651 — C = A·B over the last two dims; 2-D B shared over leading dims; if the inner
    dims disagree but A.shape[-1] == B.shape[-1], B is transposed first (:506-512)."""
    a = np.asarray(a, F)
    b = np.asarray(b, F)
    if a.shape[-1] != b.shape[-2] and a.shape[-1] == b.shape[-1]:
        b = np.swapaxes(b, -1, -2)
    return np.matmul(a, b).astype(F)


def matmul_bwd(a, b, gy):
    """core/backward_functions.cpp:94-138 — dA = dC·Bᵀ ; dB = Aᵀ·dC (flattened over leading dims for 2-D B)."""
    a, b, gy = np.asarray(a, F), np.asarray(b, F), np.asarray(gy, F)
    ga = np.matmul(gy, np.swapaxes(b, -1, -2)).astype(F)
    if b.ndim == 2 and a.ndim > 2:
        gb = a.reshape(-1, a.shape[-1]).T @ gy.reshape(-1, gy.shape[-1])
    else:
        gb = np.matmul(np.swapaxes(a, -1, -2), gy)
    return ga, gb.astype(F)


def lora_linear_fwd(x, W, bias, A, B, scale):
    """nn/lora_linear.cpp:47-108 — y = x·W (+b) + scale·(x·A)·B; A/B in either [in,r]/[r,out] or PEFT
    [r,in]/[out,r] layout (matmul auto-transposes)."""
    y = matmul(x, W)
    if bias is not None:
        y = y + np.asarray(bias, F)
    xa = matmul(x, A)
    return (y + matmul(xa, B) * F(scale)).astype(F)


def lora_linear_bwd(x, W, A, B, scale, gy, peft=False):
    """Chain of MatmulBackward/ScaleBackward/AddBackward the reference runs for LoRALinear
    (core/backward_functions.cpp:36-138,970-974). peft=False: A [in,r], B [r,out]; peft=True: A [r,in],
    B [out,r]. Returns (dx, dA, dB) with dA/dB in the layouts A/B were given in."""
    x, W, gy = np.asarray(x, F), np.asarray(W, F), np.asarray(gy, F)
    in_dim = x.shape[-1]
    A_io = np.asarray(A, F).T if peft else np.asarray(A, F)
    B_ro = np.asarray(B, F).T if peft else np.asarray(B, F)
    x2, g2 = x.reshape(-1, in_dim), gy.reshape(-1, gy.shape[-1])
    xa = x2 @ A_io
    gxab = g2 * F(scale)
    dB = xa.T @ gxab
    dxa = gxab @ B_ro.T
    dA = x2.T @ dxa
    dx = g2 @ W.T + dxa @ A_io.T
    if peft:
        dA, dB = dA.T, dB.T
    return dx.reshape(x.shape).astype(F), np.ascontiguousarray(dA, F), np.ascontiguousarray(dB, F)


def bias_add_bwd(gy, n):
    """AddBackward + sum_to_shape for a [n] bias: core/backward_functions.cpp:36-72."""
    return np.asarray(gy, F).reshape(-1, n).sum(axis=0, dtype=F)


# ----------------------------------------------------------------------------- norms / activations
def layer_norm_fwd(x, w, b, eps=1e-5):
    """core/ops.cpp:1404-1458 — biased variance, float accumulators."""
    x = np.asarray(x, F)
    mean = x.mean(axis=-1, keepdims=True, dtype=F)
    var = ((x - mean) ** 2).mean(axis=-1, keepdims=True, dtype=F)
    inv = (F(1) / np.sqrt(var + F(eps))).astype(F)
    return ((x - mean) * inv * np.asarray(w, F) + np.asarray(b, F)).astype(F)


def layer_norm_bwd(x, w, gy, eps=1e-5):
    """core/backward_functions.cpp:425-481 — returns (dx, dgamma, dbeta)."""
    x, w, gy = np.asarray(x, F), np.asarray(w, F), np.asarray(gy, F)
    D = F(x.shape[-1])
    mean = x.mean(axis=-1, keepdims=True, dtype=F)
    var = ((x - mean) ** 2).mean(axis=-1, keepdims=True, dtype=F)
    inv = (F(1) / np.sqrt(var + F(eps))).astype(F)
    xh = (x - mean) * inv
    gh = gy * w
    s1 = gh.sum(axis=-1, keepdims=True, dtype=F)
    s2 = (gh * xh).sum(axis=-1, keepdims=True, dtype=F)
    dx = (inv / D) * (D * gh - s1 - xh * s2)
    dg = (gy * xh).reshape(-1, x.shape[-1]).sum(axis=0, dtype=F)
    db = gy.reshape(-1, x.shape[-1]).sum(axis=0, dtype=F)
    return dx.astype(F), dg, db


def rms_norm_fwd(x, w, eps=1e-6, plus_one=True):
    """core/ops.cpp:1489-1537 — double Σx², rms cast to float, y = (x/rms)·(1+w)."""
    x = np.asarray(x, F)
    sq = (x.astype(np.float64) ** 2).sum(axis=-1, keepdims=True)
    rms = np.sqrt(sq / x.shape[-1] + np.float64(F(eps))).astype(F)
    sc = (F(1) + np.asarray(w, F)) if plus_one else np.asarray(w, F)
    return ((x / rms) * sc).astype(F)


def rms_norm_bwd(x, w, gy, eps=1e-6, plus_one=True):
    """core/backward_functions.cpp:483-523 — all in double, cast per element; returns (dx, dw)."""
    x64, g64 = np.asarray(x, F).astype(np.float64), np.asarray(gy, F).astype(np.float64)
    w64 = np.asarray(w, F).astype(np.float64)
    we = 1.0 + w64 if plus_one else w64
    D = x64.shape[-1]
    inv = 1.0 / np.sqrt((x64 ** 2).sum(axis=-1, keepdims=True) / D + np.float64(F(eps)))
    dot = (g64 * we * x64).sum(axis=-1, keepdims=True)
    dx = g64 * we * inv - x64 * (inv ** 3 / D) * dot
    dw = (g64 * x64 * inv).astype(F).reshape(-1, D).sum(axis=0, dtype=F)
    return dx.astype(F), dw


def gelu_fwd(x):
    """core/ops.cpp:1055-1059 (tanh form)."""
    x = np.asarray(x, F)
    t = np.tanh(F(0.7978845608) * (x + F(0.044715) * x * x * x))
    return (F(0.5) * x * (F(1) + t)).astype(F)


def gelu_bwd(x, gy):
    """core/backward_functions.cpp:155-177."""
    x = np.asarray(x, F)
    t = np.tanh(F(0.7978845608) * (x + F(0.044715) * x * x * x))
    d = F(0.5) * (F(1) + t) + F(0.5) * x * (F(1) - t * t) * F(0.7978845608) * (F(1) + F(3) * F(0.044715) * x * x)
    return (np.asarray(gy, F) * d).astype(F)


def softmax_fwd(x):
    """core/ops.cpp:1119-1152."""
    x = np.asarray(x, F)
    e = np.exp(x - x.max(axis=-1, keepdims=True))
    return (e / e.sum(axis=-1, keepdims=True, dtype=F)).astype(F)


def softmax_bwd(y, gy):
    """core/backward_functions.cpp:297-340 — row dot in double."""
    y, gy = np.asarray(y, F), np.asarray(gy, F)
    dot = (gy.astype(np.float64) * y.astype(np.float64)).sum(axis=-1, keepdims=True).astype(F)
    return (y * (gy - dot)).astype(F)


# ----------------------------------------------------------------------------- attention / rope / gqa
def causal_attention_fwd(q, k, v, scale, window=0):
    """core/memory_efficient_attention.cpp:92-162 — q,k,v [B,H,S,Hd]; j>i masked with -1e10; Z in double."""
    q, k, v = np.asarray(q, F), np.asarray(k, F), np.asarray(v, F)
    S = q.shape[-2]
    s = np.matmul(q, np.swapaxes(k, -1, -2)) * F(scale)
    i, j = np.arange(S)[:, None], np.arange(S)[None, :]
    mask = j > i
    if window > 0:
        mask = mask | (i - j >= window)
    s = np.where(mask, F(-1e10), s).astype(F)
    e = np.exp(s - s.max(axis=-1, keepdims=True))
    z = e.astype(np.float64).sum(axis=-1, keepdims=True)
    return np.matmul((e / z).astype(F), v).astype(F)


def rope(x, theta, inverse=False):
    """core/ops.cpp:2151-2208 — interleaved pairs (2d,2d+1), angle = pos / theta^(2d/Hd), float math;
    inverse = ApplyRoPEBackward (core/backward_functions.cpp:718-763)."""
    x = np.asarray(x, F)
    S, Hd = x.shape[-2], x.shape[-1]
    d = np.arange(Hd // 2, dtype=F)
    freq = (F(1) / np.power(F(theta), F(2) * d / F(Hd))).astype(F)
    ang = (np.arange(S, dtype=F)[:, None] * freq[None, :]).astype(F)
    c, s = np.cos(ang).astype(F), np.sin(ang).astype(F)
    if inverse:
        s = -s
    x0, x1 = x[..., 0::2], x[..., 1::2]
    y = np.empty_like(x)
    y[..., 0::2] = x0 * c - x1 * s
    y[..., 1::2] = x0 * s + x1 * c
    return y


def repeat_kv_fwd(x, rep):
    """core/ops.cpp:2072-2149 — head h reads kv head h // rep."""
    return np.repeat(np.asarray(x, F), rep, axis=1)


def repeat_kv_bwd(gy, rep):
    """core/backward_functions.cpp:683-716."""
    gy = np.asarray(gy, F)
    B, H, S, D = gy.shape
    return gy.reshape(B, H // rep, rep, S, D).sum(axis=2, dtype=F)


def gqa_attention_fwd_bwd(q, k, v, rep, scale, window=0, key_bias=None, dctx=None):
    """The Gemma attention chain as the reference composes it (graph/gemma_model.cpp:481-505) and its backward through the
    nodes that chain registers: repeat_kv_heads -> Q K^T -> * scale -> + mask [S,S] (0 / -1e10; causal, or causal restricted to
    i - j < window, :261-285) -> + key_bias [B,1,1,S] -> softmax -> P V; backward = MatmulBackward, SoftmaxBackward (row dot in
    double), ScaleBackward, RepeatKVHeadsBackward. q [B,H,S,Hd], k/v [B,KV,S,Hd]; returns ctx (and dq, dk, dv when dctx is given)."""
    q, k, v = np.asarray(q, F), np.asarray(k, F), np.asarray(v, F)
    S = q.shape[-2]
    kf, vf = repeat_kv_fwd(k, rep), repeat_kv_fwd(v, rep)
    s = (np.matmul(q, np.swapaxes(kf, -1, -2)) * F(scale)).astype(F)
    i, j = np.arange(S)[:, None], np.arange(S)[None, :]
    forbid = j > i
    if window > 0:
        forbid = forbid | (i - j >= window)
    s = (s + np.where(forbid, F(-1e10), F(0))[None, None]).astype(F)
    if key_bias is not None:
        s = (s + np.asarray(key_bias, F)[:, None, None, :]).astype(F)
    p = softmax_fwd(s)
    ctx = np.matmul(p, vf).astype(F)
    if dctx is None:
        return ctx
    dctx = np.asarray(dctx, F)
    dp = np.matmul(dctx, np.swapaxes(vf, -1, -2)).astype(F)
    dvf = np.matmul(np.swapaxes(p, -1, -2), dctx).astype(F)
    ds = (softmax_bwd(p, dp) * F(scale)).astype(F)
    dq = np.matmul(ds, kf).astype(F)
    dkf = np.matmul(np.swapaxes(ds, -1, -2), q).astype(F)
    return ctx, dq, repeat_kv_bwd(dkf, rep), repeat_kv_bwd(dvf, rep)


def qk_norm_rope_fwd(x, w, theta, eps=1e-6, do_norm=True):
    """rms_norm (1+w) then apply_rope on [B,H,S,Hd] (graph/gemma_model.cpp:383-420)."""
    return rope(rms_norm_fwd(x, w, eps) if do_norm else np.asarray(x, F), theta)


def qk_norm_rope_bwd(x, w, gy, theta, eps=1e-6, do_norm=True):
    """ApplyRoPEBackward then RMSNormBackward (dx only)."""
    dn = rope(gy, theta, inverse=True)
    return rms_norm_bwd(x, w, dn, eps)[0] if do_norm else dn


def embedding_gpt2(ids, wte, wpe):
    """graph/gpt2_model.cpp:346-381,500-528 — wte[id] + wpe[pos]."""
    ids = np.asarray(ids)
    return (np.asarray(wte, F)[ids] + np.asarray(wpe, F)[np.arange(ids.shape[1])][None]).astype(F)


# ----------------------------------------------------------------------------- loss
def _ce_rows(logits, labels, ignore_index):
    B, S, V = logits.shape
    tgt = np.full((B, S), -1, dtype=np.int64)
    tgt[:, :-1] = labels[:, 1:]
    valid = (tgt != ignore_index) & (tgt >= 0) & (tgt < V)
    valid[:, -1] = False
    return tgt, valid


def lm_cross_entropy_fwd(logits, labels, ignore_index=-100, reduction="mean"):
    """core/lm_loss.cpp:106-210 — internal shift, double logsumexp, mean over valid shifted labels.
    Returns (loss float32, valid_count)."""
    logits = np.asarray(logits, F)
    labels = np.asarray(labels)
    tgt, valid = _ce_rows(logits, labels, ignore_index)
    m = logits.max(axis=-1)
    denom = np.exp(logits - m[..., None]).astype(np.float64).sum(axis=-1)
    lse = m.astype(np.float64) + np.log(denom)
    picked = np.take_along_axis(logits, np.where(valid, tgt, 0)[..., None], axis=-1)[..., 0].astype(np.float64)
    nll = np.where(valid, lse - picked, 0.0)
    cnt = int(valid.sum())
    if reduction == "sum":
        return F(nll.sum()), cnt
    return (F(nll.sum() / cnt) if cnt > 0 else F(0)), cnt


def lm_cross_entropy_bwd(logits, labels, ignore_index=-100, reduction="mean", gout=1.0):
    """core/lm_loss.cpp:19-104 — (softmax - onehot)·gout/valid on valid rows, zero elsewhere."""
    logits = np.asarray(logits, F)
    labels = np.asarray(labels)
    tgt, valid = _ce_rows(logits, labels, ignore_index)
    cnt = int(valid.sum())
    p = softmax_fwd(logits)
    onehot = np.zeros_like(p)
    bi, si = np.nonzero(valid)
    onehot[bi, si, tgt[bi, si]] = 1
    scale = F(gout) if reduction == "sum" else (F(gout) / F(cnt) if cnt > 0 else F(0))
    return ((p - onehot) * scale * valid[..., None]).astype(F)


# ----------------------------------------------------------------------------- optimiser
def grad_norm(grads):
    """gpt2_lora_finetune/main.cpp:491-516 — sqrt of the double sum of squares over non-null grads."""
    return float(np.sqrt(sum((np.asarray(g, F).astype(np.float64) ** 2).sum() for g in grads if g is not None)))


def clip_scale(norm, max_norm):
    """scale = max/(norm+1e-6) if norm > max (same lines)."""
    return F(max_norm / (norm + 1e-6)) if (max_norm > 0 and norm > max_norm) else F(1)


def adam_step(p, g, m, v, t, lr, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.0):
    """optim/adam.cpp:25-90,222-228 — L2-coupled Adam, float math, bias correction via float pow.
    Returns (p, m, v); t is this tensor's 1-based step count."""
    p, g, m, v = (np.asarray(a, F).copy() for a in (p, g, m, v))
    lr, beta1, beta2, eps, wd = F(lr), F(beta1), F(beta2), F(eps), F(wd)
    if wd > 0:
        g = g + wd * p
    m = beta1 * m + (F(1) - beta1) * g
    v = beta2 * v + (F(1) - beta2) * g * g
    bc1 = F(1) - np.power(beta1, F(t), dtype=F)
    bc2 = F(1) - np.power(beta2, F(t), dtype=F)
    p = p - lr * (m / bc1) / (np.sqrt(v / bc2) + eps)
    return p.astype(F), m.astype(F), v.astype(F)

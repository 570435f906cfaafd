"""Model shapes of the BASELINE.json configs and their algorithmic FLOPs per token — pure Python.

Nothing here touches the native libraries, so tools that must not load them (bench.py's reference arm) can import it.
"""

# BASELINE.json configs: model shapes only; batch/seq are chosen by the caller.
MODEL_SHAPES = {
    "gpt2-small": dict(model=0, n_layer=12, n_embd=768, n_head=12, vocab_size=50257, n_positions=1024),
    "gpt2-medium": dict(model=0, n_layer=24, n_embd=1024, n_head=16, vocab_size=50257, n_positions=1024),
    "gemma3-270m": dict(model=1, n_layer=18, n_embd=640, n_head=4, vocab_size=262144, n_positions=32768,
                        intermediate=2048, kv_heads=1, head_dim=256, sliding_window=512),
}


def step_flops_per_token(shape, seq, full_finetune=False, rank=8, lora_targets=2):
    """Algorithmic FLOPs of one training step per token, with the reference's semantics:
    forward of every product; backward only of the products the reference's autograd graph reaches — for GPT-2 the
    attention is a stop-gradient, so nothing upstream of it (qkv, ln_1) has a backward. Attention counts the causal
    score matrix in full (4*S*C per token forward). Returns (forward, backward)."""
    s = MODEL_SHAPES[shape] if isinstance(shape, str) else shape
    L, C, V = s["n_layer"], s["n_embd"], s["vocab_size"]
    if s["model"] == 0:
        lin = 2 * C * 3 * C + 2 * C * C + 2 * 2 * C * 4 * C                  # qkv, proj, fc_in, fc_out
        attn = 4 * seq * C
        fwd = L * (lin + attn) + 2 * C * V
        if full_finetune:
            # proj dW; fc_in / fc_out dX + dW; lm_head dX + dE (wte only receives a gradient through the head)
            bwd = L * (2 * C * C + 2 * (2 * 2 * C * 4 * C)) + 2 * 2 * C * V
        else:
            lora_fwd = 2 * rank * (C + 3 * C) + 2 * rank * (C + C)           # x A, (xA) B on qkv and proj
            fwd += L * lora_fwd
            # lm_head dX; fc_out / fc_in dX; proj adapter: dxa, dA, dB
            bwd = 2 * C * V + L * (2 * 2 * C * 4 * C + 3 * 2 * rank * C)
        return float(fwd), float(bwd)
    # Gemma-3: q/k/v/o + gate/up/down with LoRA on all seven, attention with Hd*H wide heads, tied 262k head
    H, KV, Hd, I = s["n_head"], s["kv_heads"], s["head_dim"], s["intermediate"]
    QD, KD = H * Hd, KV * Hd
    lin = 2 * C * QD + 2 * 2 * C * KD + 2 * QD * C + 3 * 2 * C * I
    attn = 4 * seq * QD
    lora = 2 * rank * ((C + QD) + 2 * (C + KD) + (QD + C) + 2 * (C + I) + (I + C))
    fwd = L * (lin + attn + lora) + 2 * C * V
    bwd = L * (lin + 2 * attn + 3 * lora) + 2 * C * V                        # dX of every product, adapter grads; embedding frozen
    return float(fwd), float(bwd)

#!/usr/bin/env python
"""Synthetic model directories + WikiText-style corpora for the reference's own trainer CLIs.

No network, so nothing real can be downloaded: this writes random-init weights in the HuggingFace layouts the reference's
SafeTensorsReader / config scrapers / tokenizers expect, a small byte-level BPE vocabulary trained on the synthetic
corpus, and the corpus itself. Everything is deterministic in --seed.

  python tools/make_fixtures.py --out /tmp/fx --family gpt2 --n_layer 2 --n_embd 64 --n_head 4          # tiny, for tests
  python tools/make_fixtures.py --out /tmp/fx_small --family gpt2 --n_layer 12 --n_embd 768 --n_head 12  # GPT-2 small shape
  python tools/make_fixtures.py --out /tmp/fx_gemma --family gemma --n_layer 4 --n_embd 64

GPT-2 directory:  <out>/model/{config.json, model.safetensors, vocab.json, merges.txt}   <out>/data/wiki.{train,valid,test}.raw
Gemma directory:  <out>/model/{config.json, model.safetensors, tokenizer.json}            <out>/data/{tokens.bin, meta.json} (pretokenised)
"""
import argparse
import collections
import json
import os
import struct

import numpy as np

GPT2_EOS = 50256            # BPEConfig hard-codes it (core/tokenizer_bpe.h:27-30), so the vocabulary must reach it


def bytes_to_unicode():
    """The published GPT-2 byte alphabet (printable bytes map to themselves, the rest to 256+n)."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return {b: chr(c) for b, c in zip(bs, cs)}


def write_safetensors(path, tensors, metadata=None):
    header, off = {}, 0
    for k in sorted(tensors):
        t = np.ascontiguousarray(tensors[k], dtype=np.float32)
        header[k] = {"dtype": "F32", "shape": list(t.shape), "data_offsets": [off, off + t.nbytes]}
        off += t.nbytes
    if metadata:
        header["__metadata__"] = {k: str(v) for k, v in metadata.items()}
    blob = json.dumps(header, separators=(",", ":")).encode()
    blob += b" " * (-len(blob) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(blob)))
        f.write(blob)
        for k in sorted(tensors):
            f.write(np.ascontiguousarray(tensors[k], dtype=np.float32).tobytes())


def synth_corpus(rng, n_lines, lexicon):
    lines = []
    for i in range(n_lines):
        if i % 9 == 0:
            lines.append(" = " + " ".join(rng.choice(lexicon, 3)).title() + " = ")
            lines.append("")
            continue
        n = int(rng.integers(6, 40))
        words = list(rng.choice(lexicon, n))
        for j in range(2, n, 7):
            words[j] = words[j] + ","
        s = " ".join(words).capitalize() + " ."
        if i % 5 == 0:
            s = s.replace(" the ", " the 19" + str(int(rng.integers(10, 99))) + " ", 1)
        if i % 11 == 0:
            s = s.replace(" is ", " isn't ", 1) + " It's what they've said ."
        lines.append(" " + s + " ")
    return lines


def train_bpe(lines, n_merges):
    """Plain BPE on the byte-mapped words (the same pre-split the reference tokenizer applies: letters / digits / other)."""
    import re
    b2u = bytes_to_unicode()
    pat = re.compile(r"'s|'t|'re|'ve|'m|'ll|'d| ?[a-zA-Z]+| ?[0-9]+| ?[^\s\w]+|\s+(?!\S)|\s+")
    words = collections.Counter()
    for ln in lines:
        mapped = "".join(b2u[b] for b in ln.encode())
        for w in pat.findall(mapped):
            words[tuple(w)] += 1
    merges = []
    for _ in range(n_merges):
        pairs = collections.Counter()
        for w, c in words.items():
            for a, b in zip(w, w[1:]):
                pairs[(a, b)] += c
        if not pairs:
            break
        (a, b), cnt = max(pairs.items(), key=lambda kv: (kv[1], kv[0]))
        if cnt < 2:
            break
        merges.append((a, b))
        new = collections.Counter()
        for w, c in words.items():
            out, i = [], 0
            while i < len(w):
                if i + 1 < len(w) and w[i] == a and w[i + 1] == b:
                    out.append(a + b)
                    i += 2
                else:
                    out.append(w[i])
                    i += 1
            new[tuple(out)] += c
        words = new
    return merges


def make_gpt2(a, rng, out):
    model_dir, data_dir = os.path.join(out, "model"), os.path.join(out, "data")
    os.makedirs(model_dir, exist_ok=True)
    os.makedirs(data_dir, exist_ok=True)
    lexicon = np.array(["the", "of", "and", "in", "to", "a", "was", "is", "for", "as", "on", "with", "by", "that", "it", "from", "his", "at",
                        "were", "an", "which", "be", "this", "first", "also", "their", "had", "has", "one", "two", "game", "season", "river",
                        "album", "city", "film", "north", "south", "army", "road", "species", "century", "station", "church", "battle",
                        "storm", "song", "team", "school", "played", "released", "located", "during", "after", "between", "against"])
    for name, n in (("train", a.train_lines), ("valid", a.train_lines // 8 + 8), ("test", a.train_lines // 8 + 8)):
        with open(os.path.join(data_dir, f"wiki.{name}.raw"), "w") as f:
            f.write("\n".join(synth_corpus(rng, n, lexicon)) + "\n")
    train_lines = open(os.path.join(data_dir, "wiki.train.raw")).read().split("\n")
    b2u = bytes_to_unicode()
    merges = train_bpe(train_lines, a.merges)
    vocab = {}
    for b in range(256):
        vocab[b2u[b]] = len(vocab)
    for x, y in merges:
        if x + y not in vocab:
            vocab[x + y] = len(vocab)
    while len(vocab) < GPT2_EOS:                        # filler entries nobody produces: ids must reach the hard-coded EOS
        vocab[f"<|unused{len(vocab)}|>"] = len(vocab)
    vocab["<|endoftext|>"] = GPT2_EOS
    json.dump(vocab, open(os.path.join(model_dir, "vocab.json"), "w"), ensure_ascii=False)
    with open(os.path.join(model_dir, "merges.txt"), "w") as f:
        f.write("#version: 0.2\n")
        for x, y in merges:
            f.write(f"{x} {y}\n")
    V, C, L = GPT2_EOS + 1, a.n_embd, a.n_layer
    cfg = {"model_type": "gpt2", "vocab_size": V, "n_positions": a.n_positions, "n_ctx": a.n_positions, "n_embd": C, "n_layer": L,
           "n_head": a.n_head, "layer_norm_epsilon": 1e-5}
    json.dump(cfg, open(os.path.join(model_dir, "config.json"), "w"), indent=1)
    n = lambda *s: (rng.standard_normal(s) * 0.02).astype(np.float32)
    t = {"wte.weight": n(V, C), "wpe.weight": n(a.n_positions, C), "ln_f.weight": np.ones(C, np.float32), "ln_f.bias": n(C)}
    for i in range(L):
        p = f"h.{i}."
        t[p + "ln_1.weight"], t[p + "ln_1.bias"] = np.ones(C, np.float32), n(C)
        t[p + "ln_2.weight"], t[p + "ln_2.bias"] = np.ones(C, np.float32), n(C)
        t[p + "attn.c_attn.weight"], t[p + "attn.c_attn.bias"] = n(C, 3 * C), n(3 * C)      # Conv1D layout [in, out]
        t[p + "attn.c_proj.weight"], t[p + "attn.c_proj.bias"] = n(C, C), n(C)
        t[p + "mlp.c_fc.weight"], t[p + "mlp.c_fc.bias"] = n(C, 4 * C), n(4 * C)
        t[p + "mlp.c_proj.weight"], t[p + "mlp.c_proj.bias"] = n(4 * C, C), n(C)
    write_safetensors(os.path.join(model_dir, "model.safetensors"), t, {"format": "pt"})
    print(f"gpt2 fixture: {model_dir} (V={V} C={C} L={L}, {len(merges)} merges), corpus in {data_dir}")


def make_gemma(a, rng, out):
    model_dir, data_dir = os.path.join(out, "model"), os.path.join(out, "data")
    os.makedirs(model_dir, exist_ok=True)
    os.makedirs(data_dir, exist_ok=True)
    V, C, L = a.vocab_size, a.n_embd, a.n_layer
    H, KV, Hd, I = a.n_head, a.kv_heads, a.head_dim, a.intermediate
    layer_types = ["full_attention" if (i + 1) % 6 == 0 else "sliding_attention" for i in range(L)]
    cfg = {"model_type": "gemma3_text", "vocab_size": V, "hidden_size": C, "intermediate_size": I, "num_hidden_layers": L,
           "num_attention_heads": H, "num_key_value_heads": KV, "head_dim": Hd, "max_position_embeddings": a.n_positions,
           "rms_norm_eps": 1e-6, "rope_theta": 1000000.0, "rope_local_base_freq": 10000.0, "sliding_window": 512,
           "query_pre_attn_scalar": Hd, "layer_types": layer_types, "bos_token_id": 2, "eos_token_id": 1, "pad_token_id": 0}
    json.dump(cfg, open(os.path.join(model_dir, "config.json"), "w"), indent=1)
    n = lambda *s: (rng.standard_normal(s) * 0.02).astype(np.float32)
    t = {"model.embed_tokens.weight": n(V, C), "model.norm.weight": n(C)}
    for i in range(L):
        p = f"model.layers.{i}."
        for k in ("input_layernorm", "post_attention_layernorm", "pre_feedforward_layernorm", "post_feedforward_layernorm"):
            t[p + k + ".weight"] = n(C)
        t[p + "self_attn.q_proj.weight"], t[p + "self_attn.k_proj.weight"] = n(H * Hd, C), n(KV * Hd, C)     # HF layout [out, in]
        t[p + "self_attn.v_proj.weight"], t[p + "self_attn.o_proj.weight"] = n(KV * Hd, C), n(C, H * Hd)
        t[p + "self_attn.q_norm.weight"], t[p + "self_attn.k_norm.weight"] = n(Hd), n(Hd)
        t[p + "mlp.gate_proj.weight"], t[p + "mlp.up_proj.weight"], t[p + "mlp.down_proj.weight"] = n(I, C), n(I, C), n(C, I)
    write_safetensors(os.path.join(model_dir, "model.safetensors"), t, {"format": "pt"})
    # pretokenised stream (train_lora_gemma --pretokenized_path): zipf-ish ids, EOS between "lines"
    total = a.tokens
    ids = (rng.zipf(1.3, total) % (V - 8) + 8).astype(np.int32)
    ids[rng.integers(0, total, total // 40)] = 1
    ids.tofile(os.path.join(data_dir, "tokens.bin"))
    n_train = int(total * 0.8)
    n_valid = (total - n_train) // 2
    meta = {"total_tokens": total, "vocab_size": V, "eos_token_id": 1, "pad_token_id": 0, "bos_token_id": 2, "unk_token_id": 3,
            "insert_eos_between_lines": True,
            "splits": {"train": {"offset": 0, "length": n_train}, "valid": {"offset": n_train, "length": n_valid},
                       "test": {"offset": n_train + n_valid, "length": total - n_train - n_valid}}}
    json.dump(meta, open(os.path.join(data_dir, "meta.json"), "w"), indent=1)
    # minimal tokenizer.json so GemmaTokenizer::load() succeeds when a CLI insists on it
    vocab = {"<pad>": 0, "<eos>": 1, "<bos>": 2, "<unk>": 3}
    for b in range(256):
        vocab[f"<0x{b:02X}>"] = len(vocab)
    for w in ("▁", "▁the", "▁of", "e", "t", "a", "o", "n", "i", "s", "r", "h"):
        vocab.setdefault(w, len(vocab))
    tok = {"version": "1.0", "model": {"type": "BPE", "vocab": vocab, "merges": []},
           "added_tokens": [{"id": i, "content": c, "special": True} for c, i in list(vocab.items())[:4]]}
    json.dump(tok, open(os.path.join(model_dir, "tokenizer.json"), "w"), ensure_ascii=False)
    print(f"gemma fixture: {model_dir} (V={V} C={C} L={L}), {total} pretokenised tokens in {data_dir}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--family", default="gpt2", choices=["gpt2", "gemma"])
    ap.add_argument("--n_layer", type=int, default=2)
    ap.add_argument("--n_embd", type=int, default=64)
    ap.add_argument("--n_head", type=int, default=4)
    ap.add_argument("--n_positions", type=int, default=1024)
    ap.add_argument("--kv_heads", type=int, default=1)
    ap.add_argument("--head_dim", type=int, default=32)
    ap.add_argument("--intermediate", type=int, default=128)
    ap.add_argument("--vocab_size", type=int, default=1009, help="gemma only (gpt2 is pinned to 50257 by the tokenizer's EOS id)")
    ap.add_argument("--train_lines", type=int, default=400)
    ap.add_argument("--merges", type=int, default=300)
    ap.add_argument("--tokens", type=int, default=60000, help="gemma pretokenised stream length")
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    (make_gpt2 if a.family == "gpt2" else make_gemma)(a, rng, a.out)


if __name__ == "__main__":
    main()

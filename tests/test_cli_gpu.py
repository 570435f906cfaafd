"""The reference's own trainer CLIs, end to end.

`mobilefinetuner_b200/bin/<cli>` is the reference's main.cpp compiled from its source and linked against libmft_host (the C++
mirror over the CUDA kernels); `oracle/_ref/ref_<cli>` is the same source linked against the reference's own CPU library.
Both read the same synthetic model directory (random-init safetensors, BPE vocabulary, WikiText-style corpus written by
tools/make_fixtures.py), so tokenisation, batching, LoRA injection, the step itself, the LR schedule and the adapter
file must all agree. Neither binary nor /root/reference is needed at test time beyond these prebuilt files."""
import json
import os
import re
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "mobilefinetuner_b200", "bin")
REF = os.path.join(ROOT, "oracle", "_ref")
pytestmark = pytest.mark.gpu

LINE = re.compile(r"\(global (\d+)/\d+\) \| lr (\S+) \| loss (\S+) \| ppl \S+ \| grad_norm (\S+)")


def fixture(tmp_path, **kw):
    out = str(tmp_path / "fx")
    args = [sys.executable, os.path.join(ROOT, "tools", "make_fixtures.py"), "--out", out, "--family", "gpt2"]
    for k, v in kw.items():
        args += [f"--{k}", str(v)]
    subprocess.check_call(args, stdout=subprocess.DEVNULL)
    return out


def run_cli(exe, args, env=None, timeout=900):
    r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rows = [(int(m[1]), float(m[2]), float(m[3]), float(m[4])) for m in LINE.finditer(r.stdout)]
    return rows, r.stdout


def read_safetensors(path):
    with open(path, "rb") as f:
        n = struct.unpack("<Q", f.read(8))[0]
        header = json.loads(f.read(n))
        data = f.read()
    meta = header.pop("__metadata__", {})
    return {k: np.frombuffer(data[v["data_offsets"][0]:v["data_offsets"][1]], np.float32).reshape(v["shape"]) for k, v in header.items()}, meta


def need(*paths):
    for p in paths:
        if not os.path.exists(p):
            pytest.skip(f"{os.path.relpath(p, ROOT)} not built (needs /root/reference at build time)")


def test_gpt2_lora_finetune_cli_matches_reference_build(tmp_path):
    mine, ref = os.path.join(BIN, "gpt2_lora_finetune"), os.path.join(REF, "ref_gpt2_lora_finetune")
    need(mine, ref)
    fx = fixture(tmp_path, n_layer=2, n_embd=64, n_head=4, n_positions=128)
    common = ["--data_dir", fx + "/data", "--pretrained_dir", fx + "/model", "--steps", "12", "--batch_size", "2", "--seq_len", "32", "--lr", "1e-3",
              "--seed", "7", "--warmup_steps", "2"]
    rows_ref, _ = run_cli(ref, common + ["--lora_out", fx + "/ref.safetensors"])
    env = dict(os.environ, MFT_USE_TENSOR_CORES="0")          # exact fp32 mode: only summation order differs
    rows, out = run_cli(mine, common + ["--lora_out", fx + "/mine.safetensors"], env=env)
    assert len(rows) == len(rows_ref) == 12, out[-2000:]
    for a, b in zip(rows, rows_ref):
        assert a[0] == b[0] and a[1] == b[1]                                   # same step numbering and LR schedule
        assert abs(a[2] - b[2]) <= 2e-4 * abs(b[2]), (a, b)                    # loss (printed with 4 decimals)
        assert abs(a[3] - b[3]) <= 2e-3 * max(b[3], 1e-3) + 1e-3, (a, b)       # grad norm (3 decimals)
    ta, ma = read_safetensors(fx + "/mine.safetensors")
    tb, mb = read_safetensors(fx + "/ref.safetensors")
    assert sorted(ta) == sorted(tb) and ma == mb
    for k in tb:
        assert ta[k].shape == tb[k].shape
        assert np.abs(ta[k] - tb[k]).max() <= 1e-4 * max(np.abs(tb[k]).max(), 1e-6) + 1e-7, k
    # default (TF32 tensor-core) mode stays within the 1e-3 loss tolerance
    rows_tc, _ = run_cli(mine, common + ["--lora_out", fx + "/mine_tc.safetensors"])
    for a, b in zip(rows_tc, rows_ref):
        assert abs(a[2] - b[2]) <= 1e-3 * abs(b[2]), (a, b)


def test_gpt2_lora_cli_resume_from_saved_adapter(tmp_path):
    mine = os.path.join(BIN, "gpt2_lora_finetune")
    need(mine)
    fx = fixture(tmp_path, n_layer=2, n_embd=64, n_head=4, n_positions=128)
    common = ["--data_dir", fx + "/data", "--pretrained_dir", fx + "/model", "--batch_size", "2", "--seq_len", "32", "--lr", "1e-3", "--seed", "7"]
    run_cli(mine, common + ["--steps", "4", "--lora_out", fx + "/a.safetensors"])
    rows, out = run_cli(mine, common + ["--steps", "2", "--resume_from", fx + "/a.safetensors", "--lora_out", fx + "/b.safetensors"])
    assert len(rows) == 2 and "Loaded" in out
    ta, _ = read_safetensors(fx + "/a.safetensors")
    tb, _ = read_safetensors(fx + "/b.safetensors")
    assert sorted(ta) == sorted(tb)
    assert any(np.abs(ta[k] - tb[k]).max() > 0 for k in ta)            # training continued from the loaded factors


def test_gpt2_full_finetune_cli_matches_reference_build(tmp_path):
    mine, ref = os.path.join(BIN, "gpt2_full_finetune"), os.path.join(REF, "ref_gpt2_full_finetune")
    need(mine, ref)
    fx = fixture(tmp_path, n_layer=2, n_embd=64, n_head=4, n_positions=128)
    common = ["--data_dir", fx + "/data", "--pretrained_dir", fx + "/model", "--steps", "8", "--batch_size", "2", "--seq_len", "32", "--lr", "1e-4", "--seed", "7"]
    extra = lambda name: ["--output_path", fx + "/" + name]
    rows_ref, _ = run_cli(ref, common + extra("ref_full.safetensors"))
    rows, out = run_cli(mine, common + extra("mine_full.safetensors"), env=dict(os.environ, MFT_USE_TENSOR_CORES="0"))
    assert len(rows) == len(rows_ref) and rows, out[-2000:]
    for a, b in zip(rows, rows_ref):
        assert abs(a[2] - b[2]) <= 2e-4 * abs(b[2]), (a, b)


def test_train_lora_gemma_cli_matches_reference_build(tmp_path):
    """train_lora_gemma.cpp (unchanged, relative includes resolved through a symlink tree) + this library's GemmaLoRATrainer,
    pretokenised data: loss per step, LR schedule and the saved attention adapters against the reference build."""
    mine, ref = os.path.join(BIN, "train_lora_gemma"), os.path.join(REF, "ref_train_lora_gemma")
    need(mine, ref)
    out = str(tmp_path / "fxg")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_fixtures.py"), "--out", out, "--family", "gemma", "--n_layer", "4",
                           "--n_embd", "64", "--n_head", "4", "--head_dim", "32", "--intermediate", "128", "--n_positions", "256"], stdout=subprocess.DEVNULL)
    step = re.compile(r"\[Step (\d+)\] Loss=(\S+) PPL=\S+ LR=(\S+)")

    def run(exe, outdir, env=None):
        os.makedirs(outdir, exist_ok=True)
        r = subprocess.run([exe, "--model_dir", out + "/model", "--pretokenized_path", out + "/data/tokens.bin", "--pretokenized_meta",
                            out + "/data/meta.json", "--output_dir", outdir, "--seq_len", "32", "--batch", "2", "--epochs", "1", "--max_steps", "10",
                            "--lr", "1e-3", "--targets", "full"], capture_output=True, text=True, timeout=900, env=env)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        return [(int(m[1]), float(m[2]), float(m[3])) for m in step.finditer(r.stdout)], r.stdout

    rows_ref, _ = run(ref, out + "/ref_out")
    rows, log = run(mine, out + "/mine_out", env=dict(os.environ, MFT_USE_TENSOR_CORES="0"))
    assert len(rows) == len(rows_ref) == 10, log[-2000:]
    for a, b in zip(rows, rows_ref):
        assert a[0] == b[0] and abs(a[2] - b[2]) <= 1e-6 * abs(b[2]) + 1e-12
        assert abs(a[1] - b[1]) <= 2e-4 * abs(b[1]), (a, b)
    ta, ma = read_safetensors(out + "/mine_out/gemma_lora.safetensors")
    tb, mb = read_safetensors(out + "/ref_out/gemma_lora.safetensors")
    assert sorted(ta) == sorted(tb) and ma == mb and len(tb) == 4 * 4 * 2       # q, k, v, o of 4 layers, A and B
    for k in tb:
        # Adam divides by sqrt(v): entries whose gradients are ~1e-8 amplify summation-order noise, so the bound is relative to the
        # tensor's scale (B starts at zero and is ~2e-3 after ten warm-up steps)
        assert np.abs(ta[k] - tb[k]).max() <= 2e-3 * max(np.abs(tb[k]).max(), 1e-6) + 1e-7, k
    rows_tc, _ = run(mine, out + "/mine_tc")
    for a, b in zip(rows_tc, rows_ref):
        assert abs(a[1] - b[1]) <= 1e-3 * abs(b[1]), (a, b)


def test_eval_ppl_cli_matches_reference_build(tmp_path):
    """Evaluation forward path: eval_ppl.cpp unchanged, on a trained adapter — plain, attached and merged
    into the base weights (LoRALinear::merge_to_base) — against the reference build on the same files."""
    mine, ref, trainer = os.path.join(BIN, "eval_ppl"), os.path.join(REF, "ref_eval_ppl"), os.path.join(BIN, "gpt2_lora_finetune")
    need(mine, ref, trainer)
    fx = fixture(tmp_path, n_layer=2, n_embd=64, n_head=4, n_positions=128)
    run_cli(trainer, ["--data_dir", fx + "/data", "--pretrained_dir", fx + "/model", "--steps", "30", "--batch_size", "4", "--seq_len", "32", "--lr", "5e-3",
                      "--seed", "3", "--lora_out", fx + "/lora.safetensors"])
    res = re.compile(r"split=(\S+)\s+tokens=(\d+)\s+nll=(\S+)\s+ppl=(\S+)")

    def ppl(exe, extra, env=None):
        r = subprocess.run([exe, "--data_root", fx + "/data", "--pretrained_dir", fx + "/model", "--split", "valid", "--seq_len", "32", "--batch_size", "4"] + extra,
                           capture_output=True, text=True, timeout=900, env=env)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        m = res.search(r.stdout)
        assert m, r.stdout[-1500:]
        return int(m[2]), float(m[3])

    exact = dict(os.environ, MFT_USE_TENSOR_CORES="0")
    nlls = []
    for extra in ([], ["--lora_path", fx + "/lora.safetensors"], ["--lora_path", fx + "/lora.safetensors", "--lora_merge", "1"]):
        tok_ref, nll_ref = ppl(ref, extra)
        tok, nll = ppl(mine, extra, env=exact)
        assert tok == tok_ref and abs(nll - nll_ref) <= 2e-4 * nll_ref, (extra, nll, nll_ref)
        _, nll_tc = ppl(mine, extra)
        assert abs(nll_tc - nll_ref) <= 1e-3 * nll_ref
        nlls.append(nll)
    assert nlls[1] < nlls[0]                       # the adapter was trained on this distribution
    assert abs(nlls[1] - nlls[2]) <= 2e-4 * nlls[1]    # merged weights evaluate like attached adapters

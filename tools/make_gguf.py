#!/usr/bin/env python
"""Write a random-init GGUF v3 file of a llama-arch model (TEST TOOLING, not product code).

There is no network for checkpoints, so every model in BASELINE.json's configs is a seeded random
initialisation of the named architecture.  Tensor names and hyper-parameter keys are the ones the
reference loader reads (/root/reference/llama.cpp:2227-2276 hparams, :351-403 tensor tables,
:2756-2812 llama tensors).  The vocabulary is synthetic and character-level (see `synth_vocab`), so
neither this script nor the tests need the reference's models/ggml-vocab-*.gguf.

Two weight flavours:
  iid     every 2-D weight ~ N(0, 0.02^2): target and draft disagree, acceptance ~ 0 (exercises cancel).
  bigram  token_embd[t] = g_t, output[pi(t)] = g_t for a fixed permutation pi, block weights scaled by
          --eps so the residual stream stays ~ g_t: both models predict pi(last token), so drafts get
          accepted.  --agree A makes a draft (built with --draft-perturb) agree with pi on a fraction A.

Quantisation: tensors are written as F32 and then passed through the reference's own
`quantize --pure` (oracle/_ref/quantize) by `--quant q8_0|q4_0|q4_k` so that block contents come from
the reference quantiser, not from a re-implementation.
"""
import argparse
import os
import subprocess
import sys

import numpy as np
import gguf

N_SPECIAL = 3          # <unk> <s> </s>
N_BYTE = 256           # <0x00>..<0xFF>
CHAR0 = N_SPECIAL + N_BYTE   # first printable-character token (space == U+2581)


def synth_vocab(n_vocab):
    """3 specials, 256 byte tokens, 95 single printable ASCII characters, then filler pieces that no
    text can produce.  With it the reference SPM tokenizer (/root/reference/llama.cpp:5970-6110) maps a
    prompt one character -> one token, so tests know the ids without running a tokenizer."""
    toks, scores, types = [], [], []
    for t, ty in (("<unk>", 2), ("<s>", 3), ("</s>", 3)):
        toks.append(t); scores.append(0.0); types.append(ty)
    for b in range(256):
        toks.append("<0x%02X>" % b); scores.append(0.0); types.append(6)
    for c in range(0x20, 0x7F):
        toks.append("▁" if c == 0x20 else chr(c)); scores.append(-1.0); types.append(1)
    i = 0
    while len(toks) < n_vocab:
        toks.append("▁§%d§" % i); scores.append(-1000.0 - i); types.append(1)
        i += 1
    return toks[:n_vocab], scores[:n_vocab], types[:n_vocab]


def text_to_ids(text):
    """ids the reference tokenizer yields for printable-ASCII `text` with the synthetic vocab
    (BOS, then the SPM space prefix, then one token per character)."""
    return [1, CHAR0] + [CHAR0 + (ord(c) - 0x20) for c in text]


def perm(n_vocab, seed):
    return np.random.default_rng(seed).permutation(n_vocab)


def build(args):
    rng = np.random.default_rng(args.seed)
    E, L, H, HKV, FF, V = args.n_embd, args.n_layer, args.n_head, args.n_head_kv, args.n_ff, args.n_vocab
    hd = E // H
    falcon = args.arch == "falcon"
    w = gguf.GGUFWriter(args.out if args.quant == "f32" else args.out + ".f32.gguf", args.arch)
    w.add_name(args.name)
    w.add_context_length(args.n_ctx_train)
    w.add_embedding_length(E)
    w.add_block_count(L)
    w.add_feed_forward_length(FF)
    w.add_rope_dimension_count(hd)
    w.add_head_count(H)
    w.add_head_count_kv(HKV)
    if falcon:
        w.add_layer_norm_eps(1e-5)
    else:
        w.add_layer_norm_rms_eps(1e-5)
    w.add_file_type(0)
    toks, scores, types = synth_vocab(V)
    w.add_tokenizer_model("llama")
    w.add_token_list(toks)
    w.add_token_scores(scores)
    w.add_token_types(types)
    w.add_bos_token_id(1)
    w.add_eos_token_id(2)
    w.add_unk_token_id(0)

    def mat(rows, cols, std):
        return (rng.standard_normal((rows, cols), dtype=np.float32) * np.float32(std))

    if args.flavour == "bigram":
        g = rng.standard_normal((V, E), dtype=np.float32)
        g /= np.linalg.norm(g, axis=1, keepdims=True)
        p = perm(V, 7)
        if args.agree < 1.0:   # a draft that disagrees with the target on a deterministic token subset
            p2 = perm(V, 8)
            flip = np.random.default_rng(9).random(V) >= args.agree
            p = np.where(flip, p2, p)
        emb = (g * np.float32(args.emb_scale)).astype(np.float32)
        out = np.zeros((V, E), dtype=np.float32)
        out[p] = g * np.float32(args.emb_scale)
        std = 0.02 * args.eps
    else:
        emb = mat(V, E, 0.02)
        out = mat(V, E, 0.02)
        std = 0.02
    w.add_tensor("token_embd.weight", emb)

    def norm_w():
        return rng.uniform(0.5, 1.5, E).astype(np.float32)

    def norm_b():
        return (rng.standard_normal(E) * 0.02).astype(np.float32)

    for il in range(L if falcon else 0):
        # tensor table of the falcon architecture (reference llama.cpp:2769-2811): LayerNorm with bias, fused QKV,
        # optional second norm (Falcon-40B/180B), GELU MLP without gate
        pre = "blk.%d." % il
        w.add_tensor(pre + "attn_norm.weight", norm_w())
        w.add_tensor(pre + "attn_norm.bias", norm_b())
        if args.norm2:
            w.add_tensor(pre + "attn_norm_2.weight", norm_w())
            w.add_tensor(pre + "attn_norm_2.bias", norm_b())
        w.add_tensor(pre + "attn_qkv.weight", mat(E + 2 * hd * HKV, E, std))
        w.add_tensor(pre + "attn_output.weight", mat(E, E, std))
        w.add_tensor(pre + "ffn_down.weight", mat(E, FF, std))
        w.add_tensor(pre + "ffn_up.weight", mat(FF, E, std))
    if falcon:
        w.add_tensor("output_norm.weight", norm_w())
        w.add_tensor("output_norm.bias", norm_b())
    for il in range(0 if falcon else L):
        pre = "blk.%d." % il
        w.add_tensor(pre + "attn_norm.weight", np.ones(E, dtype=np.float32))
        w.add_tensor(pre + "attn_q.weight", mat(E, E, std))
        w.add_tensor(pre + "attn_k.weight", mat(hd * HKV, E, std))
        w.add_tensor(pre + "attn_v.weight", mat(hd * HKV, E, std))
        w.add_tensor(pre + "attn_output.weight", mat(E, E, std))
        w.add_tensor(pre + "ffn_norm.weight", np.ones(E, dtype=np.float32))
        w.add_tensor(pre + "ffn_gate.weight", mat(FF, E, std))
        w.add_tensor(pre + "ffn_down.weight", mat(E, FF, std))
        w.add_tensor(pre + "ffn_up.weight", mat(FF, E, std))
    if not falcon:
        w.add_tensor("output_norm.weight", np.ones(E, dtype=np.float32))
    w.add_tensor("output.weight", out)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    if args.quant != "f32":
        q = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref", "quantize")
        # q4_k_m / q4_k_s: the reference's mixed recipes (no --pure): output.weight Q6_K, attn_v / ffn_down partly Q6_K resp. Q5_K
        r = subprocess.run([q] + ([] if args.quant in ("q4_k_m", "q4_k_s", "q3_k_m", "q5_k_m") else ["--pure"]) + [args.out + ".f32.gguf", args.out, args.quant.upper()],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout)
            raise SystemExit("reference quantize failed")
        os.unlink(args.out + ".f32.gguf")


def direct_blocks(rng, qtype, rows, cols, std):
    """random weights written directly as quantised blocks (no F32 detour): value = d * (q - mid) with d chosen so that
    the dequantised std is `std`.  Used for the 70B-width CPU samples, where an F32 source would not fit."""
    if qtype == "q8_0":
        nb = cols // 32
        blk = rng.integers(0, 256, (rows, nb, 34), dtype=np.uint8)
        d = (rng.uniform(0.5, 1.5, (rows, nb)) * std / 74.0).astype(np.float16)
        blk[:, :, 0:2] = d.view(np.uint8).reshape(rows, nb, 2)
        return blk.reshape(rows, nb * 34), gguf.GGMLQuantizationType.Q8_0
    if qtype == "q4_0":
        nb = cols // 32
        blk = rng.integers(0, 256, (rows, nb, 18), dtype=np.uint8)
        d = (rng.uniform(0.5, 1.5, (rows, nb)) * std / 4.6).astype(np.float16)
        blk[:, :, 0:2] = d.view(np.uint8).reshape(rows, nb, 2)
        return blk.reshape(rows, nb * 18), gguf.GGMLQuantizationType.Q4_0
    nb = cols // 256
    blk = rng.integers(0, 256, (rows, nb, 144), dtype=np.uint8)
    d = (rng.uniform(0.5, 1.5, (rows, nb)) * std / (32.0 * 4.6)).astype(np.float16)
    dmin = (d.astype(np.float32) * 7.5).astype(np.float16)          # with m_j = sc_j: value = d * sc * (q - 7.5)
    blk[:, :, 0:2] = d.view(np.uint8).reshape(rows, nb, 2)
    blk[:, :, 2:4] = dmin.view(np.uint8).reshape(rows, nb, 2)
    s = rng.integers(16, 48, (rows, nb, 8), dtype=np.uint8)
    q = np.zeros((rows, nb, 12), dtype=np.uint8)
    for j in range(4):
        q[:, :, j] = s[:, :, j] | ((s[:, :, j + 4] >> 4) << 6)
        q[:, :, j + 4] = s[:, :, j] | ((s[:, :, j + 4] >> 4) << 6)
        q[:, :, j + 8] = (s[:, :, j + 4] & 0xF) | ((s[:, :, j + 4] & 0xF) << 4)
    blk[:, :, 4:16] = q
    return blk.reshape(rows, nb * 144), gguf.GGMLQuantizationType.Q4_K


def build_direct(args):
    """bigram-flavour model written directly in the quantised type (see direct_blocks)."""
    rng = np.random.default_rng(args.seed)
    E, L, H, HKV, FF, V = args.n_embd, args.n_layer, args.n_head, args.n_head_kv, args.n_ff, args.n_vocab
    hd = E // H
    w = gguf.GGUFWriter(args.out, "llama")
    w.add_name(args.name)
    w.add_context_length(args.n_ctx_train)
    w.add_embedding_length(E)
    w.add_block_count(L)
    w.add_feed_forward_length(FF)
    w.add_rope_dimension_count(hd)
    w.add_head_count(H)
    w.add_head_count_kv(HKV)
    w.add_layer_norm_rms_eps(1e-5)
    w.add_file_type(0)
    toks, scores, types = synth_vocab(V)
    w.add_tokenizer_model("llama")
    w.add_token_list(toks)
    w.add_token_scores(scores)
    w.add_token_types(types)
    w.add_bos_token_id(1)
    w.add_eos_token_id(2)
    w.add_unk_token_id(0)

    def put(name, rows, cols, std, blocks=None):
        b, qt = direct_blocks(rng, args.quant, rows, cols, std) if blocks is None else blocks
        w.add_tensor(name, b, raw_dtype=qt)
        return b, qt

    erng = np.random.default_rng(4242)                 # the same g_t for target and draft of one vocabulary / width
    emb, qt = direct_blocks(erng, args.quant, V, E, args.emb_scale)
    p = perm(V, 7)
    if args.agree < 1.0:
        flip = np.nonzero(np.random.default_rng(9).random(V) >= args.agree)[0]
        if len(flip) > 1:
            p = p.copy()
            p[flip] = np.roll(p[flip], -1)
    out = np.empty_like(emb)
    out[p] = emb
    w.add_tensor("token_embd.weight", emb, raw_dtype=qt)
    sE, sF = 0.1 / np.sqrt(E), 0.1 / np.sqrt(FF)
    for il in range(L):
        pre = "blk.%d." % il
        w.add_tensor(pre + "attn_norm.weight", np.ones(E, dtype=np.float32))
        put(pre + "attn_q.weight", E, E, sE)
        put(pre + "attn_k.weight", hd * HKV, E, sE)
        put(pre + "attn_v.weight", hd * HKV, E, sE)
        put(pre + "attn_output.weight", E, E, sE)
        w.add_tensor(pre + "ffn_norm.weight", np.ones(E, dtype=np.float32))
        put(pre + "ffn_gate.weight", FF, E, sE)
        put(pre + "ffn_down.weight", E, FF, sF)
        put(pre + "ffn_up.weight", FF, E, sE)
    w.add_tensor("output_norm.weight", np.ones(E, dtype=np.float32))
    w.add_tensor("output.weight", out, raw_dtype=qt)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--name", default="random-llama")
    ap.add_argument("--arch", default="llama", choices=["llama", "falcon"])
    ap.add_argument("--norm2", action="store_true", help="falcon: second attention norm per layer (the 40B/180B layout)")
    ap.add_argument("--n-layer", type=int, default=2)
    ap.add_argument("--n-embd", type=int, default=512)
    ap.add_argument("--n-head", type=int, default=8)
    ap.add_argument("--n-head-kv", type=int, default=8)
    ap.add_argument("--n-ff", type=int, default=1408)
    ap.add_argument("--n-vocab", type=int, default=32000)
    ap.add_argument("--n-ctx-train", type=int, default=2048)
    ap.add_argument("--quant", default="q8_0", choices=["f32", "q8_0", "q4_0", "q4_k", "q4_k_m", "q4_k_s", "q3_k_m", "q5_k_m", "q6_k"])
    ap.add_argument("--flavour", default="iid", choices=["iid", "bigram"])
    ap.add_argument("--eps", type=float, default=0.25)
    ap.add_argument("--emb-scale", type=float, default=1.0)
    ap.add_argument("--agree", type=float, default=1.0)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--direct", action="store_true", help="write random quantised blocks directly (no F32 source, no quantize)")
    a = ap.parse_args()
    if a.direct:
        build_direct(a)
    else:
        build(a)


if __name__ == "__main__":
    main()

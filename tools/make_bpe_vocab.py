#!/usr/bin/env python
"""Writes tests/golden/vocab_bpe.gguf: a small GPT-2 style byte-level BPE vocabulary in a tensor-less GGUF (what the
reference opens with vocab_only), for the tokenizer parity tests.

Falcon's real vocabulary cannot be fetched here, so the merges are TRAINED below, with the textbook BPE loop, on a
few paragraphs of mixed text; the layout follows Falcon's: eleven `>>NAME<<` markers, `<|endoftext|>` as id 11
(bos = eos = 11, llama.cpp:2456-2461), then the 256 byte symbols, then one token per merge.
"""
import collections
import os
import sys

import gguf
import regex

CORPUS = """
The quick brown fox jumps over the lazy dog. It's what they've said we'll do, isn't it? I'm sure you'd agree that he's right.
Pipelined speculation keeps every stage of the pipeline busy: while the target model verifies run 17, the draft model is
already proposing runs 18, 19 and 20.  In 2023 there were 1,048,576 tokens in 4096 batches; 3.14159 is not 22/7.
def main(argv):\n    for i in range(10):\n        print("value = %d" % (i * 2))\n    return 0\n
Ünïcödé tëxt: naïve café, Zürich, São Paulo, Москва, Ελληνικά, 東京都, 한국어, العربية, עברית. Emoji: \U0001F600 \U0001F680.
Whitespace   runs,\ttabs\tand trailing spaces   \n\n\nnew paragraphs.  ~!@#$%^&*()_+-={}[]|\\:;"'<>,.?/
the the the and and of of to to in in is is that that for for with with as as on on be be at at by by this this
"""

SPECIALS = [">>TITLE<<", ">>ABSTRACT<<", ">>INTRODUCTION<<", ">>SUMMARY<<", ">>COMMENT<<", ">>ANSWER<<", ">>QUESTION<<",
            ">>DOMAIN<<", ">>PREFIX<<", ">>SUFFIX<<", ">>MIDDLE<<", "<|endoftext|>"]

GPT2_SPLIT = regex.compile(r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+""")


def byte_alphabet():
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    table, extra = {}, 0
    for b in range(256):
        if b in keep:
            table[b] = chr(b)
        else:
            table[b] = chr(256 + extra)
            extra += 1
    return table


def train(n_merges):
    alpha = byte_alphabet()
    words = collections.Counter()
    for w in GPT2_SPLIT.findall(CORPUS):
        words[tuple(alpha[b] for b in w.encode("utf-8"))] += 1
    merges = []
    for _ in range(n_merges):
        pairs = collections.Counter()
        for w, c in words.items():
            for a, b in zip(w, w[1:]):
                pairs[(a, b)] += c
        if not pairs:
            break
        best = max(pairs.items(), key=lambda kv: (kv[1], kv[0]))[0]
        merges.append(best)
        merged = collections.Counter()
        for w, c in words.items():
            out, i = [], 0
            while i < len(w):
                if i + 1 < len(w) and (w[i], w[i + 1]) == best:
                    out.append(w[i] + w[i + 1]); i += 2
                else:
                    out.append(w[i]); i += 1
            merged[tuple(out)] += c
        words = merged
    return alpha, merges


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "vocab_bpe.gguf")
    alpha, merges = train(600)
    toks = list(SPECIALS) + [alpha[b] for b in range(256)]
    types = [3] * len(SPECIALS) + [1] * 256
    for a, b in merges:
        if a + b not in toks:
            toks.append(a + b); types.append(1)
    w = gguf.GGUFWriter(out, "falcon")
    w.add_name("bpe-vocab-fixture")
    w.add_context_length(2048)
    w.add_embedding_length(64)
    w.add_block_count(1)
    w.add_feed_forward_length(256)
    w.add_head_count(2)
    w.add_head_count_kv(1)
    w.add_layer_norm_eps(1e-5)
    w.add_file_type(0)
    w.add_tokenizer_model("gpt2")
    w.add_token_list(toks)
    w.add_token_types(types)
    w.add_token_merges(["%s %s" % m for m in merges])
    w.add_bos_token_id(11)
    w.add_eos_token_id(11)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    print("%d tokens, %d merges -> %s (%d bytes)" % (len(toks), len(merges), os.path.normpath(out), os.path.getsize(out)))


if __name__ == "__main__":
    main()

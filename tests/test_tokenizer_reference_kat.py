"""CPU-only: the reference's own tokenizer tests, run against this library.

/root/reference/tests/test-tokenizer-0-{llama,falcon}.cpp hold known-answer vectors (text -> ids, made with the original
SentencePiece / HF tokenizers) for the vocabularies shipped in /root/reference/models/ggml-vocab-{llama,falcon}.gguf;
test-tokenizer-1-{llama,bpe}.cpp check round trips over every token and every codepoint.  The vectors are parsed out of
the .cpp files where they lie and the vocabularies are opened in place (vocab_only: no GPU), so these tests run only in
the container that has /root/reference.
"""
import ctypes as C
import os
import re

import pytest

from tests.test_tokenizer_cpu import ROOT, open_vocab

REFDIR = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REFDIR, "models", "ggml-vocab-falcon.gguf")),
                                reason="needs the vocabulary files under /root/reference/models")
LIB = os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libllama_ref.so")

ESC = {"t": "\t", "n": "\n", '"': '"', "'": "'", "\\": "\\"}


def known_answers(cpp):
    src = open(os.path.join(REFDIR, "tests", cpp), encoding="utf-8").read()
    body = src[src.index("_k_tests = {"):src.index("return _k_tests;")]
    out = []
    for lit, ids in re.findall(r'\{\s*"((?:[^"\\]|\\.)*)"\s*,\s*\{([^}]*)\}', body):
        text = re.sub(r"\\(.)", lambda m: ESC[m.group(1)], lit)
        out.append((text, [int(x) for x in ids.replace(",", " ").split()]))
    assert len(out) >= 20
    return out


def ids_of(L, m, raw, add_bos, special=False):
    buf = (C.c_int32 * (len(raw) * 4 + 16))()
    n = L.llama_tokenize(m, raw, len(raw), buf, len(buf), add_bos, special)
    assert n >= 0, n
    return list(buf[:n])


def ids_of_raw(L, m, raw):
    buf = (C.c_int32 * (len(raw) * 4 + 16))()
    return L.llama_tokenize(m, raw, len(raw), buf, len(buf), False, False)


def text_of(L, m, ids):
    out, b = b"", C.create_string_buffer(256)
    for t in ids:
        n = L.llama_token_to_piece(m, t, b, 256)
        assert n >= 0
        out += b.raw[:n]
    return out


def test_llama_vocabulary_known_answers():
    L, m = open_vocab(LIB, os.path.join(REFDIR, "models", "ggml-vocab-llama.gguf"))
    for text, want in known_answers("test-tokenizer-0-llama.cpp"):
        raw = text.encode("utf-8")
        assert ids_of(L, m, raw, False) == want, text
        assert ids_of(L, m, raw, True) == [1] + want, text


def test_falcon_vocabulary_known_answers():
    L, m = open_vocab(LIB, os.path.join(REFDIR, "models", "ggml-vocab-falcon.gguf"))
    L.llama_vocab_type.argtypes = [C.c_void_p]
    assert L.llama_vocab_type(m) == 1 and L.llama_n_vocab(m) == 65024
    for text, want in known_answers("test-tokenizer-0-falcon.cpp"):
        assert ids_of(L, m, text.encode("utf-8"), False) == want, text


def utf8(cp):
    return chr(cp).encode("utf-8", "surrogatepass")


def test_falcon_round_trips():
    """test-tokenizer-1-bpe.cpp: every token's text and every codepoint survives tokenize -> detokenize."""
    L, m = open_vocab(LIB, os.path.join(REFDIR, "models", "ggml-vocab-falcon.gguf"))
    for tok in range(L.llama_n_vocab(m)):
        piece = text_of(L, m, [tok])
        try:
            piece.decode("utf-8")
        except UnicodeDecodeError:
            continue                      # a token that is a fragment of a character: the reference skips it too
        assert text_of(L, m, ids_of(L, m, piece, False)) == piece, tok
    skip = set(range(0x03, 0x06)) | {0x0b, 0x11, 0x19} | set(range(0x13, 0x18)) | set(range(0x1c, 0x1f)) | set(range(0xd800, 0xe000))
    for cp in range(0xffff):
        if cp not in skip:
            s = b" " + utf8(cp)
            assert text_of(L, m, ids_of(L, m, s, False)) == s, hex(cp)
    for cp in list(range(0x10000, 0x40000, 7)) + list(range(0xe0000, 0x10ffff, 13)):
        s = utf8(cp)
        assert text_of(L, m, ids_of(L, m, s, False)) == s, hex(cp)


def test_llama_round_trips():
    """test-tokenizer-1-llama.cpp: same for the SPM vocabulary (a detokenized text loses the leading space)."""
    L, m = open_vocab(LIB, os.path.join(REFDIR, "models", "ggml-vocab-llama.gguf"))
    for tok in range(L.llama_n_vocab(m)):
        piece = text_of(L, m, [tok])
        got = text_of(L, m, ids_of(L, m, piece, False))
        assert got in (piece, b" " + piece), tok
    for cp in list(range(0, 0xd800)) + list(range(0xe000, 0xffff)) + list(range(0x10000, 0x10ffff, 11)):
        if cp == 0x2581:
            continue                      # the escaped space itself comes back as a space (the reference excepts 9601 too)
        s = utf8(cp)
        assert text_of(L, m, ids_of(L, m, s, False)) in (s, b" " + s), hex(cp)


@pytest.mark.skipif(not os.path.exists(REF_LIB), reason="oracle/_ref not built")
@pytest.mark.parametrize("vocab", ["llama", "falcon"])
def test_real_vocabularies_agree_with_compiled_reference(vocab):
    from tests.test_tokenizer_cpu import BPE_TEXTS, fuzz_texts
    path = os.path.join(REFDIR, "models", "ggml-vocab-%s.gguf" % vocab)
    ours, m_o = open_vocab(LIB, path)
    ref, m_r = open_vocab(REF_LIB, path)
    for f in ("llama_token_bos", "llama_token_eos", "llama_token_nl", "llama_add_bos_token", "llama_vocab_type"):
        getattr(ours, f).argtypes = getattr(ref, f).argtypes = [C.c_void_p]
        assert getattr(ours, f)(m_o) == getattr(ref, f)(m_r), f
    corpus = [t for t, _ in known_answers("test-tokenizer-0-%s.cpp" % vocab)] + BPE_TEXTS + fuzz_texts(4000, 11)
    corpus += ["<s>x</s>", "<unk>", ">>COMMENT<<hello>>ANSWER<<", "a<|endoftext|>b", "\n<s>\n"]
    no_token = set("\x03\x04\x05\x0b\x11\x13\x14\x15\x16\x17\x19\x1c\x1d\x1e")      # bytes Falcon's vocabulary lacks: the reference throws
    for text in corpus:
        if vocab == "falcon" and no_token & set(text):
            assert ids_of_raw(ours, m_o, text.encode("utf-8")) == -2**31
            continue
        raw = text.encode("utf-8")
        for special in (False, True):
            assert ids_of(ours, m_o, raw, True, special) == ids_of(ref, m_r, raw, True, special), (text, special)
    a, b = C.create_string_buffer(256), C.create_string_buffer(256)
    for tok in range(ours.llama_n_vocab(m_o)):
        na, nb = ours.llama_token_to_piece(m_o, tok, a, 256), ref.llama_token_to_piece(m_r, tok, b, 256)
        assert na == nb and a.raw[:max(na, 0)] == b.raw[:max(nb, 0)], tok

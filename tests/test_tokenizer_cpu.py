"""CPU-only: llama_tokenize / llama_token_to_piece (SPM, llama_api.cu) against the compiled reference's tokenizer
(oracle/_ref/libllama_ref.so) on the fixture vocabulary.  Models are opened vocab-only: no CUDA device is touched."""
import ctypes as C
import os

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF = os.path.join(ROOT, "oracle", "_ref", "libllama_ref.so")
pytestmark = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref is built only where /root/reference exists")


class ModelParams(C.Structure):
    _fields_ = [("n_gpu_layers", C.c_int32), ("main_gpu", C.c_int32), ("tensor_split", C.c_void_p),
                ("progress_callback", C.c_void_p), ("progress_callback_user_data", C.c_void_p),
                ("vocab_only", C.c_bool), ("use_mmap", C.c_bool), ("use_mlock", C.c_bool)]


def open_vocab(path_lib, gguf):
    L = C.CDLL(path_lib)
    L.llama_model_default_params.restype = ModelParams
    L.llama_load_model_from_file.restype = C.c_void_p
    L.llama_load_model_from_file.argtypes = [C.c_char_p, ModelParams]
    L.llama_tokenize.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_bool, C.c_bool]
    L.llama_token_to_piece.argtypes = [C.c_void_p, C.c_int32, C.c_char_p, C.c_int]
    L.llama_n_vocab.argtypes = [C.c_void_p]
    mp = L.llama_model_default_params()
    mp.vocab_only = True
    m = L.llama_load_model_from_file(gguf.encode(), mp)
    assert m
    return L, m


def tokenize(L, m, text, add_bos):
    raw = text.encode("utf-8")
    buf = (C.c_int32 * (len(raw) * 4 + 16))()
    n = L.llama_tokenize(m, raw, len(raw), buf, len(buf), add_bos, True)
    assert n >= 0
    return list(buf[:n])


TEXTS = ["the quick brown fox", " leading space", "trailing space ", "two  spaces", "", "a", "Hello, World! 123",
         "tabs\tand\nnewlines", "café naïve über", "你好，世界", "emoji \U0001F600 done",
         "~!@#$%^&*()_+{}|:\"<>?", "x" * 200]


def test_tokenizer_matches_reference(golden_dir):
    gguf = os.path.join(golden_dir, "tiny_q8_0.gguf")
    ours, m_o = open_vocab(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"), gguf)
    ref, m_r = open_vocab(REF, gguf)
    assert ours.llama_n_vocab(m_o) == ref.llama_n_vocab(m_r)
    for text in TEXTS:
        for add_bos in (True, False):
            assert tokenize(ours, m_o, text, add_bos) == tokenize(ref, m_r, text, add_bos), (text, add_bos)
    a, b = C.create_string_buffer(64), C.create_string_buffer(64)
    for tok in range(ours.llama_n_vocab(m_o)):
        na, nb = ours.llama_token_to_piece(m_o, tok, a, 64), ref.llama_token_to_piece(m_r, tok, b, 64)
        assert na == nb and a.raw[:max(na, 0)] == b.raw[:max(nb, 0)], tok


# ---- byte-level BPE (Falcon's tokenizer kind) ---------------------------------------------------------------------------
BPE_TEXTS = TEXTS + [
    "It's what they've said we'll do, isn't it? I'm sure you'd agree that he's right.", "'s", "'", "a'", "'re", "x'll", "'ve'd't",
    "ends with bang!", "ab!", "123abc456", " 1", "  12  ", "a  b", "a   b   ", "\n\n\n", " \n x", "tab\t\tx", "x \t\n 7",
    "Москва Ελληνικά 東京都 한국어 العربية", "naïve café, Zürich", "1,048,576 and 3.14159", "def main(argv):\n    return 0\n",
    ">>TITLE<<abc<|endoftext|>", "<|endoftext|>", "x<|endoftext|>y>>SUFFIX<<>>MIDDLE<<z", ">>TITLE", "<|endoftext|", "  x　",
    "²³¹ ٣٤ ⅓", "\x1c\x1d x", "a\x00b",
]


def fuzz_texts(n, seed):
    import random
    rnd = random.Random(seed)
    alphabet = list("abcdehilnorstvmdTHE  '   019.,!?-\n\t") + ["'s", "'re", " the", "é", "ß", "東", "　", " ", "٣", "²", "\U0001F600",
                                                                 "<|endoftext|>", ">>TITLE<<", "'", "''", "ll", "ve"]
    return ["".join(rnd.choice(alphabet) for _ in range(rnd.randint(1, 40))) for _ in range(n)]


def test_bpe_tokenizer_matches_reference(golden_dir):
    gguf = os.path.join(golden_dir, "vocab_bpe.gguf")
    ours, m_o = open_vocab(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"), gguf)
    ref, m_r = open_vocab(REF, gguf)
    for L in (ours, ref):
        for f in ("llama_vocab_type", "llama_token_bos", "llama_token_eos", "llama_token_nl", "llama_add_bos_token"):
            getattr(L, f).argtypes = [C.c_void_p]
    for f in ("llama_n_vocab", "llama_vocab_type", "llama_token_bos", "llama_token_eos", "llama_token_nl", "llama_add_bos_token"):
        assert getattr(ours, f)(m_o) == getattr(ref, f)(m_r), f
    assert ours.llama_vocab_type(m_o) == 1
    n_multi = 0
    for text in BPE_TEXTS + fuzz_texts(20000, 7):
        for special in (True, False):
            a, b = tokenize_sp(ours, m_o, text, special), tokenize_sp(ref, m_r, text, special)
            assert a == b, (text, special, a, b)
            n_multi += len(a) < len(text.encode())
    assert n_multi > 10000         # merges happened: the vocabulary is not byte-level only
    a, b = C.create_string_buffer(64), C.create_string_buffer(64)
    for tok in range(ours.llama_n_vocab(m_o)):
        na, nb = ours.llama_token_to_piece(m_o, tok, a, 64), ref.llama_token_to_piece(m_r, tok, b, 64)
        assert na == nb and a.raw[:max(na, 0)] == b.raw[:max(nb, 0)], tok
    # too small a buffer: -needed, on both sides
    small = (C.c_int32 * 2)()
    raw = b"hello world hello"
    assert ours.llama_tokenize(m_o, raw, len(raw), small, 2, True, True) == ref.llama_tokenize(m_r, raw, len(raw), small, 2, True, True) < 0
    # text that is not UTF-8: the reference throws out of its C API; here the call reports it
    bad = b"ab\xff\xfe"
    assert ours.llama_tokenize(m_o, bad, len(bad), small, 2, False, False) == -2**31


def tokenize_sp(L, m, text, special):
    raw = text.encode("utf-8")
    buf = (C.c_int32 * (len(raw) * 4 + 16))()
    n = L.llama_tokenize(m, raw, len(raw), buf, len(buf), True, special)
    assert n >= 0
    return list(buf[:n])


def test_spm_special_tokens_match_reference(golden_dir):
    gguf = os.path.join(golden_dir, "tiny_q8_0.gguf")
    ours, m_o = open_vocab(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"), gguf)
    ref, m_r = open_vocab(REF, gguf)
    for text in ["<s>hello</s>", "a<s>b", "</s>", "<unk> x <s>", "<s", "x</s>y</s>", "<s><s> two", " <s> spaced"]:
        for special in (True, False):
            assert tokenize_sp(ours, m_o, text, special) == tokenize_sp(ref, m_r, text, special), (text, special)


def test_codepoint_classes_match_reference_tables():
    probe = os.path.join(ROOT, "oracle", "_ref", "libunicode_probe.so")
    if not os.path.exists(probe):
        pytest.skip("oracle/_ref/libunicode_probe.so not built")
    R = C.CDLL(probe)
    O = C.CDLL(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"))
    R.ref_codepoint_type.argtypes = [C.c_uint32]
    O.pib_tokenizer_codepoint_class.argtypes = [C.c_uint32]
    for cp in range(0x110000):
        r = R.ref_codepoint_type(cp)
        assert O.pib_tokenizer_codepoint_class(cp) == (r if r in (1, 2, 3) else 0), hex(cp)

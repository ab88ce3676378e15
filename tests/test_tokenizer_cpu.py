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

"""CPU-only: the host samplers of the llama.h surface (llama_sample_softmax / token_greedy / top_k / top_p / min_p / temp /
repetition_penalties; llama_api.cu and compat_ref.cu) against the compiled reference's own functions (oracle/_ref/libllama_ref.so)
on random candidate arrays.  They are plain host code: no CUDA device is touched."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF = os.path.join(ROOT, "oracle", "_ref", "libllama_ref.so")
pytestmark = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref is built only where /root/reference exists")


class TokenData(C.Structure):
    _fields_ = [("id", C.c_int32), ("logit", C.c_float), ("p", C.c_float)]


class TokenDataArray(C.Structure):
    _fields_ = [("data", C.POINTER(TokenData)), ("size", C.c_size_t), ("sorted", C.c_bool)]


def bind(L):
    P = C.POINTER(TokenDataArray)
    L.llama_sample_softmax.argtypes = [C.c_void_p, P]
    L.llama_sample_token_greedy.argtypes = [C.c_void_p, P]
    L.llama_sample_token_greedy.restype = C.c_int32
    L.llama_sample_top_k.argtypes = [C.c_void_p, P, C.c_int, C.c_size_t]
    L.llama_sample_top_p.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
    L.llama_sample_min_p.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
    L.llama_sample_temp.argtypes = [C.c_void_p, P, C.c_float]
    L.llama_sample_repetition_penalties.argtypes = [C.c_void_p, P, C.POINTER(C.c_int32), C.c_size_t, C.c_float, C.c_float, C.c_float]
    return L


@pytest.fixture(scope="module")
def libs():
    return bind(C.CDLL(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"))), bind(C.CDLL(REF))


def make(logits):
    n = len(logits)
    buf = (TokenData * n)(*[TokenData(i, float(l), 0.0) for i, l in enumerate(logits)])
    return buf, TokenDataArray(C.cast(buf, C.POINTER(TokenData)), n, False)


def snapshot(arr):
    return [(arr.data[i].id, np.float32(arr.data[i].logit).view(np.uint32), np.float32(arr.data[i].p).view(np.uint32))
            for i in range(arr.size)], bool(arr.sorted)


@pytest.mark.parametrize("seed", range(5))
def test_sampler_chain_matches_reference(libs, seed):
    ours, ref = libs
    rng = np.random.default_rng(seed)
    logits = (rng.standard_normal(2000) * 3).astype(np.float32)
    logits[rng.integers(0, 2000, 5)] = logits.max()                 # ties: first maximum wins in greedy
    last = (C.c_int32 * 64)(*rng.integers(0, 2000, 64).tolist())
    outs = []
    for L in (ours, ref):
        buf, arr = make(logits)
        steps = []
        steps.append(("greedy", L.llama_sample_token_greedy(None, C.byref(arr))))
        L.llama_sample_repetition_penalties(None, C.byref(arr), last, 64, 1.1, 0.02, 0.01)
        steps.append(("penalties", snapshot(arr)))
        L.llama_sample_top_k(None, C.byref(arr), 40, 1)
        steps.append(("top_k", snapshot(arr)))
        L.llama_sample_top_p(None, C.byref(arr), 0.9, 1)
        steps.append(("top_p", snapshot(arr)))
        L.llama_sample_min_p(None, C.byref(arr), 0.05, 1)
        steps.append(("min_p", snapshot(arr)))
        L.llama_sample_temp(None, C.byref(arr), 0.7)
        L.llama_sample_softmax(None, C.byref(arr))
        steps.append(("temp+softmax", snapshot(arr)))
        steps.append(("greedy2", L.llama_sample_token_greedy(None, C.byref(arr))))
        outs.append(steps)
    for (name, a), (_, b) in zip(*outs):
        assert a == b, name


def test_softmax_of_full_vocabulary_matches_reference(libs):
    ours, ref = libs
    rng = np.random.default_rng(9)
    logits = (rng.standard_normal(32000) * 4).astype(np.float32)
    res = []
    for L in (ours, ref):
        buf, arr = make(logits)
        L.llama_sample_softmax(None, C.byref(arr))
        res.append(snapshot(arr))
    assert res[0] == res[1]


@pytest.mark.parametrize("seed", range(4))
def test_tail_free_and_typical_match_reference(libs, seed):
    ours, ref = libs
    for L in libs:
        P = C.POINTER(TokenDataArray)
        L.llama_sample_tail_free.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
        L.llama_sample_typical.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
    rng = np.random.default_rng(100 + seed)
    logits = (rng.standard_normal(500) * 2).astype(np.float32)
    for z, p, keep in ((0.95, 0.9, 1), (0.5, 0.3, 1), (0.99, 0.95, 5), (1.0, 1.0, 1), (0.2, 0.5, 0)):
        outs = []
        for L in (ours, ref):
            steps = []
            buf, arr = make(logits)
            L.llama_sample_tail_free(None, C.byref(arr), z, keep)
            steps.append(snapshot(arr))
            buf2, arr2 = make(logits)
            L.llama_sample_typical(None, C.byref(arr2), p, keep)
            steps.append(snapshot(arr2))
            outs.append(steps)
        assert outs[0] == outs[1], (z, p, keep)


def probs_after(L, probs, call):
    buf, arr = make(np.log(np.asarray(probs, np.float32)))
    L.llama_sample_softmax(None, C.byref(arr))
    call(arr)
    return [arr.data[i].p for i in range(arr.size)]


def test_reference_known_answers(libs):
    """the known-answer cases of the reference's tests/test-sampling.cpp:134-155, with its tolerances, on this library"""
    L = libs[0]
    P = C.POINTER(TokenDataArray)
    L.llama_sample_tail_free.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
    L.llama_sample_typical.argtypes = [C.c_void_p, P, C.c_float, C.c_size_t]
    four = [0.1, 0.2, 0.3, 0.4]
    for k, want in ((1, [0.4]), (3, [0.4, 0.3, 0.2])):
        assert np.allclose(probs_after(L, four, lambda a: L.llama_sample_top_k(None, C.byref(a), k, 1)), want, atol=1e-5)
    for p, want in ((0.0, [0.4]), (0.7, [0.4, 0.3]), (0.8, [0.4, 0.3, 0.2]), (1.0, [0.4, 0.3, 0.2, 0.1])):
        assert np.allclose(probs_after(L, four, lambda a: L.llama_sample_top_p(None, C.byref(a), p, 1)), want, atol=1e-3)
    five = [0.1, 0.15, 0.2, 0.25, 0.3]
    for z, want in ((0.25, [0.3]), (0.75, [0.3, 0.25]), (0.99, [0.3, 0.25])):
        assert np.allclose(probs_after(L, five, lambda a: L.llama_sample_tail_free(None, C.byref(a), z, 1)), want, atol=1e-3)
    for probs, p, want in (([0.97, 0.01, 0.01, 0.01], 0.5, [0.97]), ([0.4, 0.2, 0.2, 0.2], 0.5, [0.2, 0.2, 0.2])):
        assert np.allclose(probs_after(L, probs, lambda a: L.llama_sample_typical(None, C.byref(a), p, 1)), want, atol=1e-3)
    flat = [0.2] * 5
    cases = [([0], [0.25, 0.25, 0.25, 0.25, 0], 50.0, 0.0, 0.0), ([0, 1, 2], [0.5, 0.5, 0, 0, 0], 50.0, 0.0, 0.0),
             ([0, 1, 2, 0, 0], [0.5, 0.5, 0, 0, 0], 50.0, 0.0, 0.0),
             ([0], [0.249997, 0.249997, 0.249997, 0.249997, 0.000011], 1.0, 5.0, 5.0),
             ([0, 1, 2], [0.499966, 0.499966, 0.000023, 0.000023, 0.000023], 1.0, 5.0, 5.0),
             ([0, 1, 2, 0, 0], [0.499977, 0.499977, 0.000023, 0.000023, 0.0], 1.0, 5.0, 5.0)]
    for last, want, rep, freq, pres in cases:
        buf, arr = make(np.log(np.asarray(flat, np.float32)))
        toks = (C.c_int32 * len(last))(*last)
        L.llama_sample_softmax(None, C.byref(arr))
        L.llama_sample_repetition_penalties(None, C.byref(arr), toks, len(last), rep, freq, pres)
        L.llama_sample_softmax(None, C.byref(arr))
        assert np.allclose([arr.data[i].p for i in range(arr.size)], want, atol=1e-3), (last, rep)

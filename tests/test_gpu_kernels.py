"""Parity of the CUDA operators (through the C ABI of libpipeinfer_b200.so) against the oracle and the golden fixtures.
Integer work (activation blocks, weight round trip, embedding rows) must be bit-exact; fp32 dot products may differ
from the reference only by the order of fp32 additions (tolerance below; north_star: logits within 1e-2 max-abs)."""
import os

import numpy as np
import pytest

from oracle import oracle_py as op

pytestmark = pytest.mark.gpu

TYPES = list(op.TYPE_NAMES.items())
REL_TOL = 3e-5     # of max|y|: fp32 reassociation over <= 896 block terms (the kernels that add the block terms in their own order)
# kernels that keep the reference's summation order (eight fp32 lane accumulators per output, block after block, then its
# reduction tree) return the oracle's bits: (weight type, variant) pairs held to np.array_equal
EXACT = {(op.TYPE_Q4_K, 6), (op.TYPE_Q4_K, 4), (op.TYPE_Q4_0, 4), (op.TYPE_Q8_0, 4), (op.TYPE_Q4_K, 7), (op.TYPE_Q4_0, 7), (op.TYPE_Q8_0, 7)}


def check(got, ref, wt, variant, what=None):
    if (wt, variant) in EXACT:
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), (what, "max|diff| %.3g" % np.abs(got - ref).max())
    else:
        assert np.abs(got - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max()), what


@pytest.fixture(scope="module")
def lib():
    from pipeinfer_b200 import load
    L = load()
    L.init(0)
    return L


@pytest.fixture(scope="module")
def kern(golden_dir):
    return np.load(os.path.join(golden_dir, "kernels.npz"))


def rand_blocks(rng, wt, N, K):
    """random but valid weight blocks generated directly (no quantiser needed on the GPU box)"""
    be, bb = op.BLOCK[wt]
    nb = K // be
    blk = rng.integers(0, 256, (N, nb, bb), dtype=np.uint8)
    d = (rng.uniform(0.5, 1.5, (N, nb)) * 0.02 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
    blk[:, :, 0:2] = d
    if wt == op.TYPE_Q4_K:
        dm = (rng.uniform(0.5, 1.5, (N, nb)) * 0.01 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
        blk[:, :, 2:4] = dm
    return blk.reshape(N, nb * bb)


@pytest.mark.parametrize("name,wt", TYPES)
def test_weight_repack_round_trip(lib, kern, name, wt):
    w = kern[name + "_w"]
    K = kern[name + "_x"].shape[1]
    h = lib.weight_upload(wt, w, K)
    assert np.array_equal(lib.weight_rows(h, wt, w.shape[0], K), w)
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("name,wt", TYPES)
def test_activation_quantiser_bit_exact_vs_golden(lib, kern, name, wt):
    assert np.array_equal(lib.quantize_act(wt, kern[name + "_x"]), kern[name + "_act"])


@pytest.mark.parametrize("name,wt", TYPES)
@pytest.mark.parametrize("K", [256, 4096, 11008 // 256 * 256])
def test_activation_quantiser_bit_exact_vs_oracle(lib, oracle, name, wt, K):
    rng = np.random.default_rng(K)
    x = (rng.standard_normal((7, K)) * rng.uniform(0.01, 30.0, (7, 1))).astype(np.float32)
    x[3, :K // 2] = 0
    x[5, 17] = -x[5, 100]          # two elements with the same |x| (Q8_K keeps the sign of the first)
    assert np.array_equal(lib.quantize_act(wt, x), oracle.quantize_act(wt, x))


@pytest.mark.parametrize("name,wt", TYPES)
def test_rmsnorm_quantise_bit_exact_vs_oracle(lib, oracle, name, wt):
    import ctypes as C
    rng = np.random.default_rng(11)
    K = 2048
    x = (rng.standard_normal((5, K)) * 3).astype(np.float32)
    w = rng.uniform(0.5, 1.5, K).astype(np.float32)
    xn = np.empty_like(x)
    for t in range(5):
        oracle.L.po_rms_norm(x[t].ctypes.data_as(C.c_void_p), xn[t].ctypes.data_as(C.c_void_p), K, C.c_float(1e-5))
    xn *= w
    assert np.array_equal(lib.quantize_act(wt, x, norm_w=w, eps=1e-5), oracle.quantize_act(wt, xn))


@pytest.mark.parametrize("name,wt", TYPES)
def test_get_rows_bit_exact(lib, kern, name, wt):
    w = kern[name + "_w"]
    K = kern[name + "_x"].shape[1]
    h = lib.weight_upload(wt, w, K)
    ids = np.array([3, 0, 11, 3], dtype=np.int32)
    got = lib.get_rows(h, ids, K)
    assert np.array_equal(got.view(np.uint32), kern[name + "_deq"][ids].view(np.uint32))
    lib.L.pib_weight_free(h)


def variants(wt):
    # 1 = dp4a (all types), 2 = the default tensor-core kernel (all types), 3 = TMA-staged persistent kernel (Q4_K only),
    # 4 = the reference's summation order restated literally, one thread per output (all types), 6 = the tensor-core kernel that
    # keeps the reference's order (Q4_K only), 8 = the balanced stream kernel (Q4_K only; what variant 2 picks for row-tile counts that
    # do not fill the machine evenly)
    return [1, 2, 3, 4, 6, 8] if wt == op.TYPE_Q4_K else [1, 2, 4]


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 6, 8])
@pytest.mark.parametrize("name,wt", TYPES)
def test_mul_mat_vs_golden(lib, kern, name, wt, variant):
    if variant not in variants(wt):
        pytest.skip("variants 3, 6 and 8 are Q4_K only")
    w, x, y = kern[name + "_w"], kern[name + "_x"], kern[name + "_y"]
    h = lib.weight_upload(wt, w, x.shape[1])
    got = lib.mul_mat(h, x, w.shape[0], variant)
    check(got, y, wt, variant)
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 6, 8])
@pytest.mark.parametrize("name,wt", TYPES)
@pytest.mark.parametrize("b", [1, 2, 5, 8, 9, 16])
def test_mul_mat_vs_oracle_batches(lib, oracle, name, wt, b, variant):
    if variant not in variants(wt):
        pytest.skip("variants 3, 6 and 8 are Q4_K only")
    rng = np.random.default_rng(100 + b)
    N, K = 96, 2048
    w = rand_blocks(rng, wt, N, K)
    x = (rng.standard_normal((b, K)) * 2).astype(np.float32)
    h = lib.weight_upload(wt, w, K)
    got = lib.mul_mat(h, x, N, variant)
    ref = oracle.mul_mat(wt, w, x)
    check(got, ref, wt, variant)
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("name,wt", TYPES)
def test_mul_mat_ragged_rows_and_long_k(lib, oracle, name, wt):
    # N not a multiple of the rows-per-CTA, K = 28672 (the 70B down projection)
    rng = np.random.default_rng(5)
    N, K = 37, 28672
    w = rand_blocks(rng, wt, N, K)
    x = (rng.standard_normal((3, K))).astype(np.float32)
    h = lib.weight_upload(wt, w, K)
    ref = oracle.mul_mat(wt, w, x)
    for variant in variants(wt):
        got = lib.mul_mat(h, x, N, variant)
        check(got, ref, wt, variant, variant)
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("name,wt", TYPES)
@pytest.mark.parametrize("N,K,b", [(2560, 2048, 3), (48, 8192, 16), (200, 1024, 1)])
def test_mul_mat_tensor_core_k_split_path(lib, oracle, N, K, b, name, wt):
    # few row tiles -> the kernel splits K over several CTAs and reduces partial sums through scratch memory in a fixed order;
    # run twice: the per-tile arrival counters must be left clean
    rng = np.random.default_rng(N)
    w = rand_blocks(rng, wt, N, K)
    x = rng.standard_normal((b, K)).astype(np.float32)
    h = lib.weight_upload(wt, w, K)
    ref = oracle.mul_mat(wt, w, x)
    for variant in variants(wt)[1:]:
        a1 = lib.mul_mat(h, x, N, variant)
        a2 = lib.mul_mat(h, x, N, variant)
        assert np.array_equal(a1, a2)                      # deterministic
        check(a1, ref, wt, variant, variant)
    lib.L.pib_weight_free(h)


# the matrices bench.py times, at their full size: 70B gate|up, down, fused qkv, lm_head (Q4_K) and Falcon-180B up (Q4_0).
# The oracle computes a sample of rows -- the first and last row tiles, the rows either side of every 1/8 of the matrix (where a
# balanced partition of the weight stream changes hands) and 1024 drawn at random -- for b = 1, 4 and 16 tokens.
@pytest.mark.parametrize("wt,N,K", [(op.TYPE_Q4_K, 57344, 8192), (op.TYPE_Q4_K, 8192, 28672), (op.TYPE_Q4_K, 10240, 8192),
                                    (op.TYPE_Q4_K, 32000, 8192), (op.TYPE_Q4_0, 118784, 14848)])
def test_mul_mat_benchmark_shapes_vs_oracle(lib, oracle, wt, N, K):
    rng = np.random.default_rng(N + K)
    w = rand_blocks(rng, wt, N, K)
    rows = set(range(32)) | set(range(N - 32, N)) | set(int(r) for r in rng.integers(0, N, 1024))
    for e in range(1, 8):
        rows |= set(range(N * e // 8 - 20, N * e // 8 + 20))
    rows = np.array(sorted(rows))
    h = lib.weight_upload(wt, w, K)
    wsub = np.ascontiguousarray(w[rows])
    del w
    for b in (1, 4, 16):
        x = rng.standard_normal((b, K)).astype(np.float32)
        ref = oracle.mul_mat(wt, wsub, x)
        got = lib.mul_mat(h, x, N, 0)
        assert np.isfinite(got).all()
        check(np.ascontiguousarray(got[:, rows]), ref, wt, 0, b)
        if wt == op.TYPE_Q4_K:          # the order-keeping kernel, both decompositions (PIB_Q4K_MODE picks; default by shape)
            ex = lib.mul_mat(h, x, N, 6)
            check(np.ascontiguousarray(ex[:, rows]), ref, wt, 6, b)
            assert np.abs(ex - got).max() <= REL_TOL * max(1.0, np.abs(got).max())
        # every row, cheaply: the default kernel against the simple dp4a kernel (same integer sums, another fp32 order)
        if wt == op.TYPE_Q4_0:
            alt = lib.mul_mat(h, x, N, 1)
            assert np.abs(got - alt).max() <= REL_TOL * max(1.0, np.abs(alt).max()), b
    lib.L.pib_weight_free(h)


def test_q6_k_lm_head_shape_vs_oracle(lib, oracle):
    """the lm_head of a 70B Q4_K_M file (32000 x 8192, Q6_K) through the tensor-core kernel: sampled rows against the oracle, b = 1, 4, 16"""
    rng = np.random.default_rng(6)
    N, K = 32000, 8192
    w = rng.integers(0, 256, (N, K // 256, 210), dtype=np.uint8)
    w[:, :, 208:210] = (rng.uniform(0.5, 1.5, (N, K // 256)) * 0.001).astype(np.float16).view(np.uint8).reshape(N, K // 256, 2)
    w = w.reshape(N, -1)
    rows = np.array(sorted(set(range(32)) | set(range(N - 32, N)) | set(int(r) for r in rng.integers(0, N, 512))))
    h = lib.weight_upload(op.TYPE_Q6_K, w, K)
    wsub = np.ascontiguousarray(w[rows])
    del w
    for b in (1, 4, 16):
        x = rng.standard_normal((b, K)).astype(np.float32)
        ref = oracle.mul_mat(op.TYPE_Q6_K, wsub, x)
        got = lib.mul_mat(h, x, N, 0)
        assert np.isfinite(got).all()
        assert np.abs(got[:, rows] - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max()), b
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("wt", [op.TYPE_Q4_0, op.TYPE_Q8_0])
@pytest.mark.parametrize("b", [1, 7, 12])
def test_mul_mat_width_not_a_multiple_of_256(lib, oracle, wt, b):
    # Falcon-7B: n_embd = 4544 = 142 blocks of 32; the tensor-core kernel pads the last group of eight blocks with zeros
    rng = np.random.default_rng(b)
    N, K = 100, 4544
    w = rand_blocks(rng, wt, N, K)
    x = rng.standard_normal((b, K)).astype(np.float32)
    h = lib.weight_upload(wt, w, K)
    ref = oracle.mul_mat(wt, w, x)
    for variant in (1, 2):
        got = lib.mul_mat(h, x, N, variant)
        assert np.abs(got - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max()), variant
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("wt,fname,bb,dpos", [(op.TYPE_Q6_K, "kernels_q6k.npz", 210, 208), (op.TYPE_Q5_K, "kernels_q5k.npz", 176, 0),
                                              (op.TYPE_Q3_K, "kernels_q3k.npz", 110, 108)])
@pytest.mark.parametrize("b", [1, 3, 8, 9, 16])
def test_q6_k_q5_k_mul_mat(lib, oracle, golden_dir, b, wt, fname, bb, dpos):
    # Q6_K / Q5_K: output.weight, attn_v and part of ffn_down of Q4_K_M files; Q3_K: the bulk of Q3_K_M files.  Block round trip, the
    # reference's golden, random raw blocks
    g = np.load(os.path.join(golden_dir, fname))
    K = g["x"].shape[1]
    h = lib.weight_upload(wt, g["w"], K)
    assert np.array_equal(lib.weight_rows(h, wt, g["w"].shape[0], K), g["w"])
    got = lib.mul_mat(h, g["x"], g["w"].shape[0], 0)
    assert np.abs(got - g["y"]).max() <= REL_TOL * max(1.0, np.abs(g["y"]).max())
    lib.L.pib_weight_free(h)
    rng = np.random.default_rng(60 + b)
    N, K = 77, 2048
    w = rng.integers(0, 256, (N, K // 256, bb), dtype=np.uint8)
    w[:, :, dpos:dpos + 2] = (rng.uniform(0.5, 1.5, (N, K // 256)) * 0.001).astype(np.float16).view(np.uint8).reshape(N, K // 256, 2)
    if wt == op.TYPE_Q5_K:
        w[:, :, 2:4] = (rng.uniform(0.5, 1.5, (N, K // 256)) * 0.0005).astype(np.float16).view(np.uint8).reshape(N, K // 256, 2)
    w = w.reshape(N, -1)
    x = (rng.standard_normal((b, K)) * 2).astype(np.float32)
    h = lib.weight_upload(wt, w, K)
    ref = oracle.mul_mat(wt, w, x)
    got = lib.mul_mat(h, x, N, 0)
    assert np.abs(got - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max())
    # variant 4 (what exact mode runs for these types): the reference's summation order, the oracle's bits
    ex = lib.mul_mat(h, x, N, 4)
    assert np.array_equal(ex.view(np.uint32), ref.view(np.uint32)), float(np.abs(ex - ref).max())
    lib.L.pib_weight_free(h)


def test_bad_arguments_are_rejected(lib):
    import ctypes as C
    assert lib.L.pib_gemv(None, None, 1, None, 0, None) < 0
    assert lib.L.pib_weight_upload(3, None, 4, 64) is None          # unsupported type
    assert lib.L.pib_quantize_act(op.TYPE_Q4_K, None, 1, 100, None, None) < 0   # K not a multiple of 256
    assert b"" != lib.L.pib_last_error()


@pytest.mark.parametrize("b", [1, 3, 8])
@pytest.mark.parametrize("pro", [0, 1, 2])
def test_prologue_fused_gemv_vs_oracle(lib, oracle, pro, b):
    """pib_gemv_fused (the matmul that quantises its own activations: 9 -> 5 launches per layer on narrow models) against the oracle's
    element-wise op + quantiser + dot product; the residual sum it writes on the side must be exact"""
    import ctypes as C
    rng = np.random.default_rng(400 + 10 * pro + b)
    N, K = 200, 2048
    w = rand_blocks(rng, op.TYPE_Q4_K, N, K)
    h = lib.weight_upload(op.TYPE_Q4_K, w, K)
    x = (rng.standard_normal((b, K)) * 2).astype(np.float32)
    x2 = rng.uniform(0.5, 1.5, K).astype(np.float32) if pro == 1 else (rng.standard_normal((b, K)) * 2).astype(np.float32)
    add = (rng.standard_normal((b, K))).astype(np.float32)
    if pro == 0:
        v = x
    elif pro == 1:
        s = x + add
        v = np.empty_like(s)
        for t in range(b):
            oracle.L.po_rms_norm(s[t].ctypes.data_as(C.c_void_p), v[t].ctypes.data_as(C.c_void_p), K, C.c_float(1e-5))
        v *= x2
    else:
        g = x.copy()
        for t in range(b):
            oracle.L.po_silu(g[t].ctypes.data_as(C.c_void_p), K)
        v = g * x2
    ref = oracle.mul_mat(op.TYPE_Q4_K, w, v)
    lib.L.pib_gemv_fused.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
    xd, x2d, ad = lib.to_device(x), lib.to_device(x2), lib.to_device(add)
    yd, sd = lib.alloc(b * N * 4), lib.alloc(b * K * 4)
    lib.check(lib.L.pib_gemv_fused(h, pro, xd, x2d if pro else None, ad if pro == 1 else None, sd if pro == 1 else None, 0, 1e-5, b, yd, None))
    got = lib.to_host(yd, (b, N), np.float32)
    assert np.abs(got - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max())
    if pro == 1:
        assert np.array_equal(lib.to_host(sd, (b, K), np.float32), x + add)
    assert lib.L.pib_gemv_fused(h, 7, xd, None, None, None, 0, 0.0, 1, yd, None) < 0       # unknown prologue
    for d in (xd, x2d, ad, yd, sd):
        lib.free(d)
    lib.L.pib_weight_free(h)


@pytest.mark.parametrize("wt,fname", [(op.TYPE_Q3_K, "kernels_q3k.npz"), (op.TYPE_Q5_K, "kernels_q5k.npz"), (op.TYPE_Q6_K, "kernels_q6k.npz")])
def test_k_quant_embedding_rows_bit_exact(lib, golden_dir, wt, fname):
    # token_embd of Q3_K_M / Q5_K_M / Q6_K files: get_rows must return the bits of the reference's dequantize_row_q{3,5,6}_K
    g = np.load(os.path.join(golden_dir, fname))
    K = g["x"].shape[1]
    h = lib.weight_upload(wt, g["w"], K)
    ids = np.array([3, 0, 11, 3], dtype=np.int32)
    got = lib.get_rows(h, ids, K)
    assert np.array_equal(got.view(np.uint32), g["deq"][ids].view(np.uint32))
    lib.L.pib_weight_free(h)

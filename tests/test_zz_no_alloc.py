"""The allocation-free GEMV entry point (pib_gemv_q4k_noalloc) against the regular one (pib_gemv) and against the CPU
oracle, through the C ABI.
Collected last on purpose: it must give the same answers in a fresh process and after every other test has used the
library (the built-in scratch array keeps the previous call's contents)."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle_py as op

pytestmark = pytest.mark.gpu
REL_TOL = 3e-5          # same envelope as tests/test_gpu_kernels.py (in practice the two entry points agree bit for bit)


@pytest.fixture(scope="module")
def lib():
    from pipeinfer_b200 import load
    L = load()
    L.init(0)
    return L


def rand_q4k(rng, N, K):
    nb = K // 256
    blk = rng.integers(0, 256, (N, nb, 144), dtype=np.uint8)
    blk[:, :, 0:2] = (rng.uniform(0.5, 1.5, (N, nb)) * 0.02 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
    blk[:, :, 2:4] = (rng.uniform(0.5, 1.5, (N, nb)) * 0.01 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
    return blk.reshape(N, nb * 144)


# (N, K, b, k splits the launcher picks on 148 SMs).  The first five go through the built-in scratch array: the 7B attention
# projection (256 row tiles, 2 splits), a 128-tile matrix (4), the TinyLlama fused qkv (160 tiles, 2), a 64-tile matrix (8
# splits of 4 super-blocks) and a ragged 13-tile one (4); the last has 512 row tiles and is not split (no scratch used).
@pytest.mark.parametrize("N,K,b,splits", [(4096, 4096, 4, 2), (2048, 11008 // 256 * 256, 16, 4), (2560, 2048, 3, 2), (1024, 8192, 16, 8),
                                          (200, 4096, 1, 4), (8192, 8192, 16, 1)])
def test_noalloc_matches_regular_entry_point_and_oracle_twice(lib, oracle, N, K, b, splits):
    rng = np.random.default_rng(N + b)
    wt = op.TYPE_Q4_K
    w = rand_q4k(rng, N, K)
    h = lib.weight_upload(wt, w, K)
    lib.L.pib_gemv_direct_splits.argtypes = [C.c_int64, C.c_int64]
    if lib.L.pib_device_sm_count() == 148:
        assert lib.L.pib_gemv_direct_splits(N, K) == splits
    y_ref_d, y_d = lib.alloc(b * N * 4), lib.alloc(b * N * 4)
    act = lib.alloc(lib.L.pib_act_bytes(wt, K, b))
    for rep in range(2):                                    # two calls in a row, fresh inputs each time
        x = (rng.standard_normal((b, K)) * (1 + rep)).astype(np.float32)
        xd = lib.to_device(x)
        lib.check(lib.L.pib_quantize_act(wt, xd, b, K, act, None))
        lib.check(lib.L.pib_gemv(h, act, b, y_ref_d, 0, None))
        ref = lib.to_host(y_ref_d, (b, N), np.float32)
        poison = np.full((b, N), np.nan, np.float32)
        lib.check(lib.L.pib_memcpy_h2d(y_d, poison.ctypes.data, poison.nbytes, None))
        lib.check(lib.L.pib_stream_sync(None))
        lib.check(lib.L.pib_gemv_q4k_noalloc(h, act, b, y_d, None))
        got = lib.to_host(y_d, (b, N), np.float32)
        assert np.isfinite(got).all()
        assert np.abs(got - ref).max() <= REL_TOL * max(1.0, np.abs(ref).max())
        cpu = oracle.mul_mat(wt, w, x)
        assert np.abs(got - cpu).max() <= REL_TOL * max(1.0, np.abs(cpu).max())
        lib.free(xd)
    for d in (y_ref_d, y_d, act):
        lib.free(d)
    lib.L.pib_weight_free(h)


def test_noalloc_rejects_bad_arguments(lib):
    rng = np.random.default_rng(1)
    h8 = lib.weight_upload(op.TYPE_Q8_0, rng.integers(0, 256, (16, 256 // 32 * 34), dtype=np.uint8), 256)
    h = lib.weight_upload(op.TYPE_Q4_K, rand_q4k(rng, 32, 512), 512)
    dummy = lib.alloc(1 << 16)
    assert lib.L.pib_gemv_q4k_noalloc(h8, dummy, 1, dummy, None) < 0         # not Q4_K
    assert lib.L.pib_gemv_q4k_noalloc(h, dummy, 17, dummy, None) < 0         # batch beyond PIB_MAX_BATCH
    assert lib.L.pib_gemv_q4k_noalloc(h, None, 1, dummy, None) < 0
    assert b"noalloc" in lib.L.pib_last_error()
    lib.free(dummy)
    lib.L.pib_weight_free(h8)
    lib.L.pib_weight_free(h)

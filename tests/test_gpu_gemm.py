"""The tcgen05 tensor-core GEMM for wide batches (prompt ingestion) against the oracle.  Not bit-exact by design: both operands
are rounded to fp16 (weights after the reference's dequantisation formula, activations after the reference's Q8_K quantisation)
and the fp32 sum runs over K in the tensor core's order.  Tolerance: 2e-3 of max|y| (fp16 operand rounding, ~5e-4 relative per
element, averaged over K)."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle_py as op

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(scope="module")
def lib():
    from pipeinfer_b200 import load
    L = load()
    L.init(0)
    L.L.pib_gemm_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return L


def rand_q4k(rng, N, K):
    nb = K // 256
    blk = rng.integers(0, 256, (N, nb, 144), dtype=np.uint8)
    blk[:, :, 0:2] = (rng.uniform(0.5, 1.5, (N, nb)) * 0.02 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
    blk[:, :, 2:4] = (rng.uniform(0.5, 1.5, (N, nb)) * 0.01 / 8).astype(np.float16).view(np.uint8).reshape(N, nb, 2)
    return blk.reshape(N, nb * 144)


@pytest.mark.parametrize("N,K,b", [(128, 256, 32), (256, 1024, 64), (1280, 1024, 208), (1024, 2048, 128), (384, 4096, 256), (200, 512, 17)])
def test_tensor_core_gemm_vs_oracle(lib, oracle, N, K, b):
    rng = np.random.default_rng(N + K + b)
    w = rand_q4k(rng, N, K)
    x = rng.standard_normal((b, K)).astype(np.float32)
    h = lib.weight_upload(op.TYPE_Q4_K, w, K)
    y = np.full((b, N), np.nan, np.float32)
    lib.check(lib.L.pib_gemm_host(h, x.ctypes.data, b, y.ctypes.data))
    ref = np.concatenate([oracle.mul_mat(op.TYPE_Q4_K, w, x[i:i + 16]) for i in range(0, b, 16)])
    err = np.abs(y - ref).max() / max(1e-9, np.abs(ref).max())
    print("N %d K %d b %d: max|gemm - oracle| / max|oracle| = %.2e" % (N, K, b, err))
    assert np.isfinite(y).all() and err <= TOL
    lib.L.pib_weight_free(h)

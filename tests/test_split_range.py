"""CPU-only: the layer partition (pib_split_range, and ggml_mpi_split_range on top of it) against the oracle's restatement of the
reference's ggml_mpi_split_range (ggml-mpi.c:523-559) on the authors' launch scripts' weights and on random ones."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


@pytest.fixture(scope="module")
def lib():
    L = C.CDLL(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"))
    L.pib_split_range.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_uint16)]
    return L


def both(lib, oracle, n_layer, weights):
    n = len(weights)
    w = (C.c_float * n)(*weights)
    got = (C.c_uint16 * (2 * n))()
    lib.pib_split_range(n, 0, n_layer - 1, w, got)
    exp = np.zeros(2 * n, np.uint16)
    oracle.L.po_split_range(n, 0, n_layer - 1, w, exp.ctypes.data)
    return list(got), exp.tolist()


def test_the_authors_launch_script_split(lib, oracle):
    # examples/speculative/scripts/dolphin_tinyllama_8_nodes.sh: --mpi-layer-split 0.1,0.15 x 6 (+ remainder), 80 layers
    got, exp = both(lib, oracle, 80, [0.1] + [0.15] * 6 + [0.0])
    assert got == exp
    assert got[0] == 0 and got[1] == 8 and got[-1] == 79          # the last stage runs to the end of the model


@pytest.mark.parametrize("seed", range(6))
def test_random_weights(lib, oracle, seed):
    rng = np.random.default_rng(seed)
    for _ in range(200):
        n = int(rng.integers(1, 9))
        n_layer = int(rng.integers(n, 200))
        w = rng.random(n)
        w = (w / w.sum() * rng.uniform(0.5, 1.2)).tolist()        # under- and over-subscribed splits
        got, exp = both(lib, oracle, n_layer, w)
        assert got == exp, (n_layer, w)
        starts, ends = got[0::2], got[1::2]
        assert starts[0] == 0 and all(starts[i + 1] == ends[i] for i in range(n - 1)) and all(e <= n_layer - 1 for e in ends)

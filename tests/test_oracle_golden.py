"""The oracle (oracle/pib_oracle.c) against what the compiled reference produced (tests/golden/, written by
tests/golden/make_goldens.py from oracle/_ref).  Bit-exact everywhere: integer blocks, fp32 dot products, logits and
KV-cell bookkeeping.  CPU only."""
import os

import numpy as np
import pytest

from oracle import oracle_py as op

TYPES = list(op.TYPE_NAMES.items())


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def kern(golden_dir):
    return np.load(os.path.join(golden_dir, "kernels.npz"))


@pytest.mark.parametrize("name,wt", TYPES)
def test_activation_quantiser_matches_reference(oracle, kern, name, wt):
    # K2: quantize_row_q8_0 (AVX2 branch) / quantize_row_q8_K, incl. all-zero blocks and an outlier block
    assert np.array_equal(oracle.quantize_act(wt, kern[name + "_x"]), kern[name + "_act"])


@pytest.mark.parametrize("name,wt", TYPES)
def test_dequantise_matches_reference(oracle, kern, name, wt):
    # K3
    K = kern[name + "_x"].shape[1]
    assert np.array_equal(bits(oracle.dequantize(wt, kern[name + "_w"], K)), bits(kern[name + "_deq"]))


@pytest.mark.parametrize("name,wt", TYPES)
def test_mul_mat_matches_reference_bit_exact(oracle, kern, name, wt):
    # K1: quantise + vec_dot, same association order as the reference's AVX2 code
    y = oracle.mul_mat(wt, kern[name + "_w"], kern[name + "_x"])
    assert np.array_equal(bits(y), bits(kern[name + "_y"]))


@pytest.mark.parametrize("name,wt", TYPES)
def test_vec_dot_error_bound_of_reference_unit_test(oracle, kern, name, wt):
    # tests/test-quantize-fns.cpp:140-161: |vec_dot - float dot| / n < 0.02 on 0.1 + 2cos(i + off), n = 4096
    y = oracle.mul_mat(wt, kern["cos_" + name + "_w"], kern["cos_b"][None, :])
    assert np.array_equal(bits(y), bits(kern["cos_" + name + "_y"]))
    exact = float(np.dot(kern["cos_a"].astype(np.float64), kern["cos_b"].astype(np.float64)))
    assert abs(float(y[0, 0]) - exact) / 4096 < 0.02


def _replay(oracle, golden_dir, model_file, stem, n_ctx=48, max_seq=8, check_logits=True):
    m = op.load_gguf_model(os.path.join(golden_dir, model_file))
    ctx = op.OracleContext(oracle, m, n_ctx, max_seq)
    lines = open(os.path.join(golden_dir, stem + ".kv.txt")).read().strip().split("\n")
    ref_logits = np.fromfile(os.path.join(golden_dir, stem + ".logits.bin"), dtype=np.float32) if check_logits else None
    toks = open(os.path.join(golden_dir, stem + ".script")).read().split()
    i = li = off = 0
    n_dec = n_kv = 0
    while i < len(toks):
        o = toks[i]
        if o == "D":
            n = int(toks[i + 1]); i += 2
            tk, ps, sq = [], [], []
            for _ in range(n):
                t, p, ns = int(toks[i]), int(toks[i + 1]), int(toks[i + 2])
                sq.append([int(x) for x in toks[i + 3:i + 3 + ns]])
                tk.append(t); ps.append(p); i += 3 + ns
            rc, lg = ctx.decode(tk, ps, sq)
            assert lines[li] == "D %d" % rc
            li += 1
            if rc == 0 and check_logits:
                ref = ref_logits[off:off + lg.size].reshape(lg.shape)
                off += lg.size
                assert np.array_equal(bits(lg), bits(ref))
            n_dec += 1
        elif o == "RM":
            ctx.seq_rm(int(toks[i + 1]), int(toks[i + 2]), int(toks[i + 3])); i += 4
        elif o == "CP":
            ctx.seq_cp(int(toks[i + 1]), int(toks[i + 2]), int(toks[i + 3]), int(toks[i + 4])); i += 5
        elif o == "KEEP":
            ctx.seq_keep(int(toks[i + 1])); i += 2
        elif o == "KV":
            used, cells = op.parse_ref_kv_line(lines[li]); li += 1; i += 1
            st = ctx.kv_state()
            assert used == st["used"]
            assert cells == op.kv_state_to_cells(st)
            n_kv += 1
        else:
            raise AssertionError(o)
    ctx.close()
    return n_dec, n_kv


# q4_k_m / q4_k_s / q3_k_m: files as the reference's quantize mixes them (Q3_K, Q4_K, Q5_K, Q6_K tensors side by side)
@pytest.mark.parametrize("q", ["q8_0", "q4_0", "q4_k", "falcon_q4_0", "falcon40_q8_0", "q4_k_m", "q4_k_s", "q3_k_m", "q5_k_m", "q6_k"])
def test_decode_trace_matches_reference_bit_exact(oracle, golden_dir, q):
    # R2/R6/R7/R8 + K1..K9: multi-sequence batches, tree-speculation masks, no-slot failure (-1), GQA
    n_dec, n_kv = _replay(oracle, golden_dir, "tiny_%s.gguf" % q, "decode_%s" % q)
    assert n_dec == 8 and n_kv >= 8


# the benchmark's shapes: head_dim 128, grouped-query attention with 8 query heads per KV head (70B / 120B) and without
# grouping (7B), 290 cells of a 1024-cell cache in use, 3002 sequence ids (94-word sets), 64- and 208-token batches
@pytest.mark.parametrize("stem,model,n_ctx,max_seq,n_dec", [("wide_q4_k", "wide_q4_k.gguf", 1024, 3002, 9), ("wide_q4_0", "wide_q4_0.gguf", 1024, 3002, 9),
                                                          ("wide_long", "wide_q4_k.gguf", 6144, 8, 3)])
def test_decode_trace_at_benchmark_shapes_matches_reference_bit_exact(oracle, golden_dir, stem, model, n_ctx, max_seq, n_dec):
    got_dec, n_kv = _replay(oracle, golden_dir, model, "decode_%s" % stem, n_ctx=n_ctx, max_seq=max_seq)
    assert got_dec == n_dec and n_kv >= 3


def test_config1_real_shape_trace_matches_reference_bit_exact(oracle, golden_dir, tmp_path):
    # BASELINE.json configs[0] at its real shape (22 layers, 2048 wide, 32000 tokens, Q8_0): the model file is regenerated from
    # its seed (1.17 GB, not committed) and must hash to what the reference was given
    import hashlib, json, subprocess, sys
    meta = json.load(open(os.path.join(golden_dir, "c1_tinyllama_q8_0.json")))
    sh = meta["shape"]
    path = str(tmp_path / "c1.gguf")
    subprocess.run([sys.executable, os.path.join(golden_dir, "..", "..", "tools", "make_gguf.py"), "--direct", "--out", path,
                    "--n-layer", str(sh["n_layer"]), "--n-embd", str(sh["n_embd"]), "--n-head", str(sh["n_head"]),
                    "--n-head-kv", str(sh["n_head_kv"]), "--n-ff", str(sh["n_ff"]), "--n-vocab", str(sh["n_vocab"]),
                    "--quant", sh["quant"], "--seed", str(sh["seed"]), "--emb-scale", str(sh["emb_scale"])], check=True)
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    assert h.hexdigest() == meta["sha256"]
    n_dec, n_kv = _replay(oracle, golden_dir, path, "decode_c1_tinyllama_q8_0", n_ctx=64, max_seq=8)
    os.unlink(path)
    assert n_dec == 8 and n_kv == 2


def test_kv_bookkeeping_random_ops_match_reference(oracle, golden_dir):
    # R6: 400 random seq_rm / seq_cp / seq_keep / decode ops, cell metadata compared after every op
    n_dec, n_kv = _replay(oracle, golden_dir, "tiny_q8_0.gguf", "kvops", check_logits=False)
    assert n_kv >= 400


def test_softmax_known_answers(oracle):
    # tests/test-sampling.cpp:134-139 known answers for llama_sample_softmax via top-k with k = n
    import ctypes as C
    logits = np.log(np.array([0.1, 0.2, 0.3, 0.4], dtype=np.float32)).astype(np.float32)
    ids = np.zeros(4, np.int32); lg = np.zeros(4, np.float32); p = np.zeros(4, np.float32)
    oracle.L.po_sample_softmax_topk(logits.ctypes.data_as(C.c_void_p), 4, 4, ids.ctypes.data_as(C.c_void_p),
                                    lg.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p))
    assert ids.tolist() == [3, 2, 1, 0]
    assert np.allclose(p, [0.4, 0.3, 0.2, 0.1], atol=1e-5)
    assert oracle.L.po_sample_greedy(logits.ctypes.data_as(C.c_void_p), 4) == 3


def test_rope_composition_property(oracle):
    # tests/test-rope.cpp:134-216: rope(p0) then rope(delta) == rope(p0 + delta), rel err < 1e-4, modes 0 and 2
    import ctypes as C
    rng = np.random.default_rng(3)
    for mode in (0, 2):
        x = rng.standard_normal((4, 64)).astype(np.float32)
        a = x.copy(); b = x.copy()
        oracle.L.po_rope(a.ctypes.data_as(C.c_void_p), 4, 64, 64, 17, mode, 10000.0, 1.0)
        oracle.L.po_rope(a.ctypes.data_as(C.c_void_p), 4, 64, 64, 5, mode, 10000.0, 1.0)
        oracle.L.po_rope(b.ctypes.data_as(C.c_void_p), 4, 64, 64, 22, mode, 10000.0, 1.0)
        assert np.abs(a - b).sum() / np.abs(b).sum() < 1e-4


def test_split_range_examples(oracle):
    # ggml-mpi.c:523-559 on the authors' 8-node split of an 80-layer model (scripts/dolphin_tinyllama_8_nodes.sh)
    import ctypes as C
    w = np.array([0.1] + [0.15] * 6 + [0.0], dtype=np.float32)
    r = np.zeros(16, dtype=np.uint16)
    oracle.L.po_split_range(8, 0, 79, w.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p))
    r = r.reshape(8, 2)
    assert r[0].tolist() == [0, 8] and r[1].tolist() == [8, 20] and r[7].tolist() == [r[6][1], 79]
    assert all(r[i][1] == r[i + 1][0] for i in range(7))


@pytest.mark.parametrize("wt,fname", [(op.TYPE_Q6_K, "kernels_q6k.npz"), (op.TYPE_Q5_K, "kernels_q5k.npz"), (op.TYPE_Q3_K, "kernels_q3k.npz")])
def test_q6_k_q5_k_match_reference_golden(oracle, golden_dir, wt, fname):
    # the types Q4_K_M files keep output.weight, attn_v and part of ffn_down in (reference ggml-quants.c:1917, 1994, 5680, 6594)
    g = np.load(os.path.join(golden_dir, fname))
    K = g["x"].shape[1]
    assert np.array_equal(oracle.dequantize(wt, g["w"], K).view(np.uint32), g["deq"].view(np.uint32))
    assert np.array_equal(oracle.quantize_act(wt, g["x"]), g["act"])
    assert np.array_equal(oracle.mul_mat(wt, g["w"], g["x"]).view(np.uint32), g["y"].view(np.uint32))

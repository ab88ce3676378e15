"""ctypes bindings for oracle/liboracle.so and (when built) oracle/_ref/libggml_ref.so.

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs.
The product package (pipeinfer_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

TYPE_F32, TYPE_Q4_0, TYPE_Q8_0, TYPE_Q4_K = 0, 2, 8, 12
TYPE_Q5_K = 13
TYPE_Q3_K = 11          # only matmul / dequantise (Q3_K_M files: the type of the authors' 70B target, scripts/dolphin_tinyllama_8_nodes.sh)
TYPE_Q6_K = 14          # only matmul / dequantise (Q4_K_M files keep output.weight and some attn_v / ffn_down in it)
TYPE_NAMES = {"q4_0": TYPE_Q4_0, "q8_0": TYPE_Q8_0, "q4_k": TYPE_Q4_K}
BLOCK = {TYPE_Q4_0: (32, 18), TYPE_Q8_0: (32, 34), TYPE_Q4_K: (256, 144), TYPE_Q5_K: (256, 176), TYPE_Q6_K: (256, 210), TYPE_Q3_K: (256, 110)}   # (elements, bytes) per block
ACT_BLOCK = {TYPE_Q4_0: (32, 34), TYPE_Q8_0: (32, 34), TYPE_Q4_K: (256, 292), TYPE_Q5_K: (256, 292), TYPE_Q6_K: (256, 292), TYPE_Q3_K: (256, 292)}  # activation format per weight type


def build(ref=True):
    """make liboracle.so (and oracle/_ref/* when /root/reference exists and ref=True)."""
    target = [] if ref else ["oracle"]
    subprocess.run(["make", "-s", "-j8", "-C", HERE] + target, check=True)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Model(C.Structure):
    _fields_ = [("n_layer", C.c_int32), ("n_embd", C.c_int32), ("n_head", C.c_int32), ("n_head_kv", C.c_int32),
                ("n_ff", C.c_int32), ("n_vocab", C.c_int32), ("n_rot", C.c_int32),
                ("rms_eps", C.c_float), ("rope_freq_base", C.c_float), ("rope_freq_scale", C.c_float),
                ("wtype", C.c_int32),
                ("tok_embd", C.c_void_p), ("output_norm", C.c_void_p), ("output", C.c_void_p),
                ("attn_norm", C.POINTER(C.c_void_p)),
                ("wq", C.POINTER(C.c_void_p)), ("wk", C.POINTER(C.c_void_p)), ("wv", C.POINTER(C.c_void_p)),
                ("wo", C.POINTER(C.c_void_p)),
                ("ffn_norm", C.POINTER(C.c_void_p)),
                ("ffn_gate", C.POINTER(C.c_void_p)), ("ffn_down", C.POINTER(C.c_void_p)),
                ("ffn_up", C.POINTER(C.c_void_p)),
                ("arch", C.c_int32), ("norm_eps", C.c_float),
                ("attn_norm_b", C.POINTER(C.c_void_p)), ("attn_norm2", C.POINTER(C.c_void_p)),
                ("attn_norm2_b", C.POINTER(C.c_void_p)), ("output_norm_b", C.c_void_p),
                ("t_wq", C.POINTER(C.c_int32)), ("t_wk", C.POINTER(C.c_int32)), ("t_wv", C.POINTER(C.c_int32)), ("t_wo", C.POINTER(C.c_int32)),
                ("t_gate", C.POINTER(C.c_int32)), ("t_down", C.POINTER(C.c_int32)), ("t_up", C.POINTER(C.c_int32)), ("t_output", C.c_int32)]


class KvMeta(C.Structure):
    _fields_ = [("size", C.c_uint32), ("head", C.c_uint32), ("used", C.c_uint32), ("n", C.c_uint32),
                ("n_words", C.c_int32), ("pos", C.POINTER(C.c_int32)), ("bits", C.POINTER(C.c_uint32))]


class Context(C.Structure):
    _fields_ = [("model", C.POINTER(Model)), ("kv", C.POINTER(KvMeta)), ("n_ctx", C.c_uint32),
                ("k", C.c_void_p), ("v", C.c_void_p)]


class Oracle:
    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = self.L = C.CDLL(path)
        L.po_fp16_to_fp32.restype = C.c_float
        L.po_fp16_to_fp32.argtypes = [C.c_uint16]
        L.po_fp32_to_fp16.restype = C.c_uint16
        L.po_fp32_to_fp16.argtypes = [C.c_float]
        for n in ("po_table_exp_f16", "po_table_silu_f16", "po_table_gelu_f16"):
            getattr(L, n).restype = C.POINTER(C.c_uint16)
        L.po_row_size.restype = C.c_size_t
        L.po_row_size.argtypes = [C.c_int, C.c_int64]
        L.po_mul_mat.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
        L.po_kv_new.restype = C.POINTER(KvMeta)
        L.po_kv_new.argtypes = [C.c_uint32, C.c_int32]
        L.po_kv_free.argtypes = [C.POINTER(KvMeta)]
        L.po_kv_clear.argtypes = [C.POINTER(KvMeta)]
        L.po_kv_cell_max.restype = C.c_int32
        L.po_kv_cell_max.argtypes = [C.POINTER(KvMeta)]
        L.po_kv_seq_rm.argtypes = [C.POINTER(KvMeta), C.c_int32, C.c_int32, C.c_int32]
        L.po_kv_seq_cp.argtypes = [C.POINTER(KvMeta), C.c_int32, C.c_int32, C.c_int32, C.c_int32]
        L.po_kv_seq_keep.argtypes = [C.POINTER(KvMeta), C.c_int32]
        L.po_kv_find_slot.restype = C.c_int
        L.po_kv_find_slot.argtypes = [C.POINTER(KvMeta), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.po_kv_mask.argtypes = [C.POINTER(KvMeta), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.po_context_new.restype = C.POINTER(Context)
        L.po_context_new.argtypes = [C.POINTER(Model), C.c_uint32, C.c_int32]
        L.po_context_free.argtypes = [C.POINTER(Context)]
        L.po_decode.restype = C.c_int
        L.po_decode.argtypes = [C.POINTER(Context), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.po_sample_greedy.restype = C.c_int32
        L.po_sample_greedy.argtypes = [C.c_void_p, C.c_int]
        L.po_sample_softmax_topk.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.po_rope.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float]
        L.po_rms_norm.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float]
        L.po_norm.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float]
        L.po_soft_max.argtypes = [C.c_void_p, C.c_int]
        L.po_silu.argtypes = [C.c_void_p, C.c_int]
        L.po_gelu.argtypes = [C.c_void_p, C.c_int]
        L.po_split_range.argtypes = [C.c_int, C.c_uint16, C.c_uint16, C.c_void_p, C.c_void_p]
        L.po_tables_init()

    # ---- quantisers / dot products on numpy arrays ----
    def quantize_act(self, wtype, x):
        """x: float32 [..., K] -> uint8 [..., K/blk*bytes] in the activation format the weight type pairs with."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        K = x.shape[-1]
        be, bb = ACT_BLOCK[wtype]
        rows = x.reshape(-1, K)
        out = np.zeros((rows.shape[0], K // be * bb), dtype=np.uint8)
        fn = self.L.po_quantize_row_q8_K if wtype in (TYPE_Q3_K, TYPE_Q4_K, TYPE_Q5_K, TYPE_Q6_K) else self.L.po_quantize_row_q8_0
        for r in range(rows.shape[0]):
            fn(_ptr(rows[r]), _ptr(out[r]), C.c_int(K))
        return out.reshape(x.shape[:-1] + (out.shape[1],))

    def dequantize(self, wtype, w, K):
        """w: uint8 [N][row_bytes] -> float32 [N][K]"""
        w = np.ascontiguousarray(w)
        N = w.shape[0]
        out = np.empty((N, K), dtype=np.float32)
        fn = {TYPE_Q4_0: self.L.po_dequantize_row_q4_0, TYPE_Q8_0: self.L.po_dequantize_row_q8_0,
              TYPE_Q4_K: self.L.po_dequantize_row_q4_K, TYPE_Q5_K: self.L.po_dequantize_row_q5_K, TYPE_Q6_K: self.L.po_dequantize_row_q6_K, TYPE_Q3_K: self.L.po_dequantize_row_q3_K}[wtype]
        for r in range(N):
            fn(_ptr(w[r]), _ptr(out[r]), C.c_int(K))
        return out

    def mul_mat(self, wtype, w, x):
        """w: uint8 [N][row_bytes], x: float32 [b][K] -> float32 [b][N]"""
        w = np.ascontiguousarray(w)
        x = np.ascontiguousarray(x, dtype=np.float32)
        b, K = x.shape
        N = w.shape[0]
        y = np.empty((b, N), dtype=np.float32)
        self.L.po_mul_mat(wtype, _ptr(w), N, K, _ptr(x), b, _ptr(y))
        return y

    def tables(self):
        n = 1 << 16
        return {k: np.ctypeslib.as_array(getattr(self.L, "po_table_%s_f16" % k)(), shape=(n,)).copy()
                for k in ("exp", "silu", "gelu")}


class Ref:
    """The compiled reference (oracle/_ref/libggml_ref.so): its own quantisers and dot products."""

    def __init__(self):
        self.L = C.CDLL(os.path.join(REF_DIR, "libggml_ref.so"))
        # ggml_init fills the fp16 tables the quantisers use (ggml.c:2185)
        class InitParams(C.Structure):
            _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]
        self.L.ggml_init.restype = C.c_void_p
        self.L.ggml_init.argtypes = [InitParams]
        self._ctx = self.L.ggml_init(InitParams(1 << 20, None, False))

    @staticmethod
    def available():
        return os.path.exists(os.path.join(REF_DIR, "libggml_ref.so"))

    def quantize_weights(self, wtype, w):
        """float32 [N][K] -> uint8 [N][row_bytes] with the reference quantize_row_q*_reference."""
        w = np.ascontiguousarray(w, dtype=np.float32)
        N, K = w.shape
        be, bb = BLOCK[wtype]
        out = np.zeros((N, K // be * bb), dtype=np.uint8)
        fn = {TYPE_Q4_0: self.L.quantize_row_q4_0_reference, TYPE_Q8_0: self.L.quantize_row_q8_0_reference,
              TYPE_Q4_K: self.L.quantize_row_q4_K_reference, TYPE_Q5_K: self.L.quantize_row_q5_K_reference, TYPE_Q6_K: self.L.quantize_row_q6_K_reference, TYPE_Q3_K: self.L.quantize_row_q3_K_reference}[wtype]
        for r in range(N):
            fn(_ptr(w[r]), _ptr(out[r]), C.c_int(K))
        return out

    def quantize_act(self, wtype, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        K = x.shape[-1]
        be, bb = ACT_BLOCK[wtype]
        rows = x.reshape(-1, K)
        out = np.zeros((rows.shape[0], K // be * bb), dtype=np.uint8)
        fn = self.L.quantize_row_q8_K if wtype in (TYPE_Q3_K, TYPE_Q4_K, TYPE_Q5_K, TYPE_Q6_K) else self.L.quantize_row_q8_0
        for r in range(rows.shape[0]):
            fn(_ptr(rows[r]), _ptr(out[r]), C.c_int(K))
        return out.reshape(x.shape[:-1] + (out.shape[1],))

    def dequantize(self, wtype, w, K):
        w = np.ascontiguousarray(w)
        N = w.shape[0]
        out = np.empty((N, K), dtype=np.float32)
        fn = {TYPE_Q4_0: self.L.dequantize_row_q4_0, TYPE_Q8_0: self.L.dequantize_row_q8_0,
              TYPE_Q4_K: self.L.dequantize_row_q4_K, TYPE_Q5_K: self.L.dequantize_row_q5_K, TYPE_Q6_K: self.L.dequantize_row_q6_K, TYPE_Q3_K: self.L.dequantize_row_q3_K}[wtype]
        for r in range(N):
            fn(_ptr(w[r]), _ptr(out[r]), C.c_int(K))
        return out

    def vec_dot(self, wtype, wrow, arow, K):
        s = C.c_float(0)
        fn = {TYPE_Q4_0: self.L.ggml_vec_dot_q4_0_q8_0, TYPE_Q8_0: self.L.ggml_vec_dot_q8_0_q8_0,
              TYPE_Q4_K: self.L.ggml_vec_dot_q4_K_q8_K, TYPE_Q5_K: self.L.ggml_vec_dot_q5_K_q8_K, TYPE_Q6_K: self.L.ggml_vec_dot_q6_K_q8_K, TYPE_Q3_K: self.L.ggml_vec_dot_q3_K_q8_K}[wtype]
        fn(C.c_int(K), C.byref(s), _ptr(wrow), _ptr(arow))
        return s.value

    def mul_mat(self, wtype, w, x):
        """the reference's mul_mat semantics assembled from ITS quantiser + ITS vec_dot (ggml.c:9388)."""
        b, K = x.shape
        act = self.quantize_act(wtype, x)
        N = w.shape[0]
        y = np.empty((b, N), dtype=np.float32)
        for t in range(b):
            for r in range(N):
                y[t, r] = self.vec_dot(wtype, w[r], act[t], K)
        return y


# --------------------------------------------------------------------------------------------------------------
# GGUF -> oracle model (reads the file with the installed `gguf` package; tensors stay memory-mapped)
def load_gguf_model(path):
    import gguf
    r = gguf.GGUFReader(path)
    kv = {}
    for k, f in r.fields.items():
        try:
            kv[k] = f.contents()
        except Exception:
            pass
    t = {x.name: x for x in r.tensors}
    arch = kv["general.architecture"]
    L = int(kv[arch + ".block_count"])
    m = Model()
    m.n_layer = L
    m.n_embd = int(kv[arch + ".embedding_length"])
    m.n_head = int(kv[arch + ".attention.head_count"])
    m.n_head_kv = int(kv.get(arch + ".attention.head_count_kv", m.n_head))
    m.n_ff = int(kv[arch + ".feed_forward_length"])
    m.n_rot = int(kv.get(arch + ".rope.dimension_count", m.n_embd // m.n_head))
    m.rms_eps = float(kv.get(arch + ".attention.layer_norm_rms_epsilon", 1e-5))
    m.rope_freq_base = float(kv.get(arch + ".rope.freq_base", 10000.0))
    m.rope_freq_scale = 1.0
    emb = t["token_embd.weight"]
    m.n_vocab = int(emb.shape[1])
    m.wtype = int(emb.tensor_type)
    keep = []

    def p(name):
        a = np.ascontiguousarray(t[name].data)
        keep.append(a)
        return a.ctypes.data

    def arr(fmt):
        a = (C.c_void_p * L)(*[p(fmt % i) for i in range(L)])
        keep.append(a)
        return C.cast(a, C.POINTER(C.c_void_p))

    m.tok_embd = p("token_embd.weight")
    m.output_norm = p("output_norm.weight")
    m.output = p("output.weight")
    if arch == "falcon":
        m.arch = 1
        m.norm_eps = float(kv.get(arch + ".attention.layer_norm_epsilon", 1e-5))
        m.output_norm_b = p("output_norm.bias")
        m.attn_norm = arr("blk.%d.attn_norm.weight")
        m.attn_norm_b = arr("blk.%d.attn_norm.bias")
        if "blk.0.attn_norm_2.weight" in t:
            m.attn_norm2 = arr("blk.%d.attn_norm_2.weight")
            m.attn_norm2_b = arr("blk.%d.attn_norm_2.bias")
        m.wq = arr("blk.%d.attn_qkv.weight")
        m.wo = arr("blk.%d.attn_output.weight")
        m.ffn_down = arr("blk.%d.ffn_down.weight")
        m.ffn_up = arr("blk.%d.ffn_up.weight")
        m._keep = keep
        m._reader = r
        return m
    m.attn_norm = arr("blk.%d.attn_norm.weight")
    m.wq = arr("blk.%d.attn_q.weight")
    m.wk = arr("blk.%d.attn_k.weight")
    m.wv = arr("blk.%d.attn_v.weight")
    m.wo = arr("blk.%d.attn_output.weight")
    m.ffn_norm = arr("blk.%d.ffn_norm.weight")
    m.ffn_gate = arr("blk.%d.ffn_gate.weight")
    m.ffn_down = arr("blk.%d.ffn_down.weight")
    m.ffn_up = arr("blk.%d.ffn_up.weight")

    def types(fmt):
        a = (C.c_int32 * L)(*[int(t[fmt % i].tensor_type) for i in range(L)])
        keep.append(a)
        return C.cast(a, C.POINTER(C.c_int32))

    # per-tensor types: files made without --pure mix them (Q4_K_M: Q6_K output.weight ...)
    m.t_wq, m.t_wk, m.t_wv, m.t_wo = types("blk.%d.attn_q.weight"), types("blk.%d.attn_k.weight"), types("blk.%d.attn_v.weight"), types("blk.%d.attn_output.weight")
    m.t_gate, m.t_down, m.t_up = types("blk.%d.ffn_gate.weight"), types("blk.%d.ffn_down.weight"), types("blk.%d.ffn_up.weight")
    m.t_output = int(t["output.weight"].tensor_type)
    m._keep = keep
    m._reader = r
    return m


def seq_ptrs(seq_lists):
    """list of per-token seq-id lists -> (n_seq_id int32 array, array of int32* , keepalive)"""
    n = len(seq_lists)
    ns = np.array([len(s) for s in seq_lists], dtype=np.int32)
    rows = [np.array(s, dtype=np.int32) for s in seq_lists]
    ptrs = (C.c_void_p * n)(*[r.ctypes.data for r in rows])
    return ns, ptrs, rows


class OracleContext:
    def __init__(self, oracle, model, n_ctx, max_seq):
        self.o, self.model = oracle, model
        self.ctx = oracle.L.po_context_new(C.byref(model), n_ctx, max_seq)
        self.n_ctx = n_ctx

    def decode(self, tokens, pos, seq_lists):
        n = len(tokens)
        tok = np.array(tokens, dtype=np.int32)
        ps = np.array(pos, dtype=np.int32)
        ns, ptrs, keep = seq_ptrs(seq_lists)
        logits = np.empty((n, self.model.n_vocab), dtype=np.float32)
        rc = self.o.L.po_decode(self.ctx, n, _ptr(tok), _ptr(ps), _ptr(ns), ptrs, _ptr(logits))
        return rc, logits

    @property
    def kv(self):
        return self.ctx.contents.kv

    def kv_state(self):
        kv = self.kv.contents
        pos = np.ctypeslib.as_array(kv.pos, shape=(kv.size,)).copy()
        bits = np.ctypeslib.as_array(kv.bits, shape=(kv.size, kv.n_words)).copy()
        return dict(head=kv.head, used=kv.used, n=kv.n, pos=pos, bits=bits)

    def seq_rm(self, s, p0, p1):
        self.o.L.po_kv_seq_rm(self.kv, s, p0, p1)

    def seq_cp(self, a, b, p0, p1):
        self.o.L.po_kv_seq_cp(self.kv, a, b, p0, p1)

    def seq_keep(self, s):
        self.o.L.po_kv_seq_keep(self.kv, s)

    def close(self):
        if self.ctx:
            self.o.L.po_context_free(self.ctx)
            self.ctx = None


def kv_state_to_cells(state):
    """{cell: (pos, [seq ids])} for non-empty cells -- the form ref_decode_dump prints."""
    out = {}
    for i in range(len(state["pos"])):
        ids = [w * 32 + b for w in range(state["bits"].shape[1]) for b in range(32) if (int(state["bits"][i, w]) >> b) & 1]
        if ids:
            out[i] = (int(state["pos"][i]), ids)
    return out


def parse_ref_kv_line(line):
    parts = line.split()
    assert parts[0] == "KV"
    used = int(parts[1])
    cells = {}
    for p in parts[2:]:
        i, pos, ids = p.split(":")
        cells[int(i)] = (int(pos), [int(x) for x in ids.split(",")])
    return used, cells

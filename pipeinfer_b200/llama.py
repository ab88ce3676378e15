"""ctypes mirror of include/pipeinfer_llama.h (runtime level): llama_decode / llama_batch / llama_kv_cache_seq_* ...
Same names and argument meaning as the reference's llama.h so that tests read like tests against the reference."""
import ctypes as C

import numpy as np

from .lib import load, PibError


class TokenData(C.Structure):
    _fields_ = [("id", C.c_int32), ("logit", C.c_float), ("p", C.c_float)]


class Batch(C.Structure):
    _fields_ = [("n_tokens", C.c_int32), ("token", C.POINTER(C.c_int32)), ("embd", C.POINTER(C.c_float)),
                ("pos", C.POINTER(C.c_int32)), ("n_seq_id", C.POINTER(C.c_int32)),
                ("seq_id", C.POINTER(C.POINTER(C.c_int32))), ("logits", C.POINTER(C.c_int8)),
                ("all_pos_0", C.c_int32), ("all_pos_1", C.c_int32), ("all_seq_id", C.c_int32),
                ("batch_id", C.c_int32), ("max_n_seq", C.c_int32)]


class ModelParams(C.Structure):
    _fields_ = [("n_gpu_layers", C.c_int32), ("main_gpu", C.c_int32), ("tensor_split", C.c_void_p),
                ("progress_callback", C.c_void_p), ("progress_callback_user_data", C.c_void_p),
                ("vocab_only", C.c_bool), ("use_mmap", C.c_bool), ("use_mlock", C.c_bool)]


class ContextParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("n_ctx", C.c_uint32), ("n_batch", C.c_uint32), ("n_threads", C.c_uint32),
                ("n_threads_batch", C.c_uint32), ("rope_scaling_type", C.c_int8), ("rope_freq_base", C.c_float),
                ("rope_freq_scale", C.c_float), ("yarn_ext_factor", C.c_float), ("yarn_attn_factor", C.c_float),
                ("yarn_beta_fast", C.c_float), ("yarn_beta_slow", C.c_float), ("yarn_orig_ctx", C.c_uint32),
                ("mul_mat_q", C.c_bool), ("f16_kv", C.c_bool), ("logits_all", C.c_bool), ("embedding", C.c_bool)]


class KvView(C.Structure):
    _fields_ = [("n_cells", C.c_int32), ("n_max_seq", C.c_int32), ("token_count", C.c_int32), ("used_cells", C.c_int32),
                ("max_contiguous", C.c_int32), ("max_contiguous_idx", C.c_int32), ("cells", C.POINTER(C.c_int32)),
                ("cells_sequences", C.POINTER(C.c_int32))]


_bound = False


def api():
    """the CDLL with argtypes/restypes of the llama_* entry points set"""
    global _bound
    L = load().L
    if _bound:
        return L
    L.llama_model_default_params.restype = ModelParams
    L.llama_context_default_params.restype = ContextParams
    L.llama_backend_init.argtypes = [C.c_bool]
    L.llama_load_model_from_file.restype = C.c_void_p
    L.llama_load_model_from_file.argtypes = [C.c_char_p, ModelParams]
    L.llama_pib_synthetic_model.restype = C.c_void_p
    L.llama_pib_synthetic_model.argtypes = [C.c_int] * 7 + [C.c_uint64, C.c_int, C.c_float]
    L.llama_free_model.argtypes = [C.c_void_p]
    L.llama_new_context_with_model.restype = C.c_void_p
    L.llama_new_context_with_model.argtypes = [C.c_void_p, ContextParams]
    L.llama_pib_new_context.restype = C.c_void_p
    L.llama_pib_new_context.argtypes = [C.c_void_p, C.c_uint32, C.c_int32]
    L.llama_free.argtypes = [C.c_void_p]
    L.llama_n_vocab.argtypes = [C.c_void_p]
    L.llama_n_ctx.argtypes = [C.c_void_p]
    L.llama_n_embd.argtypes = [C.c_void_p]
    L.llama_get_model.restype = C.c_void_p
    L.llama_get_model.argtypes = [C.c_void_p]
    L.llama_tokenize.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_bool, C.c_bool]
    L.llama_token_to_piece.argtypes = [C.c_void_p, C.c_int32, C.c_char_p, C.c_int]
    L.llama_decode.argtypes = [C.c_void_p, Batch]
    L.llama_start_async_decode.restype = C.c_void_p
    L.llama_start_async_decode.argtypes = [C.c_void_p, Batch]
    L.llama_finish_async_decode.argtypes = [C.c_void_p, Batch, C.c_void_p]
    L.llama_mpi_iprobe.restype = C.c_bool
    L.llama_mpi_iprobe.argtypes = [C.c_void_p]
    L.llama_cancel_run.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int]
    L.llama_get_logits_ith.restype = C.POINTER(C.c_float)
    L.llama_get_logits_ith.argtypes = [C.c_void_p, C.c_int32]
    L.llama_kv_cache_clear.argtypes = [C.c_void_p]
    L.llama_kv_cache_seq_rm.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.llama_kv_cache_seq_cp.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    L.llama_kv_cache_seq_keep.argtypes = [C.c_void_p, C.c_int32]
    L.llama_get_kv_cache_used_cells.argtypes = [C.c_void_p]
    L.llama_get_kv_cache_token_count.argtypes = [C.c_void_p]
    L.llama_kv_cache_view_init.restype = KvView
    L.llama_kv_cache_view_init.argtypes = [C.c_void_p, C.c_int32]
    L.llama_kv_cache_view_update.argtypes = [C.c_void_p, C.POINTER(KvView)]
    L.llama_kv_cache_view_free.argtypes = [C.POINTER(KvView)]
    L.llama_pib_get_top8.argtypes = [C.c_void_p, C.c_int32, C.POINTER(TokenData)]
    L.llama_pib_set_logit_bias.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_int]
    L.llama_pib_kv_state.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.llama_pib_kv_words.argtypes = [C.c_void_p]
    L.llama_pib_kernel_launches.restype = C.c_uint64
    L.llama_pib_kernel_launches.argtypes = [C.c_void_p]
    L.llama_pib_set_priority.argtypes = [C.c_void_p, C.c_int]
    L.llama_pib_weight_bytes.restype = C.c_size_t
    L.llama_pib_weight_bytes.argtypes = [C.c_void_p]
    _bound = True
    return L


def make_batch(tokens, pos, seq_lists, batch_id=0, max_n_seq=1):
    """llama_batch (by value) from python lists; returns (batch, keepalive)"""
    n = len(tokens)
    tok = (C.c_int32 * n)(*tokens)
    ps = (C.c_int32 * n)(*pos)
    ns = (C.c_int32 * n)(*[len(s) for s in seq_lists])
    rows = [(C.c_int32 * max(1, len(s)))(*s) for s in seq_lists]
    ptrs = (C.POINTER(C.c_int32) * n)(*[C.cast(r, C.POINTER(C.c_int32)) for r in rows])
    lg = (C.c_int8 * n)(*([1] * n))
    b = Batch()
    b.n_tokens = n
    b.token = C.cast(tok, C.POINTER(C.c_int32))
    b.pos = C.cast(ps, C.POINTER(C.c_int32))
    b.n_seq_id = C.cast(ns, C.POINTER(C.c_int32))
    b.seq_id = C.cast(ptrs, C.POINTER(C.POINTER(C.c_int32)))
    b.logits = C.cast(lg, C.POINTER(C.c_int8))
    b.batch_id = batch_id
    b.max_n_seq = max_n_seq
    return b, (tok, ps, ns, rows, ptrs, lg)


class Context:
    """thin object wrapper: a model + a context through the llama.h-compatible ABI"""

    def __init__(self, model_path=None, n_ctx=512, max_seq=64, synthetic=None):
        self.L = api()
        self.L.llama_backend_init(False)
        if synthetic is not None:
            self.model = self.L.llama_pib_synthetic_model(*synthetic)
        else:
            self.model = self.L.llama_load_model_from_file(model_path.encode(), self.L.llama_model_default_params())
        if not self.model:
            raise PibError(self.L.pib_last_error().decode())
        self.ctx = self.L.llama_pib_new_context(self.model, n_ctx, max_seq)
        if not self.ctx:
            raise PibError(self.L.pib_last_error().decode())
        self.n_vocab = self.L.llama_n_vocab(self.model)
        self.n_ctx = n_ctx
        self.max_seq = max_seq
        self._bid = 0

    def set_priority(self, high=True):
        if self.L.llama_pib_set_priority(self.ctx, 1 if high else 0) != 0:
            raise PibError("llama_pib_set_priority failed")

    def decode(self, tokens, pos, seq_lists):
        b, keep = make_batch(tokens, pos, seq_lists, self._bid, self.max_seq)
        self._bid += 1
        rc = self.L.llama_decode(self.ctx, b)
        if rc != 0:
            return rc, None
        out = np.empty((len(tokens), self.n_vocab), dtype=np.float32)
        for i in range(len(tokens)):
            p = self.L.llama_get_logits_ith(self.ctx, i)
            out[i] = np.ctypeslib.as_array(p, shape=(self.n_vocab,))
        return rc, out

    def start(self, tokens, pos, seq_lists, batch_id=None):
        if batch_id is None:
            batch_id = self._bid
            self._bid += 1
        b, keep = make_batch(tokens, pos, seq_lists, batch_id, self.max_seq)
        h = self.L.llama_start_async_decode(self.ctx, b)
        return h, b, keep

    def finish(self, h, b):
        return self.L.llama_finish_async_decode(self.ctx, b, h)

    def top8(self, i):
        arr = (TokenData * 8)()
        if self.L.llama_pib_get_top8(self.ctx, i, arr) != 0:
            return None
        return [(a.id, a.logit, a.p) for a in arr]

    def seq_rm(self, s, p0, p1):
        self.L.llama_kv_cache_seq_rm(self.ctx, s, p0, p1)

    def seq_cp(self, a, b, p0, p1):
        self.L.llama_kv_cache_seq_cp(self.ctx, a, b, p0, p1)

    def seq_keep(self, s):
        self.L.llama_kv_cache_seq_keep(self.ctx, s)

    def kv_state(self, device=False):
        W = self.L.llama_pib_kv_words(self.ctx)
        sc = np.zeros(3, np.int32)
        pos = np.zeros(self.n_ctx, np.int32)
        bits = np.zeros((self.n_ctx, W), np.uint32)
        rc = self.L.llama_pib_kv_state(self.ctx, 1 if device else 0, sc.ctypes.data, pos.ctypes.data, bits.ctypes.data)
        if rc != 0:
            raise PibError("llama_pib_kv_state failed")
        return dict(head=int(sc[0]), used=int(sc[1]), n=int(sc[2]), pos=pos, bits=bits)

    def kv_view(self):
        v = self.L.llama_kv_cache_view_init(self.ctx, self.max_seq)
        self.L.llama_kv_cache_view_update(self.ctx, C.byref(v))
        cells = {}
        for i in range(v.n_cells):
            ids = [v.cells_sequences[i * v.n_max_seq + k] for k in range(v.n_max_seq)]
            ids = [x for x in ids if x >= 0]
            if ids:
                cells[i] = (v.cells[i], ids)
        used = v.used_cells
        self.L.llama_kv_cache_view_free(C.byref(v))
        return used, cells

    def close(self):
        if self.ctx:
            self.L.llama_free(self.ctx)
            self.ctx = None
        if self.model:
            self.L.llama_free_model(self.model)
            self.model = None


# ---- scheduler level (include/pipeinfer_sched.h) ----------------------------------------------------------------------
class SpecParams(C.Structure):
    _fields_ = [("n_draft", C.c_int32), ("n_parallel", C.c_int32), ("p_accept", C.c_float), ("p_split", C.c_float),
                ("p_recovery", C.c_float), ("p_decay", C.c_float), ("n_predict", C.c_int32), ("ignore_eos", C.c_int32),
                ("max_spec_runs", C.c_int32)]


class SpecResult(C.Structure):
    _fields_ = [("n_predict", C.c_int32), ("n_drafted", C.c_int32), ("n_accept", C.c_int32), ("n_cancelled", C.c_int32),
                ("n_spec_runs", C.c_int32), ("n_itl", C.c_int32), ("t_encode_s", C.c_double), ("t_decode_s", C.c_double),
                ("decode_tok_per_s", C.c_double), ("ttft_s", C.c_double), ("itl_mean_s", C.c_double), ("itl_p50_s", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def _bind_sched(L):
    if getattr(L, "_pib_sched_bound", False):
        return
    L.pib_spec_default_params.restype = SpecParams
    L.pib_speculative_generate.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.POINTER(SpecParams),
                                           C.POINTER(C.c_int32), C.c_int, C.POINTER(SpecResult), C.POINTER(C.c_double), C.c_int]
    L.pib_speculative_orig_generate.argtypes = L.pib_speculative_generate.argtypes
    L.pib_plain_generate.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.c_int, C.POINTER(C.c_int32),
                                     C.POINTER(SpecResult), C.POINTER(C.c_double), C.c_int]
    L._pib_sched_bound = True


def spec_params(**kw):
    L = api()
    _bind_sched(L)
    p = L.pib_spec_default_params()
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def speculative_generate(tgt, dft, prompt_ids, params, classic=False):
    """tgt / dft: Context objects created with max_seq >= 1000 * n_parallel + 2.  Returns (token ids, result dict, itl_us).
    classic=True runs the synchronous speculative_orig loop instead of PipeInfer's asynchronous one."""
    L = api()
    _bind_sched(L)
    n = len(prompt_ids)
    pr = (C.c_int32 * n)(*prompt_ids)
    cap = params.n_predict + 64
    out = (C.c_int32 * cap)()
    itl = (C.c_double * cap)()
    res = SpecResult()
    fn = L.pib_speculative_orig_generate if classic else L.pib_speculative_generate
    k = fn(tgt.ctx, dft.ctx, pr, n, C.byref(params), out, cap, C.byref(res), itl, cap)
    if k < 0:
        raise PibError(L.pib_last_error().decode())
    return list(out[:min(k, cap)]), res.as_dict(), list(itl[:res.n_itl])


def plain_generate(ctx, prompt_ids, n_predict):
    L = api()
    _bind_sched(L)
    n = len(prompt_ids)
    pr = (C.c_int32 * n)(*prompt_ids)
    out = (C.c_int32 * (n_predict + 1))()
    itl = (C.c_double * (n_predict + 1))()
    res = SpecResult()
    k = L.pib_plain_generate(ctx.ctx, pr, n, n_predict, out, C.byref(res), itl, n_predict + 1)
    if k < 0:
        raise PibError(L.pib_last_error().decode())
    return list(out[:k]), res.as_dict(), list(itl[:res.n_itl])


def tokenize(ctx, text, add_bos=True):
    L = api()
    buf = (C.c_int32 * (len(text) * 4 + 8))()
    raw = text.encode()
    n = L.llama_tokenize(ctx.model, raw, len(raw), buf, len(buf), add_bos, True)
    if n < 0:
        raise PibError("llama_tokenize: " + (L.pib_last_error().decode() if n == -2**31 else "needs %d tokens" % -n))
    return list(buf[:n])


def detokenize(ctx, ids):
    L = api()
    out = b""
    b = C.create_string_buffer(64)
    for t in ids:
        n = L.llama_token_to_piece(ctx.model, t, b, 64)
        out += b.raw[:max(n, 0)]
    return out

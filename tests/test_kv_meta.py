"""CPU-only: the host side of the KV bookkeeping (KvMirror + the lazy, coalescing op-log front end in
pipeinfer_b200/csrc/kv_meta.cu) against the reference's llama_kv_cache_view dumps and against the oracle.
Uses the pib_kvtest_* hooks of libpipeinfer_b200.so, which make no CUDA call."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle_py as op
from tests.replay import replay

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
SO = os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so")
KVOP_CLEAR, KVOP_RM, KVOP_CP, KVOP_KEEP = 1, 2, 3, 4


class KvTest:
    def __init__(self, n_ctx, max_seq):
        if not os.path.exists(SO):
            import __graft_entry__
            __graft_entry__.build()
        self.L = C.CDLL(SO)
        self.L.pib_kvtest_new.restype = C.c_void_p
        self.L.pib_kvtest_new.argtypes = [C.c_uint32, C.c_int32]
        self.L.pib_kvtest_free.argtypes = [C.c_void_p]
        self.L.pib_kvtest_op.argtypes = [C.c_void_p] + [C.c_int] * 5
        self.L.pib_kvtest_pending.argtypes = [C.c_void_p]
        self.L.pib_kvtest_decode.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self.L.pib_kvtest_state.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.h = self.L.pib_kvtest_new(n_ctx, max_seq)
        self.n_ctx, self.W = n_ctx, (max_seq + 31) // 32

    def decode(self, tokens, pos, seqs):
        ps = np.array(pos, dtype=np.int32)
        ns, ptrs, keep = op.seq_ptrs(seqs)
        return self.L.pib_kvtest_decode(self.h, len(tokens), ps.ctypes.data, ns.ctypes.data, ptrs), None

    def seq_rm(self, s, p0, p1):
        self.L.pib_kvtest_op(self.h, KVOP_RM, s, 0, p0, p1)

    def seq_cp(self, a, b, p0, p1):
        self.L.pib_kvtest_op(self.h, KVOP_CP, a, b, p0, p1)

    def seq_keep(self, s):
        self.L.pib_kvtest_op(self.h, KVOP_KEEP, s, 0, 0, 0)

    def state(self):
        sc = np.zeros(3, np.int32); pos = np.zeros(self.n_ctx, np.int32); bits = np.zeros((self.n_ctx, self.W), np.uint32)
        self.L.pib_kvtest_state(self.h, sc.ctypes.data, pos.ctypes.data, bits.ctypes.data)
        return dict(head=int(sc[0]), used=int(sc[1]), n=int(sc[2]), pos=pos, bits=bits)

    def cells(self):
        st = self.state()
        return st["used"], op.kv_state_to_cells(st)


@pytest.mark.parametrize("stem", ["kvops", "decode_q8_0"])
def test_host_mirror_matches_reference_kv_view(golden_dir, stem):
    kv = KvTest(48, 8)
    n_dec, n_kv = replay(golden_dir, stem, kv.decode, kv.seq_rm, kv.seq_cp, kv.seq_keep, kv.cells)
    assert n_kv >= 8


def test_commit_loop_is_coalesced_and_identical_to_the_literal_loop(oracle):
    # speculative.cpp:603-612: for i in 1..max_seq: seq_rm(i, p0, p1); seq_cp(0, i, p0, p1)  ->  one KVOP_SWEEP
    rng = np.random.default_rng(3)
    n_ctx, max_seq = 64, 100
    for trial in range(20):
        kv = KvTest(n_ctx, max_seq)
        okv = oracle.L.po_kv_new(n_ctx, max_seq)
        pos_next = 0
        for _ in range(6):          # random population: tokens in random sequence subsets
            n = int(rng.integers(1, 7))
            seqs = [sorted(set(rng.integers(0, max_seq, int(rng.integers(1, 5))).tolist()) | ({0} if rng.random() < 0.6 else set()))
                    for _ in range(n)]
            ps = np.arange(pos_next, pos_next + n, dtype=np.int32); pos_next += n
            ns, ptrs, keep = op.seq_ptrs(seqs)
            assert kv.decode([0] * n, ps.tolist(), seqs)[0] == 0
            oracle.L.po_kv_find_slot(okv, n, ps.ctypes.data, ns.ctypes.data, ptrs)
            okv.contents.head = (okv.contents.head + n) % n_ctx
        p0, p1 = int(rng.integers(0, pos_next)), int(rng.integers(1, pos_next + 2))
        first, last = 1, int(rng.integers(1, max_seq))
        for i in range(first, last + 1):
            kv.seq_rm(i, p0, p1); kv.seq_cp(0, i, p0, p1)
            oracle.L.po_kv_seq_rm(okv, i, p0, p1); oracle.L.po_kv_seq_cp(okv, 0, i, p0, p1)
        assert kv.L.pib_kvtest_pending(kv.h) == 1            # collapsed into a single op
        st = kv.state()
        o = okv.contents
        assert (st["head"], st["used"]) == (o.head, o.used)
        assert np.array_equal(st["pos"], np.ctypeslib.as_array(o.pos, shape=(n_ctx,)))
        assert np.array_equal(st["bits"], np.ctypeslib.as_array(o.bits, shape=(n_ctx, o.n_words)))
        oracle.L.po_kv_free(okv)


def test_non_spec_run_loops_are_coalesced_and_identical(oracle):
    # speculative.cpp:1252-1268: for i in 0..max_seq: seq_rm(i, n, n+1)  ...decode...  for i in 1..max_seq: seq_cp(0, i, n, n+1)
    rng = np.random.default_rng(4)
    n_ctx, max_seq = 64, 70
    for trial in range(20):
        kv = KvTest(n_ctx, max_seq + 1)
        okv = oracle.L.po_kv_new(n_ctx, max_seq + 1)
        pos_next = 0
        for _ in range(5):
            n = int(rng.integers(1, 7))
            seqs = [sorted(set(rng.integers(0, max_seq + 1, int(rng.integers(1, 6))).tolist())) for _ in range(n)]
            ps = np.arange(pos_next, pos_next + n, dtype=np.int32); pos_next += n
            ns, ptrs, keep = op.seq_ptrs(seqs)
            assert kv.decode([0] * n, ps.tolist(), seqs)[0] == 0
            oracle.L.po_kv_find_slot(okv, n, ps.ctypes.data, ns.ctypes.data, ptrs)
            okv.contents.head = (okv.contents.head + n) % n_ctx
        n_past = int(rng.integers(0, pos_next))
        lo = int(rng.integers(0, 3)); hi = int(rng.integers(lo, max_seq + 1))
        for i in range(lo, hi + 1):
            kv.seq_rm(i, n_past, n_past + 1); oracle.L.po_kv_seq_rm(okv, i, n_past, n_past + 1)
        assert kv.L.pib_kvtest_pending(kv.h) == 1
        src = int(rng.integers(0, 2))
        for i in range(src + 1, hi + 1):
            kv.seq_cp(src, i, n_past - 2, -1); oracle.L.po_kv_seq_cp(okv, src, i, n_past - 2, -1)
        assert kv.L.pib_kvtest_pending(kv.h) <= 2
        st = kv.state()
        o = okv.contents
        assert (st["head"], st["used"]) == (o.head, o.used)
        assert np.array_equal(st["pos"], np.ctypeslib.as_array(o.pos, shape=(n_ctx,)))
        assert np.array_equal(st["bits"], np.ctypeslib.as_array(o.bits, shape=(n_ctx, o.n_words)))
        oracle.L.po_kv_free(okv)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_random_long_traces_match_the_oracle_cell_for_cell(oracle, seed):
    # differential test of the lazy, coalescing op-log front end + host mirror against the plain restatement of the reference's
    # llama_kv_cache_* (oracle, itself pinned on the reference's KV view): 1500 random ops with scheduler-sized id ranges, runs
    # of consecutive ids (which the log turns into ranged ops / sweeps), failed decodes, clears; state compared all along
    rng = np.random.default_rng(1000 + seed)
    n_ctx, max_seq = 96, 1002
    kv = KvTest(n_ctx, max_seq)
    okv = oracle.L.po_kv_new(n_ctx, max_seq)
    pos_next = 0

    def check():
        st, o = kv.state(), okv.contents
        assert (st["head"], st["used"]) == (o.head, o.used)
        assert np.array_equal(st["pos"], np.ctypeslib.as_array(o.pos, shape=(n_ctx,)))
        assert np.array_equal(st["bits"], np.ctypeslib.as_array(o.bits, shape=(n_ctx, o.n_words)))

    for it in range(1500):
        c = int(rng.integers(0, 12))
        if c < 4:                                                   # decode (with the reference's head heuristic and head advance)
            n = int(rng.integers(1, 9))
            seqs = [sorted(set(rng.integers(0, max_seq, int(rng.integers(1, 4))).tolist())) for _ in range(n)]
            ps = np.arange(pos_next, pos_next + n, dtype=np.int32)
            ns, ptrs, keep = op.seq_ptrs(seqs)
            o = okv.contents
            if o.head > o.used + 2 * n:
                o.head = 0
            ok_ref = oracle.L.po_kv_find_slot(okv, n, ps.ctypes.data, ns.ctypes.data, ptrs)
            rc, _ = kv.decode([0] * n, ps.tolist(), seqs)
            assert (rc == 0) == bool(ok_ref)
            if ok_ref:
                o.head = (o.head + n) % n_ctx if o.head + n < n_ctx else 0
                pos_next += n
        elif c < 7:                                                 # a run of consecutive ids: rm then cp (the commit pattern)
            a = int(rng.integers(0, max_seq - 40)); b = a + int(rng.integers(0, 40))
            p0 = int(rng.integers(-1, max(1, pos_next))); p1 = int(rng.integers(-1, pos_next + 2))
            src = int(rng.integers(0, 3))
            for i in range(a, b + 1):
                kv.seq_rm(i, p0, p1); oracle.L.po_kv_seq_rm(okv, i, p0, p1)
                if rng.random() < 0.8 or True:
                    kv.seq_cp(src, i, p0, p1); oracle.L.po_kv_seq_cp(okv, src, i, p0, p1)
        elif c < 9:
            s = int(rng.integers(-1, max_seq)); p0 = int(rng.integers(-1, max(1, pos_next))); p1 = int(rng.integers(-1, pos_next + 2))
            kv.seq_rm(s, p0, p1); oracle.L.po_kv_seq_rm(okv, s, p0, p1)
        elif c < 10:
            a, b = int(rng.integers(0, max_seq)), int(rng.integers(0, max_seq))
            kv.seq_cp(a, b, -1, -1); oracle.L.po_kv_seq_cp(okv, a, b, -1, -1)
        elif c < 11:
            s = int(rng.integers(0, 4))
            kv.seq_keep(s); oracle.L.po_kv_seq_keep(okv, s)
        else:
            kv.seq_rm(-1, -1, -1); oracle.L.po_kv_seq_rm(okv, -1, -1, -1)
            pos_next = 0
        if it % 7 == 0:
            check()
    check()
    oracle.L.po_kv_free(okv)


def test_ranged_operations_are_the_records_the_literal_loops_collapse_to():
    """llama_pib_kv_seq_rm_range / _cp_range / _sweep (what the scheduler issues) against the per-id loops of the reference's
    scheduler (speculative.cpp:603-612, 1252-1268) pushed one call at a time: identical op-log records, hence identical
    bookkeeping on the host mirror and on every device replica"""
    L = C.CDLL(SO)
    L.pib_kvtest_new.restype = C.c_void_p
    L.pib_kvtest_new.argtypes = [C.c_uint32, C.c_int32]
    L.pib_kvtest_op.argtypes = [C.c_void_p] + [C.c_int] * 5
    L.pib_kvtest_op_ranged.argtypes = [C.c_void_p] + [C.c_int] * 6
    L.pib_kvtest_dump_pending.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.pib_kvtest_free.argtypes = [C.c_void_p]
    max_seq, n_past, c0 = 3001, 77, 71
    lit, rng = L.pib_kvtest_new(256, max_seq + 1), L.pib_kvtest_new(256, max_seq + 1)

    def pending(h):
        out = np.zeros((64, 6), dtype=np.int32)
        n = L.pib_kvtest_dump_pending(h, out.ctypes.data, 64)
        return out[:n].tolist()

    # commit of an accepted speculative run
    for h in (lit, rng):
        L.pib_kvtest_op(h, KVOP_RM, 0, 0, c0, n_past)
        L.pib_kvtest_op(h, KVOP_CP, 1005, 0, c0, n_past)
    for i in range(1, max_seq + 1):
        L.pib_kvtest_op(lit, KVOP_RM, i, 0, c0, n_past)
        L.pib_kvtest_op(lit, KVOP_CP, 0, i, c0, n_past)
    L.pib_kvtest_op_ranged(rng, 7, 1, max_seq, 0, c0, n_past)                  # KVOP_SWEEP
    # the non-speculative run that follows: remove the position from every sequence, (decode,) copy it to every sequence
    for i in range(0, max_seq + 1):
        L.pib_kvtest_op(lit, KVOP_RM, i, 0, n_past, n_past + 1)
    L.pib_kvtest_op_ranged(rng, KVOP_RM, 0, max_seq, 0, n_past, n_past + 1)
    a, b = pending(lit), pending(rng)
    assert a == b and len(a) == 4, (a, b)
    for h in (lit, rng):
        L.pib_kvtest_free(h)
    lit, rng = L.pib_kvtest_new(256, max_seq + 1), L.pib_kvtest_new(256, max_seq + 1)
    for i in range(1, max_seq + 1):
        L.pib_kvtest_op(lit, KVOP_CP, 0, i, n_past, n_past + 1)
    L.pib_kvtest_op_ranged(rng, KVOP_CP, 0, 1, max_seq, n_past, n_past + 1)
    assert pending(lit) == pending(rng) and len(pending(rng)) == 1
    # a source inside the destination range: the literal loop skips src == dst (llama.cpp:9303-9305)
    L.pib_kvtest_op_ranged(rng, KVOP_CP, 5, 3, 8, 0, 10)
    assert [r[:3] + [r[5]] for r in pending(rng)[1:]] == [[KVOP_CP, 5, 3, 4], [KVOP_CP, 5, 6, 8]]
    for h in (lit, rng):
        L.pib_kvtest_free(h)
    # a sweep that starts below 1 (exported ABI; the scheduler never does it): the closed form assumes sequence 0 is the
    # untouched source, so id 0 must travel as its literal records -- same state as the per-id loop, on mirror and log alike
    L.pib_kvtest_state.argtypes = [C.c_void_p] * 4
    lit, rng = L.pib_kvtest_new(64, 40), L.pib_kvtest_new(64, 40)
    pos = (C.c_int32 * 6)(*range(6))
    nsq = (C.c_int32 * 6)(*([2] * 6))
    rows = [(C.c_int32 * 2)(0, 3) for _ in range(6)]
    sid = (C.POINTER(C.c_int32) * 6)(*[C.cast(r, C.POINTER(C.c_int32)) for r in rows])
    L.pib_kvtest_decode.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    for h in (lit, rng):
        assert L.pib_kvtest_decode(h, 6, pos, nsq, sid) == 0
    for i in range(0, 9):
        L.pib_kvtest_op(lit, KVOP_RM, i, 0, 2, 5)
        if i != 0:
            L.pib_kvtest_op(lit, KVOP_CP, 0, i, 2, 5)        # llama_kv_cache_seq_cp returns early for src == dst
    L.pib_kvtest_op_ranged(rng, 7, 0, 8, 0, 2, 5)
    assert all(r[0] != 7 or r[1] >= 1 for r in pending(rng))

    def state(h):
        sc, ps, bt = np.zeros(3, np.int32), np.zeros(64, np.int32), np.zeros((64, 2), np.uint32)
        L.pib_kvtest_state(h, sc.ctypes.data, ps.ctypes.data, bt.ctypes.data)
        return sc.tolist(), ps.tolist(), bt.tolist()

    assert state(lit) == state(rng)
    for h in (lit, rng):
        L.pib_kvtest_free(h)

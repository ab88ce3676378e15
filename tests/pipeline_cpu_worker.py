"""One rank of the CPU-only pipeline test (launched by tests/test_pipeline_cpu.py with torch.distributed.run, gloo,
world_size 2+).  Exercises what a multi-GPU run needs from the host: the control ring, the ggml_mpi_* shim on top of it,
the layer partition, and KV bookkeeping replicated from the head to the other stages through decode records."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from oracle import oracle_py as op          # noqa: E402
from tests.replay import replay             # noqa: E402
from tests.test_kv_meta import KvTest       # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, size = dist.get_rank(), dist.get_world_size()
    golden = os.path.join(ROOT, "tests", "golden")
    L = C.CDLL(os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so"))
    L.pib_last_error.restype = C.c_char_p
    L.pib_pipe_ring_open.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_size_t]
    L.ggml_mpi_init.restype = C.c_void_p
    L.ggml_mpi_split_range.restype = C.POINTER(C.POINTER(C.c_uint16))
    L.ggml_mpi_split_range.argtypes = [C.c_void_p, C.c_uint16, C.c_uint16, C.POINTER(C.c_float)]
    L.ggml_mpi_scatter_layers.argtypes = [C.c_void_p, C.c_void_p]
    L.ggml_mpi_layer_range.argtypes = [C.c_void_p, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16)]
    L.ggml_mpi_sync_ints_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.ggml_mpi_rank.argtypes = [C.c_void_p]
    L.ggml_mpi_size.argtypes = [C.c_void_p]
    L.ggml_mpi_size.restype = C.c_size_t
    L.ggml_mpi_next_node.argtypes = [C.c_void_p]
    L.ggml_mpi_prev_node.argtypes = [C.c_void_p]
    L.ggml_mpi_barrier.argtypes = [C.c_void_p]
    L.pib_kvtest_ship.argtypes = [C.c_void_p, C.c_int]
    L.pib_kvtest_replay.argtypes = [C.c_void_p]
    L.pib_kvtest_pause.argtypes = [C.c_void_p]

    name = [("/pib_test_%d" % os.getpid()) if rank == 0 else None]
    dist.broadcast_object_list(name, src=0)
    # a deliberately small ring (64 KiB) so that the wrap-around and back-pressure paths run
    assert L.pib_pipe_ring_open(name[0].encode(), rank, size, 64 * 1024) == 0, L.pib_last_error()
    assert L.pib_pipe_rank() == rank and L.pib_pipe_size() == size

    # ---- ggml_mpi_* shim: rank / size / neighbours / layer table / ordered int broadcast --------------------------------
    ctx = L.ggml_mpi_init()
    assert L.ggml_mpi_rank(ctx) == rank and L.ggml_mpi_size(ctx) == size
    assert L.ggml_mpi_next_node(ctx) == (rank + 1) % size and L.ggml_mpi_prev_node(ctx) == (rank - 1) % size
    n_layer = 80
    w = (C.c_float * size)(*([0.1] + [0.9 / (size - 1)] * (size - 1)))
    ranges = L.ggml_mpi_split_range(ctx, 0, n_layer - 1, w)
    assert bool(ranges) == (rank == 0)                       # only the head computes the split (ggml-mpi.c:536-538)
    L.ggml_mpi_scatter_layers(ctx, ranges if rank == 0 else None)
    lo, hi = C.c_uint16(), C.c_uint16()
    L.ggml_mpi_layer_range(ctx, C.byref(lo), C.byref(hi))
    exp = np.zeros(2 * size, np.uint16)
    op.Oracle().L.po_split_range(size, 0, n_layer - 1, w, exp.ctypes.data)
    assert (lo.value, hi.value) == (int(exp[2 * rank]), int(exp[2 * rank + 1])), (rank, lo.value, hi.value, exp)
    for tag, vals in ((13, [5]), (14, [7, 8, 9, 10, 11])):
        buf = np.array(vals if rank == 0 else [0] * len(vals), dtype=np.int32)
        L.ggml_mpi_sync_ints_pipelined(ctx, buf.ctypes.data, len(vals), tag)
        assert buf.tolist() == vals, (rank, tag, buf)
    # ggml_mpi_eval_init: the batch of a decode, head -> stages, into the stages' pre-allocated arrays
    libc = C.CDLL(None)
    libc.calloc.restype = C.c_void_p
    libc.calloc.argtypes = [C.c_size_t, C.c_size_t]
    L.ggml_mpi_eval_init.restype = C.c_bool
    L.ggml_mpi_eval_init.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(C.c_int32)),
                                     C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(C.POINTER(C.c_int32))),
                                     C.POINTER(C.POINTER(C.c_int8)), C.POINTER(C.c_int32), C.c_bool]
    L.ggml_mpi_synch_float.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int]
    cap, max_ids = 16, 4
    want_tok, want_pos = [11, 22, 33, 44, 55], [7, 8, 9, 8, 9]
    want_seq = [[1, 2], [1], [1], [2], [2, 3, 0]]
    want_lg = [0, 0, 1, 0, 1]
    tok = (C.c_int32 * cap)(*(want_tok if rank == 0 else []))
    pos = (C.c_int32 * cap)(*(want_pos if rank == 0 else []))
    nsq = (C.c_int32 * cap)(*([len(x) for x in want_seq] if rank == 0 else []))
    rows = [(C.c_int32 * max_ids)(*(want_seq[i] if rank == 0 and i < len(want_seq) else [])) for i in range(cap)]
    sid = (C.POINTER(C.c_int32) * cap)(*[C.cast(r, C.POINTER(C.c_int32)) for r in rows])
    lgp = C.cast(libc.calloc(cap, 1), C.POINTER(C.c_int8))            # malloc'd: stages free and replace it (ggml-mpi.c:270-290)
    if rank == 0:
        for i, v in enumerate(want_lg):
            lgp[i] = v
    n_tok, bid = C.c_int32(len(want_tok) if rank == 0 else 0), C.c_int32(77 if rank == 0 else 0)
    ptok, ppos, pnsq = C.cast(tok, C.POINTER(C.c_int32)), C.cast(pos, C.POINTER(C.c_int32)), C.cast(nsq, C.POINTER(C.c_int32))
    psid = C.cast(sid, C.POINTER(C.POINTER(C.c_int32)))
    assert L.ggml_mpi_eval_init(ctx, C.byref(n_tok), C.byref(ptok), C.byref(ppos), C.byref(pnsq), C.byref(psid), C.byref(lgp), C.byref(bid), False)
    assert (n_tok.value, bid.value) == (5, 77)
    assert list(tok[:5]) == want_tok and list(pos[:5]) == want_pos and list(nsq[:5]) == [len(x) for x in want_seq]
    assert [list(rows[i][:len(want_seq[i])]) for i in range(5)] == want_seq
    assert [int(lgp[i]) for i in range(5)] == want_lg
    f = C.c_float(3.25 if rank == 0 else 0.0)
    L.ggml_mpi_synch_float(ctx, C.byref(f), 0)
    assert f.value == 3.25
    L.ggml_mpi_barrier(ctx)

    # ---- KV bookkeeping replicated through decode records ----------------------------------------------------------------
    for stem, n_ctx, max_seq in (("kvops", 48, 8), ("decode_q8_0", 48, 8)):
        kv = KvTest(n_ctx, max_seq)
        if rank == 0:
            L.pib_kvtest_ship(kv.h, 1)
            replay(golden, stem, kv.decode, kv.seq_rm, kv.seq_cp, kv.seq_keep, kv.cells)      # also checks vs the reference
            assert L.pib_kvtest_pause(kv.h) == 0
        else:
            assert L.pib_kvtest_replay(kv.h) == 0, L.pib_last_error()
        st = kv.state()
        flat = np.concatenate([[st["head"], st["used"], st["n"]], st["pos"], st["bits"].view(np.int32).ravel()]).astype(np.int64)
        mine = torch.from_numpy(flat)
        alls = [torch.zeros_like(mine) for _ in range(size)]
        dist.all_gather(alls, mine)
        for r in range(size):
            assert torch.equal(alls[r], alls[0]), "%s: KV replica of rank %d differs from the head's" % (stem, r)

    # ---- a long random trace: many records through the small ring (wrap-around, consumers slower than the producer) ------
    rng = np.random.default_rng(7)
    kv = KvTest(256, 3002)                                   # -np 3 sized bitsets (94 words)
    if rank == 0:
        L.pib_kvtest_ship(kv.h, 1)
        pos_next = 0
        for it in range(300):
            n = int(rng.integers(1, 9))
            seqs = [sorted(set(rng.integers(0, 3002, int(rng.integers(1, 4))).tolist()) | {0}) for _ in range(n)]
            rc, _ = kv.decode([0] * n, list(range(pos_next, pos_next + n)), seqs)
            if rc != 0:                                      # cache full: drop everything past position 32 and go on
                kv.seq_rm(-1, 32, -1)
                pos_next = 32
                continue
            pos_next += n
            if it % 5 == 4:
                for s in range(1, 40):                       # the scheduler's commit loop (coalesced into one SWEEP)
                    kv.seq_rm(s, pos_next - 3, -1)
                    kv.seq_cp(0, s, pos_next - 3, -1)
            if it % 11 == 10:
                kv.seq_keep(0)
        assert L.pib_kvtest_pause(kv.h) == 0
    else:
        assert L.pib_kvtest_replay(kv.h) == 0, L.pib_last_error()
    st = kv.state()
    mine = torch.from_numpy(np.concatenate([[st["head"], st["used"], st["n"]], st["pos"], st["bits"].view(np.int32).ravel()]).astype(np.int64))
    alls = [torch.zeros_like(mine) for _ in range(size)]
    dist.all_gather(alls, mine)
    assert all(torch.equal(a, alls[0]) for a in alls), "random trace: replicas differ"
    assert int(alls[0][1]) > 0

    L.pib_pipe_ring_close()
    dist.barrier()
    dist.destroy_process_group()
    print("PIPELINE_CPU_OK rank %d" % rank)


if __name__ == "__main__":
    main()

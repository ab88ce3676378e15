#!/usr/bin/env python
"""Where the wall time of one bench step goes outside the decode loop: KV clears, prompt ingestion (t_encode_s), decode (t_decode_s),
and what is left (drain of the runs in flight at the end)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from pipeinfer_b200 import llama
(L, E, H, HKV, FF, V, wt), dshape, label = bench.CONFIGS["70b"]
lib = llama.api()
cap = 3002
tgt = llama.Context(None, n_ctx=1024, max_seq=cap, synthetic=(L, E, H, HKV, FF, V, wt, 1, 1, 1.0))
dft = llama.Context(None, n_ctx=1024, max_seq=cap, synthetic=dshape + (2, 1, 0.8))
dft.set_priority(True)
prompt = [1] + [3 + (i * 7919) % (V - 3) for i in range(63)]
params = llama.spec_params(n_draft=3, n_parallel=3, p_accept=0.1, p_split=0.9, n_predict=128, max_spec_runs=2)
for it in range(5):
    t0 = time.perf_counter()
    lib.llama_kv_cache_clear(tgt.ctx); lib.llama_kv_cache_clear(dft.ctx)
    t1 = time.perf_counter()
    toks, res, itl = llama.speculative_generate(tgt, dft, prompt, params)
    t2 = time.perf_counter()
    print("clear %.2f ms  generate %.1f ms = encode %.1f + decode %.1f + rest %.1f ms   ttft %.1f ms" % (
        (t1 - t0) * 1e3, (t2 - t1) * 1e3, res["t_encode_s"] * 1e3, res["t_decode_s"] * 1e3,
        (t2 - t1 - res["t_encode_s"] - res["t_decode_s"]) * 1e3, res["ttft_s"] * 1e3), flush=True)

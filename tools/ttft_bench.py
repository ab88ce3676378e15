#!/usr/bin/env python
"""Time to ingest a prompt of --n tokens (one llama_decode) with the tensor-core prompt path on and off:
  python tools/ttft_bench.py --config 70b --n 512 [--layers 80]"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="70b"); ap.add_argument("--n", type=int, default=512); ap.add_argument("--layers", type=int, default=0)
a = ap.parse_args()
(L, E, H, HKV, FF, V, wt), dshape, label = bench.CONFIGS[a.config]
if a.layers: L = a.layers
from pipeinfer_b200 import llama
toks = [3 + (i * 7919) % (V - 3) for i in range(a.n)]
for mode in ("1", "0"):
    os.environ["PIB_PROMPT_GEMM"] = mode
    ctx = llama.Context(None, n_ctx=max(1024, a.n + 16), max_seq=4, synthetic=(L, E, H, HKV, FF, V, wt, 1, 1, 1.0))
    lib = llama.api()
    best = 1e9
    for rep in range(3):
        lib.llama_kv_cache_clear(ctx.ctx)
        b, keep = llama.make_batch(toks, list(range(a.n)), [[0]] * a.n, 100 + rep, 4)
        t0 = time.perf_counter()
        assert lib.llama_decode(ctx.ctx, b) == 0
        best = min(best, time.perf_counter() - t0)
    print("%s, %d layers, %d-token prompt, %s: %.1f ms  (%.0f prompt tokens/s)" % (label, L, a.n, "tcgen05 GEMM, 256 tokens per pass" if mode == "1" else "GEMV in chunks of 16", best * 1e3, a.n / best), flush=True)
    ctx.close()

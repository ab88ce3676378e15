#!/usr/bin/env python
"""Plain (non-speculative) greedy decoding speed of one model on one GPU: python tools/plain_bench.py --which draft|target [--config 70b]"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from pipeinfer_b200 import llama
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="70b"); ap.add_argument("--which", default="draft"); ap.add_argument("-n", type=int, default=128)
a = ap.parse_args()
tshape, dshape, label = bench.CONFIGS[a.config]
shape = dshape if a.which == "draft" else tshape
ctx = llama.Context(None, n_ctx=1024, max_seq=8, synthetic=shape + (1, 1, 1.0))
prompt = [1] + [3 + (i * 7919) % (shape[5] - 3) for i in range(63)]
for rep in range(3):
    ctx.L.llama_kv_cache_clear(ctx.ctx)
    t0 = time.time()
    toks, res, itl = llama.plain_generate(ctx, prompt, a.n)
    dt = time.time() - t0
    itl = sorted(itl)
    print("%s %s: %d tokens, decode %.1f tok/s, p50 %.3f ms/token, wall %.3f s" % (a.config, a.which, len(toks), res["decode_tok_per_s"], itl[len(itl) // 2] / 1000.0, dt))

#!/usr/bin/env python
"""One target verification pass of the bench workload inside a cudaProfilerStart/Stop range, for ncu:

  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
      python tools/profile_step.py --config 70b --batch 4
  ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv -c 3 -o gpurun_out/prof \
      python tools/profile_step.py --config 70b --batch 4 --layers 2
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402
from pipeinfer_b200 import llama  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="70b")
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--layers", type=int, default=0, help="override the number of target layers (0 = config)")
ap.add_argument("--passes", type=int, default=1)
ap.add_argument("--ctx", type=int, default=16, help="tokens already in the cache (multiple of 16)")
a = ap.parse_args()
(L, E, H, HKV, FF, V, wt), dshape, label = bench.CONFIGS[a.config]
if a.layers:
    L = a.layers
lib = llama.api()
tgt = llama.Context(None, n_ctx=1024, max_seq=64, synthetic=(L, E, H, HKV, FF, V, wt, 1, 1, 1.0))
C0 = a.ctx
prompt = [1] + [3 + (i * 7919) % (V - 3) for i in range(C0 + 63)]
for c0 in range(0, C0, 16):
    rc, _ = tgt.decode(prompt[c0:c0 + 16], list(range(c0, c0 + 16)), [[0]] * 16)
    assert rc == 0
b, keep = llama.make_batch(prompt[C0:C0 + a.batch], list(range(C0, C0 + a.batch)), [[0]] * a.batch, 7, 64)
assert lib.llama_decode(tgt.ctx, b) == 0                      # warm
lib.llama_kv_cache_seq_rm(tgt.ctx, 0, C0, -1)
lib.pib_profiler_start()
for p in range(a.passes):
    b, keep = llama.make_batch(prompt[C0:C0 + a.batch], list(range(C0, C0 + a.batch)), [[0]] * a.batch, 8 + p, 64)
    assert lib.llama_decode(tgt.ctx, b) == 0
    lib.llama_kv_cache_seq_rm(tgt.ctx, 0, C0, -1)
lib.pib_profiler_stop()
print("profiled %d pass(es): %s, %d layers, batch %d" % (a.passes, label, L, a.batch))

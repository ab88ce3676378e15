#!/usr/bin/env python
"""Per-kernel timing of one target layer (CUDA events on the launching stream, nothing else running on the GPU):
  python tools/layer_bench.py --config 70b --batch 4 --passes 20 [--layers 8]"""
import argparse, ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from pipeinfer_b200 import llama
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="70b"); ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--passes", type=int, default=20); ap.add_argument("--layers", type=int, default=0)
ap.add_argument("--ctx", type=int, default=128)
a = ap.parse_args()
(L, E, H, HKV, FF, V, wt), dshape, label = bench.CONFIGS[a.config]
if a.layers: L = a.layers
lib = llama.api()
lib.llama_pib_profile_layer.argtypes = [C.c_void_p, C.c_int]
lib.llama_pib_profile_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
lib.llama_pib_matrix_shape.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
tgt = llama.Context(None, n_ctx=1024, max_seq=64, synthetic=(L, E, H, HKV, FF, V, wt, 1, 1, 1.0))
toks = [3 + (i * 7919) % (V - 3) for i in range(1024)]
pos = 0
while pos < a.ctx:                                   # fill the context
    n = min(16, a.ctx - pos)
    assert tgt.decode(toks[pos:pos + n], list(range(pos, pos + n)), [[0]] * n)[0] == 0
    pos += n
lib.llama_pib_profile_layer(tgt.ctx, L // 2)
t0 = time.time()
for p in range(a.passes):
    b, keep = llama.make_batch(toks[pos:pos + a.batch], list(range(pos, pos + a.batch)), [[0]] * a.batch, 100 + p, 64)
    assert lib.llama_decode(tgt.ctx, b) == 0
    lib.llama_kv_cache_seq_rm(tgt.ctx, 0, pos, -1)
dt = (time.time() - t0) / a.passes
n = 8192
kind, bb, ms = (C.c_int32 * n)(), (C.c_int32 * n)(), (C.c_float * n)()
cnt = lib.llama_pib_profile_read(tgt.ctx, kind, bb, ms, n)
agg = {}
for i in range(cnt):
    agg.setdefault(kind[i], []).append(ms[i])
tot = 0.0
print("%s, %d layers, batch %d, context %d: %.3f ms per pass (host wall)" % (label, L, a.batch, a.ctx, dt * 1e3))
for k in sorted(agg):
    v = sorted(agg[k]); med = v[len(v) // 2] * 1000
    extra = ""
    if k <= 4:
        N_, K_, by = C.c_int64(), C.c_int64(), C.c_uint64()
        lib.llama_pib_matrix_shape(tgt.model, L // 2, k, C.byref(N_), C.byref(K_), C.byref(by))
        extra = "  %6.0f GB/s (%d x %d)" % (by.value / med / 1e3, N_.value, K_.value)
    if k != 4: tot += med
    print("  %-22s median %7.2f us%s" % (bench.KIND[k], med, extra))
print("  layer total (sum of medians, lm_head excluded): %.1f us -> %.2f ms for %d layers" % (tot, tot * L / 1000, L))

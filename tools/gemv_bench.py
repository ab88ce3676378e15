#!/usr/bin/env python
"""Isolated timing of one quantised matmul shape through the C ABI (pib_gemv), weights generated on the device:
  python tools/gemv_bench.py --type q4_k --N 57344 --K 8192 --b 1,4 --variant 0,5 [--reps 50]
Rotates over enough copies of the matrix to exceed the 126 MB L2."""
import argparse, ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from pipeinfer_b200 import load
ap = argparse.ArgumentParser()
ap.add_argument("--type", default="q4_k"); ap.add_argument("--N", type=int, default=57344); ap.add_argument("--K", type=int, default=8192)
ap.add_argument("--b", default="1,4"); ap.add_argument("--variant", default="0"); ap.add_argument("--reps", type=int, default=50)
a = ap.parse_args()
wt = {"q4_k": 12, "q4_0": 2, "q8_0": 8, "q6_k": 14, "q5_k": 13, "q3_k": 11}[a.type]
bb = {12: 144, 2: 18, 8: 34, 14: 210, 13: 176, 11: 110}[wt]; be = {12: 256, 2: 32, 8: 32, 14: 256, 13: 256, 11: 256}[wt]
lib = load(); lib.init(0)
rng = np.random.default_rng(0)
wbytes = a.N * a.K // be * bb
ncopies = max(1, int(300e6 // wbytes) + 1)
blk = rng.integers(0, 256, (a.N, a.K // be, bb), dtype=np.uint8)
dpos = 208 if wt == 14 else 108 if wt == 11 else 0
blk[:, :, dpos:dpos + 2] = (rng.uniform(0.5, 1.5, (a.N, a.K // be)) * 0.0025).astype(np.float16).view(np.uint8).reshape(a.N, a.K // be, 2)
if wt in (12, 13):
    blk[:, :, 2:4] = (rng.uniform(0.5, 1.5, (a.N, a.K // be)) * 0.001).astype(np.float16).view(np.uint8).reshape(a.N, a.K // be, 2)
hs = [lib.weight_upload(wt, blk.reshape(a.N, -1), a.K) for _ in range(ncopies)]
lib.L.pib_gemm_scratch_bytes.restype = C.c_size_t
lib.L.pib_gemm_scratch_bytes.argtypes = [C.c_int64, C.c_int]
lib.L.pib_gemm.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
lib.L.pib_quantize_act_rows.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
for b in [int(x) for x in a.b.split(",")]:
    if b > 16:                                   # wide batches: the tcgen05 GEMM (Q4_K)
        x = rng.standard_normal((b, a.K)).astype(np.float32)
        xd = lib.to_device(x)
        act = lib.alloc(lib.L.pib_act_bytes(wt, a.K, b)); yd = lib.alloc(b * a.N * 4); sc = lib.alloc(lib.L.pib_gemm_scratch_bytes(a.K, b))
        lib.check(lib.L.pib_quantize_act_rows(wt, xd, b, a.K, act, None))
        for i in range(3):
            lib.check(lib.L.pib_gemm(hs[i % ncopies], act, b, yd, sc, None))
        lib.check(lib.L.pib_stream_sync(None))
        t0 = time.perf_counter()
        for i in range(a.reps):
            lib.L.pib_gemm(hs[i % ncopies], act, b, yd, sc, None)
        lib.check(lib.L.pib_stream_sync(None))
        dt = (time.perf_counter() - t0) / a.reps
        print("%s %dx%d b=%d tcgen05 gemm: %.1f us  %.1f TFLOP/s  (weights at %.0f GB/s)" % (a.type, a.N, a.K, b, dt * 1e6, 2.0 * a.N * a.K * b / dt / 1e12, wbytes / dt / 1e9), flush=True)
        continue
    x = rng.standard_normal((b, a.K)).astype(np.float32)
    xd = lib.to_device(x)
    awt = 12 if wt in (11, 13, 14) else wt                       # the K-quants share the Q8_K activation layout
    act = lib.alloc(lib.L.pib_act_bytes(awt, a.K, b))
    yd = lib.alloc(b * a.N * 4)
    lib.check(lib.L.pib_quantize_act(awt, xd, b, a.K, act, None))
    for v in [int(x) for x in a.variant.split(",")]:
        for i in range(5):
            lib.check(lib.L.pib_gemv(hs[i % ncopies], act, b, yd, v, None))
        lib.check(lib.L.pib_stream_sync(None))
        t0 = time.perf_counter()
        for i in range(a.reps):
            lib.L.pib_gemv(hs[i % ncopies], act, b, yd, v, None)
        lib.check(lib.L.pib_stream_sync(None))
        dt = (time.perf_counter() - t0) / a.reps
        print("%s %dx%d b=%d variant %d: %.2f us  %.0f GB/s" % (a.type, a.N, a.K, b, v, dt * 1e6, wbytes / dt / 1e9), flush=True)

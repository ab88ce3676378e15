#!/usr/bin/env python
"""Where a verification pass spends its time when it runs the way the bench runs it (CUDA graph, programmatic dependent launch):
the pass is timed whole, then again with every launch of one kind left out (PIB_DEBUG_SKIP_KINDS, a debugging switch of the
engine; the results of such a pass are garbage).  The difference is what that kind costs INSIDE the pipelined pass, which is
what event pairs around single launches cannot show.

  python tools/pass_breakdown.py --config 70b --batch 1,2,4
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)

KINDS = ["qkv", "attn_out", "gate_up", "down", "lm_head", "attn_norm_quant", "-", "attention", "attn_out_quant", "resid_ffn_norm_quant", "silu_mul_quant"]


def child(a):
    import bench
    from pipeinfer_b200 import llama
    (L, E, H, HKV, FF, V, wt), dshape, label = bench.CONFIGS[a.config]
    if a.draft:
        L, E, H, HKV, FF, V, wt = dshape
    lib = llama.api()
    tgt = llama.Context(None, n_ctx=1024, max_seq=64, synthetic=(L, E, H, HKV, FF, V, wt, 1, 1, 1.0))
    prompt = [1] + [3 + (i * 7919) % (V - 3) for i in range(255)]
    for c0 in range(0, a.ctx, 16):
        rc, _ = tgt.decode(prompt[c0:c0 + 16], list(range(c0, c0 + 16)), [[0]] * 16)
        assert rc == 0
    out = {}
    for b in a.batch:
        ts = []
        for p in range(a.passes + 3):
            bt, keep = llama.make_batch(prompt[a.ctx:a.ctx + b], list(range(a.ctx, a.ctx + b)), [[0]] * b, 7 + p, 64)
            t0 = time.perf_counter()
            assert lib.llama_decode(tgt.ctx, bt) == 0          # synchronous: returns when the pass's results are on the host
            t1 = time.perf_counter()
            lib.llama_kv_cache_seq_rm(tgt.ctx, 0, a.ctx, -1)
            if p >= 3:
                ts.append((t1 - t0) * 1e3)
        ts.sort()
        out[str(b)] = ts[len(ts) // 2]
    print("RESULT " + json.dumps(out))


ap = argparse.ArgumentParser()
ap.add_argument("--config", default="70b")
ap.add_argument("--batch", default="1,4")
ap.add_argument("--ctx", type=int, default=64, help="tokens already in the cache")
ap.add_argument("--passes", type=int, default=15)
ap.add_argument("--child", action="store_true")
ap.add_argument("--draft", action="store_true", help="time the config's draft model instead of its target")
ap.add_argument("--kinds", default="", help="with --alone: only these kinds (comma separated names)")
ap.add_argument("--alone", action="store_true", help="instead: time a pass that launches ONLY one kind (a chain of n_layer such kernels)")
a = ap.parse_args()
a.batch = [int(x) for x in a.batch.split(",")]
if a.child:
    child(a)
    sys.exit(0)


def run(mask):
    env = dict(os.environ, PIB_DEBUG_SKIP_KINDS=str(mask))
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--config", a.config, "--batch", ",".join(map(str, a.batch)),
                        "--ctx", str(a.ctx), "--passes", str(a.passes)] + (["--draft"] if a.draft else []), env=env, capture_output=True, text=True)
    for line in r.stdout.splitlines():
        if line.startswith("RESULT "):
            return json.loads(line[7:])
    raise SystemExit(r.stdout + r.stderr)


if a.alone:
    ALL = 0b11110111111
    base = run(ALL)
    print("nothing but the fixed part (embedding rows, mask, final norm, lm_head, sampling, host): %s" % base)
    for k, name in enumerate(KINDS):
        if name == "-" or (a.kinds and name not in a.kinds.split(",")):
            continue
        r = run(ALL & ~(1 << k))
        print("%-22s alone: %s" % (name, {b: "%.3f ms" % (r[b] - base[b]) for b in r}))
    sys.exit(0)

full = run(0)
print("whole pass (ms, median of %d): %s" % (a.passes, full))
rows = {}
for k, name in enumerate(KINDS):
    if name == "-":
        continue
    r = run(1 << k)
    rows[name] = {b: full[b] - r[b] for b in full}
    print("%-22s costs %s" % (name, {b: "%.3f ms" % v for b, v in rows[name].items()}))
allmm = run(0b11111)
print("all matmuls left out: %s  (what the small kernels and the launch chain cost on their own)" % allmm)
nomm = run(0b11110100000)
print("all small kernels left out: %s  (the matmuls back to back)" % nomm)
print(json.dumps({"whole": full, "marginal_ms": rows, "without_matmuls": allmm, "matmuls_only": nomm}))

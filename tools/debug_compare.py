#!/usr/bin/env python
"""Debugging aid: decode one batch with the CUDA engine and with the CPU oracle, both dumping every intermediate of the
llama graph (PIB_DUMP_DIR / PO_DUMP_DIR), and print where the two first part ways.
  python tools/debug_compare.py model.gguf [n_tokens] [n_ctx]"""
import os, sys, tempfile
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
d1, d2 = tempfile.mkdtemp(), tempfile.mkdtemp()
os.environ["PIB_DUMP_DIR"], os.environ["PO_DUMP_DIR"], os.environ["PIB_GRAPHS"], os.environ["PIB_PDL"] = d1, d2, "0", "0"
import numpy as np
from oracle import oracle_py as op
from pipeinfer_b200 import llama
path = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
n_ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 64
m = op.load_gguf_model(path)
rng = np.random.default_rng(1)
toks = rng.integers(3, m.n_vocab, n).tolist()
ctx = llama.Context(path, n_ctx=n_ctx, max_seq=8)
octx = op.OracleContext(op.Oracle(), m, n_ctx, 8)
rc, lg = ctx.decode(toks, list(range(n)), [[0]] * n)
rc2, ref = octx.decode(toks, list(range(n)), [[0]] * n)
print("logits: max|diff| %.3g (max|ref| %.3g); per row: %s" % (np.abs(lg - ref).max(), np.abs(ref).max(), " ".join("%.1e" % v for v in np.abs(lg - ref).max(1))))
names = ["inp", "q", "k", "v", "att", "o", "ffi", "gate", "up", "down"]
for il in range(m.n_layer):
    for nm in names:
        a, b = os.path.join(d1, "%s.%d.bin" % (nm, il)), os.path.join(d2, "%s.%d.bin" % (nm, il))
        if not (os.path.exists(a) and os.path.exists(b)):
            continue
        x, y = np.fromfile(a, np.float32), np.fromfile(b, np.float32)
        if x.shape != y.shape:           # the engine dumps the last 16-token chunk of a long batch
            cols = y.size // n
            y = y.reshape(n, cols)[n - x.size // cols:].reshape(-1)
        d = np.abs(x - y)
        print("layer %2d %-5s max|diff| %.3g  rel %.3g  (n differing > 1e-6 rel: %d of %d)" % (il, nm, d.max(), d.max() / max(1e-30, np.abs(y).max()), int((d > 1e-6 * np.abs(y).max()).sum()), x.size))
    if il >= int(os.environ.get("DBG_LAYERS", "3")) - 1:
        break

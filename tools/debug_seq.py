#!/usr/bin/env python
"""Debugging aid: engine vs oracle logits, (a) one batch of n tokens, (b) the same tokens one decode at a time."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import numpy as np
from oracle import oracle_py as op
from pipeinfer_b200 import llama
path, n, n_ctx = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
m = op.load_gguf_model(path)
toks = np.random.default_rng(1).integers(3, m.n_vocab, n).tolist()
O = op.Oracle()
for mode in ("batch", "single", "chunks-of-5"):
    ctx = llama.Context(path, n_ctx=n_ctx, max_seq=8)
    octx = op.OracleContext(O, m, n_ctx, 8)
    step = {"batch": n, "single": 1, "chunks-of-5": 5}[mode]
    errs = []
    for p0 in range(0, n, step):
        tk = toks[p0:p0 + step]
        ps = list(range(p0, p0 + len(tk)))
        rc, lg = ctx.decode(tk, ps, [[0]] * len(tk))
        rc2, ref = octx.decode(tk, ps, [[0]] * len(tk))
        errs += np.abs(lg - ref).max(1).tolist()
    print(mode, " ".join("%.0e" % e for e in errs))
    ctx.close(); octx.close()

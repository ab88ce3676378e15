import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from tests.replay import replay
from pipeinfer_b200 import llama
g = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests", "golden")
for q in sys.argv[1:] or ["falcon_q4_0"]:
    ctx = llama.Context(os.path.join(g, "tiny_%s.gguf" % q), n_ctx=48, max_seq=8)
    errs = []
    def check(got, ref):
        d = np.abs(got - ref)
        errs.append("%.2g(med %.2g, rows %s)" % (d.max(), np.median(d), " ".join("%.1g" % x for x in d.max(1))))
    replay(g, "decode_%s" % q, ctx.decode, ctx.seq_rm, ctx.seq_cp, ctx.seq_keep, ctx.kv_view, check)
    print(q, "\n  ".join(errs))

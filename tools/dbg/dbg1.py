import sys, os, numpy as np
sys.path.insert(0, '/root/repo')
from oracle import oracle_py as op
from pipeinfer_b200 import load
import gguf
lib = load(); lib.init(0); o = op.Oracle()
r = gguf.GGUFReader('/root/repo/tests/golden/tiny_q4_k.gguf')
t = {x.name: x for x in r.tensors}
wt = 12
rng = np.random.default_rng(0)
emb = np.ascontiguousarray(t['token_embd.weight'].data)
K = 256
rows = o.dequantize(wt, emb[:16], K)
for name in ['blk.0.attn_q.weight', 'blk.0.ffn_gate.weight', 'blk.0.ffn_down.weight', 'output.weight']:
    w = np.ascontiguousarray(t[name].data)
    N = w.shape[0]
    h = lib.weight_upload(wt, w, K)
    for b in (1, 3, 7, 16):
        x = rows[:b] * 0.7
        got = lib.mul_mat(h, x, N, 1); ref = o.mul_mat(wt, w, x)
        aq = lib.quantize_act(wt, x); ar = o.quantize_act(wt, x)
        print(name, 'b', b, 'maxdiff', np.abs(got-ref).max(), 'max', np.abs(ref).max(), 'act eq', np.array_equal(aq, ar))
    # round trip
    print(name, 'roundtrip', np.array_equal(lib.weight_rows(h, wt, N, K), w))

import sys, os, subprocess, numpy as np
sys.path.insert(0, '/root/repo')
from oracle import oracle_py as op
from pipeinfer_b200 import llama
llama.api()
o = op.Oracle()
def run(q, L, E, H, HKV, FF, ntok=5):
    path = '/tmp/dbg_%s.gguf' % q
    subprocess.run([sys.executable, '/root/repo/tools/make_gguf.py', '--out', path, '--n-layer', str(L), '--n-embd', str(E), '--n-head', str(H),
                    '--n-head-kv', str(HKV), '--n-ff', str(FF), '--n-vocab', '384', '--quant', q, '--flavour', 'bigram', '--eps', '1.0'], check=True)
    ctx = llama.Context(path, n_ctx=48, max_seq=8)
    m = op.load_gguf_model(path); oc = op.OracleContext(o, m, 48, 8)
    rng = np.random.default_rng(1)
    toks = rng.integers(3, 384, ntok).tolist()
    rc, lg = ctx.decode(toks, list(range(ntok)), [[0]]*ntok)
    rc2, ref = oc.decode(toks, list(range(ntok)), [[0]]*ntok)
    print(q, 'L', L, 'E', E, 'FF', FF, 'ntok', ntok, 'maxdiff %.3g' % np.abs(lg-ref).max(), 'max %.3g' % np.abs(ref).max())
    ctx.close(); oc.close()
for q in ('q4_k', 'q8_0'):
    pass
    run(q, 1, 256, 4, 2, 256, 1)
    run(q, 1, 256, 4, 2, 256)
    run(q, 1, 256, 4, 2, 512)
    run(q, 1, 512, 8, 2, 512)
    run(q, 2, 256, 4, 2, 256)

import sys, os, json, numpy as np
sys.path.insert(0, '/root/repo')
from pipeinfer_b200 import llama
from oracle import oracle_py as op
llama.api()
g = '/root/repo/tests/golden'
tgt = llama.Context(g + '/tiny_q8_0.gguf', n_ctx=256, max_seq=1002)
dft = llama.Context(g + '/tiny_draft_q8_0.gguf', n_ctx=256, max_seq=1002)
prompt = llama.tokenize(tgt, "the quick brown fox")
rc, lg = dft.decode(prompt, list(range(len(prompt))), [[0]] * len(prompt))
print('draft top8 last row', dft.top8(len(prompt) - 1)[:3])
o = op.Oracle(); m = op.load_gguf_model(g + '/tiny_draft_q8_0.gguf'); oc = op.OracleContext(o, m, 256, 8)
rc, ref = oc.decode(prompt, list(range(len(prompt))), [[0]] * len(prompt))
e = np.exp(ref[-1] - ref[-1].max()); p = e / e.sum(); print('oracle p max', p.max(), p.argmax(), 'maxdiff', np.abs(lg - ref).max())
dft.close(); dft = llama.Context(g + '/tiny_draft_q8_0.gguf', n_ctx=256, max_seq=1002)
for pa in (0.5, 0.05):
    params = llama.spec_params(n_draft=4, n_parallel=1, n_predict=24, p_accept=pa)
    spec, res, itl = llama.speculative_generate(tgt, dft, prompt, params)
    print(pa, res)
    tgt.close(); dft.close()
    tgt = llama.Context(g + '/tiny_q8_0.gguf', n_ctx=256, max_seq=1002)
    dft = llama.Context(g + '/tiny_draft_q8_0.gguf', n_ctx=256, max_seq=1002)

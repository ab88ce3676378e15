import sys, numpy as np
sys.path.insert(0, '/root/repo')
from pipeinfer_b200 import llama
llama.api()
toks = [5, 77, 1033, 900, 12, 2000, 1500, 31]
for wt in (8, 12, 2):
    dft = llama.Context(None, n_ctx=64, max_seq=4, synthetic=(1, 256, 4, 4, 512, 2048, wt, 2, 1, 1.0))
    rc, ld = dft.decode(toks, list(range(8)), [[0]] * 8)
    for L in (1, 4, 12, 24):
        tgt = llama.Context(None, n_ctx=64, max_seq=4, synthetic=(L, 512, 8, 2, 1024, 2048, wt, 1, 1, 1.0))
        rc, lt = tgt.decode(toks, list(range(8)), [[0]] * 8)
        srt = np.sort(lt, axis=1)
        print('wt', wt, 'L', L, 'agree', (lt.argmax(1) == ld.argmax(1)).mean(), 'top1', srt[:, -1].mean(), 'top2', srt[:, -2].mean(), 'finite', np.isfinite(lt).all())
        tgt.close()
    dft.close()

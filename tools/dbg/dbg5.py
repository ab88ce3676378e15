import sys, ctypes as C, numpy as np
sys.path.insert(0, '/root/repo')
from pipeinfer_b200 import llama
L = llama.api()
L.llama_pib_debug_read.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
toks = [5, 77, 1033, 900]
def rd(ctx, which, n):
    out = np.zeros(n, np.float32); L.llama_pib_debug_read(ctx.ctx, which, out.ctypes.data, n); return out
for wt in (8, 12):
    xs = {}
    for nl in (1, 2, 8):
        tgt = llama.Context(None, n_ctx=64, max_seq=4, synthetic=(nl, 512, 8, 2, 1024, 2048, wt, 1, 1, 1.0))
        rc, lt = tgt.decode(toks, list(range(4)), [[0]] * 4)
        x = rd(tgt, 0, 4 * 512).reshape(4, 512); att = rd(tgt, 2, 4 * 512).reshape(4, 512); proj = rd(tgt, 3, 4*512).reshape(4,512); x2 = rd(tgt, 1, 4*512).reshape(4,512)
        xs[nl] = x
        print('wt', wt, 'L', nl, '|x|', np.linalg.norm(x, axis=1), '|att|', np.linalg.norm(att, axis=1), '|down|', np.linalg.norm(proj, axis=1), 'mean x', x.mean(axis=1))
        tgt.close()
    print('  |x_L2 - x_L1|', np.linalg.norm(xs[2] - xs[1], axis=1), '|x_L8 - x_L1|', np.linalg.norm(xs[8] - xs[1], axis=1))

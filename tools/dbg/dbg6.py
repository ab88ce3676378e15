import sys, ctypes as C, numpy as np
sys.path.insert(0, '/root/repo')
from pipeinfer_b200 import llama, load
from oracle import oracle_py as op
L = llama.api(); lib = load(); o = op.Oracle()
L.llama_pib_debug_weight.restype = C.c_void_p
L.llama_pib_debug_weight.argtypes = [C.c_void_p, C.c_int, C.c_int]
tgt = llama.Context(None, n_ctx=64, max_seq=4, synthetic=(1, 512, 8, 2, 1024, 2048, 12, 1, 1, 1.0))
w = L.llama_pib_debug_weight(tgt.model, -1, 0)
rows = np.zeros((4, 512 // 256 * 144), np.uint8)
for i, r in enumerate([5, 77, 1033, 900]):
    lib.check(lib.L.pib_weight_row_blocks(w, r, rows[i].ctypes.data))
deq = o.dequantize(12, rows, 512)
print('mean', deq.mean(1), 'std', deq.std(1))
b = rows[2].reshape(2, 144)
for sb in range(2):
    d = b[sb, 0:2].view(np.float16)[0]; dm = b[sb, 2:4].view(np.float16)[0]; q = b[sb, 4:16]
    sc = []; mn = []
    for j in range(8):
        if j < 4: sc.append(q[j] & 63); mn.append(q[j + 4] & 63)
        else: sc.append((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4)); mn.append((q[j + 4] >> 4) | ((q[j] >> 6) << 4))
    qs = b[sb, 16:]
    print('sb', sb, 'd', d, 'dmin', dm, 'ratio', float(dm) / float(d), 'sc', sc, 'm', mn, 'nib mean lo', (qs & 15).mean(), 'hi', (qs >> 4).mean())
M = 0xffffffff
def h32(x):
    x &= M; x ^= x >> 16; x = (x * 0x7feb352d) & M; x ^= x >> 15; x = (x * 0x846ca68b) & M; x ^= x >> 16; return x
def h3(seed, a, b): return h32(seed ^ h32((a * 0x9E3779B9 + h32((b + 0x85ebca6b) & M)) & M))
embseed = 0xE3B0C442
for r in (1033, 5):
    for c in range(2):
        h = h3(embseed ^ 0xA5A5A5A5, r, c); u = 0.5 + (h & 0xffff) / 65536.0
        dd = np.float32(1.0 * u / (32.0 * 4.6)); hd = np.float16(dd); hm = np.float16(np.float32(7.5) * np.float32(hd))
        print('row', r, 'c', c, 'expect d', hd, 'dmin', hm, 'bytes', hd.tobytes().hex(), hm.tobytes().hex())
print('actual hdr bytes row 1033:', rows[2].reshape(2, 144)[:, :16].tobytes().hex())
print('actual hdr bytes row 5:', rows[0].reshape(2, 144)[:, :16].tobytes().hex())

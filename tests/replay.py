"""Replay of a llama.h call trace (the *.script fixtures written by tests/golden/make_goldens.py) against any object
with decode / seq_rm / seq_cp / seq_keep, comparing with what the reference returned (*.kv.txt, *.logits.bin)."""
import os

import numpy as np

from oracle import oracle_py as op


def parse_script(path):
    toks = open(path).read().split()
    ops, i = [], 0
    while i < len(toks):
        o = toks[i]
        if o == "D":
            n = int(toks[i + 1]); i += 2
            tk, ps, sq = [], [], []
            for _ in range(n):
                t, p, ns = int(toks[i]), int(toks[i + 1]), int(toks[i + 2])
                sq.append([int(x) for x in toks[i + 3:i + 3 + ns]])
                tk.append(t); ps.append(p); i += 3 + ns
            ops.append(("D", tk, ps, sq))
        elif o == "RM":
            ops.append(("RM", int(toks[i + 1]), int(toks[i + 2]), int(toks[i + 3]))); i += 4
        elif o == "CP":
            ops.append(("CP", int(toks[i + 1]), int(toks[i + 2]), int(toks[i + 3]), int(toks[i + 4]))); i += 5
        elif o == "KEEP":
            ops.append(("KEEP", int(toks[i + 1]))); i += 2
        elif o == "KV":
            ops.append(("KV",)); i += 1
        else:
            raise AssertionError(o)
    return ops


def replay(golden_dir, stem, decode, seq_rm, seq_cp, seq_keep, kv_cells, logits_check=None):
    """kv_cells() -> (used, {cell: (pos, [ids])}); decode(tokens, pos, seqs) -> (rc, logits or None);
    logits_check(got, ref) is called for every successful decode when given."""
    ops = parse_script(os.path.join(golden_dir, stem + ".script"))
    lines = open(os.path.join(golden_dir, stem + ".kv.txt")).read().strip().split("\n")
    lp = os.path.join(golden_dir, stem + ".logits.bin")
    ref_logits = np.fromfile(lp, dtype=np.float32) if (logits_check and os.path.exists(lp)) else None
    li = off = n_dec = n_kv = 0
    for o in ops:
        if o[0] == "D":
            rc, lg = decode(o[1], o[2], o[3])
            assert lines[li] == "D %d" % rc, (stem, li, lines[li], rc)
            li += 1
            if rc == 0 and ref_logits is not None:
                ref = ref_logits[off:off + lg.size].reshape(lg.shape)
                off += lg.size
                logits_check(lg, ref)
            n_dec += 1
        elif o[0] == "RM":
            seq_rm(*o[1:])
        elif o[0] == "CP":
            seq_cp(*o[1:])
        elif o[0] == "KEEP":
            seq_keep(*o[1:])
        else:
            used, cells = op.parse_ref_kv_line(lines[li]); li += 1
            got_used, got_cells = kv_cells()
            assert used == got_used, (stem, li, used, got_used)
            assert cells == got_cells, (stem, li)
            n_kv += 1
    return n_dec, n_kv

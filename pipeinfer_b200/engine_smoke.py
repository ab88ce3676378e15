"""The end-to-end slice of __graft_entry__.smoke(): decode the tiny Q4_K fixture through llama_decode on cuda:0 and
compare the logits with what the reference's own CPU build produced for the same call trace (tests/golden, written by
tests/golden/make_goldens.py), then one short speculative generation that must equal plain greedy decoding."""
import os

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def run():
    from tests.replay import replay          # test infrastructure: only smoke() and tests/ reach for it
    from . import llama
    golden = os.path.join(ROOT, "tests", "golden")
    ctx = llama.Context(os.path.join(golden, "tiny_q4_k.gguf"), n_ctx=48, max_seq=8)
    errs = []

    def check(got, ref):
        errs.append(float(np.abs(got - ref).max()))

    n_dec, n_kv = replay(golden, "decode_q4_k", ctx.decode, ctx.seq_rm, ctx.seq_cp, ctx.seq_keep, ctx.kv_view, check)
    ctx.close()
    # same bar as tests/test_gpu_engine.py: typical decodes differ by fp32 reassociation only; a decode in which one
    # activation lands on the other side of a quantisation step ("flip") stays inside the 6e-2 envelope
    assert np.median(errs) < 1e-4 and max(errs) < 6e-2, errs
    print("smoke engine: %d decodes + %d KV views against the reference; max |logit diff| per decode: median %.2g, worst %.2g"
          % (n_dec, n_kv, float(np.median(errs)), max(errs)))

    tgt = llama.Context(None, n_ctx=256, max_seq=2002, synthetic=(4, 512, 8, 2, 1024, 2048, 12, 1, 1, 1.0))
    dft = llama.Context(None, n_ctx=256, max_seq=2002, synthetic=(2, 256, 4, 2, 512, 2048, 12, 2, 1, 0.8))
    prompt = [1] + [3 + (i * 7919) % 2045 for i in range(24)]
    toks, res, _ = llama.speculative_generate(tgt, dft, prompt, llama.spec_params(n_draft=3, n_parallel=2, n_predict=24))
    lib = llama.api()
    lib.llama_kv_cache_clear(tgt.ctx)
    ref, _, _ = llama.plain_generate(tgt, prompt, 24)
    n = min(len(toks), len(ref))
    assert n >= 24 and toks[:n] == ref[:n], (toks, ref)
    print("smoke scheduler: %d tokens, speculative == plain greedy (%d drafted, %d accepted)" % (n, res["n_drafted"], res["n_accept"]))
    tgt.close()
    dft.close()

#!/usr/bin/env python
"""TEST INFRASTRUCTURE: the what-if runs behind DESIGN.md §9 (scheduler simulator).

  python tests/sched_sim/explore.py calibrate    model against the measured 8-GPU numbers in profiles/
  python tests/sched_sim/explore.py sensitivity  tokens/s in the model when the draft step / the stage time / the agreement change,
                                                 under the reference policy and under PIB_SCHED_CHAIN=1
  python tests/sched_sim/explore.py scaling      1, 2, 4, 8 stages
  python tests/sched_sim/explore.py round2       the same with round 2's constants (pass 9.39 ms + 0.34 ms per extra token, draft step 0.74 ms,
                                                 policy chain2, balanced layer split) against round 2's measurements, and what would move them

Every run also checks that the text is the greedy text.  Numbers are means over 6 prompts / model seeds.
"""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..")))
from tests.sched_sim import sim  # noqa: E402

MEASURED_8 = {"d3": 206.4, "d2": 203.7, "d5": 184.2, "d8": 163.4, "d3np1": 210.3, "classic": 182.1, "plain": 89.0}   # profiles/r1_bench_*_n8.json
MEASURED_N = {1: 97.2, 2: 142.0, 4: 194.0, 8: 206.4}


def mean_tps(cfg=None, seeds=6, **kw):
    out = []
    for seed in range(seeds):
        r = sim.run(sim.default_config(seed=seed, **(cfg or {})), prompt_seed=seed, n_prompt=64, n_predict=128, n_ctx=1024, **kw)
        assert r["tokens"] == r["truth"] and r["n"] >= 128, (cfg, kw, r["error"])
        out.append(r["result"]["decode_tok_per_s"])
    return sum(out) / len(out)


BEFORE_FIX = dict(spec_issue_extra_us=900, nonspec_issue_extra_us=500)    # host KV-mirror cost per run issue when profiles/ was measured


def calibrate():
    """the measurements in profiles/ predate the host KV-mirror fix: compare them with the model carrying that cost"""
    runs = {"d3": dict(n_draft=3, max_spec_runs=9), "d2": dict(n_draft=2, max_spec_runs=9), "d5": dict(n_draft=5, max_spec_runs=9),
            "d8": dict(n_draft=8, max_spec_runs=9), "d3np1": dict(n_draft=3, n_parallel=1, max_spec_runs=9),
            "classic": dict(mode="classic", n_draft=3), "plain": dict(mode="plain")}
    for k, kw in runs.items():
        b, a = mean_tps(BEFORE_FIX, **kw), mean_tps(**kw)
        print("%-8s measured %6.1f   model (as measured) %6.1f (%+.0f %%)   model (after the fix) %6.1f" % (k, MEASURED_8[k], b, 100 * (b / MEASURED_8[k] - 1), a))


def sensitivity():
    rows = [("today", {}), ("draft step 0.35 ms", dict(draft_us=200)), ("draft step 0.10 ms", dict(draft_us=0, enqueue_us=100)),
            ("stage 0.90 ms", dict(stage_us=900)), ("stage 0.90 ms + draft 0.35 ms", dict(stage_us=900, draft_us=200)),
            ("agreement 0.6", dict(agree=0.6)), ("agreement 0.9", dict(agree=0.9)), ("agreement 0.95", dict(agree=0.95))]
    for chain in ("0", "1"):
        os.environ["PIB_SCHED_CHAIN"] = chain
        print("PIB_SCHED_CHAIN=%s" % chain)
        for name, cfg in rows:
            print("  %-32s --draft 3: %5.0f   --draft 5: %5.0f   classic: %5.0f" %
                  (name, mean_tps(cfg, n_draft=3, max_spec_runs=9), mean_tps(cfg, n_draft=5, max_spec_runs=9), mean_tps(cfg, mode="classic", n_draft=3)))


SPLITS = {1: (0, 0), 2: (30 / 80, 50 / 80), 4: (15 / 80, 23 / 80), 8: (7 / 80, 13 / 80)}     # bench.py's default layer split: (head, tail) share


def scaling():
    """1, 2, 4, 8 stages with bench.py's layer split.  The draft shares GPU 0 with the first stage and has priority there: the
    model brackets that with share_gpu0 = 1 (the target stands still while a draft decode runs) and 0 (no interference)."""
    for stages in (1, 2, 4, 8):
        cfg = dict(n_stages=stages, stage_us=10320.0 / stages, stage_tok_us=480.0 / stages, head_frac=SPLITS[stages][0], tail_frac=SPLITS[stages][1])
        band = lambda extra, **kw: "%5.1f..%5.1f" % tuple(mean_tps(dict(cfg, share_gpu0=sh, **extra), n_draft=3, max_spec_runs=stages + 1, **kw) for sh in (1.0, 0.0))
        os.environ["PIB_SCHED_CHAIN"] = "0"
        line = "%d stage(s): measured %5.1f   model as measured %s   after the host fix %s" % (stages, MEASURED_N[stages], band(BEFORE_FIX), band({}))
        os.environ["PIB_SCHED_CHAIN"] = "2"
        line += "   with policy 2 %s" % band({})
        os.environ["PIB_SCHED_CHAIN"] = "0"
        print(line)


MEASURED_R2 = {1: 157.3, 2: 213.6, 4: 244.8, 8: 260.6}          # profiles/r2_bench_70b_n*.json
SPLITS_R2 = {1: (0, 0), 2: (30 / 80, 50 / 80), 4: (5 / 80, 27 / 80), 8: (1 / 80, 13 / 80)}


def round2():
    """round 2's engine in the model: a pass of 9.39 ms at one token + 0.343 ms per further token (profiles/r2_pass_breakdown.md),
    a draft step of 0.74 ms of which 0.15 ms is host time, policy chain2, bench.py's balanced layer split"""
    os.environ["PIB_SCHED_CHAIN"] = "2"
    base = dict(draft_us=590)
    for stages in (1, 2, 4, 8):
        cfg = dict(base, n_stages=stages, stage_us=9390.0 / stages, stage_tok_us=343.0 / stages, head_frac=SPLITS_R2[stages][0], tail_frac=SPLITS_R2[stages][1])
        band = lambda extra: "%5.1f..%5.1f" % tuple(mean_tps(dict(cfg, share_gpu0=sh, **extra), n_draft=3, max_spec_runs=stages + 1) for sh in (1.0, 0.0))
        print("%d stage(s): measured %5.1f   model %s   draft step 0.35 ms %s   agreement 0.9 %s   both %s" %
              (stages, MEASURED_R2[stages], band({}), band(dict(draft_us=200)), band(dict(agree=0.9)), band(dict(draft_us=200, agree=0.9))))
    os.environ["PIB_SCHED_CHAIN"] = "0"
    cfg = dict(base, n_stages=8, stage_us=9390.0 / 8, stage_tok_us=343.0 / 8)
    print("plain pipelined decode, 8 stages: measured 105.3   model %5.1f" % mean_tps(cfg, mode="plain"))


if __name__ == "__main__":
    {"calibrate": calibrate, "sensitivity": sensitivity, "scaling": scaling, "round2": round2}[sys.argv[1] if len(sys.argv) > 1 else "calibrate"]()

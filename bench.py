#!/usr/bin/env python
"""bench.py -- accepted tokens/s + p50 per-token latency of PipeInfer-style asynchronous pipelined speculation.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config 70b|7b|tiny]

Workload (BASELINE.json configs[2], the one the metric is quoted on): Dolphin/Llama-2-70B-architecture target in Q4_K
(38.8 GB, fits one B200) + TinyLlama-1.1B-architecture draft, random-init weights generated in HBM ("bigram" flavour:
both models predict a fixed permutation of the last token, the draft agreeing with the target on --agree of the tokens),
synthetic 64-token prompt, flags of the authors' launch script (examples/speculative/scripts/dolphin_tinyllama_8_nodes.sh:
--draft 3 -np 3 -pa 0.1 -ps 0.9 -n 128 -c 1024, greedy).  One step = one generation of 128 tokens.

value  = sum(n_predict) / sum(t_decode) over the timed steps: the reference's own metric (speculative.cpp:697), prompt
         already ingested (inputs resident in HBM).
e2e    = sum(n_predict) / wall time of the timed steps through the public C ABI (pib_speculative_generate) with host
         buffers: prompt ids in, KV clear, prompt ingestion, every descriptor / op-log H2D and result D2H included.
The roofline object times the dominant kernel (the gate|up projection GEMV of the target, 264 MB of Q4_K weights per
launch) with CUDA events on its own stream during the timed steps.
"""
import argparse
import ctypes as C
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

Q4_K, Q4_0, Q8_0 = 12, 2, 8
CONFIGS = {
    # name: (target (L, E, H, HKV, FF, V, type), draft (...), label)
    "70b": ((80, 8192, 64, 8, 28672, 32000, Q4_K), (22, 2048, 32, 4, 5632, 32000, Q4_K),
            "Llama-2-70B-arch Q4_K target + TinyLlama-1.1B-arch Q4_K draft"),
    "7b": ((32, 4096, 32, 32, 11008, 32000, Q4_0), (22, 2048, 32, 4, 5632, 32000, Q4_0),
           "Llama-2-7B-arch Q4_0 target + TinyLlama-1.1B-arch Q4_0 draft"),
    "1b": ((22, 2048, 32, 4, 5632, 32000, Q8_0), (2, 512, 8, 8, 1408, 32000, Q8_0),
           "TinyLlama-1.1B-arch Q8_0 target + 2-layer Llama-arch Q8_0 draft"),
    "120b": ((137, 8192, 64, 8, 28672, 32000, Q4_K), (32, 4096, 32, 32, 11008, 32000, Q4_K),
             "Goliath-120B-arch Q4_K target + XWin-7B-arch Q4_K draft"),
    "180b": ((80, 14848, 232, 8, 59392, 65024, Q4_0), (32, 4544, 71, 1, 18176, 65024, Q4_0),
             "Falcon-180B-arch Q4_0 target + Falcon-7B-arch Q4_0 draft"),
    "tiny": ((8, 1024, 8, 2, 2048, 4096, Q4_K), (2, 512, 8, 2, 1024, 4096, Q4_K), "8-layer 1024-wide target + 2-layer draft (smoke)"),
    "tiny-falcon": ((8, 1024, 16, 2, 4096, 4096, Q4_0), (2, 512, 8, 1, 2048, 4096, Q4_0), "8-layer falcon-arch target + 2-layer falcon-arch draft (smoke)"),
}
# synthetic-model flavour of (target, draft): bigram (1) [+ 0x10 falcon architecture, + 0x20 second attention norm]
FLAVOUR = {"180b": (0x31, 0x11), "tiny-falcon": (0x31, 0x11)}
TYPE_NAME = {Q4_K: "q4_k", Q4_0: "q4_0", Q8_0: "q8_0"}
BPW = {Q4_K: 4.5, Q4_0: 4.5, Q8_0: 8.5}
KIND = {0: "qkv", 1: "attn_out", 2: "gate_up", 3: "down", 4: "lm_head", 5: "attn_norm_quant", 6: "rope_kv_store", 7: "attention",
        8: "attn_out_quant", 9: "resid_ffn_norm_quant", 10: "silu_mul_quant"}


def nvlink_tx_kib(gpu):
    """NVLink transmit counter of one GPU in KiB (all links), or None where the driver exposes none: NVML field values first
    (NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, KiB since driver load), then the per-link table of `nvidia-smi nvlink -gt d`"""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu)
        fid = getattr(pynvml, "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX", 138)
        try:
            vals = pynvml.nvmlDeviceGetFieldValues(h, [(fid, 0xFFFFFFFF)])      # scope: all links
        except TypeError:
            vals = pynvml.nvmlDeviceGetFieldValues(h, [fid])
        v = vals[0]
        if v.nvmlReturn == 0:
            for name in ("ullVal", "ulVal", "uiVal"):
                x = getattr(v.value, name, None)
                if x:
                    return int(x)
    except Exception:
        pass
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(gpu)], capture_output=True, text=True, timeout=20).stdout
        vals = [int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out)]
        return sum(vals) if vals else None
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, gpu):
        self.rows, self.proc, self.gpu = [], None, gpu

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


# --------------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the UNMODIFIED reference `speculative` (oracle/_ref, built from /root/reference by
# oracle/Makefile) under the mini-MPI shim on this box's host cores.  A 70B-width model cannot be run at full depth on the
# host in bounded time, so the sample is depth-reduced (same width, same draft) at two depths and the per-token time is
# extrapolated linearly in the number of target layers.
def run_reference_sample(cfg_name, args, budget="small"):
    ref = os.path.join(ROOT, "oracle", "_ref")
    spec, shim = os.path.join(ref, "speculative"), os.path.join(ref, "mpirun_shim")
    if not (os.path.exists(spec) and os.path.exists(shim)):
        return None
    (L, E, H, HKV, FF, V, wt), (dL, dE, dH, dHKV, dFF, dV, dwt), label = CONFIGS[cfg_name]
    cores = os.cpu_count() or 1
    n_tgt_ranks = 2
    threads = max(1, cores // (n_tgt_ranks + 1))
    depths = (2, 4) if budget == "small" else (2, 4, 6)
    n_pred = 24 if budget == "small" else 64
    if cfg_name == "tiny":
        depths = (2, 4) if budget == "small" else (2, 3, 4)
    tmp = "/tmp/pib_ref_%d" % os.getpid()
    os.makedirs(tmp, exist_ok=True)
    mk = [sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), "--direct", "--quant", TYPE_NAME[wt]]
    dft = os.path.join(tmp, "dft.gguf")
    subprocess.run(mk + ["--out", dft, "--n-layer", str(dL), "--n-embd", str(dE), "--n-head", str(dH), "--n-head-kv", str(dHKV),
                         "--n-ff", str(dFF), "--n-vocab", str(dV), "--agree", str(args.agree), "--seed", "5"], check=True)
    prompt = os.path.join(tmp, "prompt.txt")
    open(prompt, "w").write("".join(chr(0x21 + (i * 7) % 90) for i in range(args.prompt_len - 2)))
    res = {}
    for d in depths:
        tgt = os.path.join(tmp, "tgt_%d.gguf" % d)
        subprocess.run(mk + ["--out", tgt, "--n-layer", str(d), "--n-embd", str(E), "--n-head", str(H), "--n-head-kv", str(HKV),
                             "--n-ff", str(FF), "--n-vocab", str(V), "--agree", "1.0", "--seed", "3"], check=True)
        pre = os.path.join(tmp, "run_%d" % d)
        t0 = time.time()
        subprocess.run([shim, "-np", str(n_tgt_ranks + 1), "--timeout", "900", "--out-prefix", pre, spec, "-m", tgt, "-md", dft,
                        "-f", prompt, "-n", str(n_pred), "-c", "1024", "--temp", "-1", "--repeat-last-n", "0", "--ignore-eos",
                        "--draft", str(args.draft), "-np", str(args.np), "-pa", str(args.pa), "-ps", str(args.ps), "-pr", "0.0",
                        "-t", str(threads), "-tb", str(threads), "--mpi-layer-split", "0.5,0.5/1.0"],
                       check=True, cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        err = open(pre + ".0.err", errors="replace").read()
        m = re.search(r"decoded\s+(\d+) tokens in\s+([0-9.]+) seconds", err)
        itl = re.search(r"Average inter-token latency: ([0-9.]+)", err)
        acc = re.search(r"n_accept\s+=\s+(\d+)", err)
        drf = re.search(r"n_drafted\s+=\s+(\d+)", err)
        res[d] = dict(tokens=int(m.group(1)), seconds=float(m.group(2)), itl_mean_s=float(itl.group(1)) if itl else None,
                      n_accept=int(acc.group(1)) if acc else None, n_drafted=int(drf.group(1)) if drf else None,
                      wall_s=time.time() - t0)
        os.unlink(tgt)
    os.unlink(dft)
    # seconds per accepted token as a straight line in the number of target layers (least squares over the depths run)
    xs = [float(d) for d in depths]
    ys = [res[d]["seconds"] / res[d]["tokens"] for d in depths]
    n  = len(xs)
    mx, my = sum(xs) / n, sum(ys) / n
    per_layer = max(0.0, sum((x - mx) * (y - my) for x, y in zip(xs, ys)) / sum((x - mx) ** 2 for x in xs))
    fixed = max(0.0, my - per_layer * mx)
    full = fixed + per_layer * L
    resid = max(abs(fixed + per_layer * x - y) / y for x, y in zip(xs, ys))
    sample = ("reference `speculative` (oracle/_ref) under the mini-MPI shim, 2 target ranks + 1 draft rank x %d threads, "
              "same flags, %d tokens predicted; %s-width target at %s layers (%s s per accepted token), full draft; value = "
              "1 / (%.4f + %d * %.4f) s extrapolated to %d layers, %s"
              % (threads, n_pred, "x".join(map(str, (E, FF))), "/".join(str(d) for d in depths), " / ".join("%.3f" % y for y in ys),
                 fixed, L, per_layer, L,
                 ("worst relative residual of the least-squares fit %.3f" % resid) if len(depths) > 2
                 else "two depths in this bounded leg: the line through them, no residual (`--impl reference` fits three depths over 64 tokens)"))
    return dict(value=1.0 / full if full > 0 else None, unit="tokens/s", cores=cores, kind="reference", sample=sample,
                raw={str(k): v for k, v in res.items()}, itl_est_s=full, fit_residual=resid)


def head_layers(world, tshape, dshape, n_draft):
    """How many of the target's layers the head GPU takes.  It also runs the draft, so it gets the share that evens out the work per
    run: n_draft draft steps + L0 layers on the head against (L - L0) / (N - 1) layers on every other stage.  Costs from the shapes: a
    target layer streams its weights at ~5 TB/s plus ~25 us of small kernels; a draft step is a chain of nine dependent kernels per
    layer (~28 us) plus its bytes.  Returns (L0, the drafting of one run expressed in target layers)."""
    def layer_bytes(e, h, hkv, ff, t):
        return (e * (e + 2 * e // h * hkv) + e * e + 3 * e * ff) * ({Q4_K: 4.5, 2: 4.5, 8: 8.5}.get(t, 4.5) / 8.0)
    L, E, H, HKV, FF, V, wt = tshape
    dL, dE, dH, dHKV, dFF, dV, dwt = dshape
    t_layer = layer_bytes(E, H, HKV, FF, wt) / 5.0e12 + 25e-6
    t_draft = dL * (layer_bytes(dE, dH, dHKV, dFF, dwt) / 5.0e12 + 28e-6) + 60e-6
    r = n_draft * t_draft / t_layer
    l0 = int(round((L - r * (world - 1)) / world))
    return max(1, min(L // world, l0)), r


# --------------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="70b", choices=sorted(CONFIGS))
    ap.add_argument("--agree", type=float, default=0.8, help="fraction of tokens on which the random-init draft predicts the target's token")
    ap.add_argument("--n-predict", type=int, default=128)
    ap.add_argument("--prompt-len", type=int, default=64)
    ap.add_argument("--draft", type=int, default=3)
    ap.add_argument("--n-parallel", dest="np", type=int, default=3)
    ap.add_argument("--pa", type=float, default=0.1)
    ap.add_argument("--ps", type=float, default=0.9)
    ap.add_argument("--max-spec-runs", type=int, default=-1,
                    help="bound on target runs in flight: 0 = the reference's behaviour (unbounded), -1 = number of pipeline stages + 1")
    ap.add_argument("--mode", default="speculative", choices=["speculative", "plain", "classic"],
                    help="plain = non-speculative greedy decoding through the same pipeline (the reference's `main`): the second denominator")
    ap.add_argument("--sched-policy", default="chain2", choices=["reference", "chain1", "chain2", "chain3"],
                    help="scheduler policy: chain2 (the library's default: text-identical to the reference's policy, checked in the first "
                         "warm-up step), the reference's own, or the other chain-continuation policies (PIB_SCHED_CHAIN, DESIGN.md §9)")
    ap.add_argument("--no-draft-priority", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-plain", action="store_true", help="skip the non-speculative generation after the timed steps (plain_tok_s, plain_tokens_sha)")
    ap.add_argument("--layer-split", default="", help="comma-separated per-stage weights, as the reference's --mpi-layer-split")
    ap.add_argument("--head-share", type=float, default=None, help="share of the target's layers on the head GPU (which also runs the draft)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.max_spec_runs < 0:
        args.max_spec_runs = world + 1
    (L, E, H, HKV, FF, V, wt), dshape, label = CONFIGS[args.config]
    wbytes = L * (E * (E + 2 * (E // H) * HKV) + E * E + 3 * E * FF) * BPW[wt] / 8
    l2_policy = ("inputs larger than L2: every pass streams the target's weights (%.1f GB >> 126 MB)" % (wbytes / 1e9)) if wbytes > 1e9 else \
                ("smoke configuration: the target's weights (%.0f MB) fit in the 126 MB L2 -- not a bandwidth measurement" % (wbytes / 1e6))
    config = {"workload": "%s, random-init (bigram flavour, draft agreement %.2f), %d-token synthetic prompt, greedy, "
                          "--draft %d -np %d -pa %g -ps %g -n %d -c 1024, at most %s runs in flight, layer split over %d GPU(s)"
                          % (label, args.agree, args.prompt_len, args.draft, args.np, args.pa, args.ps, args.n_predict,
                             args.max_spec_runs or "unbounded", args.gpus),
              "l2_policy": l2_policy}
    # scheduler policy (PIB_SCHED_CHAIN is read by the scheduler at every generation).  A non-reference policy is checked on
    # the GPU before it is timed: the first warm-up step runs the reference policy, the others the chosen one, and any
    # difference in the generated text falls back to the reference policy (the text must be the greedy text under any policy)
    os.environ["PIB_SCHED_CHAIN"] = "0"
    config["sched_policy"] = args.sched_policy

    if args.impl == "reference":
        if rank != 0:
            return 0
        t0 = time.time()
        r = run_reference_sample(args.config, args, budget="large")
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/speculative not built (needs /root/reference at build time)"}))
            return 0
        line = {"impl": "reference", "metric": "accepted tokens/sec", "value": r["value"], "unit": "tokens/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * args.n_predict / r["value"] if r["value"] else None,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int8 x int4/int8 blocks, fp32 accumulate",
                "data": "synthetic", "config": config, "latency_p50_ms": None, "latency_mean_ms": 1000.0 * r["itl_est_s"],
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": r["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "fit_residual": r.get("fit_residual"), "raw": r["raw"], "wall_s": time.time() - t0}
        print(json.dumps(line))
        return 0

    import torch
    from pipeinfer_b200 import llama, pipeline
    lib = llama.api()
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("cpu:gloo,cuda:nccl")
        pipeline.open_pipeline(rank, world, dev)

    def allreduce(vals, op="sum"):
        if world == 1:
            return list(vals)
        t = torch.tensor(list(vals), dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return t.tolist()

    lib.llama_pib_io_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.llama_pib_profile_layer.argtypes = [C.c_void_p, C.c_int]
    lib.llama_pib_profile_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.llama_pib_matrix_shape.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
    cap = 1000 * args.np + 2
    t_load = time.time()
    head_share = args.head_share
    if head_share is None and world > 1:
        l0, r = head_layers(world, (L, E, H, HKV, FF, V, wt), dshape, args.draft)
        head_share = (l0 + 0.25) / L
        config["head_layers_rule"] = "draft work per run ~ %.1f target layers -> %d layer(s) on the head" % (r, l0)
    weights = [float(x) for x in args.layer_split.split(",")] if args.layer_split else pipeline.default_weights(world, head_share)
    fl_t, fl_d = FLAVOUR.get(args.config, (1, 1))
    ranges = pipeline.split_layers(L, weights) if world > 1 else [(0, L)]
    config["layer_ranges"] = ranges
    if world > 1:
        tgt = pipeline.StageContext(rank, world, ranges, 1024, cap, synthetic=(L, E, H, HKV, FF, V, wt, 1, fl_t, 1.0))
    else:
        tgt = llama.Context(None, n_ctx=1024, max_seq=cap, synthetic=(L, E, H, HKV, FF, V, wt, 1, fl_t, 1.0))
    dft = llama.Context(None, n_ctx=1024, max_seq=cap, synthetic=dshape + (2, fl_d, args.agree)) if rank == 0 else None
    if dft is not None and not args.no_draft_priority:
        dft.set_priority(True)                           # the draft sits on the scheduler's critical path
    ctxs = [c for c in (tgt, dft) if c is not None]
    t_load = time.time() - t_load
    prompt = [1] + [3 + (i * 7919) % (V - 3) for i in range(args.prompt_len - 1)]
    params = llama.spec_params(n_draft=args.draft, n_parallel=args.np, p_accept=args.pa, p_split=args.ps, n_predict=args.n_predict, max_spec_runs=args.max_spec_runs)

    def io(ctx):
        a, b = C.c_uint64(), C.c_uint64()
        lib.llama_pib_io_bytes(ctx.ctx, C.byref(a), C.byref(b))
        return a.value, b.value

    def step():
        if rank != 0:                                   # a stage: serve until the head pauses the pipeline
            tgt.serve()
            return None, None, None
        lib.llama_kv_cache_clear(tgt.ctx)
        lib.llama_kv_cache_clear(dft.ctx)
        if args.mode == "plain":
            out = llama.plain_generate(tgt, prompt, args.n_predict)
            out[1].update(n_drafted=0, n_accept=0, n_cancelled=0, n_spec_runs=0)
        else:
            out = llama.speculative_generate(tgt, dft, prompt, params, classic=args.mode == "classic")
        if world > 1:
            tgt.pause()
        return out

    def launches():
        return sum(lib.llama_pib_kernel_launches(c.ctx) for c in ctxs)

    first = None
    policy_env = "0" if args.sched_policy == "reference" or args.mode != "speculative" else args.sched_policy[-1]
    for w in range(args.warmup):
        os.environ["PIB_SCHED_CHAIN"] = policy_env if (w > 0 or args.warmup == 1) else "0"
        toks, res, itl = step()
        first = first or toks
        if rank == 0 and policy_env != "0" and toks[:len(first)] != first[:len(toks)]:
            sys.stderr.write("bench: scheduler policy %s changed the text; falling back to the reference policy\n" % args.sched_policy)
            policy_env = "0"
            config["sched_policy"] = "reference (fallback: %s changed the text)" % args.sched_policy
    os.environ["PIB_SCHED_CHAIN"] = policy_env
    if policy_env != "0":
        config["workload"] += ", scheduler policy %s (additive, text-identical; DESIGN.md §9)" % args.sched_policy
    b0, e0l = ranges[rank] if world > 1 else (0, L)
    prof_layer = (b0 + e0l) // 2
    lib.llama_pib_profile_layer(tgt.ctx, prof_layer)
    k0 = launches()
    io0 = [io(c) for c in ctxs]
    tr0 = pipeline.traffic()
    nv0 = nvlink_tx_kib(dev) if world > 1 else None
    clocks = ClockSampler(dev)
    clocks.start()
    allreduce([0.0])                                    # barrier
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    results, itls = [], []
    for _ in range(args.steps):
        toks, res, itl = step()
        if rank == 0:
            assert toks[:len(first)] == first[:len(toks)], "non-deterministic generation"
            results.append(res)
            itls += itl
    torch.cuda.synchronize()
    e1.record()
    e1.synchronize()
    wall_ms = allreduce([e0.elapsed_time(e1)], "max")[0]        # barrier + max over ranks
    clk = clocks.stop()
    lib.llama_pib_profile_layer(tgt.ctx, -1)
    k1 = launches()
    # outside the timed region: one NON-SPECULATIVE greedy generation through the same pipeline (the reference's `main`): the second
    # denominator of north_star, measured in the same run, and the text every speculative step must reproduce
    plain = None
    if args.mode == "speculative" and not args.no_plain:
        if rank != 0:
            tgt.serve()
        else:
            lib.llama_kv_cache_clear(tgt.ctx)
            ptoks, pres, pitl = llama.plain_generate(tgt, prompt, args.n_predict)
            if world > 1:
                tgt.pause()
            plain = {"tokens": ptoks, "tok_s": pres["n_predict"] / pres["t_decode_s"]}
    io1 = [io(c) for c in ctxs]
    tr1 = pipeline.traffic()
    nv1 = nvlink_tx_kib(dev) if world > 1 else None
    h2d_all, d2h_all, launches_all, wbytes_all = allreduce([
        sum(io1[i][0] - io0[i][0] for i in range(len(ctxs))), sum(io1[i][1] - io0[i][1] for i in range(len(ctxs))), k1 - k0,
        sum(lib.llama_pib_weight_bytes(c.model) for c in ctxs)])
    hops = None
    if world > 1:
        nv = float((nv1 - nv0) * 1024) if (nv0 is not None and nv1 is not None) else -1.0
        per_rank = [torch.zeros(9, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(per_rank, torch.tensor([tr1[i] - tr0[i] for i in range(3)] + [nv] + list(pipeline.cancel_stats()), dtype=torch.float64))
        hops = [{"rank": r, "activation_bytes_to_next_stage": int(t[0]), "messages": int(t[1]), "result_bytes_to_head": int(t[2]),
                 "nvlink_tx_bytes_counter": (int(t[3]) if t[3] >= 0 else None),
                 "cancel_records": int(t[4]), "cancel_to_stage_host_us": {"mean": float(t[5]), "max": float(t[6])},
                 "cancel_to_device_flag_us": ({"mean": float(t[7]), "max": float(t[8])} if os.environ.get("PIB_CANCEL_TIMING") == "1" else None)}
                for r, t in enumerate(per_rank)]
        if rank == 0:
            tgt.shutdown()
        else:
            assert tgt.serve() == 1
        pipeline.close_pipeline()
        dist.destroy_process_group()
        if rank != 0:
            return 0

    n_tok = sum(r["n_predict"] for r in results)
    t_dec = sum(r["t_decode_s"] for r in results)
    itls_ms = sorted(x / 1000.0 for x in itls)
    p50 = itls_ms[len(itls_ms) // 2] if itls_ms else None

    # roofline of the dominant kernel from the per-launch events
    n = 8192
    kind, bb, ms = (C.c_int32 * n)(), (C.c_int32 * n)(), (C.c_float * n)()
    cnt = lib.llama_pib_profile_read(tgt.ctx, kind, bb, ms, n)
    peak, peak_src = measured_peaks()
    per_kind = {}
    for i in range(cnt):
        per_kind.setdefault(kind[i], []).append((bb[i], ms[i]))
    kernels = {}
    others = {}
    for k, lst in list(per_kind.items()):
        if k >= 5:
            others[KIND[k]] = {"launches": len(lst), "mean_ms": sum(x[1] for x in lst) / len(lst)}
            del per_kind[k]
    for k, lst in per_kind.items():
        N_, K_, by = C.c_int64(), C.c_int64(), C.c_uint64()
        lib.llama_pib_matrix_shape(tgt.model, prof_layer, k, C.byref(N_), C.byref(K_), C.byref(by))
        mean_b = sum(x[0] for x in lst) / len(lst)
        mean_ms = sum(x[1] for x in lst) / len(lst)
        alg = by.value + mean_b * K_.value * 4 + mean_b * N_.value * 4        # weights + F32 activations in + F32 out
        kernels[KIND[k]] = {"launches": len(lst), "N": N_.value, "K": K_.value, "mean_batch": round(mean_b, 2), "mean_ms": mean_ms,
                            "algorithmic_bytes": alg, "achieved_gbs": alg / mean_ms / 1e6, "frac": alg / mean_ms / 1e6 / peak}
    dom = kernels.get("gate_up")
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic, traffic_src = tj.get("gate_up_dram_bytes_per_launch"), tj.get("source")
    roofline = None
    if dom:
        roofline = {"bound": "hbm", "kernel": "dequant-GEMV %s, gate|up projection %dx%d, layer %d of the target" % (TYPE_NAME[wt], dom["N"], dom["K"], prof_layer),
                    "achieved": dom["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": dom["frac"], "traffic": traffic,
                    "peak_source": peak_src, "frac_of_8TBs_nominal": dom["achieved_gbs"] / 8000.0, "traffic_source": traffic_src,
                    "note": "event pairs around single launches of layer %d, one verification decode in eight, launched directly and without programmatic "
                            "dependent launch (so the short kernels carry their launch ramp: in the pipelined pass they cost less, profiles/r2_pass_breakdown.md); "
                            "the peak is a COPY bandwidth (half reads, half writes), a read-only weight stream can exceed it: frac > 1 means above the copy "
                            "figure, see frac_of_8TBs_nominal" % prof_layer,
                    "all_matmuls": kernels, "other_kernels_of_the_layer": others}

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        try:
            r = run_reference_sample(args.config, args, budget="small")
            if r:
                cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:      # the baseline is informative; never lose the GPU line over it
            cpu = {"value": None, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "reference", "sample": "failed: %r" % (e,)}

    h2d = int(h2d_all) + 4 * len(prompt) * args.steps
    d2h = int(d2h_all)
    acc = sum(r["n_accept"] for r in results) / max(1, sum(r["n_drafted"] for r in results))
    if args.mode == "plain":
        config["workload"] = config["workload"].replace("greedy,", "greedy, NON-SPECULATIVE pipelined decode (reference `main`),")
    if args.mode == "classic":
        config["workload"] = config["workload"].replace("greedy,", "greedy, CLASSIC synchronous speculation (reference `speculative_orig`),")
    line = {"metric": "accepted tokens/sec", "value": n_tok / t_dec, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": wall_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": {Q4_K: "int8 x int4 blocks (Q4_K x Q8_K), fp32 accumulate", Q4_0: "int8 x int4 blocks (Q4_0 x Q8_0), fp32 accumulate", Q8_0: "int8 x int8 blocks (Q8_0 x Q8_0), fp32 accumulate"}[wt], "data": "synthetic", "config": config,
            "latency_p50_ms": p50, "latency_mean_ms": sum(itls_ms) / len(itls_ms) if itls_ms else None,
            "ttft_ms": 1000.0 * results[-1]["ttft_s"], "accept_rate": acc,
            "spec": {k: sum(r[k] for r in results) / len(results) for k in ("n_predict", "n_drafted", "n_accept", "n_cancelled", "n_spec_runs")},
            "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": n_tok / (wall_ms / 1000.0), "unit": "tokens/s", "h2d_bytes_per_step": h2d // args.steps,
                    "d2h_bytes_per_step": d2h // args.steps},
            "gpu_launches": int(launches_all), "weights_gb": wbytes_all / 1e9, "model_build_s": t_load}
    # the generated ids, hashed: the same value at every N (and equal to the plain-greedy hash) shows that the N-stage pipeline and
    # the speculative scheduler produce the text of single-GPU greedy decoding
    import hashlib

    def sha(ids):
        return hashlib.sha256(b"".join(int(t).to_bytes(4, "little", signed=True) for t in ids)).hexdigest()[:16]

    line["tokens_sha"] = sha(toks[:args.n_predict])
    if plain is not None:
        line["plain_tokens_sha"] = sha(plain["tokens"][:args.n_predict])
        line["plain_tok_s"] = plain["tok_s"]
        line["speedup_over_plain"] = line["value"] / plain["tok_s"]
        line["text_equals_plain_greedy"] = toks[:args.n_predict] == plain["tokens"][:args.n_predict]
    if hops is not None:
        line["nvlink_hops"] = {"per_rank_over_timed_steps": hops,
                               "formula": "n_tokens * n_embd * 4 B of fp32 activations per hop (%d B per token), stored by the sending stage's last kernel into the "
                                          "next stage's memory; 80 B per verification batch (device-side acceptance walk) back to the head" % (E * 4),
                               "nvlink_tx_bytes_counter": "not exposed on this driver: NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX returns NVML_ERROR_NOT_SUPPORTED and "
                                                          "`nvidia-smi nvlink -gt d` prints N/A (tools/nvlink_probe.py); the byte counts above are counted at the "
                                                          "point where each transfer is enqueued" if all(h["nvlink_tx_bytes_counter"] is None for h in hops) else "KiB counters read around the timed steps"}
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""TEST INFRASTRUCTURE: Python side of the scheduler simulator (see sim_llama.cpp).

build() links the scheduler's own object file (pipeinfer_b200/csrc/build/scheduler.o, the one inside libpipeinfer_b200.so)
and the KV bookkeeping (kv_meta.o) with the virtual-clock engine model into tests/sched_sim/libsched_sim.so.
run(...) plays one generation and returns tokens, the true greedy continuation, the scheduler's result and counters.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
BUILD = os.path.join(ROOT, "pipeinfer_b200", "csrc", "build")
SO = os.path.join(HERE, "libsched_sim.so")


class Config(C.Structure):
    _fields_ = [("n_stages", C.c_int32), ("stage_us", C.c_double), ("stage_tok_us", C.c_double), ("hop_us", C.c_double),
                ("ret_us", C.c_double), ("skip_us", C.c_double), ("draft_us", C.c_double), ("draft_tok_us", C.c_double),
                ("enqueue_us", C.c_double), ("call_us", C.c_double), ("agree", C.c_double), ("p_top", C.c_double),
                ("p_second", C.c_double), ("jitter", C.c_double), ("share_gpu0", C.c_double), ("n_vocab", C.c_int32),
                ("seed", C.c_uint64), ("spec_issue_extra_us", C.c_double), ("nonspec_issue_extra_us", C.c_double),
                ("head_frac", C.c_double), ("tail_frac", C.c_double)]


class SpecParams(C.Structure):
    _fields_ = [("n_draft", C.c_int32), ("n_parallel", C.c_int32), ("p_accept", C.c_float), ("p_split", C.c_float),
                ("p_recovery", C.c_float), ("p_decay", C.c_float), ("n_predict", C.c_int32), ("ignore_eos", C.c_int32),
                ("max_spec_runs", C.c_int32)]


class SpecResult(C.Structure):
    _fields_ = [("n_predict", C.c_int32), ("n_drafted", C.c_int32), ("n_accept", C.c_int32), ("n_cancelled", C.c_int32),
                ("n_spec_runs", C.c_int32), ("n_itl", C.c_int32), ("t_encode_s", C.c_double), ("t_decode_s", C.c_double),
                ("decode_tok_per_s", C.c_double), ("ttft_s", C.c_double), ("itl_mean_s", C.c_double), ("itl_p50_s", C.c_double)]


def build():
    objs = [os.path.join(BUILD, "scheduler.o"), os.path.join(BUILD, "kv_meta.o")]
    src = os.path.join(HERE, "sim_llama.cpp")
    if not all(os.path.exists(o) for o in objs):
        subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "pipeinfer_b200", "csrc")], check=True)
    if not os.path.exists(SO) or any(os.path.getmtime(f) > os.path.getmtime(SO) for f in objs + [src]):
        cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I" + os.path.join(cuda, "include"), src] + objs +
                       ["-L" + os.path.join(cuda, "lib64"), "-lcudart", "-o", SO], check=True)
    return SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.sim_new_context.restype = C.c_void_p
        L.sim_new_context.argtypes = [C.c_uint32, C.c_int32, C.c_int]
        L.sim_free_context.argtypes = [C.c_void_p]
        L.sim_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.sim_configure.argtypes = [C.POINTER(Config)]
        L.sim_default_config.argtypes = [C.POINTER(Config)]
        L.sim_truth.argtypes = [C.POINTER(C.c_int32), C.c_int, C.POINTER(C.c_int32), C.c_int]
        L.sim_last_error.restype = C.c_char_p
        L.sim_now_us.restype = C.c_double
        gen = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.POINTER(SpecParams), C.POINTER(C.c_int32), C.c_int,
               C.POINTER(SpecResult), C.POINTER(C.c_double), C.c_int]
        L.pib_speculative_generate.argtypes = gen
        L.pib_speculative_orig_generate.argtypes = gen
        L.pib_plain_generate.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(SpecResult),
                                         C.POINTER(C.c_double), C.c_int]
        L.pib_spec_default_params.restype = SpecParams
        _lib = L
    return _lib


def default_config(**kw):
    cfg = Config()
    lib().sim_default_config(C.byref(cfg))
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


def run(cfg, mode="speculative", n_prompt=24, n_predict=128, n_ctx=1024, n_draft=3, n_parallel=3, p_accept=0.1, p_split=0.9,
        p_recovery=0.0, p_decay=0.0, max_spec_runs=0, prompt_seed=5, prompt=None):
    L = lib()
    L.sim_configure(C.byref(cfg))
    max_seq = 1000 * max(1, n_parallel) + 2 if mode == "speculative" else max(1, n_parallel) + 2
    tgt = L.sim_new_context(n_ctx, max_seq, 1)
    dft = L.sim_new_context(n_ctx, max_seq, 0)
    if prompt is not None:
        n_prompt = len(prompt)
    else:
        prompt = [1] + [3 + (i * 7919 + prompt_seed * 104729) % 30000 for i in range(n_prompt - 1)]
    prompt = (C.c_int32 * n_prompt)(*prompt)
    cap = n_predict + 64
    out = (C.c_int32 * cap)()
    res = SpecResult()
    itl = (C.c_double * 4096)()
    sp = L.pib_spec_default_params()
    sp.n_draft, sp.n_parallel, sp.p_accept, sp.p_split = n_draft, n_parallel, p_accept, p_split
    sp.p_recovery, sp.p_decay, sp.n_predict, sp.ignore_eos, sp.max_spec_runs = p_recovery, p_decay, n_predict, 1, max_spec_runs
    if mode == "speculative":
        n = L.pib_speculative_generate(tgt, dft, prompt, n_prompt, C.byref(sp), out, cap, C.byref(res), itl, 4096)
    elif mode == "classic":
        n = L.pib_speculative_orig_generate(tgt, dft, prompt, n_prompt, C.byref(sp), out, cap, C.byref(res), itl, 4096)
    else:
        n = L.pib_plain_generate(tgt, prompt, n_prompt, n_predict, out, C.byref(res), itl, 4096)
    err = L.sim_last_error().decode()
    truth = (C.c_int32 * cap)()
    L.sim_truth(prompt, n_prompt, truth, cap)
    st_t, st_d = (C.c_int64 * 6)(), (C.c_int64 * 6)()
    L.sim_stats(tgt, st_t)
    L.sim_stats(dft, st_d)
    names = ("decodes", "refused", "cancelled_runs", "bad_context_rows", "read_cancelled_rows", "cells_used")
    L.sim_free_context(tgt)
    L.sim_free_context(dft)
    return {"n": n, "tokens": list(out[:max(n, 0)]), "truth": list(truth[:max(n, 0)]), "error": err,
            "result": {f: getattr(res, f) for f, _ in SpecResult._fields_},
            "target": dict(zip(names, st_t)), "draft": dict(zip(names, st_d))}


if __name__ == "__main__":
    import json
    import sys
    mode = sys.argv[1] if len(sys.argv) > 1 else "speculative"
    r = run(default_config(), mode=mode, max_spec_runs=9 if mode == "speculative" else 0)
    r["ok"] = r["tokens"] == r["truth"]
    del r["tokens"], r["truth"]
    print(json.dumps(r, indent=1))


# ---- the reference's own scheduler on the same model (oracle/_ref/speculative_sim, built by oracle/Makefile) ----------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "speculative_sim")
REF_BIN_CLASSIC = os.path.join(ROOT, "oracle", "_ref", "speculative_orig_sim")     # examples/speculative_orig


def run_reference(cfg_fields, text, n_predict=128, n_ctx=1024, n_draft=3, n_parallel=3, p_accept=0.1, p_split=0.9, p_recovery=0.0,
                  p_decay=0.0, trace=None, workdir="/tmp", classic=False):
    """plays examples/speculative/speculative.cpp (unmodified) against the simulator; returns tokens, counters, tok/s"""
    import json
    import tempfile
    d = tempfile.mkdtemp(dir=workdir)
    env = dict(os.environ, SIM_CFG=",".join("%s=%s" % kv for kv in cfg_fields.items()), SIM_OUT=os.path.join(d, "out.json"))
    if trace:
        env["SIM_TRACE"] = trace
    args = [REF_BIN_CLASSIC if classic else REF_BIN, "-m", "sim:target", "-md", "sim:draft", "-p", text, "-n", str(n_predict), "-c", str(n_ctx), "--draft", str(n_draft),
            "-np", str(n_parallel), "-pa", str(p_accept), "-ps", str(p_split), "-pr", str(p_recovery), "-pd", str(p_decay), "--temp", "-1",
            "--repeat-last-n", "0", "--ignore-eos", "--mpi-layer-split", "1.0/1.0"]
    p = subprocess.run(args, cwd=d, env=env, capture_output=True, text=True, timeout=int(os.environ.get("SIM_REF_TIMEOUT", "120")))
    if p.returncode != 0:
        raise RuntimeError("speculative_sim failed: %s" % p.stderr[-2000:])
    toks = [int(x) for line in p.stdout.splitlines() if not line.startswith("Size of") for x in line.split()]
    stats = json.load(open(os.path.join(d, "out.json")))
    row = open(os.path.join(d, "results.csv")).read().strip().splitlines()[-1].split(",")
    return {"tokens": toks, "stats": stats, "decode_tok_per_s": float(row[1])}


def text_prompt(text):
    return [1] + [3 + b for b in text.encode()]


def truth_of(cfg, prompt, n):
    L = lib()
    L.sim_configure(C.byref(cfg))
    P = (C.c_int32 * len(prompt))(*prompt)
    T = (C.c_int32 * n)()
    L.sim_truth(P, len(prompt), T, n)
    return list(T)

#!/bin/bash
# bash tools/scale_run.sh N TAG [extra bench args]: one bench line at N GPUs into gpurun_out/bench_TAG_nN.json
N=$1; TAG=$2; shift 2
O=gpurun_out; mkdir -p $O
if [ "$N" = "1" ]; then
  python bench.py --gpus 1 "$@" > $O/bench_${TAG}_n1.json 2> $O/bench_${TAG}_n1.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" 2> $O/bench_${TAG}_n$N.err | tail -1 > $O/bench_${TAG}_n$N.json
fi
python - <<PY
import json
try:
    d = json.loads(open("$O/bench_${TAG}_n$N.json").readline())
    print("$TAG N=$N", {k: d.get(k) for k in ("value", "latency_p50_ms", "latency_mean_ms", "accept_rate", "ttft_ms")}, "e2e", d["e2e"]["value"], "frac", (d.get("roofline") or {}).get("frac"), d["config"].get("layer_ranges"))
except Exception as e:
    print("$TAG N=$N FAILED", e); print(open("$O/bench_${TAG}_n$N.err").read()[-1500:])
PY

set -u
O=gpurun_out/exp8; mkdir -p $O
run() { local name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 8 --no-cpu-baseline --no-plain "$@" 2>/dev/null | grep "^{" | tail -1 > $O/$name.json; python -c "
import json; d=json.load(open('$O/$name.json')); print('$name', round(d['value'],1), round(d['e2e']['value'],1), d['latency_p50_ms'], d['spec'], d['config']['layer_ranges'])"; }
run hs003 --head-share 0.03
run hs015 --head-share 0.15
run msr5 --max-spec-runs 5
run msr16 --max-spec-runs 16
run nodp --no-draft-priority

#!/bin/bash
# final bench lines of round 2 for N GPUs (called with N as $1): 70B headline, then the other BASELINE.json configurations that fit N
set -u
N=${1:-1}
O=gpurun_out/r2final
mkdir -p $O
run() {   # name, extra args...
    local name=$1; shift
    if [ "$N" = 1 ]; then timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err
    else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > $O/$name.json 2> $O/$name.err; fi
    tail -c 300 $O/$name.json; echo
}
case $N in
  1) run bench_70b_n1
     run bench_1b_n1 --config 1b --no-cpu-baseline
     timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1300 -c 1600 --csv --log-file $O/launches_bench_cmd.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-plain > $O/ncu_bench.log 2>&1 ;;
  2) run bench_70b_n2 --no-cpu-baseline
     run bench_7b_n2 --config 7b --no-cpu-baseline ;;
  4) run bench_70b_n4 --no-cpu-baseline ;;
  8) run bench_70b_n8 --no-cpu-baseline
     PIB_CANCEL_TIMING=1 run bench_70b_n8_cancel_timing --no-cpu-baseline --no-plain
     run bench_120b_n8 --config 120b --draft 8 --no-cpu-baseline
     run bench_180b_n8 --config 180b --no-cpu-baseline ;;
esac
ls -la $O

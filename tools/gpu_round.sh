#!/bin/bash
# One GPU-box visit: the default bench line, the reference arm, and the ncu evidence kept under profiles/.
#   gpurun --timeout 2400 -- 'bash tools/gpu_round.sh r1'
R=${1:-r1}
O=gpurun_out
mkdir -p $O
python bench.py > $O/bench_${R}_n1.json 2> $O/bench_${R}_n1.err; echo "bench rc=$?"; tail -c 600 $O/bench_${R}_n1.json
python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_${R}_ref.json 2> $O/bench_${R}_ref.err; echo "reference arm rc=$?"; tail -c 400 $O/bench_${R}_ref.json
# launch list of the bench command itself (first decode launches after model generation and most of the prompt)
PIB_GRAPHS=0 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 4000 -c 6000 --csv --log-file $O/launches_${R}_bench.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --n-predict 8 > $O/ncu_bench_${R}.log 2>&1; echo "ncu bench rc=$?"
# one verification pass (batch 4) of the 70B target, every kernel
PIB_GRAPHS=0 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_${R}_pass_b4.csv \
    python tools/profile_step.py --config 70b --batch 4 > $O/ncu_pass_${R}.log 2>&1; echo "ncu pass rc=$?"
# full sections for the matmul kernels of two layers
PIB_GRAPHS=0 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_q4k_direct -c 8 -f -o $O/prof_${R}_gemv \
    python tools/profile_step.py --config 70b --batch 4 --layers 2 > $O/ncu_full_${R}.log 2>&1; echo "ncu full rc=$?"
ls -la $O | tail -12

#!/bin/bash
# Round-2 evidence on one B200: bench line, launch lists, one full ncu capture of the dominant kernel, isolated GEMV / GEMM timings and
# the in-situ cost of every kernel kind.  Everything lands in gpurun_out/r2/ ; the summaries are copied into profiles/ by hand.
set -u
O=gpurun_out/r2
mkdir -p $O
timeout 600 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err
tail -c 400 $O/bench_n1.json
# launch list of the bench command itself (first 600 launches), then of one verification pass of the target at batch 2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_bench_cmd.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-plain > $O/ncu_bench.log 2>&1
PIB_GRAPHS=0 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_target_pass_b2.csv python tools/profile_step.py --config 70b --batch 2 --ctx 128 > $O/ncu_pass.log 2>&1
# full capture of the gate|up launches (grid 3584) of two layers
PIB_GRAPHS=0 timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_q4k_ring -c 8 -o $O/prof_ring python tools/profile_step.py --config 70b --batch 2 --ctx 128 --layers 2 > $O/ncu_ring.log 2>&1
# isolated timings
for shape in "57344 8192" "8192 28672" "10240 8192" "8192 8192" "32000 8192" "2560 2048" "2048 2048" "11264 2048" "2048 5632" "32000 2048"; do set -- $shape; timeout 100 python tools/gemv_bench.py --N $1 --K $2 --b 1,2,4,8,16 --variant 0 2>&1 | grep q4_k; done > $O/gemv_isolated.txt
for shape in "57344 8192" "8192 28672" "10240 8192" "8192 8192"; do set -- $shape; timeout 100 python tools/gemv_bench.py --N $1 --K $2 --b 64,128,256 2>&1 | grep q4_k; done > $O/gemm_tcgen05.txt
timeout 100 python tools/gemv_bench.py --type q4_0 --N 118784 --K 14848 --b 1,4 --variant 0 2>&1 | grep q4_0 >> $O/gemv_isolated.txt
# in-situ costs
timeout 400 python tools/pass_breakdown.py --config 70b --batch 1,2,4 --ctx 128 > $O/pass_breakdown_target.txt 2>&1
timeout 400 python tools/pass_breakdown.py --config 70b --batch 1,2,4 --ctx 128 --alone > $O/pass_alone_target.txt 2>&1
timeout 400 python tools/pass_breakdown.py --config 70b --draft --batch 1 --ctx 128 --passes 25 > $O/pass_breakdown_draft.txt 2>&1
timeout 400 python tools/pass_breakdown.py --config 70b --draft --batch 1 --ctx 128 --passes 25 --alone > $O/pass_alone_draft.txt 2>&1
timeout 200 python tools/ttft_bench.py > $O/ttft.txt 2>&1
ls -la $O

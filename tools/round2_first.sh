#!/bin/bash
# First 8-GPU visit of the next round: what this round could not afford to re-measure.
#   gpurun --gpus 8 --timeout 600 -- 'bash tools/round2_first.sh'
# Each line is one bench run (--no-cpu-baseline, 3 + 3 steps, ~20 s): the default policy after the host KV-mirror fix, the two
# experimental chain policies, the reference's cancel behaviour, and the same at a draft agreement of 0.9.
set -u
A="--no-cpu-baseline --steps 3 --warmup 3"
bash tools/scale_run.sh 8 r2_default $A
bash tools/scale_run.sh 8 r2_chain1 $A --sched-policy chain1
bash tools/scale_run.sh 8 r2_chain2 $A --sched-policy chain2
PIB_REFERENCE_HAZARDS=1 bash tools/scale_run.sh 8 r2_refcancel $A
bash tools/scale_run.sh 8 r2_default_a90 $A --agree 0.9
bash tools/scale_run.sh 8 r2_chain1_a90 $A --agree 0.9 --sched-policy chain1
bash tools/scale_run.sh 8 r2_chain2_a90 $A --agree 0.9 --sched-policy chain2
# one GPU (the driver's headline bench): the model predicts 94 -> 115 tok/s for chain2
# gpurun --timeout 300 -- 'bash tools/scale_run.sh 1 r2_n1 --no-cpu-baseline; bash tools/scale_run.sh 1 r2_n1_chain2 --no-cpu-baseline --sched-policy chain2; python -m pytest tests/test_gpu_scheduler.py -q -m gpu'
#   (and PIB_SCHED_CHAIN=2 python -m pytest tests/test_gpu_scheduler.py tests/test_gpu_cli.py -q -m gpu  before making it the default)
# model: 102 -> 157 tok/s on one GPU, 168 -> 221 on two (gpurun --gpus 2 ... scale_run.sh 2 ...), 245 -> 260 on four

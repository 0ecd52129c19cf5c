#!/bin/bash
# ncu --set full of one kernel of the bench step: bash scripts/gpu_ncu_one.sh <kernel regex> [skip]
K=$1; S=${2:-3}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_$K.log 2>&1
tail -3 gpurun_out/ncu_$K.log

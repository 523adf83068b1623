#!/bin/bash
# one ncu --set full capture of kernel regex $1 (after a plain run of the same command)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$1 -s ${2:-2} -c 1 -o gpurun_out/prof_$1 $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log

#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:radix_pass -s 14 -c 1 -o gpurun_out/prof_radix_pass $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu1 exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:partition_entries -s 1 -c 1 -o gpurun_out/prof_partition_entries $CMD > gpurun_out/ncu_full2.log 2>&1
echo "ncu2 exit $?"

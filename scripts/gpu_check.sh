#!/bin/bash
# parity tests + bench (+ optional ncu launch list when NCU=1)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?" | tee -a gpurun_out/bench.log
tail -3 gpurun_out/bench.log
if [ "$NCU" = "1" ]; then
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
fi
if [ "$NCUFULL" != "" ]; then
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$NCUFULL -s 2 -c 1 -o gpurun_out/prof_$NCUFULL $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
fi

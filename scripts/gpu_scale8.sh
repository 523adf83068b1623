#!/bin/bash
# 8-GPU box: multi-GPU parity, then the bench at N=4 and N=8 (default = balanced owners + p2p), and
# N=8 with the previous scheme (partition mod ranks + NCCL fixed slots) for comparison
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29601 tests/mp_gpu_parity.py > gpurun_out/parity8.log 2>&1; echo "parity8 exit $?"; grep -c "parity ok" gpurun_out/parity8.log
timeout 300 $TR --nproc-per-node 8 --master-port 29602 bench.py --gpus 8 --steps 20 --warmup 3 2> gpurun_out/bench_n8.err | tail -1 > gpurun_out/bench_n8.json; echo "n8 exit $?"
if [ "$OLD" = "1" ]; then
timeout 300 $TR --nproc-per-node 8 --master-port 29603 bench.py --gpus 8 --steps 20 --warmup 3 --shuffle nccl --mod-owners 2> gpurun_out/bench_n8_old.err | tail -1 > gpurun_out/bench_n8_old.json; echo "n8 old exit $?"
fi
timeout 300 $TR --nproc-per-node 2 --master-port 29605 bench.py --gpus 2 --steps 20 --warmup 3 2> gpurun_out/bench_n2.err | tail -1 > gpurun_out/bench_n2.json; echo "n2 exit $?"
timeout 300 $TR --nproc-per-node 4 --master-port 29604 bench.py --gpus 4 --steps 20 --warmup 3 2> gpurun_out/bench_n4.err | tail -1 > gpurun_out/bench_n4.json; echo "n4 exit $?"
python - <<PY
import json
for f in ("bench_n8", "bench_n4", "bench_n2"):
    try:
        j = json.loads(open(f"gpurun_out/{f}.json").read())
        print(f, round(j["value"] / 1e9, 1), "G tok/s", round(j["ms_per_step"], 4), "ms; e2e", round(j["e2e"]["ms_per_step"], 2), "ms;", j["roofline"]["map_ms"], j["roofline"]["reduce_ms"], j["gpu_launches"])
    except Exception as e:
        print(f, "failed", e)
PY
tail -5 gpurun_out/parity8.log

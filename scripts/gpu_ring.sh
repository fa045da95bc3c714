#!/bin/bash
# N-GPU check + bench of the NCCL ring (config 5).  usage: gpurun --gpus N -- ./scripts/gpu_ring.sh N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
timeout 120 $TR tests/ring_nccl_check.py 8192 2>&1 | grep -E "rank|Error|error" | tee gpurun_out/ring_check_n$N.log
timeout 150 $TR bench.py --gpus $N --workload ring --steps 5 --warmup 3 2>/dev/null | tail -1 | tee gpurun_out/ring_bench_n$N.json
timeout 150 $TR bench.py --gpus $N --steps 20 --warmup 3 2>/dev/null | tail -1 | tee gpurun_out/bench_n$N.json

#!/bin/bash
# gpu_multi.sh N [STEPS] : the driver's N-GPU launch of bench.py (strong scaling of config 3 + config5_ring +
# weak figure in one line), then the real-NCCL ring check at T = 131072 (4 heads) on the same N GPUs
N=${1:-2}; STEPS=${2:-10}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
timeout 900 $TR bench.py --gpus $N --steps $STEPS --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -c 2500 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
timeout 600 $TR tests/ring_nccl_check.py 131072 2>&1 | grep -E "rank|Error|error" | tee gpurun_out/ring_check_n${N}_T131072.log

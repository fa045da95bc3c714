#!/bin/bash
# ring bench at N GPUs: NCCL's SM-kernel P2P against NCCL_P2P_USE_CUDA_MEMCPY=1 (copy engines)
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
for mode in default memcpy; do
  if [ $mode = memcpy ]; then export NCCL_P2P_USE_CUDA_MEMCPY=1; else unset NCCL_P2P_USE_CUDA_MEMCPY; fi
  echo -n "p2p=$mode " | tee -a gpurun_out/ring_ce_n$N.log
  timeout 120 $TR bench.py --gpus $N --workload ring --steps 5 --warmup 3 2>gpurun_out/ring_ce_$mode.err | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(round(j['value'],1), 'TFLOP/s', round(j['ms_per_step'],2), 'ms', j['clocks'])" | tee -a gpurun_out/ring_ce_n$N.log
done

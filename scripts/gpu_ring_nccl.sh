#!/bin/bash
# ring bench at N GPUs with different NCCL channel limits (NCCL P2P copies run as SM kernels and
# take SMs from the attention kernels)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
for ch in default 2 4 8; do
  if [ $ch = default ]; then unset NCCL_MAX_NCHANNELS NCCL_MIN_NCHANNELS; else export NCCL_MAX_NCHANNELS=$ch NCCL_MIN_NCHANNELS=$ch; fi
  echo -n "channels=$ch " | tee -a gpurun_out/ring_nccl_channels_n$N.log
  timeout 300 $TR bench.py --gpus $N --workload ring --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(round(j['value'],1), 'TFLOP/s', round(j['ms_per_step'],2), 'ms', j['clocks'])" | tee -a gpurun_out/ring_nccl_channels_n$N.log
done

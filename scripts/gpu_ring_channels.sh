#!/bin/bash
# gpu_ring_channels.sh N CH... : config-5 ring (bench.py --workload ring) on N GPUs with NCCL_MAX_NCHANNELS = each CH
# ("default" = unset): how many SMs NCCL's P2P copy kernels take from the one-CTA-per-SM attention kernels
N=$1; shift
mkdir -p gpurun_out; : > gpurun_out/ring_channels.log
for ch in "$@"; do
  if [ "$ch" = default ]; then unset NCCL_MAX_NCHANNELS; else export NCCL_MAX_NCHANNELS=$ch; fi
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 \
    bench.py --gpus $N --workload ring --steps 3 --warmup 2 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('NCCL_MAX_NCHANNELS=$ch', 'ring %.0f TFLOP/s  %.1f ms per step  %.0f per GPU' % (d['value'], d['ms_per_step'], d['value'] / d['n_gpus']))" | tee -a gpurun_out/ring_channels.log
done

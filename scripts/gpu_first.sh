#!/bin/bash
# first GPU contact: building-block probes, then the C-ABI test binary case by case
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/first.log 2>&1
for m in 0 1 2 3 4; do timeout 60 build/umma_probe $m >> gpurun_out/first.log 2>&1; echo "probe $m exit $?" >> gpurun_out/first.log; done
for f in "L=6 " "L=256 D=128 chunks=2x2 fp32" "tile G=1 L=128 D=128 chunks=1x1 bf16 full" "tile G=1 L=128 D=128 chunks=1x1 bf16 causal" "tile G=2 L=256 D=128" "tile G=2 L=512 D=64" "tile G=3 L=1024" "ragged" "L=2048"; do
  timeout 120 build/tests "$f" >> gpurun_out/first.log 2>&1; echo "tests '$f' exit $?" >> gpurun_out/first.log
done
tail -80 gpurun_out/first.log

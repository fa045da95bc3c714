#!/bin/bash
# CTA-pair (cta_group::2) probes: layout discovery + exactness (modes 0-3), issue rates under
# shared-memory store traffic (modes 10-14).  One process per mode: a trap kills only that mode.
mkdir -p gpurun_out; L=gpurun_out/umma2.log; : > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader >> $L
for m in 0 1 2 3 4; do
  echo "=== mode $m" >> $L
  timeout 60 build/umma2_probe $m >> $L 2>&1; echo "exit $?" >> $L
done
for t in 0 2 8; do
  for m in 10 11 12 13 14; do
    timeout 60 build/umma2_probe $m $t >> $L 2>&1 || echo "mode $m traffic $t exit $?" >> $L
  done
done
cat $L
if [ -n "$WITH_TESTS" ]; then
  timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
fi

#!/bin/bash
# issue rates of the 1-CTA / 2-CTA MMA forms under heavier shared-memory store traffic
mkdir -p gpurun_out; L=gpurun_out/umma2_rate.log; : > $L
for t in 32 128 512; do
  for m in 10 11 13 14; do
    timeout 60 build/umma2_probe $m $t >> $L 2>&1 || echo "mode $m traffic $t exit $?" >> $L
  done
done
cat $L

#!/bin/bash
# C-ABI test binary case by case (a hang or trap in one case does not block the rest)
mkdir -p gpurun_out; : > gpurun_out/ctests.log
for f in "tile G=1 L=128 D=128 chunks=1x1 bf16 full" "tile G=1 L=128 D=128 chunks=1x1 bf16 causal" "tile G=2 L=256 D=128" "tile G=2 L=512 D=64" "tile G=3 L=1024" "ragged" "L=2048" "fp32"; do
  timeout 120 build/tests "$f" >> gpurun_out/ctests.log 2>&1; echo "tests '$f' exit $?" >> gpurun_out/ctests.log
done
grep -E "FAILED|exit|timeout|error" gpurun_out/ctests.log | head -40

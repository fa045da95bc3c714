#!/bin/bash
# gpu_variant_cycles.sh NAME... : SM cycles of the backward kernel under ncu for each build_exp/NAME (no tests)
mkdir -p gpurun_out; L=gpurun_out/variant_cycles.log; : > $L
cp build/libflash_attention.so /tmp/default.so
for n in "$@"; do
  cp build_exp/$n/libflash_attention.so build/libflash_attention.so
  echo "== $n" | tee -a $L
  ncu --metrics sm__cycles_elapsed.max,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
      --clock-control none -k regex:fa_bwd_sm100 -s 2 -c 1 --csv --log-file /tmp/v.csv \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  grep -E 'fa_bwd' /tmp/v.csv | awk -F'","' '{print $(NF-2), $NF}' | tee -a $L
done
cp /tmp/default.so build/libflash_attention.so

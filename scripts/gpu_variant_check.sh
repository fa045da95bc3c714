#!/bin/bash
# gpu_variant_check.sh NAME : swap build_exp/NAME/libflash_attention.so in, run the bf16/fp16 parity
# subset (the tcgen05 kernels), then SM cycles of the two hot kernels under ncu.  Restores the default.
n=$1; mkdir -p gpurun_out; L=gpurun_out/variant_$n.log; : > $L
cp build/libflash_attention.so /tmp/default.so
cp build_exp/$n/libflash_attention.so build/libflash_attention.so
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q \
  -k "packed_vs_oracle or golden or ragged or grouped_query or long_sequence or chunk_hints" > gpurun_out/pytest_$n.log 2>&1
echo "pytest exit $?" | tee -a $L; tail -3 gpurun_out/pytest_$n.log | tee -a $L
ncu --metrics sm__cycles_elapsed.max,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:fa_bwd_sm100 -s 2 -c 2 --csv --log-file /tmp/v.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
grep -E 'fa_bwd' /tmp/v.csv | awk -F'","' '{print $(NF-2), $NF}' | tee -a $L
cp /tmp/default.so build/libflash_attention.so

#!/bin/bash
# gpu_cycles.sh [LABEL] : SM cycles / duration / tensor-pipe activity of the two hot kernels under ncu
# for the library in build/ (one config-3 bench step after warm-up; cycle counts repeat to 0.2 %)
n=${1:-current}; mkdir -p gpurun_out; L=gpurun_out/cycles_$n.log; : > $L
ncu --metrics sm__cycles_elapsed.max,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:fa_.*_sm100 -s 6 -c 2 --csv --log-file /tmp/v.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-small > /dev/null 2>&1
grep -E 'fa_(fwd|bwd)_sm100' /tmp/v.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | sed 's/(.*)//' | tee -a $L

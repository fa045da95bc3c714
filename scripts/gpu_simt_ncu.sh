#!/bin/bash
# per-kernel times + one full capture of each tiled SIMT kernel on the nanoGPT fp32 shape
mkdir -p gpurun_out
CMD="python bench.py --workload nanogpt --steps 2 --warmup 3"
$CMD > gpurun_out/plain_simt.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_simt.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_simt.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:fa_tiled -s 6 -c 3 -o gpurun_out/prof_simt -f $CMD > gpurun_out/ncu_simt.log 2>&1
tail -1 gpurun_out/plain_simt.log | cut -c1-200

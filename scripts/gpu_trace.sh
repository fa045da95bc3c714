#!/bin/bash
# gpu_trace.sh NAME : run one bench step with build_exp/NAME (an FA_BWD_TRACE / FA_FWD_TRACE build) and keep the trace lines
n=${1:-trace}
cp build/libflash_attention.so /tmp/default.so; cp build_exp/$n/libflash_attention.so build/libflash_attention.so
python bench.py --steps 1 --warmup 3 --no-cpu-baseline 2>&1 | grep trace | tail -8 | tee gpurun_out/trace_$n.log
cp /tmp/default.so build/libflash_attention.so

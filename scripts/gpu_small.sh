#!/bin/bash
mkdir -p gpurun_out
for w in nanogpt gpt2 bert; do timeout 200 python bench.py --workload $w --steps 20 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_$w.json; done

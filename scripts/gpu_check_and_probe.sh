#!/bin/bash
# GPU suite, then the power-cap probe of the exp-split variants (build_exp/*), then one bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/gputests.log
timeout 600 bash scripts/gpu_power_probe.sh default bp0 bp6 bp16 fp0 fp2 fp3 noreduce
timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 3000 gpurun_out/bench_n1.json

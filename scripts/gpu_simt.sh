#!/bin/bash
# tiled fp32 path: parity subset + the nanoGPT-shape bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "fp32 or tiled or d2 or nanogpt" 2>&1 | tail -15 | tee gpurun_out/simt_tests.log
timeout 200 python bench.py --workload nanogpt --steps 20 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_nanogpt.json

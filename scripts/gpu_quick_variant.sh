#!/bin/bash
# gpu_quick_variant.sh NAME : shortest possible look at build_exp/NAME: 3 bench steps (fwd / bwd ms), then one parity case
cp build/libflash_attention.so /tmp/default.so; cp build_exp/$1/libflash_attention.so build/libflash_attention.so
python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant $1', d['value'], d['roofline']['ms_fwd'], d['roofline']['ms_bwd'], d['clocks'])" | tee gpurun_out/quick_$1.log
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "packed_vs_oracle and bf16" 2>&1 | tail -1 | tee -a gpurun_out/quick_$1.log
cp /tmp/default.so build/libflash_attention.so

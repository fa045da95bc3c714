#!/bin/bash
# GPT-2 / BERT shaped bench lines (CUDA-graph replay time) for the default build and every build_exp/*/ variant
mkdir -p gpurun_out; : > gpurun_out/small_variants.log
cp build/libflash_attention.so /tmp/default.so
for d in default build_exp/*/; do
  if [ "$d" != default ]; then cp $d/libflash_attention.so build/libflash_attention.so; else cp /tmp/default.so build/libflash_attention.so; fi
  for w in gpt2 bert; do
    echo -n "$d $w " >> gpurun_out/small_variants.log
    timeout 120 python bench.py --workload $w --steps 30 --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(round(j['ms_per_step']*1000,1), 'us', round(j['value'],1), 'TFLOP/s')" >> gpurun_out/small_variants.log 2>&1
  done
done
cp /tmp/default.so build/libflash_attention.so
cat gpurun_out/small_variants.log

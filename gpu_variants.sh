#!/bin/bash
# run the parity tests once on the default build, then a short bench for every variant in build_exp/
mkdir -p gpurun_out; : > gpurun_out/variants.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/variants.log
tail -3 gpurun_out/pytest_gpu.log
cp build/libflash_attention.so /tmp/default.so
for d in default build_exp/*/; do
  if [ "$d" != default ]; then cp $d/libflash_attention.so build/libflash_attention.so; else cp /tmp/default.so build/libflash_attention.so; fi
  for rep in 1 2; do
  echo "== $d" >> gpurun_out/variants.log
  timeout 300 python bench.py --steps ${STEPS:-20} --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    r = j['roofline']; print('total %.1f fwd %.1f bwd %.1f ms_fwd %.3f ms_bwd %.3f clk %s' % (j['value'], r['fwd_tflops'], r['bwd_tflops'], r['ms_fwd'], r['ms_bwd'], j['clocks']))
" >> gpurun_out/variants.log
  done
done
cp /tmp/default.so build/libflash_attention.so
cat gpurun_out/variants.log

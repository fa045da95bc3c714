#!/bin/bash
# gpu_power_probe.sh NAME... : sustained (>= 1.5 s) forward / backward legs of bench.py for build_exp/NAME variants:
# TFLOP/s, ms and the SM clock the backward kernel reports -- what a change does AT THE POWER CAP
mkdir -p gpurun_out; L=gpurun_out/power_probe.log; : > $L
cp build/libflash_attention.so /tmp/default.so
for n in "$@"; do
  [ "$n" = default ] || cp build_exp/$n/libflash_attention.so build/libflash_attention.so
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-small 2>/dev/null | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); r = d['roofline']
print('$n', 'fwd %.1f TFLOP/s %.3f ms | bwd %.1f TFLOP/s %.3f ms | SM clock under load %.0f MHz' % (r['fwd_tflops'], r['ms_fwd'], r['bwd_tflops'], r['ms_bwd'], r['sm_mhz_effective']))" | tee -a $L
  cp /tmp/default.so build/libflash_attention.so
done

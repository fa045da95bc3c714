#!/bin/bash
# gpu_bias_probe.sh D NAME... : tools/bias_cost.py (masked / dropout / plain fwd+bwd and forward-only times) per variant
D=$1; shift
cp build/libflash_attention.so /tmp/default.so
for n in "$@"; do
  [ "$n" = default ] || cp build_exp/$n/libflash_attention.so build/libflash_attention.so
  echo "== $n (D=$D)"; timeout 200 python tools/bias_cost.py $D 2>&1 | grep -E "^non-causal|float bias|forward only"
  cp /tmp/default.so build/libflash_attention.so
done

#!/bin/bash
# gpu_small_probe.sh NAME... : GPT-2-small and BERT-large attention steps (bench.small_config: eager and graph-replayed
# ms per fwd+bwd, L2 flushed between iterations) for build_exp/NAME variants of the library
mkdir -p gpurun_out; L=gpurun_out/small_probe.log; : > $L
cp build/libflash_attention.so /tmp/default.so
for n in "$@"; do
  [ "$n" = default ] || cp build_exp/$n/libflash_attention.so build/libflash_attention.so
  python - <<PY 2>/dev/null | tee -a $L
import torch, bench
dev = torch.device("cuda:0")
r = {k: bench.small_config(k, dev, steps=60) for k in ("gpt2", "bert")}
print("$n", " | ".join("%s eager %.4f ms graph %.4f ms (%.0f TFLOP/s)" % (k, v["ms_eager"], v["ms_graph"], v["tflops_graph"]) for k, v in r.items()))
PY
  cp /tmp/default.so build/libflash_attention.so
done

#!/bin/bash
# Builds (locally, no GPU) the backward variants that were written after round 1's GPU budget ended
# and prints the gpurun command that checks them: parity subset + SM cycles under ncu each
# (default build: 7.92 M cycles, no-drain floor 5.66 M; DESIGN.md sections 4.6 and 7).
set -e
cd "$(dirname "$0")/.."
rm -rf build_exp
bash scripts/build_variant.sh drain_pipe  -DFA_BWD_DRAIN_PIPE=1 &
bash scripts/build_variant.sh drain_regs  -DFA_BWD_DRAIN_REGS=1 &
bash scripts/build_variant.sh stats_regs  -DFA_BWD_STATS_MMA=1 -DFA_BWD_DRAIN_REGS=1 &
wait
make -s build/umma2_probe
cat <<'MSG'
built: build_exp/{drain_pipe,drain_regs,stats_regs}, build/umma2_probe.  Now (about 3 GPU-minutes):
  gpurun --timeout 400 -- 'for v in drain_pipe drain_regs stats_regs; do bash scripts/gpu_variant_check.sh $v; done; timeout 30 build/umma2_probe 4 | tee gpurun_out/umma2_mode4.log'
A variant below 7.7 M cycles with "pytest exit 0": run the full suite with it
  (cp build_exp/NAME/libflash_attention.so build/; gpurun -- python -m pytest tests -m gpu -x -q)
and then make its flag the default in csrc/fa_bwd_sm100.cuh.
MSG

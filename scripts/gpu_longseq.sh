#!/bin/bash
# cycles per KV step at T=32768 (per-CTA prologue/epilogue amortised 4x better than at T=8192)
ncu --metrics sm__cycles_elapsed.max --clock-control none -k regex:fa_.*_sm100 -c 4 --csv --log-file /tmp/v.csv \
  python bench.py --workload ring --ring-seq 32768 --steps 1 --warmup 3 > /dev/null 2>&1
grep -E "fa_(fwd|bwd)" /tmp/v.csv | awk -F'","' '{print $5, $NF}' | cut -c1-120

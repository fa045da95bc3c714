#!/bin/bash
# per-kernel times of the small BASELINE configs (GPT-2 small, BERT-large shapes)
mkdir -p gpurun_out
for w in gpt2 bert; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 > /dev/null 2>&1
done

#!/bin/bash
# programmatic dependent launch on / off: GPU suite with it on, then the bench line (config 3 + small configs) both ways
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/gputests_pdl.log
for pdl in 1 0 1 0; do
  FA_B200_PDL=$pdl timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_pdl$pdl.json
  python - <<PY | tee -a gpurun_out/pdl_compare.log
import json
d = json.loads(open('gpurun_out/bench_pdl$pdl.json').read())
o = d.get('other_configs', {})
print('PDL=$pdl', 'config3 %.1f TFLOP/s %.3f ms | e2e %.2f ms |' % (d['value'], d['ms_per_step'], d['e2e']['ms_per_step']),
      ' | '.join('%s eager %.4f ms graph %.4f ms' % (k, v['ms_eager'], v['ms_graph']) for k, v in o.items()))
PY
done

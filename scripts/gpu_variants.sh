#!/bin/bash
# Clock-independent comparison of kernel variants: SM cycles of the two hot kernels under ncu
# (power-cap clock wander makes wall-time comparisons of a few % meaningless on this pool).
# The default build is tested for parity first; every build_exp/*/ variant is then measured.
mkdir -p gpurun_out; : > gpurun_out/variants.log
if [ -z "$SKIP_TESTS" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/variants.log
tail -3 gpurun_out/pytest_gpu.log
fi
cp build/libflash_attention.so /tmp/default.so
for d in default build_exp/*/; do
  [ -d "$d" ] || [ "$d" = default ] || continue
  if [ "$d" != default ]; then cp $d/libflash_attention.so build/libflash_attention.so; else cp /tmp/default.so build/libflash_attention.so; fi
  echo "== $d" >> gpurun_out/variants.log
  ncu --metrics sm__cycles_elapsed.max,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
      --clock-control none -k regex:fa_.*_sm100 -s 4 -c 6 --csv --log-file /tmp/v.csv \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  python - >> gpurun_out/variants.log <<'PY'
import csv
rows = [r for r in csv.reader(open('/tmp/v.csv')) if len(r) > 10]
h = rows[0]; ik = h.index('Kernel Name'); im = h.index('Metric Name'); iv = h.index('Metric Value')
agg = {}
for r in rows[1:]:
    k = 'fwd' if 'fwd' in r[ik] else 'bwd'
    agg.setdefault((k, r[im]), []).append(float(r[iv].replace(',', '')))
for (k, m), v in sorted(agg.items()):
    print(f"  {k} {m:70s} min {min(v):14.1f} mean {sum(v)/len(v):14.1f} n={len(v)}")
PY
done
cp /tmp/default.so build/libflash_attention.so
cat gpurun_out/variants.log

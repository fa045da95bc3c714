#!/bin/bash
# gpurun_retry.sh [gpurun options] -- 'command' : repeat a gpurun call while the pod answers "transient" / busy
# (nothing is charged for those), at most 20 times, 3 minutes apart.  Output of the final call on stdout.
for attempt in $(seq 1 20); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient\|no box or slot"; then sleep 180; continue; fi
  [ $rc -eq 3 ] && { sleep 180; continue; }
  echo "$out"; echo "[gpurun_retry] attempt $attempt rc $rc"; exit $rc
done
echo "$out"; echo "[gpurun_retry] gave up after 20 attempts"; exit 3

#!/bin/bash
# gpu_contention.sh : tests/cuda/smem_contention.cu sweep (what kind of shared-memory traffic slows tcgen05.mma)
L=gpurun_out/smem_contention.log; : > $L
for ts in 0 1; do
  timeout 60 build/smem_contention 0 256 $ts | tee -a $L
  for kind in 1 2 3 4 5 6 7; do
    for gap in 1 128 256 512 1024; do
      timeout 60 build/smem_contention $kind $gap $ts | tee -a $L
    done
  done
done

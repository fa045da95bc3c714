#!/bin/bash
# TMA store / reduce throughput per SM: how it depends on the number of SMs storing, the spread of the targets, and on MMAs running
L=gpurun_out/smem_contention2.log; : > $L
for kind in 4 5 6; do
  for grid in 148 74 37 8; do
    for region in 64 1024 4096; do
      for mma in 0 1; do
        timeout 60 build/smem_contention $kind 1 1 $grid $region $mma | tee -a $L
      done
    done
  done
done

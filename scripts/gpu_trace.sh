#!/bin/bash
cp build/libflash_attention.so /tmp/default.so; cp build_exp/trace/libflash_attention.so build/libflash_attention.so
python bench.py --steps 1 --warmup 3 --no-cpu-baseline 2>&1 | grep trace | tail -8
cp /tmp/default.so build/libflash_attention.so

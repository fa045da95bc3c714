#!/bin/bash
# build_variant.sh NAME [-DFLAG ...] : compile an experimental variant of the library into build_exp/NAME/
set -e
name=$1; shift
mkdir -p build_exp/$name
nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC "$@" -shared \
  -o build_exp/$name/libflash_attention.so flash_attention_ipu_b200/csrc/fa_api.cu
echo "$@" > build_exp/$name/flags.txt

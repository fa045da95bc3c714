#!/bin/bash
# compute-sanitizer memcheck (out-of-bounds / misaligned global and shared accesses, leaks of
# device-side errors) over a small slice of the GPU suite: both persistent tcgen05 kernels, ragged
# lengths, GQA, the deterministic mode and the SIMT family.  NOTE (round 2): compute-sanitizer is closed on
# this GPU pool (the wrapper answers "closed ... runs under it have left GPUs needing a reset", exit 86), so this
# script is kept for other machines only.  (racecheck does not model the async
# proxy -- TMA writes, tcgen05 reads -- and reports every operand tile; not run.)
mkdir -p gpurun_out
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q \
  -k "golden and bf16 or ragged or test_c_abi_tensor_family_direct or grouped_query and 512 or deterministic and 2048 or tiled_simt and 331 or strongly_negative or bwd_dq_accum or dropout_matches or key_padding or noalloc" tests/test_zz_no_alloc.py \
  > gpurun_out/sanitize_memcheck.log 2>&1
echo "memcheck exit $?" | tee -a gpurun_out/sanitize_memcheck.log
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds|misaligned" gpurun_out/sanitize_memcheck.log | tail -8

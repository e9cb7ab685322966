#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
for tail in 0 70; do
  echo "== B200_GEMM_TAIL=$tail"
  B200_GEMM_TRACE=1 B200_GEMM_TAIL=$tail timeout 600 python tools/gemm_bench.py 512 q4_0 4 > $O/trace_tail$tail.log 2>&1; echo "rc=$?"
  grep -A9 "M=11008" $O/trace_tail$tail.log | tail -10
  grep -A9 "M=14336" $O/trace_tail$tail.log | tail -10
done

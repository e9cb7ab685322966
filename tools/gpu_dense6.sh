#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python tools/gemm_bench.py 512 ${1:-q4_0} > $O/gemm_bench.log 2>&1; echo "gemm_bench rc=$?"; tail -${2:-5} $O/gemm_bench.log
B200_GEMM_TRACE=1 timeout 600 python tools/gemm_bench.py 512 q4_0 2 > $O/trace.log 2>&1
grep -A9 "M=4096" $O/trace.log | tail -10; grep -A9 "M=11008" $O/trace.log | tail -10
for i in 1 2; do timeout 300 python tools/pp_model.py llama2-7b q4_0 512 10 2>&1 | tail -1; done

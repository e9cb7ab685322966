#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
export B200_GEMM_TAIL=0
echo "== BN=128"
B200_GEMM_BN=128 timeout 600 python tools/gemm_bench.py 512 q4_0 > $O/bn128.log 2>&1; tail -5 $O/bn128.log
echo "== BN=256"
timeout 600 python tools/gemm_bench.py 512 q4_0 > $O/bn256.log 2>&1; tail -5 $O/bn256.log
echo "== trace 4096x4096 (BN=256, split 2)"
B200_GEMM_TRACE=1 timeout 600 python tools/gemm_bench.py 512 q4_0 1 > $O/trace_4096.log 2>&1
grep -A9 "M=4096" $O/trace_4096.log | tail -10
echo "== trace BN=128"
B200_GEMM_BN=128 B200_GEMM_TRACE=1 timeout 600 python tools/gemm_bench.py 512 q4_0 2 > $O/trace_bn128.log 2>&1
grep -A9 "M=4096" $O/trace_bn128.log | tail -10
grep -A9 "M=11008" $O/trace_bn128.log | tail -10

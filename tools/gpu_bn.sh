#!/bin/bash
O=gpurun_out; mkdir -p $O
echo "== default (BN=256 + split-K)"; timeout 300 python tools/gemm_bench.py 512 q4_0 3 2>&1 | tail -3
echo "== BN=128"; B200_GEMM_BN=128 timeout 300 python tools/gemm_bench.py 512 q4_0 3 2>&1 | tail -3
echo "== BN=128 trace"; B200_GEMM_BN=128 B200_GEMM_TRACE=1 timeout 300 python tools/gemm_bench.py 512 q4_0 1 2>&1 | tail -9

#!/bin/bash
O=gpurun_out; mkdir -p $O
for d in 0 1 2 4; do echo "== B200_GEMM_DBG=$d"; B200_GEMM_TAIL=0 B200_GEMM_DBG=$d timeout 300 python tools/gemm_bench.py 512 q4_0 4 2>&1 | tail -2; done
B200_GEMM_TAIL=0 B200_GEMM_DBG=0 B200_GEMM_TRACE=1 timeout 300 python tools/gemm_bench.py 512 q4_0 4 2>&1 | tail -12

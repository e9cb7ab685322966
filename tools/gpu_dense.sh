#!/bin/bash
# dense GEMM: parity of both kernels (+ the exact GEMM tests, which share the split-K reduce), the N = 512 shape sweep, a phase trace
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest dense rc=$?"; tail -3 $O/pytest_dense.log
B200_GEMM_CTA=2 timeout 300 python tools/gemm_bench.py 512 q4_0,q6_K > $O/gemm_bench_pair.log 2>&1; echo "pair rc=$?"; cat $O/gemm_bench_pair.log | tail -12
cp $O/gemm_bench.json $O/gemm_bench_pair.json
B200_GEMM_CTA=1 timeout 300 python tools/gemm_bench.py 512 q4_0 > $O/gemm_bench_1cta.log 2>&1; echo "1cta rc=$?"; cat $O/gemm_bench_1cta.log | tail -6
bash tools/gpu_trace.sh

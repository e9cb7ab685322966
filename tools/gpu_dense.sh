#!/bin/bash
# dense GEMM: parity of both kernels (+ the exact GEMM tests, which share the split-K reduce), the N = 512 shape sweep, whole bench
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest dense rc=$?"; tail -8 $O/pytest_dense.log
bash tools/gpu_bench2.sh

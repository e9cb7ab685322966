#!/bin/bash
# quick loop: dense parity tests (fail fast), per-shape times for one type, whole-model pp512
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_gemm_dense.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_dense.log
timeout 600 python tools/gemm_bench.py 512 ${1:-q4_0} > $O/gemm_bench.log 2>&1; echo "gemm_bench rc=$?"; tail -${2:-5} $O/gemm_bench.log
for i in 1 2; do timeout 300 python tools/pp_model.py llama2-7b q4_0 512 10 2>&1 | tail -1; done

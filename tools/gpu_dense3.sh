#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest_dense.log
timeout 600 python tools/gemm_bench.py 512 > $O/gemm_bench.log 2>&1; echo "gemm_bench rc=$?"; tail -30 $O/gemm_bench.log
timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 > $O/pp_model.log 2>&1; echo "pp_model rc=$?"; tail -5 $O/pp_model.log

#!/bin/bash
# dense GEMM: parity of both kernels (+ the exact GEMM tests, which share the split-K reduce), whole-model prefill, whole bench
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest dense rc=$?"; tail -8 $O/pytest_dense.log
timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 2>&1 | tail -1
B200_GEMM_NARROW=0 timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 2>&1 | tail -1
timeout 300 python tools/pp_model.py llama3-8b q4_K_M 512 5 2>&1 | tail -1
B200_GEMM_NARROW=0 timeout 300 python tools/pp_model.py llama3-8b q4_K_M 512 5 2>&1 | tail -1

#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest dense rc=$?"; tail -8 $O/pytest_dense.log
timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 2>&1 | tail -1
B200_FUSE=0 timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 2>&1 | tail -1
timeout 300 python tools/pp_model.py llama3-8b q4_K_M 512 5 2>&1 | tail -1

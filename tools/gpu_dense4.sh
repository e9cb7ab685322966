#!/bin/bash
# dense GEMM: parity tests, per-shape times and the whole-model pp512 step, with and without the stream-K tail
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_gemm_dense.py tests/test_gpu_gemm.py -m gpu -x -q > $O/pytest_dense.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest_dense.log
for tail in 0 70 100; do
  echo "== B200_GEMM_TAIL=$tail"
  B200_GEMM_TAIL=$tail timeout 600 python tools/gemm_bench.py 512 ${1:-q4_0} > $O/gemm_bench_tail$tail.log 2>&1; echo "gemm_bench rc=$?"; tail -6 $O/gemm_bench_tail$tail.log
  B200_GEMM_TAIL=$tail timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 > $O/pp_model_tail$tail.log 2>&1; echo "pp_model rc=$?"; tail -2 $O/pp_model_tail$tail.log
done

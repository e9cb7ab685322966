#!/bin/bash
# the reference's test-backend-ops on the ops graph_ops.cu adds, then llama-bench (tools/gpu_llama_bench.sh)
HERE=$(cd "$(dirname "$0")/.." && pwd)
TAG=${1:-r02_j}
OUT=$HERE/gpurun_out; mkdir -p $OUT
V=$(cd $HERE/oracle && python -c 'import pyref; print(pyref.best_variant())')
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
cd $HERE/oracle/_ref/$V
for op in ${OPS:-MUL SILU ROPE CPY CONT DUP GET_ROWS MUL_MAT ADD RMS_NORM SOFT_MAX}; do
  timeout 600 ./test-backend-ops test -o $op > $OUT/${TAG}_tbo_$op.log 2>&1
  echo "$op rc=$? ok=$(grep -c 'OK' $OUT/${TAG}_tbo_$op.log) fail=$(grep -c 'FAIL' $OUT/${TAG}_tbo_$op.log) notsup=$(grep -c 'not supported' $OUT/${TAG}_tbo_$op.log)"
  grep 'FAIL' $OUT/${TAG}_tbo_$op.log | head -5
done
cd $HERE
[ "${SKIP_LB:-0}" = 1 ] || SKIP_CPU=1 NKVO="${NKVO:-0 1}" bash tools/gpu_llama_bench.sh $TAG

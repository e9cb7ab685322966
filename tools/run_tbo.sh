#!/bin/bash
# Runs the reference's own test-backend-ops (built unmodified into oracle/_ref) against libggml-b200.so.
# usage: tools/run_tbo.sh <mode: test|perf> <op>
HERE=$(cd "$(dirname "$0")/.." && pwd)
V=${REF_VARIANT:-avx2}
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
cd $HERE/oracle/_ref/$V && ./test-backend-ops "$1" -o "$2"

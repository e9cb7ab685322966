#!/bin/bash
# decode sweep of the GEMV tile knobs on the K-quant mix (Llama-3-8B, q4_K / q6_K)
run() { echo -n "$* : "; env "$@" timeout 200 python tools/pp_model.py llama3-8b q4_K_M 1 30 2>&1 | tail -1; }
run B200_QX=1
run B200_GEMV5_MIN_MB=8
run B200_GEMV5_MIN_MB=4
run B200_GEMV5_MIN_MB=2
run B200_GEMV5_MIN_MB=2 B200_GEMV5_ROWS=4
run B200_GEMV5_MIN_MB=2 B200_GEMV5_ROWS=16
run B200_GEMV_ROWS=8
run B200_GEMV_ROWS=2

#!/bin/bash
# decode sweep of the GEMV tile knobs with quantise-on-exit (Llama-2-7B q4_0, one column)
run() { echo -n "$* : "; env "$@" timeout 200 python tools/pp_model.py llama2-7b q4_0 1 30 2>&1 | tail -1; }
run B200_QX=1
run B200_GEMV5_MIN_MB=8
run B200_GEMV5_MIN_MB=4
run B200_GEMV5_ROWS=16
run B200_GEMV5_ROWS=4
run B200_GEMV5_ROWS=16 B200_GEMV5_MIN_MB=4
run B200_GEMV_ROWS=8 
run B200_GEMV_ROWS=2
run B200_QX=0

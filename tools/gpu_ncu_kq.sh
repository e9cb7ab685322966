#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python tools/pp_model.py llama3-8b q4_K_M 1 5 > $O/kq_plain.log 2>&1; echo "plain rc=$?"; tail -1 $O/kq_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv --launch-skip 600 --launch-count 7 -f -o $O/kq_gemv_full \
    python tools/pp_model.py llama3-8b q4_K_M 1 5 > $O/kq_ncu.log 2>&1; echo "full rc=$?"; tail -2 $O/kq_ncu.log

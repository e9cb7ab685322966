#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 300 python tools/pp_model.py llama2-7b q4_0 512 2 > $O/plain_pp.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_dense2 -s 8 -c 3 -f -o $O/dense3_full \
    python tools/pp_model.py llama2-7b q4_0 512 2 > $O/ncu_dense3.log 2>&1
echo "ncu rc=$?"; tail -3 $O/ncu_dense3.log

#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 > $O/pp_model.log 2>&1; echo "pp rc=$?"; tail -2 $O/pp_model.log
B200_PDL=0 timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 2>&1 | tail -1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_dense2|dequant_act|add_f32|reduce_splits" -s 1200 -c 800 --csv --log-file $O/pp_launches.csv python tools/pp_model.py llama2-7b q4_0 512 3 > $O/ncu_pp.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py $O/pp_launches.csv | head -20

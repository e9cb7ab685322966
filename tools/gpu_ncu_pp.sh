#!/bin/bash
# launch list (per-kernel durations, serialised, cold cache) of one whole-model pp512 step
set -u
O=gpurun_out; mkdir -p $O
export B200_GEMM_TAIL=${TAIL:-0}
PP_NOGRAPH=1 timeout 300 python tools/pp_model.py llama2-7b q4_0 512 2 > $O/pp_plain.log 2>&1; echo "plain rc=$?"; tail -1 $O/pp_plain.log
PP_NOGRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/pp_launches.csv python tools/pp_model.py llama2-7b q4_0 512 1 > $O/pp_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py $O/pp_launches.csv | head -20

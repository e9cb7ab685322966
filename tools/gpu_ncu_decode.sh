#!/bin/bash
# decode step (Llama-2-7B q4_0, one column): launch list of the run, and a --set full capture of the LAST layer's 4 GEMV launches + the output matrix
# (129 GEMV launches per step: skip 5 steps + 31 layers)
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python tools/pp_model.py llama2-7b q4_0 1 3 > $O/decode_plain.log 2>&1; echo "plain rc=$?"; tail -1 $O/decode_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/decode_launches.csv python tools/pp_model.py llama2-7b q4_0 1 3 > $O/decode_ncu_list.log 2>&1; echo "list rc=$?"
python tools/launch_summary.py $O/decode_launches.csv | head -12
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv --launch-skip 769 --launch-count 5 -f -o $O/decode_gemv_full \
    python tools/pp_model.py llama2-7b q4_0 1 3 > $O/decode_ncu_full.log 2>&1; echo "full rc=$?"; tail -2 $O/decode_ncu_full.log
ls -la $O

#!/bin/bash
O=gpurun_out; mkdir -p $O
PP_NOGRAPH=1 B200_GEMM_TRACE=1 timeout 300 python tools/pp_model.py llama2-7b q4_0 512 1 > $O/pp_trace.log 2> $O/pp_trace.err; echo "rc=$?"
python - <<'PY'
t=open('gpurun_out/pp_trace.err').read().split('[gemm_dense2 trace]')
for b in t[-9:-5]: print(b)
PY

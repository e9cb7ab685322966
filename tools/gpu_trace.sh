#!/bin/bash
O=gpurun_out; mkdir -p $O
B200_GEMM_TRACE=1 timeout 300 python tools/gemm_bench.py 512 q4_0 > $O/gemm_trace.log 2> $O/gemm_trace.err; echo "trace rc=$?"
# the last trace block of each shape
python - <<'PY'
import re
t=open('gpurun_out/gemm_trace.err').read().split('[gemm_dense2 trace]')
seen={}
for b in t[1:]:
    seen[b.split('\n')[0]]=b
for k,b in seen.items(): print(b)
PY

#!/bin/bash
# round 2: the whole GPU suite (plugin tests included), then the default bench line
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS:-} > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_gpu.log
if [ "${SKIPBENCH:-0}" != "1" ]; then
  timeout 900 python bench.py ${BENCH_ARGS:-} > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
  python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
    print('value', d['value'], 'frac', d['roofline']['frac'], 'e2e', json.dumps(d['e2e'])[:400])
    print('e2e_op_layer', d.get('e2e_op_layer', {}).get('value'), 'rowsplit_n1', d.get('rowsplit_n1'), 'add', d.get('add'))
    print('pp512 model', d.get('pp512', {}).get('model', {}).get('tokens_per_s'), 'cpu', d.get('cpu_baseline'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/bench.err').read()[-3000:])
PY
fi

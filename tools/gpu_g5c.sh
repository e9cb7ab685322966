#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_gemv6.py tests/test_gpu_parity.py tests/test_gpu_rowsplit.py -m gpu -x -q > $O/pytest_g5c.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_g5c.log
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-shapes --no-plugin > $O/bench_$name.json 2> $O/bench_$name.err
  python - <<PY
import json
try:
    d=json.loads(open('$O/bench_$name.json').read().strip().splitlines()[-1])
    print('$name', d['value'], 'tok/s', d['ms_per_step'], 'ms frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
except Exception as e:
    print('$name parse failed', e); print(open('$O/bench_$name.err').read()[-800:])
PY
}
run off B200_GEMV5C=0
for cfg in "$@"; do run "$(echo $cfg | tr ' =' '__')" $cfg; done

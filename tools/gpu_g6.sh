#!/bin/bash
# gemv6 bring-up: parity tests of the new kernel, the timeline, and the step time under a few knob settings (each its own process)
set -u
O=gpurun_out; mkdir -p $O
if [ "${SKIPTESTS:-0}" != "1" ]; then timeout 900 python -m pytest tests/test_gpu_gemv6.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_g6.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_g6.log; fi
if [ "${SKIPTESTS:-0}" != "1" ]; then timeout 300 env B200_GEMV6=1 python tools/g6_trace.py llama2-7b 64 9 > $O/g6_trace.log 2>&1; echo "trace rc=$?"; cat $O/g6_trace.log | tail -20; fi
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-shapes > $O/bench_$name.json 2> $O/bench_$name.err
  python - <<PY
import json
try:
    d=json.loads(open('$O/bench_$name.json').read().strip().splitlines()[-1])
    print('$name', d['value'], 'tok/s', d['ms_per_step'], 'ms frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
except Exception as e:
    print('$name parse failed', e); print(open('$O/bench_$name.err').read()[-800:])
PY
}
run g6off B200_GEMV6=0
run g6def B200_GEMV6=1
for cfg in "$@"; do
  run "$(echo $cfg | tr ' =' '__')" B200_GEMV6=1 $cfg
done

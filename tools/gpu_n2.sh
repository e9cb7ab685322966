#!/bin/bash
O=gpurun_out; mkdir -p $O; N=${NGPU:-2}
if [ "${FULLTESTS:-0}" = "1" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu_n$N.log
  timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
  exit 0
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
    print(d['value'], d['e2e']['value'], d.get('pp512_rowsplit'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/bench_n$N.err').read()[-1500:])
PY

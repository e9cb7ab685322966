#!/bin/bash
# multi-GPU check on one box with $NGPU GPUs: row-split parity tests (every world size that fits), then the bench at N = 2 .. $NGPU
set -u
O=gpurun_out; mkdir -p $O; N=${NGPU:-2}
timeout 900 python -m pytest tests/test_gpu_rowsplit.py -m gpu -q > $O/pytest_rowsplit_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_rowsplit_n$N.log
for n in ${BENCH_NS:-$N}; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps ${STEPS:-10} --warmup 3 --no-cpu > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "bench n$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
    print('n=$n value', d['value'], 'e2e', d['e2e']['value'], 'parity', d.get('rowsplit_parity', d.get('parity')), 'eff', d.get('strong_scaling_eff'), 'pp', d.get('pp512_rowsplit'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/bench_n$n.err').read()[-1500:])
PY
done

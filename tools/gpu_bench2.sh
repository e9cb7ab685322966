#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --no-cpu > $O/bench2.json 2> $O/bench2.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench2.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'])
print(json.dumps(d['pp512']['q4_0'])[:900])
print(json.dumps(d['pp512'].get('model')))
print(json.dumps(d['kquant'])[:1500])
PY

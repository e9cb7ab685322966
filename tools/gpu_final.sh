#!/bin/bash
# final pass of a round on one B200: whole GPU suite, reference arm, bench line, smoke, test-backend-ops logs (compute-sanitizer is closed on this pool)
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $O/rc.log; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.log; tail -1 $O/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench_ref rc=$?" | tee -a $O/rc.log
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.log
timeout 300 tools/run_tbo.sh test MUL_MAT > $O/tbo_mul_mat.log 2>&1; echo "tbo mm rc=$?" | tee -a $O/rc.log
timeout 120 tools/run_tbo.sh test ADD > $O/tbo_add.log 2>&1; echo "tbo add rc=$?" | tee -a $O/rc.log
# the ops of graph_ops.cu under the reference's harness, then llama-bench itself (CPU backend, plugin with the KV cache on the host, plugin fully offloaded)
SKIP_LB=1 OPS="MUL SILU ROPE CPY CONT DUP GET_ROWS SOFT_MAX RMS_NORM" bash tools/gpu_graph_ops.sh final | tee -a $O/rc.log
NKVO="0 1" bash tools/gpu_llama_bench.sh final | tee -a $O/rc.log
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'frac', d['roofline']['frac'], 'traffic', d['roofline']['traffic'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
print('pp512', d.get('pp512', {}).get('model', {}).get('tokens_per_s'), d.get('pp512', {}).get('model', {}).get('frac_of_peaks'))
print('llama_bench', d.get('llama_bench'))
print('ref', open('gpurun_out/bench_ref.json').read().strip()[-700:])
PY

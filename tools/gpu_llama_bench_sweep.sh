#!/bin/bash
# robustness sweep: llama-bench configurations beside the headline one (longer prompts and generations, small ubatches, flash attention and a
# quantised KV cache -- ops the plugin declines, which the scheduler then keeps on ggml-cpu); every line must end rc=0 with sane rates
HERE=$(cd "$(dirname "$0")/.." && pwd)
TAG=${1:-r02_u}
OUT=$HERE/gpurun_out; mkdir -p $OUT
V=$(cd $HERE/oracle && python -c 'import pyref; print(pyref.best_variant())')
BIN=$HERE/oracle/_ref/$V/llama-bench
M=/dev/shm/l7b.gguf
python $HERE/tools/make_gguf.py $M --preset 7b 2> /dev/null || exit 1
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
run() { name=$1; shift; timeout 600 $BIN -m $M -ngl 99 -r 2 -o jsonl "$@" > $OUT/${TAG}_sweep_$name.jsonl 2> $OUT/${TAG}_sweep_$name.err; echo "$name rc=$? $(python - <<PY
import json
o = []
for l in open("$OUT/${TAG}_sweep_$name.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    o.append(("pp%d" % d["n_prompt"] if d["n_prompt"] else "tg%d" % d["n_gen"]) + " %.1f" % d["avg_ts"])
print("; ".join(o))
PY
)"; grep -i 'abort\|error\|assert' $OUT/${TAG}_sweep_$name.err | head -3 | cut -c1-250; }
run pp2048 -p 2048 -n 32
run tg1100 -p 0 -n 1100
run ub32 -p 128 -n 16 -b 64 -ub 32
run ub8 -p 64 -n 8 -b 8 -ub 8
run pg -p 0 -n 0 -pg 512,64
run fa1 -p 64 -n 16 -fa 1
run nkvo_q8 -p 64 -n 16 -nkvo 1 -ctk q8_0
# the cDSP admission table alone (MUL_MAT / ADD / RMS_NORM / SOFT_MAX offloaded, the rest on ggml-cpu, KV cache on the host)
B200_GRAPH_OPS=0 run cdsp_table -p 512 -n 128 -nkvo 1
# Llama-3.1-style frequency factors: ROPE with a third operand stays on the device
python $HERE/tools/make_gguf.py $M --preset 7b --rope-freqs 2> /dev/null || exit 1
run ropefreqs -p 512 -n 64
rm -f $M

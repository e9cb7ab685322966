#!/bin/bash
# llama-bench with the layers split over all visible B200s by llama.cpp itself (-sm layer): the scheduler moves the activations between the
# plugin's devices through cpy_tensor_async + events.  usage (gpurun --gpus N): bash tools/gpu_llama_bench_multi.sh [tag]
HERE=$(cd "$(dirname "$0")/.." && pwd)
TAG=${1:-r02_r}
OUT=$HERE/gpurun_out; mkdir -p $OUT
V=$(cd $HERE/oracle && python -c 'import pyref; print(pyref.best_variant())')
BIN=$HERE/oracle/_ref/$V/llama-bench
M=/dev/shm/l7b.gguf
python $HERE/tools/make_gguf.py $M --preset ${GGUF_PRESET:-7b} --type ${GGUF_TYPE:-q4_0} ${GGUF_EXTRA:-} 2> /dev/null || exit 1
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
nvidia-smi -L | head -8
timeout 600 $BIN -m $M -p 512 -n 128 -ngl 99 -sm layer -r 3 -o jsonl > $OUT/${TAG}_llama_bench_sm_layer.jsonl 2> $OUT/${TAG}_llama_bench_sm_layer.err
echo "sm layer rc=$?"; tail -5 $OUT/${TAG}_llama_bench_sm_layer.err | cut -c1-300
timeout 600 $BIN -m $M -p 512 -n 128 -ngl 99 -sm none -mg 1 -r 3 -o jsonl > $OUT/${TAG}_llama_bench_mg1.jsonl 2> $OUT/${TAG}_llama_bench_mg1.err
echo "main-gpu 1 rc=$?"; tail -3 $OUT/${TAG}_llama_bench_mg1.err | cut -c1-300
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/${TAG}_llama_bench_*.jsonl")):
    for l in open(f):
        try: d = json.loads(l)
        except Exception: continue
        print(f.split("/")[-1], "pp" if d["n_prompt"] else "tg", d["n_prompt"] or d["n_gen"], "%.2f t/s" % d["avg_ts"], d.get("backends"), d.get("split_mode"), d.get("gpu_info"))
PY
rm -f $M

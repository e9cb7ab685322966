#!/bin/bash
# the reference's own llama-bench (built unmodified into oracle/_ref/<variant>/ by `make -f oracle/Makefile.ref llama`)
# on a synthetic Llama-2-7B Q4_0 GGUF (tools/make_gguf.py), once on the reference's CPU backend and once with libggml-b200.so
# loaded through GGML_BACKEND_PATH, the ggml scheduler splitting the graph between the two.
# usage (on the GPU box): bash tools/gpu_llama_bench.sh [tag] [extra llama-bench args for the plugin run]
HERE=$(cd "$(dirname "$0")/.." && pwd)
TAG=${1:-r02_i}; shift
OUT=$HERE/gpurun_out; mkdir -p $OUT
V=${REF_VARIANT:-$(cd $HERE/oracle && python -c 'import pyref; print(pyref.best_variant())')}
BIN=$HERE/oracle/_ref/$V/llama-bench
M=/dev/shm/l7b.gguf
python $HERE/tools/make_gguf.py $M --preset ${GGUF_PRESET:-7b} --type ${GGUF_TYPE:-q4_0} ${GGUF_EXTRA:-} 2> $OUT/${TAG}_gguf.log || exit 1
NT=$(nproc)
if [ "${SKIP_CPU:-0}" != 1 ]; then
  timeout 900 $BIN -m $M -p ${CPU_PP:-512} -n ${CPU_TG:-32} -ngl 0 -t $NT -r 1 -o jsonl > $OUT/${TAG}_llama_bench_cpu.jsonl 2> $OUT/${TAG}_llama_bench_cpu.err
  echo "cpu rc=$?" >> $OUT/${TAG}_llama_bench_cpu.err
fi
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
for KVO in ${NKVO:-1 0}; do
  timeout 900 $BIN -m $M -p 512 -n 128 -ngl 99 -nkvo $KVO -t $NT -r ${REPS:-3} -o jsonl "$@" > $OUT/${TAG}_llama_bench_b200_nkvo$KVO.jsonl 2> $OUT/${TAG}_llama_bench_b200_nkvo$KVO.err
  echo "b200 nkvo=$KVO rc=$?" >> $OUT/${TAG}_llama_bench_b200_nkvo$KVO.err
done
tail -n 3 $OUT/${TAG}_llama_bench_*.err
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/${TAG}_llama_bench_*.jsonl")):
    for l in open(f):
        try: d = json.loads(l)
        except Exception: continue
        print(f.split("/")[-1], "pp" if d["n_prompt"] else "tg", d["n_prompt"] or d["n_gen"], "%.2f t/s" % d["avg_ts"], d.get("backends"), "ngl", d.get("n_gpu_layers"))
PY
rm -f $M

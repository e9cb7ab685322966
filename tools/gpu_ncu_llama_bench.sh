#!/bin/bash
# per-kernel durations of a few llama-bench decode steps (ncu serialises launches: the SHARE per kernel is what to read)
HERE=$(cd "$(dirname "$0")/.." && pwd)
TAG=${1:-r02_j}
OUT=$HERE/gpurun_out; mkdir -p $OUT
V=$(cd $HERE/oracle && python -c 'import pyref; print(pyref.best_variant())')
BIN=$HERE/oracle/_ref/$V/llama-bench
M=/dev/shm/l7b.gguf
python $HERE/tools/make_gguf.py $M --preset 7b 2> /dev/null || exit 1
export GGML_BACKEND_PATH=$HERE/ggml-hexagon_b200/lib/libggml-b200.so
ARGS="-m $M -p ${PP:-0} -n ${TG:-6} -ngl 99 -nkvo 0 -r 1 -o jsonl"
timeout 300 $BIN $ARGS > $OUT/${TAG}_ncu_plain.log 2>&1 || { echo plain run failed; tail -5 $OUT/${TAG}_ncu_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_llama_bench_launches.csv $BIN $ARGS > $OUT/${TAG}_ncu_run.log 2>&1
echo ncu rc=$?
python - <<PY
import csv, collections
rows = [r for r in csv.reader(l for l in open("$OUT/${TAG}_llama_bench_launches.csv") if l.startswith('"'))]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try: v = float(r[vi].replace(",", ""))
    except Exception: continue
    k = r[ki].split("(")[0][:60]; tot[k][0] += 1; tot[k][1] += v
s = sum(v[1] for v in tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])[:25]:
    print("%-62s n=%6d  total %10.1f us  %5.1f%%  avg %8.2f us" % (k, v[0], v[1] / 1e3, 100 * v[1] / s, v[1] / v[0] / 1e3))
print("total", s / 1e3, "us over", sum(v[0] for v in tot.values()), "launches")
PY
rm -f $M

#!/bin/bash
# Development sweep of the q4_0 N=1 GEMV variants (run on the GPU box): prints tokens/s and per-shape GB/s per config.
out=gpurun_out/tune.txt; : > $out
run() { echo "== $*" >> $out; env "$@" python bench.py --steps 20 --warmup 3 --no-cpu 2>>$out | python -c '
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print("tok/s", j["value"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], "launches", j["gpu_launches"], "finite", j["logits_finite"])
    print("  ", " | ".join("%s %.2fus %.0f" % (s["shape"].split()[0], s["us"], s["GB/s"]) for s in j.get("shapes", [])))
' >> $out; }
run B200_FUSE=0 B200_PDL=0
run B200_FUSE=1 B200_PDL=0
run B200_FUSE=0 B200_PDL=1
for r in 2 4 8; do run B200_GEMV_ROWS=$r; done
cat $out

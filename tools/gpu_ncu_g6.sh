#!/bin/bash
# one ncu --set full capture of the decode GEMV (4 consecutive launches = one layer) after a plain run of the same command
O=gpurun_out; mkdir -p $O
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-shapes > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:b200_gemv6 -s 260 -c 4 -f -o $O/g6_full \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-shapes > $O/ncu_g6.log 2>&1
echo "ncu rc=$?"; tail -3 $O/ncu_g6.log

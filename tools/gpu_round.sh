#!/bin/bash
# One GPU-box pass: parity tests, both bench arms, reference test-backend-ops, ncu launch list + full captures.
# Every profiler run follows a plain run of the same command that exited 0.
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench_ref rc=$?" | tee -a $O/rc.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; rc=$?; echo "bench rc=$rc" | tee -a $O/rc.log
timeout 300 tools/run_tbo.sh test MUL_MAT > $O/tbo_mul_mat.log 2>&1; echo "tbo mm rc=$?" | tee -a $O/rc.log
timeout 120 tools/run_tbo.sh test ADD > $O/tbo_add.log 2>&1; echo "tbo add rc=$?" | tee -a $O/rc.log
timeout 300 python tools/gemm_bench.py 512 > $O/gemm_bench.log 2>&1; echo "gemm_bench rc=$?" | tee -a $O/rc.log
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-shapes > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?" | tee -a $O/rc.log
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv5_kernel -s 200 -c 4 -f -o $O/gemv5_full \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-shapes > $O/ncu_gemv5.log 2>&1; echo "ncu gemv5 rc=$?" | tee -a $O/rc.log
  timeout 300 python tools/pp_model.py llama2-7b q4_0 512 5 > $O/pp_model.log 2>&1; echo "pp_model rc=$?" | tee -a $O/rc.log
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 800 --csv --log-file $O/pp_launches.csv \
      python tools/pp_model.py llama2-7b q4_0 512 3 > $O/ncu_pp_launch.log 2>&1; echo "ncu pp launches rc=$?" | tee -a $O/rc.log
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_dense2 -s 300 -c 4 -f -o $O/gemm_dense2_full \
      python tools/pp_model.py llama2-7b q4_0 512 2 > $O/ncu_dense.log 2>&1; echo "ncu dense2 rc=$?" | tee -a $O/rc.log
fi
tail -3 $O/pytest_gpu.log; cat $O/bench.json | cut -c1-600

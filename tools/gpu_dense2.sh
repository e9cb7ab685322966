#!/bin/bash
O=gpurun_out; mkdir -p $O
B200_GEMM_BN=128 timeout 300 python tools/gemm_bench.py 512 q4_0 > $O/gemm_bench_pair_bn128.log 2>&1; echo "pair bn128 rc=$?"; cat $O/gemm_bench_pair_bn128.log | tail -6
timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__t_sectors_srcunit_tex_op_read.sum,dram__bytes_read.sum --clock-control none -s 40 -c 200 --csv --log-file $O/dense_pair_launches.csv python tools/gemm_bench.py 512 q4_0 > $O/ncu_pair.log 2>&1; echo "ncu rc=$?"

cd /root/repo
SKIP_LB=1 OPS="CPY CONT DUP SOFT_MAX MUL_MAT" bash tools/gpu_graph_ops.sh r02_n
python -m pytest tests/test_gpu_graph_ops.py tests/test_gpu_norm_ops.py -x -q -m gpu 2>&1 | tail -5
SKIP_CPU=1 NKVO=0 bash tools/gpu_llama_bench.sh r02_n
V=$(cd oracle && python -c 'import pyref; print(pyref.best_variant())')
python tools/make_gguf.py /dev/shm/l7b.gguf --preset 7b 2>/dev/null
B200_GRAPH_TIME=1 GGML_BACKEND_PATH=$PWD/ggml-hexagon_b200/lib/libggml-b200.so oracle/_ref/$V/llama-bench -m /dev/shm/l7b.gguf -p 512 -n 32 -r 2 -ngl 99 > gpurun_out/r02_n_graph_time.log 2>&1
grep 'graph replay' gpurun_out/r02_n_graph_time.log | sort | uniq -c | sort -rn | head -6

cd /root/repo
V=$(cd oracle && python -c 'import pyref; print(pyref.best_variant())')
python tools/make_gguf.py /dev/shm/l7b.gguf --preset 7b 2>/dev/null
B200_GRAPH_TIME=1 GGML_BACKEND_PATH=$PWD/ggml-hexagon_b200/lib/libggml-b200.so oracle/_ref/$V/llama-bench -m /dev/shm/l7b.gguf -p 0 -n 64 -r 1 -ngl 99 > gpurun_out/r02_l_graph_time.log 2>&1
grep 'graph replay' gpurun_out/r02_l_graph_time.log | sort | uniq -c | sort -rn | head -5
grep 'graph replay' gpurun_out/r02_l_graph_time.log | awk '{s+=$(NF-1); n++} END {print "mean ms", s/n, "n", n}'
tail -3 gpurun_out/r02_l_graph_time.log

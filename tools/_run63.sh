cd /root/repo
V=$(cd oracle && python -c 'import pyref; print(pyref.best_variant())')
python tools/make_gguf.py /dev/shm/l7b.gguf --preset 7b 2>/dev/null
GGML_B200_DUMP=1 GGML_BACKEND_PATH=$PWD/ggml-hexagon_b200/lib/libggml-b200.so oracle/_ref/$V/llama-bench -m /dev/shm/l7b.gguf -p 0 -n 4 -r 1 -ngl 99 > gpurun_out/r02_k_dump.log 2>&1
grep -c node gpurun_out/r02_k_dump.log

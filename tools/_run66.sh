cd /root/repo
python -m pytest tests/test_gpu_graph_ops.py tests/test_gpu_plugin.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -5
SKIP_CPU=1 NKVO=0 bash tools/gpu_llama_bench.sh r02_m
PP=512 TG=0 bash tools/gpu_ncu_llama_bench.sh r02_m_pp

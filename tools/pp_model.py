#!/usr/bin/env python
"""Whole-model prefill (bench.py's pp512 'model' leg alone): every MUL_MAT + residual ADD of the model over N columns as one
b200_graph_compute per step.  usage: pp_model.py [model=llama2-7b] [quant=q4_0] [N=512] [steps=5]  -- for profiler runs."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ggml-hexagon_b200"))
import bench
import b200ops as B

def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b"
    quant = sys.argv[2] if len(sys.argv) > 2 else "q4_0"
    N = int(sys.argv[3]) if len(sys.argv) > 3 else 512
    steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    with B.Session(0) as s:
        pw = bench.DecodeWorkload(B, s, model, quant=quant, ncols=N)
        if os.environ.get("PP_NOGRAPH"):          # plain launches instead of a CUDA graph replay (for B200_GEMM_TRACE)
            pw.step = lambda: s.graph_compute(pw._nodes, use_cuda_graph=False)
        for _ in range(3):
            pw.step()
        s.sync()
        l0 = s.launch_count()
        e0, e1 = s.event(), s.event()
        s.record(e0)
        for _ in range(steps):
            pw.step()
        s.record(e1)
        ms = s.elapsed_ms(e0, e1) / steps
        print("pp%d %s %s: %.3f ms per step, %.1f tokens/s, %d kernels per step" % (N, model, quant, ms, N / ms * 1e3, (s.launch_count() - l0) // steps))
        pw.free()

main()

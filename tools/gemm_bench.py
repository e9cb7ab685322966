#!/usr/bin/env python
"""Times prefill MUL_MATs (many activation columns) through the C-ABI: per shape us, TOP/s (2*M*N*K), device-resident."""
import sys, os, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-hexagon_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ctypes as C
import b200ops as B
import pyoracle as O

def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    types = [(O.Q8_0, "q8_0"), (O.Q4_0, "q4_0"), (O.Q4_K, "q4_K"), (O.Q6_K, "q6_K")]
    if len(sys.argv) > 2:
        types = [x for x in types if x[1] in sys.argv[2].split(",")]
    shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (14336, 4096), (4096, 14336)]
    if len(sys.argv) > 3:
        shapes = shapes[:int(sys.argv[3])]
    rng = np.random.default_rng(0)
    out = []
    with B.Session(0) as s:
        for t, name in types:
            for M, K in shapes:
                if K % 256 and t in (O.Q4_K, O.Q6_K):
                    continue
                raw = O.random_blocks(t, 128, K, rng)
                raw = np.tile(raw, (M // 128, 1))
                w = s.weights(t, raw, K)
                x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
                xb, db = s.to_device(x), s.malloc(4 * N * M)
                s0, s1, d = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(db.ptr, [M, N], op=B.OP_MUL_MAT)
                for _ in range(3):
                    B._ck(B.lib().b200_mulmat(s.h, C.byref(s0), C.byref(s1), C.byref(d)))
                s.sync()
                e0, e1 = s.event(), s.event()
                iters = 20
                s.record(e0)
                for _ in range(iters):
                    B._ck(B.lib().b200_mulmat(s.h, C.byref(s0), C.byref(s1), C.byref(d)))
                s.record(e1)
                us = s.elapsed_ms(e0, e1) * 1e3 / iters
                out.append({"type": name, "M": M, "K": K, "N": N, "us": round(us, 1), "TOPS": round(2.0 * M * N * K / us / 1e6, 1)})
                print(out[-1], flush=True)
                for b in (xb, db): b.free()
                w.free()
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gemm_bench.json"), "w"))

main()

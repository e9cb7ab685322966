#!/usr/bin/env python
"""bench.py — measures BASELINE.json's metric: tokens/s of the Llama-shaped Q4_0 decode MUL_MAT/ADD sweep on B200,
with the MUL_MAT fraction of the HBM roofline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one decoded token's worth of the hot path: every MUL_MAT of every layer (batch 1) + the residual ADDs,
submitted as ONE b200_graph_compute call (CUDA-graph replay) through the C-ABI (include/ggml_b200_ops.h).
N = 1 : Llama-2-7B shapes (BASELINE.json configs[1]).        weights 3.7 GB >> L2, streamed from HBM every step.
N > 1 : Llama-3-70B shapes row-split over N GPUs (configs[4]), one process per GPU (torchrun), NCCL all-gather of the
        output slices after every MUL_MAT; strong scaling (total work fixed).
--impl reference : the reference's own CPU implementation (oracle/_ref ggml-cpu, all host threads) on the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "ggml-hexagon_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

Q4_0 = 2
METRIC = "tokens/s Llama Q4_0 decode (tg, MUL_MAT+ADD path, batch 1)"


# ------------------------------------------------------------------------------------------------ workloads
def llama_shapes(model):
    """(name, M, K) of one layer's MUL_MATs, n_layers, (output M, K), hidden"""
    if model == "llama2-7b":
        h, f, kv, L, V = 4096, 11008, 4096, 32, 32000
    elif model == "llama3-70b":
        h, f, kv, L, V = 8192, 28672, 1024, 80, 128256
    elif model == "llama3-8b":
        h, f, kv, L, V = 4096, 14336, 1024, 32, 128256
    else:
        raise ValueError(model)
    layer = [("wq", h, h), ("wk", kv, h), ("wv", kv, h), ("wo", h, h), ("gate", f, h), ("up", f, h), ("down", h, f)]
    return layer, L, ("output", V, h), h, f


def gemv_bytes(M, K, N=1, blk=(32, 18)):
    """algorithmic bytes of one MUL_MAT: weights once + f32 activations in + f32 out"""
    return M * (K // blk[0]) * blk[1] + N * K * 4 + N * M * 4


def random_q4_0(rows, K, rng):
    """random-init q4_0 weights in the wire format: random nibbles, fp16 scale ~U(0.001,0.003) with random sign
    (|w| ~ 0.01, keeps a 32..80 layer chain finite)"""
    nb = K // 32
    raw = rng.integers(0, 256, size=(rows, nb, 18), dtype=np.uint8)
    d = (rng.uniform(0.001, 0.003, size=(rows, nb)) * rng.choice([-1.0, 1.0], size=(rows, nb))).astype(np.float16).view(np.uint16)
    raw[:, :, 0] = (d & 0xFF).astype(np.uint8)
    raw[:, :, 1] = (d >> 8).astype(np.uint8)
    return raw.reshape(rows, nb * 18)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
                  "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------ B200 arm
class DecodeWorkload:
    """All MUL_MAT/ADD nodes of one decoded token, device resident.  Row r-th slice of every weight when world > 1."""

    def __init__(self, B, sess, model, rank=0, world=1, seed=1234):
        import ctypes as C
        self.B, self.s, self.C = B, sess, C
        self.rank, self.world = rank, world
        layer, L, out, h, f = llama_shapes(model)
        self.model, self.h, self.f, self.L = model, h, f, L
        rng = np.random.default_rng(seed)            # same seed on every rank: identical full matrices, each rank keeps its rows
        self.weights = []                            # [layer][name] -> Weights
        self.alg_bytes = 0
        self.n_mulmat = 0
        host = {}
        for name, M, K in layer + [out]:
            assert M % world == 0, (name, M, world)
            full_rows = M // world
            # one random host matrix per shape; every layer gets its own HBM copy (distinct addresses, no L2 reuse)
            host[name] = random_q4_0(full_rows, K, rng)
        for _ in range(L):
            self.weights.append({name: sess.weights(Q4_0, host[name], K) for name, M, K in layer})
        self.w_out = sess.weights(Q4_0, host[out[0]], out[2])
        self.layer, self.out = layer, out
        for name, M, K in layer:
            self.alg_bytes += L * gemv_bytes(M, K)
            self.n_mulmat += L
        self.alg_bytes += gemv_bytes(out[1], out[2])
        self.n_mulmat += 1
        self.weight_bytes_local = sum(w.nbytes for lw in self.weights for w in lw.values()) + self.w_out.nbytes
        # activations (f32, device)
        kv = dict((n, m) for n, m, _ in layer)["wk"]
        mk = lambda n: sess.malloc(4 * n)
        self.x = mk(h); self.q = mk(h); self.k = mk(kv); self.v = mk(kv); self.o = mk(h); self.hres = mk(h)
        self.g = mk(f); self.u = mk(f); self.dn = mk(h); self.x2 = mk(h); self.logits = mk(out[1])
        self.x_host = np.random.default_rng(seed + 1).uniform(-1, 1, h).astype(np.float32)
        sess.h2d(self.x, self.x_host)
        self._nodes = None
        self._build()

    def _mm(self, w, src, K, dst, M):
        B = self.B
        n = B.Node()
        n.src0 = w.tensor()
        n.src1 = B.f32_tensor(src.ptr, [K, 1])
        n.dst = B.f32_tensor(dst.ptr, [M, 1], op=B.OP_MUL_MAT)
        return n

    def _add(self, a, b, dst, n_el):
        B = self.B
        n = B.Node()
        n.src0, n.src1, n.dst = B.f32_tensor(a.ptr, [n_el]), B.f32_tensor(b.ptr, [n_el]), B.f32_tensor(dst.ptr, [n_el], op=B.OP_ADD)
        return n

    def _build(self):
        h, f = self.h, self.f
        dims = dict((n, (m, k)) for n, m, k in self.layer)
        nodes = []
        xin, xout = self.x, self.x2
        for lw in self.weights:
            # attention block: q,k,v share src1 (quantised once); attention itself is not on this path -> q stands in
            nodes.append(self._mm(lw["wq"], xin, h, self.q, dims["wq"][0]))
            nodes.append(self._mm(lw["wk"], xin, h, self.k, dims["wk"][0]))
            nodes.append(self._mm(lw["wv"], xin, h, self.v, dims["wv"][0]))
            nodes.append(self._mm(lw["wo"], self.q, h, self.o, h))
            nodes.append(self._add(xin, self.o, self.hres, h))
            # MLP block: gate, up share src1; down consumes an f-vector (gate stands in for silu(gate)*up)
            nodes.append(self._mm(lw["gate"], self.hres, h, self.g, f))
            nodes.append(self._mm(lw["up"], self.hres, h, self.u, f))
            nodes.append(self._mm(lw["down"], self.g, f, self.dn, h))
            nodes.append(self._add(self.hres, self.dn, xout, h))
            xin, xout = xout, (self.x if xout is self.x2 else self.x2)
        nodes.append(self._mm(self.w_out, xin, h, self.logits, self.out[1]))
        self.final_x = xin
        self.node_list = nodes
        self._nodes = (self.B.Node * len(nodes))(*nodes)

    def step(self):
        B, C = self.B, self.C
        if self.world == 1:
            self.s.graph_compute(self._nodes, use_cuda_graph=True)
            return
        L = B.lib()
        for n in self.node_list:
            if n.dst.op == B.OP_MUL_MAT:
                # n.src0 is this rank's row slice; dst is the full vector on every rank
                d = B.Tensor.from_buffer_copy(n.dst)
                B._ck(L.b200_mulmat_rowsplit(self.s.h, C.byref(n.src0), C.byref(n.src1), C.byref(d)))
            else:
                d = B.Tensor.from_buffer_copy(n.dst)
                B._ck(L.b200_add(self.s.h, C.byref(n.src0), C.byref(n.src1), C.byref(d)))


def time_shapes(B, s, wl, iters=3):
    """Per-shape GEMV time: each shape's launches cycle over the L distinct per-layer copies (L x matrix bytes >> L2 for
    all but wk/wv of 70B), back to back on the stream, CUDA events around the lot."""
    import ctypes as C
    out = []
    for name, M, K in wl.layer:
        rows = M // wl.world
        xin = wl.g if K == wl.f else wl.x
        dst = s.malloc(4 * rows)
        nodes = []
        for lw in wl.weights:
            n = B.Node()
            n.src0, n.src1, n.dst = lw[name].tensor(), B.f32_tensor(xin.ptr, [K, 1]), B.f32_tensor(dst.ptr, [rows, 1], op=B.OP_MUL_MAT)
            nodes.append(n)
        arr = (B.Node * len(nodes))(*nodes)
        s.graph_compute(arr, True)
        s.sync()
        e0, e1 = s.event(), s.event()
        s.record(e0)
        for _ in range(iters):
            s.graph_compute(arr, True)
        s.record(e1)
        ms = s.elapsed_ms(e0, e1)
        per = ms * 1e-3 / (iters * len(nodes))
        # the first node quantises, the others reuse it (same src1): time is GEMV-dominated
        by = gemv_bytes(rows, K)
        out.append({"shape": "%s %dx%d" % (name, rows, K), "us": round(per * 1e6, 3), "GB/s": round(by / per / 1e9, 1)})
        dst.free()
    return out

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
import rowsplit as RS          # host-side partition helpers of the row split (ggml-hexagon_b200/rowsplit.py)

Q4_0 = 2
METRIC = "tokens/s Llama Q4_0 decode (tg, MUL_MAT+ADD path, batch 1)"
def measured_traffic():
    """dram__bytes_read + dram__bytes_write per MUL_MAT launch of the decode step, parsed from the ncu --set full capture of THIS bench
    command that tools/ncu_traffic.py summarised into profiles/ (with the commit it was taken at); None when there is none"""
    p = os.path.join(ROOT, "profiles", "r02_ncu_decode_traffic.json")
    try:
        d = json.load(open(p))
        return int(d["dram_bytes_per_mulmat"]), "profiles/r02_ncu_decode_traffic.json (ncu --set full, commit %s)" % d.get("commit", "?")
    except (OSError, KeyError, ValueError):
        return None, None


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


BLK = {2: (32, 18), 8: (32, 34), 12: (256, 144), 14: (256, 210)}      # ggml type -> (elements, bytes) per block
Q8_0, Q4_K, Q6_K = 8, 12, 14


def random_weights(t, rows, K, rng):
    """random-init weights of ggml type t in the wire format: random quant bytes, small fp16 scales (|w| << 1 so that chains
    of layers stay finite).  q4_K: 6-bit scales/mins are whatever the random bytes say; d, dmin positive."""
    if t == Q4_0:
        return random_q4_0(rows, K, rng)
    bl, by = BLK[t]
    nb = K // bl
    raw = rng.integers(0, 256, size=(rows, nb, by), dtype=np.uint8)

    def put(off, lo, hi, signed):
        v = rng.uniform(lo, hi, size=(rows, nb))
        if signed:
            v = v * rng.choice([-1.0, 1.0], size=(rows, nb))
        d = v.astype(np.float16).view(np.uint16)
        raw[:, :, off] = (d & 0xFF).astype(np.uint8)
        raw[:, :, off + 1] = (d >> 8).astype(np.uint8)

    if t == Q8_0:
        put(0, 0.00002, 0.00004, True)
    elif t == Q4_K:
        put(0, 0.000004, 0.000008, False)
        put(2, 0.000004, 0.000008, False)
    elif t == Q6_K:
        put(208, 0.0000005, 0.000001, True)
    else:
        raise ValueError(t)
    return raw.reshape(rows, nb * by)


def quant_mix(quant):
    """weight type per matrix name.  q4_K_M-like: llama.cpp's Q4_K_M recipe in outline (q6_K for attn_v, ffn_down and the output
    matrix, q4_K elsewhere)"""
    if quant == "q4_0":
        return lambda name: Q4_0
    if quant == "q4_K_M":
        return lambda name: Q6_K if name in ("wv", "down", "output") else Q4_K
    raise ValueError(quant)


def random_q4_0(rows, K, rng):
    """random-init q4_0 weights in the wire format: random nibbles, fp16 scale ~U(0.0002,0.0004) with random sign
    (|w| ~ 0.0014: the per-layer map stays a contraction, so an 80-layer chain of identical matrices stays finite)"""
    nb = K // 32
    raw = rng.integers(0, 256, size=(rows, nb, 18), dtype=np.uint8)
    d = (rng.uniform(0.0002, 0.0004, size=(rows, nb)) * rng.choice([-1.0, 1.0], size=(rows, nb))).astype(np.float16).view(np.uint16)
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

    def __init__(self, B, sess, model, rank=0, world=1, seed=1234, mode="p2p", quant="q4_0", ncols=1, share=None, heap_cols=0, host_override=None):
        """ncols > 1: the same node list over ncols activation columns (prefill); share: another workload of the same model and
        quant whose HBM weights (and, under row split, symmetric heap) are used instead of making a second copy; heap_cols: room in
        the symmetric heap for a later workload of that many columns sharing it"""
        import ctypes as C
        self.mode = mode
        self.ncols, self.share = ncols, share
        wt = quant_mix(quant)
        self.quant = quant
        self.B, self.s, self.C = B, sess, C
        self.rank, self.world = rank, world
        layer, L, out, h, f = llama_shapes(model)
        self.model, self.h, self.f, self.L = model, h, f, L
        rng = np.random.default_rng(seed)            # same seed on every rank: identical full matrices, each rank keeps its rows
        self.weights = []                            # [layer][name] -> Weights
        self.alg_bytes = 0
        self.n_mulmat = 0
        host = {}
        self.host_full = {}                          # world > 1: the FULL matrices (rank 0 checks sampled rows of every rank's slice)
        for name, M, K in layer + [out]:
            assert M % world == 0, (name, M, world)
            # one random host matrix per shape; every layer gets its own HBM copy (distinct addresses, no L2 reuse).
            # Row split: every rank generates the same full matrix (same seed) and keeps ITS rows [rank * M / world, (rank + 1) * M / world)
            if share is None:
                full = host_override[name] if host_override is not None else random_weights(wt(name), M, K, rng)
                if world > 1:
                    lo, hi = RS.row_range(M, rank, world)
                    host[name] = np.ascontiguousarray(RS.slice_rows(full, rank, world))
                    assert host[name].shape[0] == hi - lo
                    if rank == 0:
                        self.host_full[name] = full
                else:
                    host[name] = full
        if share is not None:
            self.host_full = share.host_full
        if share is not None:
            self.weights, self.w_out = share.weights, share.w_out
        else:
            for _ in range(L):
                self.weights.append({name: sess.weights(wt(name), host[name], K) for name, M, K in layer})
            self.w_out = sess.weights(wt(out[0]), host[out[0]], out[2])
        self.layer, self.out = layer, out
        for name, M, K in layer:
            self.alg_bytes += L * gemv_bytes(M, K, blk=BLK[wt(name)])
            self.n_mulmat += L
        self.alg_bytes += gemv_bytes(out[1], out[2], blk=BLK[wt(out[0])])
        self.n_mulmat += 1
        self.weight_bytes_local = sum(w.nbytes for lw in self.weights for w in lw.values()) + self.w_out.nbytes
        # activations (f32, device)
        kv = dict((n, m) for n, m, _ in layer)["wk"]
        # row split: activations live in the symmetric heap (same offsets on every rank) so that the GEMVs can store
        # their slices straight into every peer's copy over NVLink
        act_floats = 8 * h + 2 * kv + 2 * f + out[1]
        if world > 1:
            self.heap = share.heap if share is not None else B.SymHeap(sess, 4 * act_floats * (ncols + heap_cols) + 2 * 64 * 256)
        else:
            self.heap = None
        mk = (lambda n: self.heap.malloc(4 * n * ncols)) if world > 1 else (lambda n: sess.malloc(4 * n * ncols))
        self.x = mk(h); self.q = mk(h); self.k = mk(kv); self.v = mk(kv); self.o = mk(h); self.hres = mk(h)
        self.g = mk(f); self.u = mk(f); self.dn = mk(h); self.x2 = mk(h); self.x3 = mk(h); self.logits = mk(out[1])
        self.x_host = np.random.default_rng(seed + 1).uniform(-1, 1, h * ncols).astype(np.float32)
        sess.h2d(self.x, self.x_host)
        self._nodes = None
        self._build()

    def free(self):
        if self.share is None:
            for lw in self.weights:
                for w in lw.values():
                    w.free()
            self.w_out.free()
        if self.heap is None:
            for b in (self.x, self.q, self.k, self.v, self.o, self.hres, self.g, self.u, self.dn, self.x2, self.x3, self.logits):
                b.free()

    def _mm(self, w, src, K, dst, M):
        B = self.B
        n = B.Node()
        n.src0 = w.tensor()
        n.src1 = B.f32_tensor(src.ptr, [K, self.ncols])
        n.dst = B.f32_tensor(dst.ptr, [M, self.ncols], op=B.OP_MUL_MAT)
        return n

    def _add(self, a, b, dst, n_el):
        B = self.B
        n = B.Node()
        ne = [n_el] if self.ncols == 1 else [n_el, self.ncols]
        n.src0, n.src1, n.dst = B.f32_tensor(a.ptr, ne), B.f32_tensor(b.ptr, ne), B.f32_tensor(dst.ptr, ne, op=B.OP_ADD)
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
            # self.x (the step's input) is never written: every step computes the same token from the same input
            xin, xout = xout, (self.x3 if xout is self.x2 else self.x2)
        nodes.append(self._mm(self.w_out, xin, h, self.logits, self.out[1]))
        self.final_x = xin
        self.node_list = nodes
        self._nodes = (self.B.Node * len(nodes))(*nodes)

    def step(self):
        B, C = self.B, self.C
        if self.world == 1:
            self.s.graph_compute(self._nodes, use_cuda_graph=True)
            return
        if self.mode == "p2p":
            self.s.graph_compute_rowsplit(self._nodes, use_cuda_graph=True)
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


def time_prefill(B, s, model, N=512, mixes=(("q8_0", lambda name: Q8_0), ("q4_0", lambda name: Q4_0)), iters=5):
    """BASELINE.json configs[2] (and [3] with a K-quant mix): the layer MUL_MATs with N = 512 activation columns through the
    tensor-core GEMM (b200_mulmat, device resident).  One HBM copy per (shape, type), timed back to back with CUDA events;
    pp tokens/s = N / (layers * sum of the layer's MUL_MAT times + output matrix) -- MUL_MAT only, as for decode."""
    import ctypes as C
    layer, L, out, h, f = llama_shapes(model)
    rng = np.random.default_rng(7)
    res = {}
    for name, wt in mixes:
        per_shape, tot_s, flops = [], 0.0, 0.0
        seen = {}
        for nm, M, K in layer + [out]:
            t = wt(nm)
            if (M, K, t) not in seen:
                raw = np.tile(random_weights(t, 128, K, rng), ((M + 127) // 128, 1))[:M]
                w = s.weights(t, raw, K)
                x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
                xb, db = s.to_device(x), s.malloc(4 * N * M)
                s0, s1, d = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(db.ptr, [M, N], op=B.OP_MUL_MAT)
                for _ in range(2):
                    B._ck(B.lib().b200_mulmat(s.h, C.byref(s0), C.byref(s1), C.byref(d)))
                s.sync()
                e0, e1 = s.event(), s.event()
                s.record(e0)
                for _ in range(iters):
                    B._ck(B.lib().b200_mulmat(s.h, C.byref(s0), C.byref(s1), C.byref(d)))
                s.record(e1)
                sec = s.elapsed_ms(e0, e1) * 1e-3 / iters
                seen[(M, K, t)] = sec
                per_shape.append({"shape": "%dx%d" % (M, K), "type": {2: "q4_0", 8: "q8_0", 12: "q4_K", 14: "q6_K"}[t],
                                  "us": round(sec * 1e6, 1), "TOPS": round(2.0 * M * N * K / sec / 1e12, 1)})
                for b in (xb, db):
                    b.free()
                w.free()
            reps = 1 if nm == out[0] else L
            tot_s += reps * seen[(M, K, t)]
            flops += reps * 2.0 * M * N * K
        res[name] = {"tokens_per_s": round(N / tot_s, 1), "ms": round(tot_s * 1e3, 3), "TOPS": round(flops / tot_s / 1e12, 1),
                     "frac_of_int8_nominal_4500": round(flops / tot_s / 4.5e15, 4), "shapes": per_shape}
    mode = os.environ.get("B200_GEMM_MODE", "dense")
    res["mode"] = mode
    res["note"] = ("N=%d columns, MUL_MAT only, per shape back to back on one matrix (operand preparation inside the timed region); " % N +
                   ("dense fp16 tcgen05.mma.cta_group::2 over the whole K (256 x 256 tiles per CTA pair) with the block scales folded into the "
                    "operands, weights from their cached fp16 shadow; a single matrix alone pays wave quantisation and split-K, the whole-model "
                    "leg ('model') groups q/k/v and gate/up into one launch (DESIGN.md section 6)" if mode != "exact" else
                    "exact per-block tcgen05.mma sums + fp32 scale-accumulate: bound by the epilogue's issue slots, not the tensor pipe "
                    "(DESIGN.md section 6)"))
    return res


def with_env(name, value, fn):
    old = os.environ.get(name)
    os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None:
            del os.environ[name]
        else:
            os.environ[name] = old


def time_prefill_model(B, s, wl, N=512, steps=5):
    """pp512 as llama-bench runs it, restricted to this path: ALL MUL_MAT + residual ADD nodes of the model over N = 512 columns
    as one b200_graph_compute, every layer with its own weights (no L2 reuse between layers or steps).  The first call is timed
    separately: it builds the fp16 weight shadows of the dense GEMM (2 bytes per weight beside the native blocks)."""
    pw = DecodeWorkload(B, s, wl.model, quant=wl.quant, ncols=N, share=wl)
    s.sync()
    t0 = time.perf_counter()
    pw.step()
    s.sync()
    first_s = time.perf_counter() - t0
    for _ in range(2):
        pw.step()
    s.sync()
    l0 = s.launch_count()
    e0, e1 = s.event(), s.event()
    s.record(e0)
    for _ in range(steps):
        pw.step()
    s.record(e1)
    sec = s.elapsed_ms(e0, e1) * 1e-3 / steps
    launches = (s.launch_count() - l0) // steps
    flops = 2.0 * N * (sum(m * k for _, m, k in wl.layer) * wl.L + wl.out[1] * wl.out[2])
    tc_peak = None
    try:
        tc_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except (OSError, KeyError, ValueError):
        pass
    pw.free()
    umma = {}                                  # tools/ubench_umma.cu on a B200 of this pool: back-to-back tcgen05.mma on shared-memory operands, no loads
    try:
        for ln in open(os.path.join(ROOT, "profiles", "r02_e_ubench_umma.jsonl")):
            d = json.loads(ln)
            umma[d["kind"]] = d["TOPS_best"]
    except (OSError, ValueError, KeyError):
        pass
    tf = flops / sec / 1e12
    vs_peaks = {"bf16_cublas_sustained_measured": round(tf / tc_peak, 4) if tc_peak else None,
                "fp16_mma_issue_peak_measured": round(tf / umma["f16"], 4) if "f16" in umma else None,
                "int8_mma_issue_peak_measured": round(tf / umma["i8"], 4) if "i8" in umma else None,
                "fp16_nominal_2250": round(tf / 2250.0, 4), "int8_nominal_4500": round(tf / 4500.0, 4),
                "peaks": {"fp16_mma_TFLOPS": umma.get("f16"), "int8_mma_TOPS": umma.get("i8"), "source": "profiles/r02_e_ubench_umma.jsonl (tools/ubench_umma.cu)"}}
    try:
        vs_peaks["bf16_cublas_burst_measured"] = round(tf / float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]), 4)
    except (OSError, KeyError, ValueError):
        pass
    return {"tokens_per_s": round(N / sec, 1), "ms": round(sec * 1e3, 3), "TFLOPS": round(flops / sec / 1e12, 1), "frac_of_peaks": vs_peaks,
            "frac_of_measured_16bit_tensor_peak": round(flops / sec / 1e12 / tc_peak, 4) if tc_peak else None, "tensor_peak_TFLOPS": tc_peak,
            "roofline": {"bound": "tensor", "achieved": round(flops / sec / 1e12, 1), "peak": tc_peak, "unit": "TFLOP/s",
                         "frac": round(flops / sec / 1e12 / tc_peak, 4) if tc_peak else None, "traffic": None,
                         "kernel": "gemm_dense2_kernel<256, q4_0, WTMA> (tcgen05.mma.cta_group::2 kind::f16; native q4_0 blocks by TMA, dequantised by converter warps)",
                         "note": "2*M*N*K of every MUL_MAT of the step / device time of the whole step (conversions, reduces, ADDs included)"},
            "kernels_per_step": int(launches), "first_call_ms": round(first_s * 1e3, 1),
            "note": "whole node list (%d MUL_MAT + %d ADD) over N=%d columns, one graph_compute per step, distinct weights per layer; "
                    "peak = MEASURED_PEAKS.json bf16_tflops_sustained (fp16 and bf16 run at the same tensor rate)" % (wl.n_mulmat, 2 * wl.L, N)}


def rowsplit_parity(B, s, wl, nsample=256, seed=99):
    """Rank 0, after a row-split step: every kind of op of the LAST layer (and the output matrix) is recomputed on this GPU alone for
    `nsample` rows drawn over the WHOLE matrix (= every rank's slice) from the full host copy, with the op's actual input vector as it
    stands in the symmetric heap, and compared with what the row-split step left in the output vector.  A permuted, stale or missing
    slice of any rank shows up here (NMSE bar 1e-9: same integers and scales, float summation order may differ)."""
    rng = np.random.default_rng(seed)
    dims = dict((n, (m, k)) for n, m, k in wl.layer)
    last_in = wl.x3 if wl.final_x is wl.x2 else wl.x2          # input of the last layer (the ping-pong partner of its output)
    if wl.L == 1:
        last_in = wl.x
    ops = [("wq", last_in, wl.q), ("wk", last_in, wl.k), ("wv", last_in, wl.v), ("wo", wl.q, wl.o), ("gate", wl.hres, wl.g),
           ("up", wl.hres, wl.u), ("down", wl.g, wl.dn), (wl.out[0], wl.final_x, wl.logits)]
    worst, detail = 0.0, {}
    for name, src, dst in ops:
        M, K = (wl.out[1], wl.out[2]) if name == wl.out[0] else dims[name]
        rows = np.sort(rng.choice(M, size=min(nsample, M), replace=False))
        w = s.weights(Q4_0, np.ascontiguousarray(wl.host_full[name][rows]), K)
        x = s.to_host(src, (1, K), np.float32)
        ref = s.mulmat(w, x).reshape(-1)
        got = s.to_host(dst, (M,), np.float32)[rows]
        e = float(((got.astype(np.float64) - ref) ** 2).sum() / max(float((ref.astype(np.float64) ** 2).sum()), 1e-30))
        detail[name] = e
        worst = max(worst, e)
        w.free()
    return {"parity_ok": bool(worst <= 1e-9), "nmse_max": worst, "rows_sampled_per_op": int(nsample), "per_op": detail,
            "what": "rank 0 recomputes sampled rows of every op of the last layer + the output matrix on one GPU from the full host matrices"}


def time_decode(B, s, wl, steps):
    for _ in range(3):
        wl.step()
    s.sync()
    e0, e1 = s.event(), s.event()
    s.record(e0)
    for _ in range(steps):
        wl.step()
    s.record(e1)
    return s.elapsed_ms(e0, e1) * 1e-3 / steps


def time_add(B, s, iters=50):
    """GGML_OP_ADD as KERNEL time: 64 ADD nodes over rotating, distinct buffers (no L2 reuse at 8 MB each) submitted as one
    captured CUDA graph, device time / 64.  ne=[4096]: launch-latency bound (report us); [4096, 512]: HBM-bound (3 x nbytes / time)"""
    out = []
    for ne in ([4096], [4096, 512]):
        n = int(np.prod(ne))
        nn = 64
        bufs = [s.malloc(4 * n) for _ in range(3 * nn)]
        nodes = []
        for i in range(nn):
            nd = B.Node()
            a, b, d = bufs[3 * i], bufs[3 * i + 1], bufs[3 * i + 2]
            nd.src0, nd.src1, nd.dst = B.f32_tensor(a.ptr, ne), B.f32_tensor(b.ptr, ne), B.f32_tensor(d.ptr, ne, op=B.OP_ADD)
            nodes.append(nd)
        arr = (B.Node * nn)(*nodes)
        for _ in range(3):
            s.graph_compute(arr, True)
        s.sync()
        e0, e1 = s.event(), s.event()
        s.record(e0)
        for _ in range(iters):
            s.graph_compute(arr, True)
        s.record(e1)
        us = s.elapsed_ms(e0, e1) * 1e3 / (iters * nn)
        out.append({"ne": ne, "us": round(us, 3), "GB/s": round(3 * 4 * n / us / 1e3, 1), "how": "64 ADD nodes per captured CUDA graph, distinct buffers"})
        for x in bufs:
            x.free()
    return out


def time_kquant(B, s, steps):
    """BASELINE.json configs[3]: Llama-3-8B shapes with a Q4_K_M-like q4_K / q6_K mix: decode step (same node list shape as the
    headline workload) and pp512"""
    wl = DecodeWorkload(B, s, "llama3-8b", quant="q4_K_M", seed=4321)
    for _ in range(3):
        wl.step()
    s.sync()
    e0, e1 = s.event(), s.event()
    s.record(e0)
    for _ in range(steps):
        wl.step()
    s.record(e1)
    step_s = s.elapsed_ms(e0, e1) * 1e-3 / steps
    peak, _ = measured_peak_hbm()
    out = {"workload": "llama3-8b-q4_K_M-like-decode (q6_K: wv, down, output; q4_K: the rest)", "tokens_per_s": round(1.0 / step_s, 2),
           "ms_per_step": round(step_s * 1e3, 4), "GB/s": round(wl.alg_bytes / step_s / 1e9, 1),
           "hbm_frac": round(wl.alg_bytes / step_s / 1e9 / peak, 4), "weight_bytes": wl.weight_bytes_local}
    out["pp512"] = time_prefill(B, s, "llama3-8b", mixes=(("q4_K_M", quant_mix("q4_K_M")),))["q4_K_M"]
    out["pp512"]["model"] = time_prefill_model(B, s, wl)
    wl.free()
    return out


def plugin_e2e(model, steps, x_host):
    """The headline e2e: the SAME node list through the drop-in seat, libggml-b200.so, called the way a ggml host calls a backend
    (the reference's public C API from its own libggml, tests/ggmlapi.py): per step ggml_backend_tensor_set(x, host buffer) ->
    ggml_backend_graph_compute (= graph_compute_async + synchronize, ggml-backend.cpp:327-335) -> ggml_backend_tensor_get(logits).
    The host library is the caller, not the thing measured; it comes from oracle/_ref because the plugin API needs a ggml host."""
    import ctypes as C
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ggmlapi as GA
    G = GA.lib()
    reg, devs = GA.load_plugin()
    be = G.ggml_backend_dev_init(devs[0], None)
    layer, L, out, h, f = llama_shapes(model)
    ctx = GA.Ctx(L * 20 + 16)
    gf = G.ggml_new_graph(ctx.ctx)
    x = ctx.tensor(GA.F32, h, 1)
    weights, cur = [], x
    dims = dict((n, (m, k)) for n, m, k in layer)

    def W(name):
        m, k = dims[name] if name != out[0] else (out[1], out[2])
        t = ctx.tensor(GA.Q4_0, k, m)
        weights.append((name, t))
        return t

    def node(t):
        G.ggml_build_forward_expand(gf, t)
        return t
    for _ in range(L):
        q = node(G.ggml_mul_mat(ctx.ctx, W("wq"), cur))
        node(G.ggml_mul_mat(ctx.ctx, W("wk"), cur))
        node(G.ggml_mul_mat(ctx.ctx, W("wv"), cur))
        o = node(G.ggml_mul_mat(ctx.ctx, W("wo"), q))
        hres = node(G.ggml_add(ctx.ctx, cur, o))
        g = node(G.ggml_mul_mat(ctx.ctx, W("gate"), hres))
        node(G.ggml_mul_mat(ctx.ctx, W("up"), hres))
        dn = node(G.ggml_mul_mat(ctx.ctx, W("down"), g))
        cur = node(G.ggml_add(ctx.ctx, hres, dn))
    logits_t = node(G.ggml_mul_mat(ctx.ctx, W(out[0]), cur))
    buf = G.ggml_backend_alloc_ctx_tensors(ctx.ctx, be)
    assert buf, "alloc_ctx_tensors failed"
    rng = np.random.default_rng(1234)
    host = {}
    for name, m, k in layer + [out]:
        host[name] = random_q4_0(m, k, rng)
    for name, t in weights:
        a = host[name]
        G.ggml_backend_tensor_set(t, a.ctypes.data_as(C.c_void_p), 0, a.nbytes)
    xh = np.ascontiguousarray(x_host[:h], dtype=np.float32)
    logits = np.empty(out[1], dtype=np.float32)

    def step():
        G.ggml_backend_tensor_set(x, xh.ctypes.data_as(C.c_void_p), 0, xh.nbytes)
        rc = G.ggml_backend_graph_compute(be, gf)
        assert rc == 0, rc
        G.ggml_backend_tensor_get(logits_t, logits.ctypes.data_as(C.c_void_p), 0, logits.nbytes)
    for _ in range(5):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    sec = (time.perf_counter() - t0) / steps
    res = {"value": round(1.0 / sec, 3), "unit": "tokens/s", "ms_per_step": round(sec * 1e3, 5), "h2d_bytes_per_step": int(xh.nbytes),
           "d2h_bytes_per_step": int(logits.nbytes), "logits_finite": bool(np.isfinite(logits).all()),
           "call": "ggml_backend_tensor_set(x) -> ggml_backend_graph_compute(libggml-b200.so, %d nodes) -> ggml_backend_tensor_get(logits); pageable host "
                   "buffers, host perf_counter" % (9 * L + 1)}
    G.ggml_backend_buffer_free(buf)
    ctx.free()
    G.ggml_backend_free(be)
    return res, logits


def llama_bench(use_plugin, n_prompt=512, n_gen=128, reps=3):
    """BASELINE.json's literal metric on its literal harness: the reference's own llama-bench, built unmodified into
    oracle/_ref by `make -f oracle/Makefile.ref llama`, on a synthetic Llama-2-7B Q4_0 GGUF (tools/make_gguf.py; there are no model files
    on the box), the ggml scheduler placing the graph.  use_plugin: libggml-b200.so through GGML_BACKEND_PATH, all layers offloaded;
    otherwise the reference's CPU backend with every host core."""
    import subprocess, tempfile
    import pyref
    d = pyref.ref_dir()
    exe = os.path.join(d or "", "llama-bench")
    if not d or not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/<variant>/llama-bench not built (make -f oracle/Makefile.ref llama)"}
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
    gguf = os.path.join(tmp, "b200_bench_l7b_%d.gguf" % os.getpid())
    try:
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), gguf, "--preset", "7b"], check=True, capture_output=True, timeout=300)
        env = dict(os.environ)
        cmd = [exe, "-m", gguf, "-p", str(n_prompt), "-n", str(n_gen), "-r", str(reps), "-t", str(os.cpu_count() or 1), "-o", "jsonl"]
        if use_plugin:
            env["GGML_BACKEND_PATH"] = os.path.join(ROOT, "ggml-hexagon_b200", "lib", "libggml-b200.so")
            cmd += ["-ngl", "99"]
        else:
            env.pop("GGML_BACKEND_PATH", None)
            cmd += ["-ngl", "0"]
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
        out = {"harness": "reference llama-bench, unmodified (oracle/_ref/%s)" % os.path.basename(d), "model": "synthetic llama 7B Q4_0 GGUF (3.56 GiB, random blocks)",
               "backend": None, "reps": reps}
        for l in r.stdout.splitlines():
            try:
                j = json.loads(l)
            except ValueError:
                continue
            out["backend"] = j.get("backends")
            key = ("pp%d" % j["n_prompt"]) if j["n_prompt"] else ("tg%d" % j["n_gen"])
            out[key] = {"tokens_per_s": round(j["avg_ts"], 2), "stddev": round(j["stddev_ts"], 2)}
        if r.returncode != 0 or not any(k.startswith(("pp", "tg")) for k in out):
            out["unavailable"] = "llama-bench rc=%d: %s" % (r.returncode, (r.stderr or "")[-300:])
        return out
    except (subprocess.SubprocessError, OSError) as e:
        return {"unavailable": "%s: %s" % (type(e).__name__, e)}
    finally:
        if os.path.exists(gguf):
            os.remove(gguf)


def cpu_baseline(model, budget_s):
    """reference ggml-cpu (oracle/_ref) on the host cores; falls back to the single-threaded C port only if _ref is absent"""
    import pyref
    layer, L, out, h, f = llama_shapes(model)
    cores = os.cpu_count() or 1
    try:
        R = pyref.Ref()
        per_tok, desc = pyref.bench_decode(R, layer, out, L, 8, Q4_0, cores, budget_s=budget_s)
        return {"value": round(1.0 / per_tok, 4), "unit": "tokens/s", "cores": cores, "kind": "reference",
                "sample": desc + "; oracle/_ref/%s ggml-cpu, llamafile on, OpenMP off" % R.variant}
    except (FileNotFoundError, OSError):
        import pyoracle as O
        rng = np.random.default_rng(0)
        M, K = 512, 4096
        w = random_q4_0(M, K, rng)
        x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
        t0 = time.perf_counter(); n = 0
        while time.perf_counter() - t0 < budget_s:
            O.mul_mat(O.Q4_0, w, x, 0); n += 1
        per_row = (time.perf_counter() - t0) / n / M / K
        tot = sum(m * k for _, m, k in layer) * L + out[1] * out[2]
        return {"value": round(1.0 / (per_row * tot), 5), "unit": "tokens/s", "cores": 1, "kind": "port",
                "sample": "oracle C port, %dx%d q4_0 GEMV x%d, extrapolated by weight elements" % (M, K, n)}


def run_reference(args, rank):
    if rank != 0:
        return
    model = "llama2-7b" if args.gpus == 1 else "llama3-70b"
    vals = []
    cb = None
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(model, budget_s=max(2.0, 40.0 / (args.warmup + args.steps)))
        if i >= args.warmup:
            vals.append(cb["value"])
    v = float(np.mean(vals))
    cb["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": None, "dtype": "q4_0 x q8_0 int8 dot, f32 accumulate", "data": "synthetic",
            "config": {"workload": model + "-q4_0-decode", "batch": 1}, "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    if args.gpus == 1:
        line["llama_bench"] = llama_bench(False, n_gen=32, reps=1)
    print(json.dumps(line))


def run_b200(args, rank, world):
    import ctypes as C
    import b200ops as B
    # libraries (NCCL's version banner) write to fd 1: keep stdout clean for the ONE JSON line
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist = None
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    s = B.Session(local)
    model = "llama2-7b" if world == 1 else "llama3-70b"
    if world > 1 and args.rowsplit == "nccl":
        import torch
        idb = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = (C.c_uint8 * 128)()
            B._ck(B.lib().b200_rowsplit_unique_id(buf))
            idb = torch.tensor(list(buf), dtype=torch.uint8)
        idb = idb.cuda()
        dist.broadcast(idb, 0)
        idh = (C.c_uint8 * 128)(*idb.cpu().tolist())
        B._ck(B.lib().b200_rowsplit_init(s.h, idh, rank, world))
    rs_prefill_cols = 512 if world > 1 and args.rowsplit == "p2p" else 0
    wl = DecodeWorkload(B, s, model, rank, world, mode=args.rowsplit, heap_cols=rs_prefill_cols)
    if world > 1:
        import torch
        mine = torch.tensor(list(wl.heap.handle), dtype=torch.uint8, device="cuda")
        allh = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine)
        wl.heap.attach(b"".join(bytes(t.cpu().tolist()) for t in allh), rank, world)
        dist.barrier()

    def barrier():
        s.sync()
        if dist is not None:
            dist.barrier()
            import torch
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        wl.step()
    barrier()
    # clocks: nvidia-smi needs ~0.1 s per sample and the timed region is K milliseconds, so the same steps keep running without a gap
    # before and after it (0.6 s each, untimed) while the sampler runs: the samples are of this workload under load, timed region included
    sampler = ClockSampler(local)
    sampler.start()

    def roll(seconds):
        if world > 1:                              # row split: every rank must run the same number of steps (they meet in every exchange)
            for _ in range(150):
                wl.step()
            s.sync()
            return
        t_end = time.perf_counter() + seconds
        while time.perf_counter() < t_end:
            for _ in range(20):
                wl.step()
            s.sync()
    roll(0.6)
    l0 = s.launch_count()
    e0, e1 = s.event(), s.event()
    s.record(e0)
    for _ in range(args.steps):
        wl.step()
    s.record(e1)
    ms = s.elapsed_ms(e0, e1)
    launches = s.launch_count() - l0
    roll(0.6)
    barrier()
    if dist is not None:
        import torch
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    step_s = ms * 1e-3 / args.steps
    # e2e: host buffers in, host buffers out, through the C-ABI, every step (pinned x -> HBM, graph, logits -> pinned)
    hx, hy = C.c_void_p(), C.c_void_p()
    B._ck(B.lib().b200_host_alloc(s.h, 4 * wl.h, C.byref(hx)))
    B._ck(B.lib().b200_host_alloc(s.h, 4 * wl.out[1], C.byref(hy)))
    C.memmove(hx, wl.x_host.ctypes.data, 4 * wl.h)
    def e2e_step():
        B._ck(B.lib().b200_memcpy_h2d_async(s.h, wl.x.ptr, hx, 4 * wl.h))
        wl.step()
        B._ck(B.lib().b200_memcpy_d2h_async(s.h, hy, wl.logits.ptr, 4 * wl.out[1]))
        s.sync()                                       # the step's result is in host memory here
    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    if dist is not None:
        import torch
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    clocks = sampler.stop()
    logits = np.ctypeslib.as_array(C.cast(hy, C.POINTER(C.c_float)), shape=(wl.out[1],)).copy()
    parity = None
    single = None
    if world > 1:
        # (rank 0 works alone here; the other ranks wait in the barrier below)
        wl.step()              # one more step on every rank: the vectors the parity check reads are this step's
        barrier()
        if rank == 0:
            parity = rowsplit_parity(B, s, wl)
            if not args.no_shapes:
                # the SAME model on ONE GPU (39 GB of q4_0 fit): the denominator of the strong-scaling efficiency of this line
                one = DecodeWorkload(B, s, model, 0, 1, host_override=wl.host_full)
                t1 = time_decode(B, s, one, max(5, args.steps // 2))
                single = {"tokens_per_s": round(1.0 / t1, 3), "ms_per_step": round(t1 * 1e3, 4),
                          "note": "llama3-70b-q4_0 decode node list on rank 0's GPU alone (same kernels, no split), measured in this run"}
                one.free()
        barrier()
    plugin = plugin_logits = None
    if world == 1 and rank == 0 and not args.no_plugin:
        try:
            plugin, plugin_logits = plugin_e2e(model, args.steps, wl.x_host)
            # same weights (same generator, same order), same kernels: the seat must reproduce the op layer bit for bit
            plugin["matches_op_layer_bitwise"] = bool(np.array_equal(plugin_logits, logits))
        except (FileNotFoundError, OSError, RuntimeError, AssertionError) as e:
            plugin = {"unavailable": "%s: %s" % (type(e).__name__, e)}
    shapes = time_shapes(B, s, wl) if world == 1 and not args.no_shapes else None
    prefill = None
    if world == 1 and not args.no_shapes and rank == 0:
        prefill = time_prefill(B, s, model)
        prefill["exact_mode"] = with_env("B200_GEMM_MODE", "exact", lambda: time_prefill(B, s, model, iters=3))
        prefill["model"] = time_prefill_model(B, s, wl)
    rs_prefill = None
    if rs_prefill_cols:
        # BASELINE.json configs[4] with N = 512 columns: every rank runs the dense GEMM on its row slices and pushes them into the
        # peers' dst over NVLink (every rank takes part: the flag exchange per op is collective)
        pw = DecodeWorkload(B, s, model, rank, world, mode=args.rowsplit, ncols=rs_prefill_cols, share=wl)
        for _ in range(2):
            pw.step()
        barrier()
        pe0, pe1 = s.event(), s.event()
        s.record(pe0)
        for _ in range(3):
            pw.step()
        s.record(pe1)
        pms = s.elapsed_ms(pe0, pe1) / 3
        barrier()
        import torch
        t = torch.tensor([pms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pms = float(t.item())
        pflops = 2.0 * rs_prefill_cols * (sum(m * k for _, m, k in wl.layer) * wl.L + wl.out[1] * wl.out[2])
        rs_prefill = {"tokens_per_s": round(rs_prefill_cols / pms * 1e3, 1), "ms": round(pms, 3), "TFLOPS_all_gpus": round(pflops / pms / 1e9, 1),
                      "note": "pp%d of the same model under row split: dense GEMM on each rank's row slices + slice push over NVLink + flag "
                              "exchange per op, one CUDA graph per rank; max over ranks" % rs_prefill_cols}
    kquant = None
    adds = time_add(B, s) if world == 1 and not args.no_shapes and rank == 0 else None
    rowsplit_n1 = None
    if world == 1 and not args.no_shapes and rank == 0:
        wl.free()
        kquant = time_kquant(B, s, args.steps)
        # the N >= 2 lines run Llama-3-70B under row split; its single-GPU number belongs to the N = 1 line as well (the driver's own
        # 1 -> N efficiency divides by the 7B headline, which is another workload)
        w70 = DecodeWorkload(B, s, "llama3-70b", seed=1234)
        t70 = time_decode(B, s, w70, max(5, args.steps // 2))
        peak70, _ = measured_peak_hbm()
        rowsplit_n1 = {"workload": "llama3-70b-q4_0-decode on one GPU (the N >= 2 workload, unsplit)", "tokens_per_s": round(1.0 / t70, 3),
                       "ms_per_step": round(t70 * 1e3, 4), "GB/s": round(w70.alg_bytes / t70 / 1e9, 1), "hbm_frac": round(w70.alg_bytes / t70 / 1e9 / peak70, 4),
                       "weight_bytes": w70.weight_bytes_local}
        w70.free()
    if rank != 0:
        s.close()
        return
    peak, peak_src = measured_peak_hbm()
    # roofline of the dominant kernel (gemv_q_kernel<q4_0>): algorithmic bytes of all MUL_MATs of one rank's step over the
    # device time of the WHOLE step (quantise + add kernels included -> conservative); per launch = / n_mulmat
    alg = wl.alg_bytes if world == 1 else sum(gemv_bytes(M // world, K) for _, M, K in wl.layer) * wl.L + gemv_bytes(wl.out[1] // world, wl.out[2])
    achieved = alg / step_s / 1e9
    line = {"metric": METRIC, "value": round(1.0 / step_s, 3), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(step_s * 1e3, 5), "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
            "dtype": "q4_0 x q8_1 int8 dp4a, f32 accumulate", "data": "synthetic (random-init q4_0 blocks, seeded)",
            "config": {"workload": model + "-q4_0-decode" + ("-rowsplit%d" % world if world > 1 else ""), "batch": 1,
                       "layers": wl.L, "mul_mat_per_step": wl.n_mulmat, "add_per_step": 2 * wl.L,
                       "weight_bytes_per_gpu": wl.weight_bytes_local, "l2": "inputs larger than L2 (weights streamed once per step)",
                       "submit": "one b200_graph_compute (CUDA graph + PDL)" if world == 1 else
                       ("one b200_graph_compute_rowsplit (CUDA graph; GEMV stores {value, seq} slots into peer HBM over NVLink, polled by a gather kernel per op; B200_RS_PROTO=flag: plain stores + flag exchange)"
                        if args.rowsplit == "p2p" else "per-node b200_mulmat_rowsplit + ncclAllGather")},
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "frac_of_nominal_8000": round(achieved / 8000.0, 4),
                         "traffic": measured_traffic()[0] if world == 1 else None, "traffic_source": measured_traffic()[1] if world == 1 else None, "peak_source": peak_src, "kernel": "b200_gemv5_kernel<q4_0,8,1> (launches >= 12 MB: q/k/v, gate/up, down, output); b200_gemv4_kernel<q4_0,1,4> (wo)",
                         "note": "algorithmic MUL_MAT bytes of the step / device time of the whole step; per launch: %d B avg" % (alg // wl.n_mulmat)},
            "gpu_launches": int(launches), "clocks": clocks, "logits_finite": bool(np.isfinite(logits).all())}
    e2e_ops = {"value": round(1.0 / e2e_s, 3), "unit": "tokens/s", "h2d_bytes_per_step": 4 * wl.h, "d2h_bytes_per_step": 4 * wl.out[1],
               "call": "op-layer C-ABI (libb200ops.so): b200_memcpy_h2d_async from pinned memory + b200_graph_compute + b200_memcpy_d2h_async + b200_synchronize",
               "timer": "host perf_counter around K steps"}
    if plugin and "value" in plugin:
        # headline e2e = the reference-facing plugin call (ggml_backend_graph_compute on libggml-b200.so); the op-layer figure explains it
        line["e2e"] = dict(plugin, through="libggml-b200.so (ggml backend plugin API)")
        line["e2e_op_layer"] = e2e_ops
        line["e2e"]["ratio_to_op_layer"] = round(plugin["value"] / e2e_ops["value"], 4)
    else:
        line["e2e"] = e2e_ops
        if plugin:
            line["e2e_plugin"] = plugin
    if parity is not None:
        line["parity_ok"] = parity["parity_ok"]
        line["parity"] = parity
    if single is not None:
        line["single_gpu_same_model"] = single
        line["strong_scaling_eff"] = round((1.0 / step_s) / (world * single["tokens_per_s"]), 4)
    if rowsplit_n1 is not None:
        line["rowsplit_n1"] = rowsplit_n1
    if shapes:
        line["shapes"] = shapes
    if prefill:
        line["pp512"] = prefill
    if rs_prefill is not None:
        line["pp512_rowsplit"] = rs_prefill
    if kquant:
        line["kquant"] = kquant
    if adds:
        line["add"] = adds
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(model, budget_s=15.0)
    if world == 1 and rank == 0 and not args.no_plugin:
        s.sync()
        line["llama_bench"] = llama_bench(True)          # (its own process and CUDA context; this one is idle meanwhile)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    s.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rowsplit", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: p2p = GEMV stores into peer memory + flag exchange (default); nccl = ncclAllGather per MUL_MAT")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-shapes", action="store_true", help="skip the per-shape GEMV table")
    ap.add_argument("--no-plugin", action="store_true", help="skip the e2e leg through libggml-b200.so (needs oracle/_ref as the ggml host)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus and not (world == 1 and args.gpus == 1):
        if world == 1 and args.gpus > 1:
            sys.exit("launch with torchrun --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    run_b200(args, rank, world)


if __name__ == "__main__":
    main()

"""GPU tests (pytest -m gpu) of the drop-in seat itself: libggml-b200.so, driven the way a ggml host drives a backend plugin.

  * the reference's own parity harness, tests/test-backend-ops.cpp (built unmodified into oracle/_ref), against the plugin loaded
    through GGML_BACKEND_PATH: MUL_MAT and ADD vs ggml-cpu at its own bars (NMSE 5e-4 / 1e-7), also with the tensor-core GEMM
    paths forced down to the small batches that harness uses (dense fp16 path and exact int8 path)
  * in THIS process, through the reference's public C API (tests/ggmlapi.py): ggml_backend_load -> device -> buffer ->
    ggml_backend_graph_compute, compared with the same graph on the reference's CPU backend: decode chains (one column),
    prefill (64 columns: dense GEMM through the plugin), weight re-upload and buffer free / re-allocation (stale caches),
    supports_op answers, events across two backends, async tensor copies
"""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import ggmlapi as GA
import pyoracle as O

pytestmark = pytest.mark.gpu
ROOT = GA.ROOT


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


def _need_ref():
    if GA.ref_dir() is None:
        pytest.skip("oracle/_ref not built")


# ------------------------------------------------------------------------------------------------ test-backend-ops
def run_tbo(op, env_extra=None, mode="test"):
    _need_ref()
    d = GA.ref_dir()
    env = dict(os.environ, GGML_BACKEND_PATH=GA.PLUGIN)
    env.update(env_extra or {})
    r = subprocess.run([os.path.join(d, "test-backend-ops"), mode, "-o", op], cwd=d, env=env, capture_output=True, text=True, timeout=900)
    out = re.sub(r"\x1b\[[0-9;]*m", "", r.stdout)
    return r.returncode, out, r.stderr


def tbo_counts(out):
    ok = [l for l in out.splitlines() if l.rstrip().endswith("OK") and "(" in l]
    fail = [l for l in out.splitlines() if "FAIL" in l and "(" in l]
    return ok, fail


def test_tbo_mul_mat_through_plugin():
    rc, out, err = run_tbo("MUL_MAT")
    assert "load_backend: loaded b200 backend" in err + out or "B200_0" in out, (out[-2000:], err[-2000:])
    ok, fail = tbo_counts(out)
    assert rc == 0 and not fail, (fail[:5], out[-1500:])
    assert len(ok) >= 138, len(ok)           # (138 cases + the two summary lines = the 140 'OK' of the round-1 logs)
    for t in ("q4_0", "q8_0", "q4_K", "q6_K", "f32"):
        n = sum(1 for l in ok if "type_a=%s," % t in l)
        assert n >= 10, (t, n)
    assert re.search(r"Backend B200_0: .*OK", out.replace("\n", " ")), out[-800:]


def test_tbo_add_through_plugin():
    rc, out, err = run_tbo("ADD")
    ok, fail = tbo_counts(out)
    assert rc == 0 and not fail, (fail[:5], out[-1500:])
    assert len(ok) >= 26, len(ok)            # (26 cases + the two summary lines)


@pytest.mark.parametrize("mode", ["dense", "exact"])
def test_tbo_mul_mat_gemm_paths(mode):
    """test-backend-ops' quantised MUL_MAT cases have n <= 16 columns (tests/test-backend-ops.cpp:4134-4197), which the GEMV kernels
    take; B200_GEMM_MIN_N=8 sends its n = 8, 9, 16 cases through the tcgen05 GEMM kernels instead (default dense fp16 contraction,
    and B200_GEMM_MODE=exact = per-block exact int8 sums), still compared with ggml-cpu by the reference's own harness"""
    env = {"B200_GEMM_MIN_N": "8"}
    if mode == "exact":
        env["B200_GEMM_MODE"] = "exact"
    rc, out, err = run_tbo("MUL_MAT", env)
    ok, fail = tbo_counts(out)
    assert rc == 0 and not fail, (fail[:5], out[-1500:])
    assert len(ok) >= 138, len(ok)


@pytest.mark.parametrize("op,min_ok", [("RMS_NORM", 2), ("SOFT_MAX", 20), ("POOL_2D", 20)])
def test_tbo_stub_ops_through_plugin(op, min_ok):
    """the ops ggml-hexagon lists without a kernel (ggml-hexagon.cpp:619, 639, 649), vs ggml-cpu by the reference's own harness"""
    rc, out, err = run_tbo(op)
    ok, fail = tbo_counts(out)
    assert rc == 0 and not fail, (fail[:5], out[-1500:])
    assert len(ok) >= min_ok, (len(ok), out[-1500:])


# ------------------------------------------------------------------------------------------------ in-process, public ggml API
@pytest.fixture(scope="module")
def plug():
    _need_ref()
    G = GA.lib()
    reg, devs = GA.load_plugin()
    assert devs, "the plugin registered no device"
    be = G.ggml_backend_dev_init(devs[0], None)
    assert be
    cpu = G.ggml_backend_init_by_type(GA.DEV_CPU, None)
    assert cpu
    yield G, reg, devs, be, cpu
    G.ggml_backend_free(be)
    G.ggml_backend_free(cpu)


def _set(G, t, arr):
    arr = np.ascontiguousarray(arr)
    assert G.ggml_nbytes(t) == arr.nbytes, (G.ggml_nbytes(t), arr.nbytes)
    G.ggml_backend_tensor_set(t, arr.ctypes.data_as(C.c_void_p), 0, arr.nbytes)


def _get(G, t, shape, dtype=np.float32):
    out = np.empty(shape, dtype=dtype)
    assert G.ggml_nbytes(t) == out.nbytes
    G.ggml_backend_tensor_get(t, out.ctypes.data_as(C.c_void_p), 0, out.nbytes)
    return out


class Chain:
    """x -> [W_q, W_k (grouped: same src1)] -> W_o(q) -> + x -> [gate, up] -> down(gate) -> + h   on one backend"""

    def __init__(self, G, backend, t, K, F, N, n_layers=2):
        self.G, self.ctx = G, GA.Ctx(32 * n_layers + 8)
        c = self.ctx
        self.x = c.tensor(GA.F32, K, N)
        self.w, cur = [], self.x
        gf = G.ggml_new_graph(c.ctx)
        for _ in range(n_layers):
            wq, wk, wo = c.tensor(t, K, K), c.tensor(t, K, K // 4), c.tensor(t, K, K)
            wg, wu, wd = c.tensor(t, K, F), c.tensor(t, K, F), c.tensor(t, F, K)
            q = G.ggml_mul_mat(c.ctx, wq, cur)
            k = G.ggml_mul_mat(c.ctx, wk, cur)
            o = G.ggml_mul_mat(c.ctx, wo, q)
            h = G.ggml_add(c.ctx, cur, o)
            g = G.ggml_mul_mat(c.ctx, wg, h)
            u = G.ggml_mul_mat(c.ctx, wu, h)
            dn = G.ggml_mul_mat(c.ctx, wd, g)
            cur = G.ggml_add(c.ctx, h, dn)
            for node in (k, u, cur):
                G.ggml_build_forward_expand(gf, node)
            self.w += [wq, wk, wo, wg, wu, wd]
            self.k, self.u = k, u
        self.out, self.gf = cur, gf
        self.buf = G.ggml_backend_alloc_ctx_tensors(c.ctx, backend)
        assert self.buf
        self.backend, self.K, self.N = backend, K, N

    def run(self, weights, x):
        G = self.G
        for t, w in zip(self.w, weights):
            _set(G, t, w)
        _set(G, self.x, x)
        assert G.ggml_backend_graph_compute(self.backend, self.gf) == 0
        return _get(G, self.out, (self.N, self.K)), _get(G, self.u, (self.N, G.ggml_nbytes(self.u) // 4 // self.N))

    def free(self):
        self.G.ggml_backend_buffer_free(self.buf)
        self.ctx.free()


def ref_quantize(t, x):
    """the reference's own weight quantiser (ggml_quantize_chunk, ggml-quants.c) -> uint8 [rows, row bytes]"""
    G = GA.lib()
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows, k = x.shape
    out = np.empty((rows, k // GA.BLOCK[t][0] * GA.BLOCK[t][1]), dtype=np.uint8)
    n = G.ggml_quantize_chunk(t, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), 0, rows, k, None)
    assert n == out.nbytes
    return out


def _weights(t, K, F, n_layers, rng, scale=0.02):
    ws = []
    for _ in range(n_layers):
        for (m, k) in ((K, K), (K // 4, K), (K, K), (F, K), (F, K), (K, F)):
            ws.append(ref_quantize(t, (rng.standard_normal((m, k)) * scale).astype(np.float32)))
    return ws


@pytest.mark.parametrize("t", [GA.Q4_0, GA.Q8_0, GA.Q4_K, GA.Q6_K])
@pytest.mark.parametrize("N", [1, 64])
def test_plugin_graph_compute_vs_ggml_cpu(plug, t, N):
    """MUL_MAT / ADD chains through ggml_backend_graph_compute of the plugin vs the reference CPU backend, same graph, same data:
    one column = the decode GEMVs (grouped launches, fused residual ADD), 64 columns = the tensor-core GEMM (grouped, split-K + ADD)"""
    G, reg, devs, be, cpu = plug
    K, F = 2048, 4096
    rng = np.random.default_rng(7 + t + N)
    ws = _weights(t, K, F, 2, rng)
    x = rng.standard_normal((N, K)).astype(np.float32)
    a, b = Chain(G, be, t, K, F, N), Chain(G, cpu, t, K, F, N)
    got, got_u = a.run(ws, x)
    ref, ref_u = b.run(ws, x)
    assert np.isfinite(got).all()
    assert nmse(got, ref) <= 5e-4, nmse(got, ref)
    assert nmse(got_u, ref_u) <= 5e-4
    # a second call replays the cached CUDA graph of the same node list: same result bit for bit
    got2, _ = a.run(ws, x)
    assert np.array_equal(got, got2)
    a.free(); b.free()


def test_plugin_weight_reupload_and_realloc(plug):
    """anything cached per weight address (CUDA graphs, converted operands) must follow ggml_backend_tensor_set
    of new weights into the same tensor, and a buffer freed and re-allocated at the same address"""
    G, reg, devs, be, cpu = plug
    K, F, N, t = 2048, 4096, 64, GA.Q4_0
    rng = np.random.default_rng(99)
    x = rng.standard_normal((N, K)).astype(np.float32)
    b = Chain(G, cpu, t, K, F, N)
    for rnd in range(3):
        a = Chain(G, be, t, K, F, N)                    # (round 2, 3: same sizes -> the allocator hands back the same addresses)
        for upl in range(2):
            ws = _weights(t, K, F, 2, rng)
            got, _ = a.run(ws, x)
            ref, _ = b.run(ws, x)
            assert nmse(got, ref) <= 5e-4, (rnd, upl, nmse(got, ref))
        a.free()
    b.free()


def test_plugin_supports_op_answers(plug):
    G, reg, devs, be, cpu = plug
    dev = devs[0]
    c = GA.Ctx(64)
    w = c.tensor(GA.Q4_0, 4096, 64)
    x = c.tensor(GA.F32, 4096, 3)
    mm = G.ggml_mul_mat(c.ctx, w, x)
    assert G.ggml_backend_dev_supports_op(dev, mm)
    assert G.ggml_backend_dev_supports_op(dev, w)                        # GGML_OP_NONE leaf (tests call it on leaves)
    xh = c.tensor(GA.F16, 4096, 3)
    wf = c.tensor(GA.F16, 4096, 64)
    assert G.ggml_backend_dev_supports_op(dev, G.ggml_mul_mat(c.ctx, wf, x))          # f16 src0 = the KV cache in attention (graph_ops.cu)
    assert not G.ggml_backend_dev_supports_op(dev, G.ggml_mul_mat(c.ctx, wf, xh))     # f16 src1: declined, never emulated
    a, b = c.tensor(GA.F32, 4096, 5), c.tensor(GA.F32, 4096, 1)
    assert G.ggml_backend_dev_supports_op(dev, G.ggml_add(c.ctx, a, b))
    assert not G.ggml_backend_dev_supports_op(dev, G.ggml_add(c.ctx, xh, xh))
    # offload_op: batched MUL_MAT is worth pulling weights over for, a single column is not
    big = G.ggml_mul_mat(c.ctx, w, c.tensor(GA.F32, 4096, 64))
    assert G.ggml_backend_dev_offload_op(dev, big) and not G.ggml_backend_dev_offload_op(dev, mm)
    p = GA.DevProps()
    G.ggml_backend_dev_get_props(dev, C.byref(p))
    assert p.type == GA.DEV_GPU and p.caps.async_ and p.caps.events and not p.caps.host_buffer
    assert p.memory_total > (100 << 30)
    buft = G.ggml_backend_dev_buffer_type(dev)
    assert G.ggml_backend_dev_supports_buft(dev, buft)
    assert G.ggml_backend_reg_get_proc_address(reg, b"ggml_backend_set_n_threads")
    c.free()


def test_plugin_events_and_async_copies(plug):
    """event_new / record / wait / synchronize and set/get_tensor_async (ggml-backend-impl.h:87-117, 170-178): producer backend
    computes y = a + a and records an event; a second backend instance (second GPU when there is one) waits for it, copies y
    device-to-device (cpy_tensor_async) and adds again; the host synchronises on a second event only"""
    G, reg, devs, be, cpu = plug
    dev2 = devs[1] if len(devs) > 1 else devs[0]
    be2 = G.ggml_backend_dev_init(dev2, None)
    n = 1 << 20
    a_host = np.random.default_rng(3).standard_normal(n).astype(np.float32)
    c1, c2 = GA.Ctx(8), GA.Ctx(8)
    a = c1.tensor(GA.F32, n)
    y = G.ggml_add(c1.ctx, a, a)
    g1 = G.ggml_new_graph(c1.ctx); G.ggml_build_forward_expand(g1, y)
    b1 = G.ggml_backend_alloc_ctx_tensors(c1.ctx, be)
    yc = c2.tensor(GA.F32, n)
    z = G.ggml_add(c2.ctx, yc, yc)
    g2 = G.ggml_new_graph(c2.ctx); G.ggml_build_forward_expand(g2, z)
    b2 = G.ggml_backend_alloc_ctx_tensors(c2.ctx, be2)
    ev1, ev2 = G.ggml_backend_event_new(devs[0]), G.ggml_backend_event_new(dev2)
    assert ev1 and ev2
    out = np.empty(n, np.float32)
    for rep in range(3):
        src = a_host * (rep + 1)
        G.ggml_backend_tensor_set_async(be, a, src.ctypes.data_as(C.c_void_p), 0, src.nbytes)
        assert G.ggml_backend_graph_compute_async(be, g1) == 0
        G.ggml_backend_event_record(ev1, be)
        G.ggml_backend_event_wait(be2, ev1)
        G.ggml_backend_tensor_copy_async(be, be2, y, yc)
        assert G.ggml_backend_graph_compute_async(be2, g2) == 0
        G.ggml_backend_tensor_get_async(be2, z, out.ctypes.data_as(C.c_void_p), 0, out.nbytes)
        G.ggml_backend_event_record(ev2, be2)
        G.ggml_backend_event_synchronize(ev2)
        assert np.array_equal(out, 4 * src), rep
    G.ggml_backend_event_free(ev1); G.ggml_backend_event_free(ev2)
    G.ggml_backend_buffer_free(b1); G.ggml_backend_buffer_free(b2)
    c1.free(); c2.free()
    G.ggml_backend_free(be2)


def test_plugin_quantised_set_get_roundtrip_and_views(plug):
    """buffer set_tensor / get_tensor with the SoA repack underneath: whole tensors, block-aligned sub-ranges, memset / clear"""
    G, reg, devs, be, cpu = plug
    c = GA.Ctx(8)
    rng = np.random.default_rng(5)
    ts = {t: c.tensor(t, 1024, 9) for t in (GA.Q4_0, GA.Q8_0, GA.Q4_K, GA.Q6_K)}
    f = c.tensor(GA.F32, 1000)
    buf = G.ggml_backend_alloc_ctx_tensors(c.ctx, be)
    for t, ten in ts.items():
        raw = rng.integers(0, 256, G.ggml_nbytes(ten), dtype=np.uint8)
        _set(G, ten, raw)
        assert np.array_equal(_get(G, ten, raw.shape, np.uint8), raw)
        bb = GA.BLOCK[t][1]
        part = rng.integers(0, 256, 3 * bb, dtype=np.uint8)
        G.ggml_backend_tensor_set(ten, part.ctypes.data_as(C.c_void_p), 5 * bb, part.nbytes)
        raw[5 * bb:8 * bb] = part
        assert np.array_equal(_get(G, ten, raw.shape, np.uint8), raw)
    G.ggml_backend_tensor_memset(f, 0x3F, 0, 4000)
    assert np.array_equal(_get(G, f, (4000,), np.uint8), np.full(4000, 0x3F, np.uint8))
    G.ggml_backend_buffer_clear(buf, 0)
    assert not _get(G, f, (1000,)).any()
    G.ggml_backend_buffer_free(buf)
    c.free()

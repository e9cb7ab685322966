"""The ops between the MUL_MATs of a Llama layer (ggml-hexagon_b200/csrc/graph_ops.cu) through the ggml backend plugin, against the
reference's CPU backend on the same graph and data:
  * the reference's own harness (test-backend-ops, built unmodified) on each op;
  * a Llama layer stack as llama.cpp builds it (src/llama-model.cpp:4191-4358, src/llama-graph.cpp:1126-1216, 1306-1420): norm * weight,
    q / k / v, RoPE, KV-cache store (f16; V transposed), K^T q, soft_max with mask, V kq, head merge, output projection, residual, SwiGLU FFN;
    one column (decode: vector kernels, fused producers that also write q8_1) and 48 columns (prefill: tile kernel, dense GEMM);
  * the fused producers (RMS_NORM*w, SILU*up, CONT, each + q8_1) must give the bits of the unfused kernels;
  * B200_TENSOR_DYNAMIC: one captured CUDA graph, two cache positions."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import ggmlapi as GA
import b200ops as B
import pyoracle as O

pytestmark = pytest.mark.gpu


def nmse(a, b):
    a = a.astype(np.float64).ravel(); b = b.astype(np.float64).ravel()
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


def _need_ref():
    if GA.ref_dir() is None:
        pytest.skip("oracle/_ref not built")


@pytest.mark.parametrize("op,min_ok", [("MUL", 20), ("SILU", 2), ("ROPE", 150), ("CPY", 30), ("CONT", 6), ("DUP", 6), ("GET_ROWS", 5)])
def test_tbo_graph_ops_through_plugin(op, min_ok):
    _need_ref()
    d = GA.ref_dir()
    env = dict(os.environ, GGML_BACKEND_PATH=GA.PLUGIN)
    r = subprocess.run([os.path.join(d, "test-backend-ops"), "test", "-o", op], cwd=d, env=env, capture_output=True, text=True, timeout=900)
    out = re.sub(r"\x1b\[[0-9;]*m", "", r.stdout)
    ok = [l for l in out.splitlines() if l.rstrip().endswith("OK") and "(" in l]
    fail = [l for l in out.splitlines() if "FAIL" in l and "(" in l]
    assert r.returncode == 0 and not fail, (fail[:5], out[-1500:])
    assert len(ok) >= min_ok, (len(ok), out[-1500:])


def test_tbo_mul_mat_f16_cases_run():
    """f16 src0 (the KV cache) is admitted now: the harness' f16 x f32 cases run on the device instead of printing 'not supported'"""
    _need_ref()
    d = GA.ref_dir()
    env = dict(os.environ, GGML_BACKEND_PATH=GA.PLUGIN)
    r = subprocess.run([os.path.join(d, "test-backend-ops"), "test", "-o", "MUL_MAT"], cwd=d, env=env, capture_output=True, text=True, timeout=900)
    out = re.sub(r"\x1b\[[0-9;]*m", "", r.stdout)
    ok16 = [l for l in out.splitlines() if "type_a=f16,type_b=f32" in l and l.rstrip().endswith("OK")]
    assert r.returncode == 0 and "FAIL" not in out, out[-1500:]
    assert len(ok16) >= 40, len(ok16)


@pytest.fixture(scope="module")
def plug():
    _need_ref()
    G = GA.lib()
    reg, devs = GA.load_plugin()
    be = G.ggml_backend_dev_init(devs[0], None)
    cpu = G.ggml_backend_init_by_type(GA.DEV_CPU, None)
    assert be and cpu
    yield G, be, cpu
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


I32 = 26


class LlamaStack:
    """n_layers Llama layers over N tokens at cache position `head`, attending to n_kv cache rows, on one backend"""

    def __init__(self, G, backend, t, n_embd, n_head, n_ff, N, head, n_kv, n_ctx, n_layers=2):
        self.G, self.ctx = G, GA.Ctx(80 * n_layers + 16)
        c, x = self.ctx, self.ctx.ctx
        hd = n_embd // n_head
        self.inp = c.tensor(GA.F32, n_embd, N)
        self.pos = c.tensor(I32, N)
        self.mask = c.tensor(GA.F32, n_kv, (N + 31) // 32 * 32)
        self.w, self.nw, self.kv = [], [], []
        gf = G.ggml_new_graph(x)
        cur = self.inp
        for _ in range(n_layers):
            an, fn = c.tensor(GA.F32, n_embd), c.tensor(GA.F32, n_embd)
            wq, wk, wv, wo = (c.tensor(t, n_embd, n_embd) for _ in range(4))
            wg, wu, wd = c.tensor(t, n_embd, n_ff), c.tensor(t, n_embd, n_ff), c.tensor(t, n_ff, n_embd)
            kc, vc = c.tensor(GA.F16, n_embd * n_ctx), c.tensor(GA.F16, n_embd * n_ctx)
            self.w += [wq, wk, wv, wo, wg, wu, wd]; self.nw += [an, fn]; self.kv += [kc, vc]
            inp_sa = cur
            cur = G.ggml_mul(x, G.ggml_rms_norm(x, cur, 1e-5), an)
            q = G.ggml_reshape_3d(x, G.ggml_mul_mat(x, wq, cur), hd, n_head, N)
            k = G.ggml_reshape_3d(x, G.ggml_mul_mat(x, wk, cur), hd, n_head, N)
            v = G.ggml_reshape_3d(x, G.ggml_mul_mat(x, wv, cur), hd, n_head, N)
            rope = lambda a: G.ggml_rope_ext(x, a, self.pos, None, hd, 0, 4096, 10000.0, 1.0, 0.0, 1.0, 32.0, 1.0)
            q, k = rope(q), rope(k)
            for node in (q, k, v):
                G.ggml_build_forward_expand(gf, node)
            # KV-cache store (llama-graph.cpp:1333-1364): K rows appended, V transposed
            k_view = G.ggml_view_1d(x, kc, N * n_embd, 2 * n_embd * head)
            G.ggml_build_forward_expand(gf, G.ggml_cpy(x, k, k_view))
            v2 = G.ggml_transpose(x, G.ggml_reshape_2d(x, v, n_embd, N))
            v_view = G.ggml_view_2d(x, vc, N, n_embd, 2 * n_ctx, 2 * head)
            G.ggml_build_forward_expand(gf, G.ggml_cpy(x, v2, v_view))
            qp = G.ggml_permute(x, q, 0, 2, 1, 3)
            kk = G.ggml_view_3d(x, kc, hd, n_kv, n_head, 2 * n_embd, 2 * hd, 0)
            vv = G.ggml_view_3d(x, vc, n_kv, hd, n_head, 2 * n_ctx, 2 * n_ctx * hd, 0)
            kq = G.ggml_soft_max_ext(x, G.ggml_mul_mat(x, kk, qp), self.mask, 1.0 / np.sqrt(hd), 0.0)
            kqv = G.ggml_permute(x, G.ggml_mul_mat(x, vv, kq), 0, 2, 1, 3)
            att = G.ggml_cont_2d(x, kqv, n_embd, N)
            ffn_inp = G.ggml_add(x, G.ggml_mul_mat(x, wo, att), inp_sa)
            cur = G.ggml_mul(x, G.ggml_rms_norm(x, ffn_inp, 1e-5), fn)
            up = G.ggml_mul_mat(x, wu, cur)
            gate = G.ggml_silu(x, G.ggml_mul_mat(x, wg, cur))
            cur = G.ggml_add(x, G.ggml_mul_mat(x, wd, G.ggml_mul(x, gate, up)), ffn_inp)
            G.ggml_build_forward_expand(gf, cur)
        self.out, self.gf = cur, gf
        self.buf = G.ggml_backend_alloc_ctx_tensors(x, backend)
        assert self.buf
        G.ggml_backend_buffer_clear(self.buf, 0)
        self.backend, self.N, self.n_embd, self.n_ctx = backend, N, n_embd, n_ctx

    def load(self, weights, norm_w, kv):
        for t, w in zip(self.w, weights):
            _set(self.G, t, w)
        for t, w in zip(self.nw, norm_w):
            _set(self.G, t, w)
        for t, w in zip(self.kv, kv):
            _set(self.G, t, w)

    def run(self, x, pos, mask):
        G = self.G
        _set(G, self.inp, x); _set(G, self.pos, pos); _set(G, self.mask, mask)
        assert G.ggml_backend_graph_compute(self.backend, self.gf) == 0
        return _get(G, self.out, (self.N, self.n_embd)), [_get(G, t, (self.n_embd * self.n_ctx,), np.float16) for t in self.kv]

    def free(self):
        self.G.ggml_backend_buffer_free(self.buf)
        self.ctx.free()


def ref_quantize(t, x):
    G = GA.lib()
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows, k = x.shape
    out = np.empty((rows, k // GA.BLOCK[t][0] * GA.BLOCK[t][1]), dtype=np.uint8)
    assert G.ggml_quantize_chunk(t, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), 0, rows, k, None) == out.nbytes
    return out


def _stack_data(t, n_embd, n_ff, n_ctx, head, n_layers, rng):
    ws, nws, kvs = [], [], []
    for _ in range(n_layers):
        for (m, k) in ((n_embd, n_embd),) * 4 + ((n_ff, n_embd),) * 2 + ((n_embd, n_ff),):
            ws.append(ref_quantize(t, (rng.standard_normal((m, k)) / np.sqrt(k)).astype(np.float32)))
        nws += [rng.uniform(0.5, 1.5, n_embd).astype(np.float32) for _ in range(2)]
        for _ in range(2):          # a cache that already holds `head` positions (rows beyond are zero, as after buffer_clear)
            kvs.append((rng.standard_normal(n_embd * n_ctx) * 0.5).astype(np.float16))
    return ws, nws, kvs


def _mask(N, head, n_kv):
    m = np.full(((N + 31) // 32 * 32, n_kv), -np.inf, np.float32)
    for i in range(N):
        m[i, :head + i + 1] = 0.0
    return m


@pytest.mark.parametrize("t", [GA.Q4_0, GA.Q6_K])
@pytest.mark.parametrize("N,head,n_kv", [(1, 37, 64), (48, 0, 64), (5, 20, 32), (64, 0, 64)])
def test_llama_stack_vs_ggml_cpu(plug, monkeypatch, t, N, head, n_kv):
    G, be, cpu = plug
    n_embd, n_head, n_ff, n_ctx, L = 1024, 8, 2816, 128, 2
    rng = np.random.default_rng(3 + t + N)
    ws, nws, kvs = _stack_data(t, n_embd, n_ff, n_ctx, head, L, rng)
    x = rng.standard_normal((N, n_embd)).astype(np.float32)
    pos = np.arange(head, head + N, dtype=np.int32)
    mask = _mask(N, head, n_kv)
    a = LlamaStack(G, be, t, n_embd, n_head, n_ff, N, head, n_kv, n_ctx, L)
    b = LlamaStack(G, cpu, t, n_embd, n_head, n_ff, N, head, n_kv, n_ctx, L)
    a.load(ws, nws, kvs); b.load(ws, nws, kvs)
    got, got_kv = a.run(x, pos, mask)
    ref, ref_kv = b.run(x, pos, mask)
    assert np.isfinite(got).all()
    assert nmse(got, ref) <= 5e-4, nmse(got, ref)
    # the cache of the FIRST layer was written from bit-identical inputs up to the q/k/v MUL_MAT tolerance; compare loosely, and exactly where untouched
    for g, r in zip(got_kv[:2], ref_kv[:2]):
        assert nmse(g.astype(np.float32), r.astype(np.float32)) <= 5e-4
    k0 = got_kv[0].reshape(n_ctx, n_embd)
    assert np.array_equal(k0[head + N:], kvs[0].reshape(n_ctx, n_embd)[head + N:])          # rows past the store are untouched
    # same graph again: the second sighting is captured as a CUDA graph and replayed -> same bits
    a.load(ws, nws, kvs)
    got2, _ = a.run(x, pos, mask)
    a.load(ws, nws, kvs)
    got3, _ = a.run(x, pos, mask)
    assert np.array_equal(got, got2) and np.array_equal(got, got3)
    # fused producers (RMS_NORM*w, SILU*up, CONT; + q8_1 at one column) = the unfused kernels bit for bit
    monkeypatch.setenv("B200_GRAPH_FUSE", "0")
    c = LlamaStack(G, be, t, n_embd, n_head, n_ff, N, head, n_kv, n_ctx, L)
    c.load(ws, nws, kvs)
    plain, _ = c.run(x, pos, mask)
    assert np.array_equal(got, plain)
    a.free(); b.free(); c.free()


def test_dynamic_destination_one_graph_two_positions():
    """B200_TENSOR_DYNAMIC: the CPY destination is read from a device table at launch time, so the CUDA graph captured for one position of
    the cache stores correctly at another (b200_graph_compute_keyed, same key)"""
    K = 1024
    rng = np.random.default_rng(1)
    with B.Session(0) as s:
        x = rng.standard_normal(K).astype(np.float32)
        xb, cache, y = s.malloc(4 * K), s.malloc(2 * K * 8), s.malloc(4 * K)
        s.h2d(xb, x)
        s.h2d(cache, np.zeros(K * 8, np.float16))

        def nodes_at(row):
            cp = B.Node()
            cp.src0 = B.f32_tensor(xb.ptr, [K])
            d = B.f32_tensor(cache.ptr + 2 * K * row, [K], nb=[2, 2 * K, 2 * K, 2 * K], op=B.OP_CPY)
            d.type = B.F16; d.flags = B.TENSOR_DYNAMIC; d.data_len = 2 * K
            cp.src1 = d; cp.dst = d
            ad = B.Node()
            ad.src0, ad.src1, ad.dst = B.f32_tensor(xb.ptr, [K]), B.f32_tensor(xb.ptr, [K]), B.f32_tensor(y.ptr, [K], op=B.OP_ADD)
            return [cp, ad]
        for row in (0, 3, 3, 5, 1):                 # first call captures; the others replay the same graph with another destination
            s.graph_compute_keyed(nodes_at(row), key=77, use_cuda_graph=True)
        s.sync()
        got = s.to_host(cache, (8, K), np.float16)
        want = x.astype(np.float16)
        for row in range(8):
            assert np.array_equal(got[row], want if row in (0, 1, 3, 5) else np.zeros(K, np.float16)), row
        assert np.array_equal(s.to_host(y, (K,), np.float32), x + x)


@pytest.mark.parametrize("M,N,K,nh,nhk", [(512, 512, 128, 32, 8), (128, 512, 512, 32, 32), (100, 70, 50, 3, 1), (64, 17, 96, 2, 2), (40, 5, 128, 4, 2)])
def test_mul_mat_f16_shapes_vs_numpy(M, N, K, nh, nhk):
    """the f16 x f32 MUL_MAT kernels on their own (node list through the C-ABI): 128-wide tiles (>= 2 CTAs per SM), 64-wide tiles, edge tiles
    and unaligned rows (scalar operand loads), the one-warp-per-row kernels at few columns; grouped-query broadcast (nh / nhk query heads per
    cache head).  Reference: f32 matmul of the f16 weights with src1 rounded to f16, as ggml-cpu computes an f16 src0."""
    rng = np.random.default_rng(M + N + K)
    w = (rng.standard_normal((nhk, M, K)) * 0.5).astype(np.float16)
    x = rng.standard_normal((nh, N, K)).astype(np.float32)
    want = np.einsum("hmk,hnk->hnm", np.repeat(w, nh // nhk, axis=0).astype(np.float32), x.astype(np.float16).astype(np.float32), dtype=np.float32)
    with B.Session(0) as s:
        wb, xb, yb = s.to_device(w), s.to_device(x), s.malloc(4 * nh * N * M)
        nd = B.Node()
        a = B.f32_tensor(wb.ptr, [K, M, nhk], nb=[2, 2 * K, 2 * K * M, 2 * K * M * nhk]); a.type = B.F16; a.data_len = w.nbytes
        nd.src0 = a
        nd.src1 = B.f32_tensor(xb.ptr, [K, N, nh])
        nd.dst = B.f32_tensor(yb.ptr, [M, N, nh], op=B.OP_MUL_MAT)
        s.graph_compute([nd], use_cuda_graph=False)
        got = s.to_host(yb, (nh, N, M), np.float32)
    assert nmse(got, want) <= 1e-6, nmse(got, want)

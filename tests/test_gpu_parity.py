"""GPU parity tests (pytest -m gpu): the CUDA path, called through the C-ABI (libb200ops.so), against the oracle.

Bars (BASELINE.json north_star):
  * quantize_q8_1 output bytes == quantize_row_q8_1_ref                       (bit-exact)
  * per-group integer dot products == oracle_block_dots                        (bit-exact)
  * weight upload/download round trip                                          (bit-exact)
  * MUL_MAT f32 output vs ggml-cpu semantics (oracle mode 0): NMSE <= 5e-4     (tests/test-backend-ops.cpp:1982-1984)
    and vs the q8_1 algebra (oracle mode 1): NMSE <= 1e-9 (only float summation order differs)
  * ADD == a + b                                                               (bit-exact, IEEE f32)
"""
import os
import numpy as np
import pytest

import pyoracle as O
import b200ops as B

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_scalar.npz"))
WTYPES = [(O.Q4_0, "q4_0"), (O.Q8_0, "q8_0"), (O.Q4_K, "q4_K"), (O.Q6_K, "q6_K")]
MUL_MAT_NMSE = 5e-4


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.fixture(scope="module")
def sess():
    s = B.Session(0)      # raises if the extension or the GPU is missing: no fallback
    yield s
    s.close()


# ------------------------------------------------------------------ quantize_q8_1
def test_quantize_q8_1_golden_bit_exact(sess):
    assert np.array_equal(sess.quantize_q8_1(G["act_x"]), G["act_q8_1"])


@pytest.mark.parametrize("rows,k,scale", [(1, 32, 1.0), (3, 4096, 1.0), (2, 11008, 30.0), (5, 14336, 1e-5), (512, 4096, 1.0)])
def test_quantize_q8_1_bit_exact(sess, rows, k, scale):
    rng = np.random.default_rng(rows * 7 + k)
    x = (rng.standard_normal((rows, k)) * scale).astype(np.float32)
    x[0, :32] = 0
    if k >= 64:
        x[0, 32:64] = 0.5 * scale          # .5 ties after scaling
    assert np.array_equal(sess.quantize_q8_1(x), O.quantize(O.Q8_1, x))


def test_quantize_q8_1_empty(sess):
    assert sess.quantize_q8_1(np.zeros((0, 64), np.float32)).shape == (0, 72)


# ------------------------------------------------------------------ weights: HBM layout round trip
@pytest.mark.parametrize("t,name", WTYPES)
def test_weight_upload_download_roundtrip(sess, t, name):
    rng = np.random.default_rng(5)
    raw = rng.integers(0, 256, size=(7, O.row_size(t, 1024)), dtype=np.uint8)
    w = sess.weights(t, raw, 1024)
    assert np.array_equal(w.download(), raw)
    # partial range
    import ctypes as C
    bl, by = B.BLOCK[t]
    part = np.empty(3 * by, dtype=np.uint8)
    assert B.lib().b200_weights_download(sess.h, t, w.buf.ptr, w.blocks, 5, 3, part.ctypes.data_as(C.c_void_p)) == 0
    assert np.array_equal(part, raw.reshape(-1)[5 * by:8 * by])
    w.free()


# ------------------------------------------------------------------ exact integers
@pytest.mark.parametrize("t,name", WTYPES)
@pytest.mark.parametrize("K", [256, 4096])
def test_block_dots_bit_exact(sess, t, name, K):
    rng = np.random.default_rng(K + t)
    raw = rng.integers(0, 256, size=(3, O.row_size(t, K)), dtype=np.uint8)    # every bit pattern, incl. extreme nibbles
    a = rng.integers(-127, 128, size=K, dtype=np.int8)
    a[:32] = 127
    a[32:64] = -127
    w = sess.weights(t, raw, K)
    for r in range(3):
        assert np.array_equal(sess.block_dots(w, r, a), O.block_dots(t, K, raw[r], a))
    w.free()


# ------------------------------------------------------------------ MUL_MAT
@pytest.mark.parametrize("t,name", WTYPES)
def test_mulmat_golden(sess, t, name):
    w = sess.weights(t, G["w_" + name], 512)
    got = sess.mulmat(w, G["mm_x"])
    assert nmse(got, G["mm_" + name]) <= MUL_MAT_NMSE            # vs ggml-cpu (reference run, scalar build)
    assert nmse(got, O.mul_mat(t, G["w_" + name], G["mm_x"], 1)) <= 1e-9
    w.free()


@pytest.mark.parametrize("t,name", WTYPES)
@pytest.mark.parametrize("M,K,N", [(1, 256, 1), (16, 256, 1), (16, 256, 9), (33, 1024, 2), (130, 4096, 1),
                                   (64, 11008 if True else 0, 3), (40, 2048, 8), (24, 512, 17)])
def test_mulmat_shapes_vs_oracle(sess, t, name, M, K, N):
    if K % B.BLOCK[t][0]:
        pytest.skip("K not a multiple of the block size")
    rng = np.random.default_rng(M * 31 + K + N + t)
    if t in (O.Q4_0, O.Q8_0):
        raw = O.quantize(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    else:
        raw = O.random_blocks(t, M, K, rng)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    w = sess.weights(t, raw, K)
    got = sess.mulmat(w, x)
    assert got.shape == (N, M)
    assert nmse(got, O.mul_mat(t, raw, x, 0)) <= MUL_MAT_NMSE
    assert nmse(got, O.mul_mat(t, raw, x, 1)) <= 1e-9
    w.free()


def test_mulmat_q4_0_4096_config0(sess):
    # BASELINE.json configs[0]: MUL_MAT q4_0 x f32 4096x4096, n=1
    rng = np.random.default_rng(4096)
    raw = O.quantize(O.Q4_0, rng.uniform(-1, 1, (4096, 4096)).astype(np.float32))
    x = rng.uniform(-1, 1, (1, 4096)).astype(np.float32)
    w = sess.weights(O.Q4_0, raw, 4096)
    got = sess.mulmat(w, x)
    assert nmse(got, O.mul_mat(O.Q4_0, raw, x, 0)) <= MUL_MAT_NMSE
    w.free()


def test_mulmat_f32_weights(sess):
    rng = np.random.default_rng(9)
    wf = rng.uniform(-1, 1, (37, 1000)).astype(np.float32)
    x = rng.uniform(-1, 1, (3, 1000)).astype(np.float32)
    w = sess.weights(B.F32, wf, 1000)
    got = sess.mulmat(w, x)
    assert nmse(got, O.mul_mat(O.F32, wf, x, 0)) <= 1e-10
    w.free()


@pytest.mark.parametrize("t,name", WTYPES)
def test_mulmat_host_pointer_call(sess, t, name):
    # FastRPC-style call: all three tensors in host memory (B200_TENSOR_HOST)
    got = sess.mulmat_host(t, G["w_" + name], G["mm_x"], 512)
    assert nmse(got, G["mm_" + name]) <= MUL_MAT_NMSE


def test_mulmat_linearity_full_size(sess):
    # size-independent property at a BASELINE shape (11008 x 4096 q4_0): W(a x) = a W x exactly for a = 2^k
    # (a power-of-two scaling of the activations changes only the fp16 scale exponent), and row order independence
    rng = np.random.default_rng(77)
    M, K = 11008, 4096
    raw = O.random_blocks(O.Q4_0, M, K, rng)
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    w = sess.weights(O.Q4_0, raw, K)
    y1, y2 = sess.mulmat(w, x), sess.mulmat(w, 4.0 * x)
    assert np.array_equal(4.0 * y1, y2)
    rows = rng.choice(M, 64, replace=False)
    ref = O.mul_mat(O.Q4_0, raw[rows], x, 1)
    assert nmse(y1[:, rows], ref) <= 1e-9
    w.free()


def test_mulmat_rejects_bad_arguments(sess):
    import ctypes as C
    raw = O.random_blocks(O.Q4_0, 4, 256, np.random.default_rng(0))
    w = sess.weights(O.Q4_0, raw, 256)
    xb = sess.malloc(4 * 512)
    db = sess.malloc(4 * 4)
    s0, s1, d = w.tensor(), B.f32_tensor(xb.ptr, [512, 1]), B.f32_tensor(db.ptr, [4, 1], op=B.OP_MUL_MAT)
    assert B.lib().b200_mulmat(sess.h, C.byref(s0), C.byref(s1), C.byref(d)) == -2      # K mismatch
    assert b"K mismatch" in B.lib().b200_last_error()
    w.free()


# ------------------------------------------------------------------ ADD
def test_add_golden(sess):
    assert np.array_equal(sess.add(G["add_a"], G["add_b"]), G["add_out"])


@pytest.mark.parametrize("sa,sb", [((4096,), (4096,)), ((512, 4096), (512, 4096)), ((512, 4096), (1, 4096)),
                                   ((3, 5, 7), (3, 5, 7)), ((2, 3, 4, 10), (1, 3, 1, 10)), ((6, 10), (2, 5)),
                                   ((1,), (1,)), ((4, 16), (4, 1))])
def test_add_shapes(sess, sa, sb):
    rng = np.random.default_rng(len(sa) * 100 + sa[-1])
    a = rng.standard_normal(sa).astype(np.float32)
    b = rng.standard_normal(sb).astype(np.float32)
    assert np.array_equal(sess.add(a, b), O.add_f32(a, b))


def test_add_special_values(sess):
    a = np.array([np.inf, -np.inf, np.nan, 0.0, -0.0, 1e38, 1e-45, 3.0], dtype=np.float32)
    b = np.array([1.0, np.inf, 1.0, -0.0, 0.0, 3e38, 1e-45, -3.0], dtype=np.float32)
    got = sess.add(a, b)
    ref = a + b
    assert np.array_equal(got.view(np.uint32)[~np.isnan(ref)], ref.view(np.uint32)[~np.isnan(ref)])
    assert np.isnan(got[2]) and np.isnan(got[1])


# ------------------------------------------------------------------ graph_compute (CUDA graph replay == direct launches)
def test_graph_compute_matches_direct(sess):
    import ctypes as C
    rng = np.random.default_rng(21)
    K, M = 2048, 2048
    ws = [sess.weights(t, O.random_blocks(t, M, K, rng) if t in (O.Q4_K, O.Q6_K) else
                       O.quantize(t, rng.uniform(-1, 1, (M, K)).astype(np.float32)), K) for t, _ in WTYPES]
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    xb = sess.to_device(x)
    bufs = [sess.malloc(4 * M) for _ in range(6)]
    nodes = []
    cur = xb
    for i, w in enumerate(ws):                      # chain: y_i = W_i y_{i-1}; then residual add
        n = B.Node()
        n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(cur.ptr, [K, 1]), B.f32_tensor(bufs[i].ptr, [M, 1], op=B.OP_MUL_MAT)
        nodes.append(n)
        cur = bufs[i]
    n = B.Node()
    n.src0, n.src1, n.dst = B.f32_tensor(bufs[3].ptr, [M]), B.f32_tensor(xb.ptr, [M]), B.f32_tensor(bufs[4].ptr, [M], op=B.OP_ADD)
    nodes.append(n)
    sess.graph_compute(nodes, use_cuda_graph=False)
    direct = sess.to_host(bufs[4], (M,), np.float32)
    for _ in range(3):                              # capture, then two replays: same kernels, so bit-equal
        sess.h2d(bufs[4], np.zeros(M, np.float32))
        sess.graph_compute(nodes, use_cuda_graph=2)
        assert np.array_equal(sess.to_host(bufs[4], (M,), np.float32), direct)
    # oracle chain (mode 1) for the value itself
    y = x
    for (t, _), w in zip(WTYPES, ws):
        y = O.mul_mat(t, w.download(), y, 1)
    assert nmse(direct, (y + x).reshape(-1)) <= 1e-6
    assert sess.launch_count() > 0


def test_async_copies_are_stream_ordered():
    """b200_memcpy_h2d_async -> op -> b200_memcpy_d2h_async -> b200_synchronize: the seat of set/get_tensor_async"""
    import ctypes as C
    with B.Session(0) as s:
        n = 1 << 16
        hx, hy = C.c_void_p(), C.c_void_p()
        B._ck(B.lib().b200_host_alloc(s.h, 4 * n, C.byref(hx)))
        B._ck(B.lib().b200_host_alloc(s.h, 4 * n, C.byref(hy)))
        a = np.random.default_rng(0).standard_normal(n).astype(np.float32)
        C.memmove(hx, a.ctypes.data, 4 * n)
        da, dd = s.malloc(4 * n), s.malloc(4 * n)
        B._ck(B.lib().b200_memcpy_h2d_async(s.h, da.ptr, hx, 4 * n))
        t0, t1, td = B.f32_tensor(da.ptr, [n]), B.f32_tensor(da.ptr, [n]), B.f32_tensor(dd.ptr, [n], op=B.OP_ADD)
        B._ck(B.lib().b200_add(s.h, C.byref(t0), C.byref(t1), C.byref(td)))
        B._ck(B.lib().b200_memcpy_d2h_async(s.h, hy, dd.ptr, 4 * n))
        s.sync()
        got = np.ctypeslib.as_array(C.cast(hy, C.POINTER(C.c_float)), shape=(n,)).copy()
        assert np.array_equal(got, a + a)
        B._ck(B.lib().b200_host_free(s.h, hx)); B._ck(B.lib().b200_host_free(s.h, hy))


@pytest.mark.parametrize("rows", [4, 8, 16])
@pytest.mark.parametrize("t", [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K])
@pytest.mark.parametrize("M,K", [(301, 4096), (64, 2048), (130, 14336), (40, 28672)])
def test_gemv5_tma_tiles(monkeypatch, t, rows, M, K):
    """the TMA-staged GEMV (gemv5: one bulk copy per SoA field into shared memory) is normally used from 12 MB per launch up;
    forced here for small matrices, ragged last tiles (M % rows != 0), two slices per warp (K = 28672) and every tile height.
    Same integer dots and scales as gemv4 -> same bar against the oracle."""
    monkeypatch.setenv("B200_GEMV5_MIN_MB", "0")
    monkeypatch.setenv("B200_GEMV6", "0")          # (the kernels that quantise in-kernel would take every one-column MUL_MAT)
    monkeypatch.setenv("B200_GEMV5_ROWS", str(rows))
    rng = np.random.default_rng(M + K + t + rows)
    raw = O.random_blocks(t, M, K, rng)
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    with B.Session(0) as s:
        w = s.weights(t, raw, K)
        got = s.mulmat(w, x)
        ref1 = O.mul_mat(t, raw, x, 1)
        assert nmse(got, ref1) <= 1e-9, nmse(got, ref1)
        assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
        w.free()


@pytest.mark.parametrize("t", [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K])
@pytest.mark.parametrize("M,K,N", [(200, 2048, 2), (301, 4096, 5), (64, 4096, 13), (130, 14336, 8), (96, 28672, 31)])
def test_small_batch_columns(monkeypatch, t, M, K, N):
    """2 .. 31 activation columns: one launch over shared-memory weight tiles (gemv5n), all columns from one copy of the tile;
    must equal the single-column results column by column (same integer dots and scales; float sums in the same order)."""
    monkeypatch.setenv("B200_GEMV5N", "2")        # also for 2 and 3 columns (default: from 4 up)
    rng = np.random.default_rng(M + K + N + t)
    raw = O.random_blocks(t, M, K, rng)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    with B.Session(0) as s:
        w = s.weights(t, raw, K)
        got = s.mulmat(w, x)
        one = np.concatenate([s.mulmat(w, x[i:i + 1]) for i in range(N)])
        assert got.shape == (N, M)
        assert nmse(got, O.mul_mat(t, raw, x, 1)) <= 1e-9
        assert nmse(got, one) <= 1e-12
        w.free()


# ------------------------------------------------------------------ quantise on exit (decode chains without quantise launches)
def _chain_nodes(ws, x, bufs, K, Ms):
    """wq / wk / wv over x (group) -> wo(q) + x (ADD fused into the GEMV) -> gate / up over the sum (group) -> down(gate) + sum -> out(sum2)"""
    def mm(w, src, dst, M):
        n = B.Node()
        n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(src.ptr, [w.K, 1]), B.f32_tensor(dst.ptr, [M, 1], op=B.OP_MUL_MAT)
        return n

    def add(a, b, d, M):
        n = B.Node()
        n.src0, n.src1, n.dst = B.f32_tensor(a.ptr, [M]), B.f32_tensor(b.ptr, [M]), B.f32_tensor(d.ptr, [M], op=B.OP_ADD)
        return n
    q, k, v, o, h, g, u, dn, h2, out = bufs
    return [mm(ws[0], x, q, Ms[0]), mm(ws[1], x, k, Ms[1]), mm(ws[2], x, v, Ms[2]), mm(ws[3], q, o, Ms[3]), add(x, o, h, K),
            mm(ws[4], h, g, Ms[4]), mm(ws[5], h, u, Ms[5]), mm(ws[6], g, dn, Ms[6]), add(h, dn, h2, K), mm(ws[7], h2, out, Ms[7])]


@pytest.mark.parametrize("K,F", [(2048, 5504), (4096, 11008), (1024, 1056)])
@pytest.mark.parametrize("t", [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K])
def test_quantise_on_exit_chain(monkeypatch, t, K, F):
    """a decoder-layer-shaped node list: every MUL_MAT but the first finds its q8_1 activations written by the GEMV in front of it
    (gemv_quant_on_exit): same bits as with the quantise launches (B200_QX=0), fewer launches, and on the oracle; gemv4 and gemv5
    tiles, groups of 3 and 2, fused ADDs, CUDA graph replay"""
    if t in (O.Q4_K, O.Q6_K) and (K % 256 or F % 256):
        pytest.skip("K-quants want multiples of 256")
    monkeypatch.setenv("B200_GEMV6", "0")
    rng = np.random.default_rng(77 + t + K)
    Ms = [K, K // 4, K // 4, K, F, F, K, 1000]
    Ks = [K, K, K, K, K, K, F, K]
    raws = [O.random_blocks(t, m, k, rng, 0.05) for m, k in zip(Ms, Ks)]
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    with B.Session(0) as s:
        ws = [s.weights(t, r, k) for r, k in zip(raws, Ks)]
        xb = s.to_device(x)
        sizes = [K, K // 4, K // 4, K, K, F, F, K, K, 1000]
        bufs = [s.malloc(4 * n) for n in sizes]
        nodes = _chain_nodes(ws, xb, bufs, K, Ms)
        res = {}
        for qx in ("0", "1"):
            monkeypatch.setenv("B200_QX", qx)
            for b, n in zip(bufs, sizes):
                s.h2d(b, np.zeros(n, np.float32))
            l0 = s.launch_count()
            s.graph_compute(nodes, use_cuda_graph=False)
            s.sync()
            res[qx] = ([s.to_host(b, (n,), np.float32) for b, n in zip(bufs, sizes)], s.launch_count() - l0)
        (plain, l_plain), (fused, l_fused) = res["0"], res["1"]
        for a, b in zip(plain, fused):
            assert np.array_equal(a, b)
        assert l_plain == 5 + 5 and l_fused == 5 + 1, (l_plain, l_fused)      # 5 GEMV launches (q/k/v, wo + ADD, gate/up, down + ADD, out), each behind a quantise launch / only the first
        for _ in range(3):                                                      # capture + replays
            s.graph_compute(nodes, use_cuda_graph=2)
        s.sync()
        for a, b, n in zip(fused, bufs, sizes):
            assert np.array_equal(a, s.to_host(b, (n,), np.float32))
        mm = lambda i, v: O.mul_mat(t, raws[i], v.reshape(1, -1), 1).reshape(-1)
        # every op against the oracle on the input the GPU itself fed it (a chained oracle drifts: each op quantises its own input)
        fq, fo, fh, fg, fdn, fh2 = fused[0], fused[3], fused[4], fused[5], fused[7], fused[8]
        assert np.array_equal(fh, x.reshape(-1) + fo) and np.array_equal(fh2, fh + fdn)
        for got, ref in ((fq, mm(0, x)), (fo, mm(3, fq)), (fg, mm(4, fh)), (fdn, mm(6, fg)), (fused[9], mm(7, fh2))):
            assert nmse(got, ref) <= 1e-8

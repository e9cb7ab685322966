"""GPU parity tests of the default prefill path (gemm_dense.cu: block scales folded into fp16 operands, the quantised weights
dequantised inside the kernel by converter warps, the whole K accumulated in TMEM by tcgen05.mma.cta_group::2 kind::f16), called
through b200_mulmat with many activation columns, against the oracle.
Bars: against the oracle's q8_1 block algebra (mode 1) only the fp16 rounding of the two dequantised operands differs
(relative 2^-12 per element -> NMSE ~4e-8; bar 1e-6); against ggml-cpu semantics (mode 0) the test-backend-ops bar
NMSE <= 5e-4 (tests/test-backend-ops.cpp:1982-1984)."""
import os
import numpy as np
import pytest

import pyoracle as O
import b200ops as B

pytestmark = pytest.mark.gpu

ALL = [(O.Q8_0, "q8_0"), (O.Q4_0, "q4_0"), (O.Q4_K, "q4_K"), (O.Q6_K, "q6_K")]
DENSE_BAR = 1e-6


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.fixture(autouse=True)
def _dense_from_16_columns(monkeypatch):
    monkeypatch.setenv("B200_GEMM_MIN_N", "16")
    monkeypatch.delenv("B200_GEMM_MODE", raising=False)
    monkeypatch.delenv("B200_GEMM_BN", raising=False)


@pytest.fixture(params=["tma", "ring"])
def wpath(request, monkeypatch):
    """how the native blocks reach the converter warps: TMA boxes (default) or per-thread cp.async rings (rows that are not 16-byte aligned)"""
    monkeypatch.setenv("B200_GEMM_WTMA", "1" if request.param == "tma" else "0")
    return request.param


@pytest.fixture(scope="module")
def sess():
    s = B.Session(0)
    yield s
    s.close()


@pytest.mark.parametrize("t,name", ALL)
@pytest.mark.parametrize("M,K,N", [(128, 256, 128), (16, 256, 32), (200, 1024, 77), (384, 4096, 512), (130, 512, 33), (257, 11008, 129)])
def test_dense_vs_oracle(sess, wpath, t, name, M, K, N):
    rng = np.random.default_rng(M + 3 * K + 7 * N + t)
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    before = sess.launch_count()
    got = sess.mulmat(w, x)
    assert sess.launch_count() > before
    assert got.shape == (N, M)
    e1 = nmse(got, O.mul_mat(t, raw, x, 1))
    assert e1 <= DENSE_BAR, (name, M, K, N, e1)
    assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
    again = sess.mulmat(w, x)                               # deterministic: same tiles, same order
    assert np.array_equal(got, again)
    w.free()


@pytest.mark.parametrize("t,name", [(O.Q8_0, "q8_0"), (O.Q4_0, "q4_0")])
def test_dense_ragged_k(sess, t, name):
    """K = 96: the last 64-element stage is half out of range (TMA zero fill on both operands)"""
    rng = np.random.default_rng(41 + t)
    M, K, N = 130, 96, 33
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    assert nmse(sess.mulmat(w, x), O.mul_mat(t, raw, x, 1)) <= DENSE_BAR
    w.free()


@pytest.mark.parametrize("bn", ["128", "256"])
@pytest.mark.parametrize("t,name", ALL)
def test_dense_both_tile_widths(sess, monkeypatch, wpath, bn, t, name):
    monkeypatch.setenv("B200_GEMM_BN", bn)
    rng = np.random.default_rng(43 + t)
    M, K, N = 300, 1280, 300
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    assert nmse(sess.mulmat(w, x), O.mul_mat(t, raw, x, 1)) <= DENSE_BAR
    w.free()


def test_dense_integer_product_is_exact(sess):
    """integers that fp16 holds exactly (|w| <= 15 with d = 1, |x| <= 127 with d_a = 1): every product and every partial sum
    is an integer below 2^24, so the f32 accumulation in TMEM must reproduce the int64 matrix product bit for bit"""
    rng = np.random.default_rng(11)
    M, K, N = 256, 2048, 96
    wq = rng.integers(-15, 16, size=(M, K // 32, 32), dtype=np.int8)
    raw = np.zeros((M, K // 32, 34), dtype=np.uint8)
    raw[:, :, 0:2] = np.frombuffer(np.float16(1.0).tobytes(), dtype=np.uint8)
    raw[:, :, 2:] = wq.view(np.uint8)
    raw = raw.reshape(M, -1)
    x = rng.integers(-127, 128, size=(N, K // 32, 32)).astype(np.float32)
    x[:, :, 5] = 127.0 * rng.choice([-1.0, 1.0], size=(N, K // 32))
    x = x.reshape(N, K)
    w = sess.weights(O.Q8_0, raw, K)
    got = sess.mulmat(w, x)
    want = x.astype(np.int64) @ wq.reshape(M, K).astype(np.int64).T
    assert np.array_equal(got, want.astype(np.float32))
    w.free()


@pytest.mark.parametrize("K", [8192, 2304])
@pytest.mark.parametrize("t,name", ALL)
def test_dense_split_k_small_m(sess, wpath, t, name, K):
    """few output tiles (M = 256, N = 96 -> 2 tiles): the K range is split over work items and reduced in a second kernel;
    K = 2304: 36 slabs in 4 splits of 9 -> the splits begin and end inside the 4-slab raw stages (super-blocks) of the weight loads"""
    rng = np.random.default_rng(23 + t)
    M, N = 256, 96
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    assert nmse(sess.mulmat(w, x), O.mul_mat(t, raw, x, 1)) <= DENSE_BAR
    w.free()


def test_dense_follows_a_new_upload(sess):
    """nothing derived from the weights is cached: uploading other weights into the same buffer is seen by the next call"""
    import ctypes as C
    rng = np.random.default_rng(47)
    M, K, N = 256, 1024, 64
    raw1, raw2 = O.random_blocks(O.Q4_0, M, K, rng), O.random_blocks(O.Q4_0, M, K, rng)
    w = sess.weights(O.Q4_0, raw1, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    assert nmse(sess.mulmat(w, x), O.mul_mat(O.Q4_0, raw1, x, 1)) <= DENSE_BAR
    raw2 = np.ascontiguousarray(raw2)
    B._ck(B.lib().b200_weights_upload(sess.h, O.Q4_0, w.buf.ptr, w.blocks, 0, w.blocks, B._p(raw2)))
    assert nmse(sess.mulmat(w, x), O.mul_mat(O.Q4_0, raw2, x, 1)) <= DENSE_BAR
    w.free()


def test_dense_inside_a_captured_graph(monkeypatch):
    """node list with dense MUL_MATs submitted with use_cuda_graph: the scratch is sized before the capture starts"""
    rng = np.random.default_rng(31)
    M, K, N = 384, 2048, 96
    with B.Session(0) as sess:
        ws, xs = [], rng.uniform(-1, 1, (N, K)).astype(np.float32)
        xb = sess.to_device(xs)
        nodes, outs, raws = [], [], []
        for t in (O.Q8_0, O.Q4_0, O.Q4_K, O.Q6_K):
            raw = O.random_blocks(t, M, K, rng)
            w = sess.weights(t, raw, K)
            d = sess.malloc(4 * M * N)
            n = B.Node()
            n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(d.ptr, [M, N], op=B.OP_MUL_MAT)
            nodes.append(n); ws.append(w); outs.append(d); raws.append((t, raw))
        arr = (B.Node * len(nodes))(*nodes)
        for _ in range(3):                                    # capture, then replays
            sess.graph_compute(arr, 2)
        sess.sync()
        for (t, raw), d in zip(raws, outs):
            got = sess.to_host(d, (N, M), np.float32)
            assert nmse(got, O.mul_mat(t, raw, xs, 1)) <= DENSE_BAR
        for b in ws + outs + [xb]:
            b.free()


@pytest.mark.parametrize("use_graph", [0, 2])
@pytest.mark.parametrize("N", [300, 64])
def test_dense_grouped_launch(use_graph, N):
    """consecutive MUL_MATs over the same activations (q / k / v, gate / up) go out as one launch of the pair kernel: members of
    different row counts (ragged against the 256-row tiles) and weight types, every dst checked against the oracle"""
    rng = np.random.default_rng(67 + N)
    K = 1536
    members = [(O.Q4_K, 700), (O.Q4_K, 130), (O.Q4_K, 1024), (O.Q6_K, 256), (O.Q6_K, 257)]      # groups of 3 + 2 (a group shares one weight type)
    with B.Session(0) as sess:
        xs = rng.standard_normal((N, K)).astype(np.float32)
        xb = sess.to_device(xs)
        nodes, keep, outs = [], [], []
        for t, M in members:
            raw = O.random_blocks(t, M, K, rng)
            w = sess.weights(t, raw, K)
            d = sess.malloc(4 * M * N)
            n = B.Node()
            n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(d.ptr, [M, N], op=B.OP_MUL_MAT)
            nodes.append(n); keep.append(w); outs.append((t, raw, M, d))
        arr = (B.Node * len(nodes))(*nodes)
        sess.graph_compute(arr, use_graph)                     # first call: sizes the scratch (and captures the graph)
        sess.sync()
        before = sess.launch_count()
        sess.graph_compute(arr, use_graph)
        sess.sync()
        assert sess.launch_count() - before <= 3              # one activation conversion + two grouped GEMMs
        for t, raw, M, d in outs:
            got = sess.to_host(d, (N, M), np.float32)
            assert nmse(got, O.mul_mat(t, raw, xs, 1)) <= DENSE_BAR, (t, M)
        for b in keep + [d for _, _, _, d in outs] + [xb]:
            b.free()


@pytest.mark.parametrize("t,name", ALL)
@pytest.mark.parametrize("M,what", [(10000, "80 pair tiles: 6 left after a wave of 74 -> 48 chunks of 2 slabs"),
                                    (13203, "104 pair tiles: 30 left -> chunks of 7 slabs that cross tile boundaries; M odd: unaligned dst columns"),
                                    (9500, "76 pair tiles: 2 left -> 32 chunks of one slab, 42 idle pairs"),
                                    (5632, "44 tiles on 74 pairs, no full wave: everything is tail")])
def test_dense_stream_k_tail_of_the_last_wave(sess, monkeypatch, t, name, M, what):
    """more tiles than CTA pairs: the tiles left over after the full waves are cut into one chunk of consecutive (tile, slab) steps per
    pair; partial tiles meet in gd_tail_fixup_kernel.  Deterministic, and the same matrix product as without the tail."""
    monkeypatch.setenv("B200_GEMM_TAIL", "70")             # (default 25: only waves up to a quarter full)
    rng = np.random.default_rng(71 + M + t)
    K, N = 1024, 300
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    l0 = sess.launch_count()
    got = sess.mulmat(w, x)
    assert sess.launch_count() - l0 == 3, what            # activations -> fp16, GEMM, tail fix-up
    ref = O.mul_mat(t, raw, x, 1)
    assert nmse(got, ref) <= DENSE_BAR, what
    assert np.array_equal(got, sess.mulmat(w, x))
    monkeypatch.setenv("B200_GEMM_TAIL", "0")
    l0 = sess.launch_count()
    plain = sess.mulmat(w, x)
    assert sess.launch_count() - l0 == 2
    assert nmse(plain, ref) <= DENSE_BAR and nmse(got, plain) <= 1e-10
    w.free()


@pytest.mark.parametrize("model_shape", [("llama2-7b gate+up", O.Q4_0, [11008, 11008], 4096), ("llama2-7b q+k+v", O.Q4_0, [4096, 4096, 4096], 4096),
                                         ("llama3-8b q+k+v", O.Q4_K, [4096, 1024, 1024], 4096), ("llama3-8b gate+up", O.Q4_K, [14336, 14336], 4096),
                                         ("llama2-7b down", O.Q8_0, [4096], 11008)])
def test_dense_full_size_against_exact_mode(monkeypatch, model_shape):
    """BASELINE.json's prefill shapes at full size (N = 512), as the node list submits them (grouped launches, narrow units of the last
    wave, split-K): the oracle is too slow here, so the check is against the exact per-block GEMM (gemm_i8.cu, itself oracle-checked at
    small sizes): only the fp16 rounding of the operands may differ"""
    name, t, Ms, K = model_shape
    N = 512
    rng = np.random.default_rng(len(name) + K)
    with B.Session(0) as sess:
        xs = rng.standard_normal((N, K)).astype(np.float32)
        xb = sess.to_device(xs)
        nodes, keep, outs = [], [], []
        for M in Ms:
            raw = np.tile(O.random_blocks(t, 256, K, rng), (M // 256, 1))
            w = sess.weights(t, raw, K)
            d = sess.malloc(4 * M * N)
            n = B.Node()
            n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(d.ptr, [M, N], op=B.OP_MUL_MAT)
            nodes.append(n); keep.append(w); outs.append((M, d))
        arr = (B.Node * len(nodes))(*nodes)
        sess.graph_compute(arr, 0)
        sess.sync()
        dense = [sess.to_host(d, (N, M), np.float32) for M, d in outs]
        monkeypatch.setenv("B200_GEMM_MODE", "exact")
        sess.graph_compute(arr, 0)
        sess.sync()
        for (M, d), got in zip(outs, dense):
            exact = sess.to_host(d, (N, M), np.float32)
            assert np.isfinite(got).all()
            assert nmse(got, exact) <= DENSE_BAR, (name, M)
        for b in keep + [d for _, d in outs] + [xb]:
            b.free()


@pytest.mark.parametrize("in_place", [False, True])
@pytest.mark.parametrize("use_graph", [0, 2])
def test_dense_residual_add_fused_into_the_split_k_reduce(in_place, use_graph):
    """MUL_MAT (few tiles -> split-K) followed by the ADD that consumes it (wo / down + residual): with the pair kernel the reduce
    writes both results; both must equal the oracle's MUL_MAT and the IEEE sum bit for bit"""
    rng = np.random.default_rng(83)
    M, K, N = 512, 4096, 128
    with B.Session(0) as sess:
        raw = O.random_blocks(O.Q4_0, M, K, rng)
        w = sess.weights(O.Q4_0, raw, K)
        xs = rng.standard_normal((N, K)).astype(np.float32)
        res = rng.standard_normal((N, M)).astype(np.float32)
        xb, rb = sess.to_device(xs), sess.to_device(res)
        y, z = sess.malloc(4 * M * N), sess.malloc(4 * M * N)
        zt = rb if in_place else z
        # a second MUL_MAT consumes the sum (K2 = M): its activations were converted by the reduce of the first
        M2 = 384
        raw2 = O.random_blocks(O.Q6_K, M2, M, rng)
        w2 = sess.weights(O.Q6_K, raw2, M)
        o2 = sess.malloc(4 * M2 * N)
        mm, ad, m2 = B.Node(), B.Node(), B.Node()
        mm.src0, mm.src1, mm.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(y.ptr, [M, N], op=B.OP_MUL_MAT)
        ad.src0, ad.src1, ad.dst = B.f32_tensor(rb.ptr, [M, N]), B.f32_tensor(y.ptr, [M, N]), B.f32_tensor(zt.ptr, [M, N], op=B.OP_ADD)
        m2.src0, m2.src1, m2.dst = w2.tensor(), B.f32_tensor(zt.ptr, [M, N]), B.f32_tensor(o2.ptr, [M2, N], op=B.OP_MUL_MAT)
        arr = (B.Node * 3)(mm, ad, m2)
        sess.graph_compute(arr, use_graph)
        sess.sync()
        before = sess.launch_count()
        if not in_place:
            sess.graph_compute(arr, use_graph)              # (in place the residual has already moved on)
            sess.sync()
            # conversion, GEMM, reduce (+ ADD + the second MUL_MAT's conversion), GEMM (8 k-stages: no split)
            assert sess.launch_count() - before == 4
        got_y, got_z = sess.to_host(y, (N, M), np.float32), sess.to_host(zt, (N, M), np.float32)
        got_o = sess.to_host(o2, (N, M2), np.float32)
        assert nmse(got_y, O.mul_mat(O.Q4_0, raw, xs, 1)) <= DENSE_BAR
        assert np.array_equal(got_z, res + got_y)
        assert nmse(got_o, O.mul_mat(O.Q6_K, raw2, got_z, 1)) <= DENSE_BAR
        for b in (xb, rb, y, z, o2):
            b.free()
        w.free(); w2.free()


def test_dense_close_to_exact_mode(sess, monkeypatch):
    """the two prefill kernels on the same inputs: they differ by the fp16 rounding of the operands only"""
    rng = np.random.default_rng(59)
    M, K, N = 512, 4096, 128
    raw = O.random_blocks(O.Q4_0, M, K, rng)
    w = sess.weights(O.Q4_0, raw, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    dense = sess.mulmat(w, x)
    monkeypatch.setenv("B200_GEMM_MODE", "exact")
    exact = sess.mulmat(w, x)
    assert nmse(dense, exact) <= DENSE_BAR and not np.array_equal(dense, exact)
    w.free()


# ------------------------------------------------------------------------------------------------ operand ranges
def _col_nmse(got, ref):
    """worst column: columns of very different magnitude must each be right, not only the largest"""
    num = ((got.astype(np.float64) - ref) ** 2).sum(axis=1)
    den = np.maximum((ref.astype(np.float64) ** 2).sum(axis=1), 1e-300)
    return float((num / den).max())


@pytest.mark.parametrize("N", [64, 512])
@pytest.mark.parametrize("scale", [1e5, 3e6, 1e-3, 1e-6, "mixed"])
@pytest.mark.parametrize("t,name", ALL)
def test_dense_activation_range_vs_oracle(sess, t, name, scale, N):
    """activations far outside what fp16(d_a q_a) could hold unscaled (|x| up to 3e6: the reference's own limit is d_a = amax / 127 <= 65504)
    and far below (1e-6: the reference's fp16 d_a is zero there and so is ours); 'mixed': every column at its own magnitude, 1e-5 .. 3e6.
    The per-column power-of-two scale of the conversion kernel keeps the dense path on the oracle for all of them."""
    rng = np.random.default_rng(900 + t + N)
    M, K = 300, 1024
    raw = O.random_blocks(t, M, K, rng)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    if scale == "mixed":
        x *= (10.0 ** rng.uniform(-5, 6.4, (N, 1))).astype(np.float32)
    else:
        x *= np.float32(scale)
    w = sess.weights(t, raw, K)
    got = sess.mulmat(w, x)
    ref = O.mul_mat(t, raw, x, 1)
    assert np.isfinite(got).all()
    if scale == 1e-6:
        assert not ref.any() and not got.any()             # d_a = amax / 127 rounds to zero in fp16: quantize_row_q8_1_ref, and the oracle, say 0
    else:
        assert _col_nmse(got, ref) <= DENSE_BAR, (name, scale, N, _col_nmse(got, ref))
        if scale != 1e-3:                                     # (1e-3: ggml-cpu keeps an f32 d for its q8_K activations, q8_1's fp16 d is subnormal there)
            assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
    w.free()


def _with_scale(t, raw, M, K, d16, rows=slice(None), blocks=slice(None)):
    """overwrite the fp16 block scale d (q4_K: d and dmin) of the given rows / blocks of a wire-format matrix"""
    bpb, nb = B.BLOCK[t][1], K // B.BLOCK[t][0]
    r = raw.reshape(M, nb, bpb).copy()
    h = np.frombuffer(np.float16(d16).tobytes(), dtype=np.uint8)
    off = 208 if t == O.Q6_K else 0
    r[rows, blocks, off:off + 2] = h
    if t == O.Q4_K:
        r[rows, blocks, 2:4] = h
    return r.reshape(M, nb * bpb)


@pytest.mark.parametrize("N", [64, 512])
@pytest.mark.parametrize("case", ["all 1e-7", "every other row 1e-7", "first block 1e-7, the rest 1e-5", "first block 6e-8, the rest 2e-3"])
@pytest.mark.parametrize("t,name", ALL)
def test_dense_tiny_weight_scales_vs_oracle(sess, t, name, case, N):
    """block scales in fp16's subnormal range: q4_0 / q8_0 products d_w q_w stay exact (integers times 2^-24), K-quant rows are scaled up by a
    power of two in the converters (taken from the first super-block) and scaled back in the epilogue"""
    rng = np.random.default_rng(950 + t + N)
    M, K = 300, 2048
    raw = O.random_blocks(t, M, K, rng)
    if case == "all 1e-7":
        raw = _with_scale(t, raw, M, K, 1e-7)
    elif case == "every other row 1e-7":
        raw = _with_scale(t, raw, M, K, 1e-7, rows=slice(0, None, 2))
    elif case == "first block 1e-7, the rest 1e-5":
        raw = _with_scale(t, _with_scale(t, raw, M, K, 1e-5), M, K, 1e-7, blocks=slice(0, 1))
    else:
        raw = _with_scale(t, _with_scale(t, raw, M, K, 2e-3), M, K, 6e-8, blocks=slice(0, 1))
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    w = sess.weights(t, raw, K)
    got = sess.mulmat(w, x)
    ref = O.mul_mat(t, raw, x, 1)
    assert np.isfinite(got).all()
    rows_err = ((got.astype(np.float64) - ref) ** 2).sum(axis=0) / np.maximum((ref.astype(np.float64) ** 2).sum(axis=0), 1e-300)
    assert float(rows_err.max()) <= DENSE_BAR, (name, case, N, float(rows_err.max()))      # every dst row (= weight row) on its own
    assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
    w.free()

"""GPU parity tests of the exact tensor-core prefill path (B200_GEMM_MODE=exact, gemm_i8.cu: int8 tcgen05.mma per 32-element block, fp32
scale-accumulate), called through b200_mulmat with many activation columns, against the oracle:
mode 1 = the same q8_1 block algebra (only float rounding of the running sum can differ), mode 0 = ggml-cpu
semantics with the test-backend-ops bar NMSE <= 5e-4 (tests/test-backend-ops.cpp:1982-1984)."""
import numpy as np
import pytest

import pyoracle as O
import b200ops as B

pytestmark = pytest.mark.gpu


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.fixture(autouse=True)
def _gemm_from_16_columns(monkeypatch):
    """the tensor-core path normally takes over at 36 (K-quants) / 52 columns; these tests want it for the small shapes too"""
    monkeypatch.setenv("B200_GEMM_MIN_N", "16")
    monkeypatch.setenv("B200_GEMM_MODE", "exact")          # this file: the exact-block-sum kernel; tests/test_gpu_gemm_dense.py: the default


@pytest.fixture(scope="module")
def sess():
    s = B.Session(0)
    yield s
    s.close()


@pytest.mark.parametrize("t,name", [(O.Q8_0, "q8_0"), (O.Q4_0, "q4_0")])
@pytest.mark.parametrize("M,K,N", [(128, 128, 128), (16, 256, 32), (200, 1024, 77), (384, 4096, 512), (130, 96, 33), (257, 11008, 129)])
def test_gemm_vs_oracle(sess, t, name, M, K, N):
    rng = np.random.default_rng(M + 3 * K + 7 * N + t)
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = sess.mulmat(w, x)
    assert got.shape == (N, M)
    ref1 = O.mul_mat(t, raw, x, 1)
    assert nmse(got, ref1) <= 1e-10, (name, M, K, N, nmse(got, ref1))
    assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
    w.free()


@pytest.mark.parametrize("t,name", [(O.Q4_K, "q4_K"), (O.Q6_K, "q6_K")])
@pytest.mark.parametrize("M,K,N", [(128, 256, 128), (16, 256, 32), (200, 1024, 77), (384, 4096, 512), (130, 512, 33), (257, 11008, 129)])
def test_gemm_kquants_vs_oracle(sess, t, name, M, K, N):
    """K-quants through the same tensor-core kernel: q4_K as 32-element sub-blocks with f32 scales d*sc and the block-minimum
    term folded into the epilogue, q6_K with every 16-element scale group padded to its own 32-byte MMA block"""
    rng = np.random.default_rng(M + 3 * K + 7 * N + t)
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = sess.mulmat(w, x)
    assert got.shape == (N, M)
    ref1 = O.mul_mat(t, raw, x, 1)
    assert nmse(got, ref1) <= 1e-10, (name, M, K, N, nmse(got, ref1))
    assert nmse(got, O.mul_mat(t, raw, x, 0)) <= 5e-4
    w.free()


@pytest.mark.parametrize("t", [O.Q4_K, O.Q6_K])
def test_gemm_kquants_match_gemv_columns(sess, t):
    rng = np.random.default_rng(5 + t)
    M, K, N = 256, 2048, 40
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    big = sess.mulmat(w, x)
    small = np.concatenate([sess.mulmat(w, x[i:i + 8]) for i in range(0, N, 8)])
    assert nmse(big, small) <= 1e-10
    w.free()


def test_gemm_matches_gemv_columns(sess):
    """the same matrix through the GEMV (8 columns at a time) and through the GEMM: same integers, same scales"""
    rng = np.random.default_rng(3)
    M, K, N = 512, 2048, 64
    raw = O.random_blocks(O.Q8_0, M, K, rng)
    w = sess.weights(O.Q8_0, raw, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    big = sess.mulmat(w, x)
    small = np.concatenate([sess.mulmat(w, x[i:i + 8]) for i in range(0, N, 8)])
    assert nmse(big, small) <= 1e-10
    w.free()


def test_gemm_block_sums_are_exact_integers(sess):
    """q8_0 weights with d = 1 and integer activations whose blocks all contain +-127 (so d_a = 1 and the quants are the
    activations themselves): every block sum and the whole result are integers below 2^24, and the tensor-core path must
    reproduce the int64 matrix product bit for bit -- whichever MMA kind computes the block sums."""
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
    assert np.abs(want).max() < 2 ** 24
    assert np.array_equal(got.astype(np.int64), want) and np.array_equal(got, want.astype(np.float32))
    w.free()


@pytest.mark.parametrize("kind", ["i8", "f16"])
@pytest.mark.parametrize("t", [O.Q8_0, O.Q4_0, O.Q4_K, O.Q6_K])
def test_gemm_both_mma_kinds(sess, monkeypatch, kind, t):
    """kind::i8 (int8 operands, s32 sums) and kind::f16 (the same integers as fp16, f32 sums) give the same block sums:
    the two results may differ only in float rounding of the scale-accumulate, and both match the oracle"""
    monkeypatch.setenv("B200_GEMM_KIND", kind)
    rng = np.random.default_rng(17 + t)
    M, K, N = 200, 1280, 70
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = sess.mulmat(w, x)
    assert nmse(got, O.mul_mat(t, raw, x, 1)) <= 1e-10
    w.free()


@pytest.mark.parametrize("t", [O.Q8_0, O.Q4_K, O.Q6_K])
def test_gemm_split_k_small_m(sess, t):
    """few output tiles (M = 256, N = 96 -> 2 tiles): the K range is split over work items and reduced in a second kernel"""
    rng = np.random.default_rng(23 + t)
    M, K, N = 256, 8192, 96
    raw = O.random_blocks(t, M, K, rng)
    w = sess.weights(t, raw, K)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = sess.mulmat(w, x)
    assert nmse(got, O.mul_mat(t, raw, x, 1)) <= 1e-10
    w.free()


def test_gemm_inside_a_captured_graph(sess):
    """a node list with tensor-core MUL_MATs submitted with use_cuda_graph: the GEMM scratch is sized before the stream
    capture starts (allocations are not allowed inside one), the replay gives the same result as the direct launch"""
    import ctypes as C
    rng = np.random.default_rng(31)
    M, K, N = 384, 2048, 96
    ws, xs = [], rng.uniform(-1, 1, (N, K)).astype(np.float32)
    xb = sess.to_device(xs)
    nodes, outs, raws = [], [], []
    for t in (O.Q8_0, O.Q4_K, O.Q6_K):
        raw = O.random_blocks(t, M, K, rng)
        w = sess.weights(t, raw, K)
        d = sess.malloc(4 * M * N)
        n = B.Node()
        n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(d.ptr, [M, N], op=B.OP_MUL_MAT)
        nodes.append(n); ws.append(w); outs.append(d); raws.append((t, raw))
    arr = (B.Node * len(nodes))(*nodes)
    for _ in range(2):                                    # capture, then replay
        sess.graph_compute(arr, 2)
    sess.sync()
    for (t, raw), d in zip(raws, outs):
        got = sess.to_host(d, (N, M), np.float32)
        assert nmse(got, O.mul_mat(t, raw, xs, 1)) <= 1e-10
    for b in ws + outs + [xb]:
        b.free()


def test_captured_graph_survives_scratch_growth():
    """a captured graph bakes the scratch addresses in; when a later, bigger op moves the scratch, the old graph must not be
    replayed (the cache key carries a scratch generation) -- the same node list is captured again and stays correct"""
    rng = np.random.default_rng(37)
    with B.Session(0) as s:
        M, K, N = 128, 1024, 64
        raw = O.random_blocks(O.Q4_0, M, K, rng)
        w = s.weights(O.Q4_0, raw, K)
        x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        xb, d = s.to_device(x), s.malloc(4 * M * N)
        n = B.Node()
        n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(xb.ptr, [K, N]), B.f32_tensor(d.ptr, [M, N], op=B.OP_MUL_MAT)
        arr = (B.Node * 1)(n)
        s.graph_compute(arr, 2); s.graph_compute(arr, 2)
        ref = O.mul_mat(O.Q4_0, raw, x, 1)
        assert nmse(s.to_host(d, (N, M), np.float32), ref) <= 1e-10
        big = O.random_blocks(O.Q4_K, 1024, 4096, rng)
        wb = s.weights(O.Q4_K, big, 4096)
        s.mulmat(wb, rng.uniform(-1, 1, (256, 4096)).astype(np.float32))        # grows every GEMM scratch buffer
        s.h2d(d, np.zeros((N, M), np.float32))
        s.graph_compute(arr, 2)
        assert nmse(s.to_host(d, (N, M), np.float32), ref) <= 1e-10

"""GPU parity tests of the tensor-core prefill path (gemm_i8.cu: int8 tcgen05.mma per 32-element block, fp32
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

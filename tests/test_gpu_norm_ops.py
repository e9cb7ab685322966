"""GPU parity tests (pytest -m gpu): RMS_NORM, SOFT_MAX, POOL_2D — the ops ggml-hexagon's cDSP table lists beside ADD / MUL_MAT and
leaves as empty stubs (ggml-hexagon.cpp:619, 639, 649; kernels/ggml-dsp.c:1353-1372) — through the C-ABI against the oracle and the
reference-generated fixtures.  Bars: test-backend-ops' default NMSE 1e-7 (tests/test-backend-ops.cpp:325-327); in fact RMS_NORM and
POOL_2D come out bit-equal (same IEEE operations, double sum of squares / same tap order), SOFT_MAX within a few ulp (expf)."""
import os

import numpy as np
import pytest

import b200ops as B
import pyoracle as O

pytestmark = pytest.mark.gpu
GO = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_scalar_ops.npz"))


def nmse(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


@pytest.fixture(scope="module")
def sess():
    s = B.Session(0)
    yield s
    s.close()


def test_rms_norm_golden(sess):
    assert nmse(sess.rms_norm(GO["rms_x"], 1e-6), GO["rms_out_eps1e-6"]) <= 1e-12
    assert nmse(sess.rms_norm(GO["rms_x"][1:], 0.0), GO["rms_out_eps0"]) <= 1e-12
    got = sess.rms_norm(GO["rms_xl"], 1e-5)
    assert nmse(got, GO["rms_outl"]) <= 1e-12
    assert np.abs(got - GO["rms_outl"]).max() <= 2.4e-7 * np.abs(GO["rms_outl"]).max()


@pytest.mark.parametrize("shape,eps", [((1, 4096), 1e-5), ((512, 4096), 1e-6), ((3, 5, 7, 33), 1e-6), ((2, 11008), 1e-5), ((4, 1), 0.0), ((2, 131072), 1e-6)])
def test_rms_norm_vs_oracle(sess, shape, eps):
    rng = np.random.default_rng(sum(shape))
    x = (rng.standard_normal(shape) * rng.uniform(0.01, 30)).astype(np.float32)
    got, ref = sess.rms_norm(x, eps), O.rms_norm(x, eps)
    assert nmse(got, ref) <= 1e-12
    assert np.isfinite(got).all()


def test_rms_norm_zero_row_and_bad_args(sess):
    x = np.zeros((2, 64), np.float32)
    assert np.array_equal(sess.rms_norm(x, 1e-6), O.rms_norm(x, 1e-6))
    with pytest.raises(B.B200Error):
        sess.rms_norm(np.ones((1, 64), np.float32), -1.0)


def test_soft_max_golden(sess):
    for got, key in ((sess.soft_max(GO["sm_x"]), "sm_plain"), (sess.soft_max(GO["sm_x"], GO["sm_mask"], 0.1, 0.0), "sm_mask_scale"),
                     (sess.soft_max(GO["sm_x"], GO["sm_mask"], 0.1, 8.0), "sm_alibi"), (sess.soft_max(GO["sm_xl"], None, 0.125, 0.0), "sm_outl"),
                     (sess.soft_max(GO["sm_x"], GO["sm_mask"].astype(np.float16).astype(np.float32), 0.1, 8.0), None)):
        if key is None:
            ref = O.soft_max(GO["sm_x"], GO["sm_mask"].astype(np.float16).astype(np.float32), 0.1, 8.0)
        else:
            ref = GO[key]
        assert nmse(got, ref) <= 1e-10, key
        assert np.allclose(got.sum(-1), 1.0, atol=1e-5)


@pytest.mark.parametrize("shape,mask,scale,max_bias", [((1, 1, 1, 4096), False, 1.0, 0.0), ((1, 32, 7, 129), True, 0.125, 0.0), ((2, 8, 16, 1024), True, 0.1, 8.0),
                                                      ((1, 5, 64, 77), False, 1.0, 0.0), ((1, 12, 3, 40000), True, 0.05, 4.0), ((1, 1, 2, 1), False, 1.0, 0.0)])
@pytest.mark.parametrize("mdtype", [np.float32, np.float16])
def test_soft_max_vs_oracle(sess, shape, mask, scale, max_bias, mdtype):
    rng = np.random.default_rng(sum(shape))
    x = (rng.standard_normal(shape) * 4).astype(np.float32)
    m = None
    if mask:
        m = rng.standard_normal((shape[2] + 1, shape[3])).astype(np.float32)        # (a mask may have more rows than src0)
        m[m > 1.2] = -np.inf
        m[:, 0] = 0.0                                                                # (never a fully masked row)
        m = m.astype(mdtype)
    elif mdtype == np.float16:
        pytest.skip("no mask")
    got = sess.soft_max(x, m, scale, max_bias)
    ref = O.soft_max(x, None if m is None else m.astype(np.float32), scale, max_bias)
    assert nmse(got, ref) <= 1e-10
    assert np.allclose(got.sum(-1), 1.0, atol=1e-4)


def test_pool_2d_golden_bit_exact(sess):
    for name in ("max_3311", "avg_3322", "avg_2321", "max_1122"):
        a = [int(v) for v in GO["poolargs_" + name]]
        assert np.array_equal(sess.pool_2d(GO["pool_x"], *a), GO["pool_" + name]), name


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
@pytest.mark.parametrize("op", [0, 1])
@pytest.mark.parametrize("k0,k1,s0,s1,p0,p1", [(1, 1, 1, 1, 0, 0), (3, 3, 1, 1, 0, 0), (3, 3, 2, 2, 1, 1), (2, 3, 2, 1, 0, 1), (5, 2, 3, 2, 2, 1)])
def test_pool_2d_vs_oracle(sess, dtype, op, k0, k1, s0, s1, p0, p1):
    rng = np.random.default_rng(k0 * 7 + s1 + p0)
    x = rng.standard_normal((2, 3, 10, 13)).astype(dtype)
    got = sess.pool_2d(x, op, k0, k1, s0, s1, p0, p1)
    assert np.array_equal(got, O.pool_2d(x.astype(np.float32), op, k0, k1, s0, s1, p0, p1))


def test_new_ops_in_a_node_list(sess):
    """RMS_NORM -> MUL_MAT -> SOFT_MAX through b200_graph_compute, direct and as a captured CUDA graph"""
    import ctypes as C
    rng = np.random.default_rng(8)
    K, M = 1024, 512
    raw = O.quantize(O.Q8_0, (rng.standard_normal((M, K)) * 0.05).astype(np.float32))
    w = sess.weights(O.Q8_0, raw, K)
    x = rng.standard_normal((1, K)).astype(np.float32)
    xb, nb, yb, pb = sess.to_device(x), sess.malloc(4 * K), sess.malloc(4 * M), sess.malloc(4 * M)
    n1, n2, n3 = B.Node(), B.Node(), B.Node()
    n1.src0, n1.dst = B.f32_tensor(xb.ptr, [K, 1]), B.f32_tensor(nb.ptr, [K, 1], op=B.OP_RMS_NORM)
    n1.dst.op_params[0] = int(np.float32(1e-5).view(np.int32))
    n2.src0, n2.src1, n2.dst = w.tensor(), B.f32_tensor(nb.ptr, [K, 1]), B.f32_tensor(yb.ptr, [M, 1], op=B.OP_MUL_MAT)
    n3.src0, n3.dst = B.f32_tensor(yb.ptr, [M, 1]), B.f32_tensor(pb.ptr, [M, 1], op=B.OP_SOFT_MAX)
    n3.dst.op_params[0] = int(np.float32(1.0).view(np.int32))
    ref = O.soft_max(O.mul_mat(O.Q8_0, raw, O.rms_norm(x, 1e-5), 1).reshape(1, 1, 1, M), None, 1.0, 0.0).reshape(-1)
    for mode in (0, 2, 2):
        sess.h2d(pb, np.zeros(M, np.float32))
        sess.graph_compute([n1, n2, n3], use_cuda_graph=mode)
        got = sess.to_host(pb, (M,), np.float32)
        assert nmse(got, ref) <= 1e-9
    w.free()
    for b in (xb, nb, yb, pb):
        b.free()

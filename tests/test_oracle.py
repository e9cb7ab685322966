"""CPU tests: the oracle (oracle/ggml_oracle.c) pinned against the reference's own outputs.

(a) tests/golden/ref_scalar.npz — produced by tests/golden/make_golden.py from the unmodified reference
    (scalar build of ggml-base/ggml-cpu); always available, also on the GPU box.
(b) oracle/_ref/* live, when the prebuilt reference libraries are present.
Integer/byte results must be bit-exact; float results of the scalar build are bit-exact too because the
oracle restates the scalar code path statement by statement.
"""
import os
import numpy as np
import pytest

import pyoracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_scalar.npz"))
WTYPES = [(O.Q4_0, "q4_0"), (O.Q8_0, "q8_0"), (O.Q4_K, "q4_K"), (O.Q6_K, "q6_K")]


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.mark.parametrize("t,name", [(O.Q8_0, "q8_0"), (O.Q8_1, "q8_1"), (O.Q8_K, "q8_K")])
def test_activation_quantisers_golden(t, name):
    got = O.quantize(t, G["act_x"])
    ref = G["act_" + name]
    if t == O.Q8_K:
        # the reference leaves bsums of an all-zero block untouched (uninitialised in its buffer); compare the rest
        assert np.array_equal(got[:, :260], ref[:, :260])
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("t,name", WTYPES)
def test_dequantize_golden(t, name):
    assert np.array_equal(O.dequantize(t, G["w_" + name], 512), G["deq_" + name])


@pytest.mark.parametrize("t,name", WTYPES)
def test_mul_mat_cpu_mode_golden_bit_exact(t, name):
    got = O.mul_mat(t, G["w_" + name], G["mm_x"], mode=0)
    assert np.array_equal(got, G["mm_" + name])


@pytest.mark.parametrize("t,name", WTYPES)
def test_mul_mat_q8_1_mode_within_backend_ops_tolerance(t, name):
    # tests/test-backend-ops.cpp:1982-1984: MUL_MAT passes at NMSE <= 5e-4
    got = O.mul_mat(t, G["w_" + name], G["mm_x"], mode=1)
    assert nmse(got, G["mm_" + name]) <= 5e-4


def test_weight_quantisers_simple_types_golden():
    assert np.array_equal(O.quantize(O.Q4_0, G["mm_wf"]), G["w_q4_0"])
    assert np.array_equal(O.quantize(O.Q8_0, G["mm_wf"]), G["w_q8_0"])


def test_add_golden():
    assert np.array_equal(O.add_f32(G["add_a"], G["add_b"]), G["add_out"])


@pytest.mark.parametrize("t,name", WTYPES)
def test_block_dots_are_the_integers_inside_vec_dot(t, name):
    # reassemble the float result from the exact integers and the scales; must reproduce mode-1 mul_mat
    rng = np.random.default_rng(3)
    K = 512
    w = O.random_blocks(t, 4, K, rng)
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    a = O.quantize(O.Q8_1, x)[0].reshape(K // 32, 36)
    qs = a[:, 4:].view(np.int8).reshape(-1)
    wd = O.dequantize(t, w, K)
    for r in range(4):
        dots = O.block_dots(t, K, w[r], qs)
        assert dots.dtype == np.int32 and dots.size == (K // 16 if t == O.Q6_K else K // 32)
        # brute force with numpy ints from the dequantised weights is not exact; recompute directly instead
    if t in (O.Q4_0, O.Q8_0):
        bs = 18 if t == O.Q4_0 else 34
        blk = w[0].reshape(K // 32, bs)
        if t == O.Q4_0:
            lo = (blk[:, 2:] & 0xF).astype(np.int32) - 8
            hi = (blk[:, 2:] >> 4).astype(np.int32) - 8
            wi = np.concatenate([lo, hi], axis=1)
        else:
            wi = blk[:, 2:].view(np.int8).astype(np.int32)
        ref = (wi * qs.reshape(K // 32, 32).astype(np.int32)).sum(axis=1)
        assert np.array_equal(O.block_dots(t, K, w[0], qs), ref.astype(np.int32))
    assert wd.shape == (4, K)


def test_empty_and_minimum_sizes():
    assert O.quantize(O.Q8_1, np.zeros((0, 32), np.float32)).shape == (0, 36)
    one = O.quantize(O.Q8_1, np.zeros((1, 32), np.float32))
    assert not one.any()
    out = O.mul_mat(O.Q8_0, O.quantize(O.Q8_0, np.ones((1, 32), np.float32)), np.ones((1, 32), np.float32), 0)
    assert out.shape == (1, 1) and abs(out[0, 0] - 32.0) < 0.1


# ------------------------------------------------------------------ live reference, when oracle/_ref is present
def _ref(variant):
    import pyref
    try:
        return pyref.Ref(variant)
    except (FileNotFoundError, OSError):
        pytest.skip("oracle/_ref/%s not built" % variant)


@pytest.mark.parametrize("t,name", WTYPES)
def test_oracle_vs_live_scalar_reference_bit_exact(t, name):
    R = _ref("scalar")
    rng = np.random.default_rng(11)
    K, M, N = 1024, 16, 2
    w = R.quantize_chunk(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    assert np.array_equal(O.dequantize(t, w, K), R.dequantize(t, w, K))
    assert np.array_equal(O.mul_mat(t, w, x, 0), R.mul_mat(t, w, x))
    wr = O.random_blocks(t, M, K, rng)       # bytes made directly in the wire format
    assert np.array_equal(O.dequantize(t, wr, K), R.dequantize(t, wr, K))
    assert np.array_equal(O.mul_mat(t, wr, x, 0), R.mul_mat(t, wr, x))


@pytest.mark.parametrize("t,name", WTYPES)
def test_oracle_vs_live_simd_reference_nmse(t, name):
    import pyref
    v = pyref.best_variant()
    if v in (None, "scalar"):
        pytest.skip("no SIMD reference build")
    R = _ref(v)
    rng = np.random.default_rng(12)
    K, M, N = 2048, 32, 3
    w = R.quantize_chunk(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    ref = R.mul_mat(t, w, x, n_threads=2)
    assert nmse(O.mul_mat(t, w, x, 0), ref) <= 1e-6
    assert nmse(O.mul_mat(t, w, x, 1), ref) <= 5e-4


def test_q8_1_vs_live_reference():
    R = _ref("scalar")
    rng = np.random.default_rng(13)
    x = (rng.standard_normal((8, 4096)) * rng.choice([1e-4, 1.0, 50.0], (8, 1))).astype(np.float32)
    assert np.array_equal(O.quantize(O.Q8_1, x), R.quantize_row_ref(O.Q8_1, x))


# ------------------------------------------------------------------ RMS_NORM / SOFT_MAX / POOL_2D
GO = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_scalar_ops.npz"))


def _nmse(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


def test_oracle_rms_norm_golden_bit_exact():
    assert np.array_equal(O.rms_norm(GO["rms_x"], 1e-6), GO["rms_out_eps1e-6"])
    assert np.array_equal(O.rms_norm(GO["rms_x"][1:], 0.0), GO["rms_out_eps0"])
    assert np.array_equal(O.rms_norm(GO["rms_xl"], 1e-5), GO["rms_outl"])


def test_oracle_soft_max_golden():
    # the reference's x86 builds run their own polynomial expf (ggml_v_expf, ggml-cpu.c:2181-2209) on groups of 4 and libm expf on
    # the tail; the restatement follows the scalar tail.  Floating point: the bar is test-backend-ops' NMSE (1e-7), far from binding
    for got, key in ((O.soft_max(GO["sm_x"]), "sm_plain"), (O.soft_max(GO["sm_x"], GO["sm_mask"], 0.1, 0.0), "sm_mask_scale"),
                     (O.soft_max(GO["sm_x"], GO["sm_mask"], 0.1, 8.0), "sm_alibi"), (O.soft_max(GO["sm_xl"], None, 0.125, 0.0), "sm_outl")):
        assert _nmse(got, GO[key]) <= 1e-12, key
        assert np.abs(got - GO[key]).max() <= 3e-7, key
        assert np.allclose(got.sum(-1), 1.0, atol=1e-5)


def test_oracle_pool_2d_golden_bit_exact():
    for name in ("max_3311", "avg_3322", "avg_2321", "max_1122"):
        a = [int(v) for v in GO["poolargs_" + name]]
        assert np.array_equal(O.pool_2d(GO["pool_x"], *a), GO["pool_" + name]), name


def test_oracle_new_ops_vs_live_reference():
    import pyref
    try:
        R = pyref.Ref("scalar")
    except FileNotFoundError:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(3)
    x = rng.standard_normal((2, 3, 7, 96)).astype(np.float32)
    assert np.array_equal(O.rms_norm(x, 1e-6), pyref.ref_rms_norm(R, x, 1e-6))
    m = rng.standard_normal((7, 96)).astype(np.float32)
    assert _nmse(O.soft_max(x, m, 0.5, 4.0), pyref.ref_soft_max(R, x, m, 0.5, 4.0)) <= 1e-12
    p = rng.standard_normal((1, 2, 9, 13)).astype(np.float32)
    assert np.array_equal(O.pool_2d(p, 1, 3, 2, 2, 1, 1, 0), pyref.ref_pool_2d(R, p, 1, 3, 2, 2, 1, 1, 0))

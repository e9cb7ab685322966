"""GPU test (pytest -m gpu) of b200_mulmat_noalloc, the MUL_MAT entry point that allocates nothing (its q8_1 activations live in a
static device array of B200_NOALLOC_SCRATCH_BYTES, include/ggml_b200_ops.h): called through the C-ABI on the caller's own stream
and checked against the existing entry point b200_mulmat on the same inputs, two calls in a row (the second call finds the first
call's activations in the shared array), and an oversize request must be refused with the sizes in the message.
Collected last (test_zz_*)."""
import ctypes as C

import numpy as np
import pytest

import pyoracle as O
import b200ops as B

pytestmark = pytest.mark.gpu
SCRATCH = 1 << 20           # B200_NOALLOC_SCRATCH_BYTES


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


class Stream:
    """the caller's own stream, made with the CUDA runtime the library itself links (libcudart.so.12 is in the process already)"""

    def __init__(self):
        self.rt = C.CDLL("libcudart.so.12")
        self.s = C.c_void_p()
        assert self.rt.cudaStreamCreate(C.byref(self.s)) == 0

    def sync(self):
        assert self.rt.cudaStreamSynchronize(self.s) == 0

    def close(self):
        self.rt.cudaStreamDestroy(self.s)


@pytest.mark.parametrize("N", [1, 3, 11])
@pytest.mark.parametrize("t", [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K])
def test_noalloc_matches_mulmat_on_two_calls_in_a_row(t, N):
    rng = np.random.default_rng(500 + t + N)
    M, K = 1003, 4096
    raw = O.random_blocks(t, M, K, rng)
    xs = [rng.uniform(-1, 1, (N, K)).astype(np.float32), (rng.standard_normal((N, K)) * 3).astype(np.float32)]
    with B.Session(0) as s:
        w = s.weights(t, raw, K)
        want = [s.mulmat(w, x) for x in xs]                  # the existing entry point
        st = Stream()
        xb = [s.to_device(x) for x in xs]
        db = [s.malloc(4 * N * M) for _ in xs]
        for b in db:
            s.h2d(b, np.full(N * M, np.nan, np.float32))
        s.sync()
        for i in (0, 1):                                     # back to back on one stream: the second call reuses the static scratch
            s.mulmat_noalloc(w, xb[i], db[i], N, stream=st.s.value)
        st.sync()
        for i in (0, 1):
            got = s.to_host(db[i], (N, M), np.float32)
            assert np.isfinite(got).all()
            if N == 1:
                assert np.array_equal(got, want[i]), i       # the same two kernels as b200_mulmat's one-column path
            assert nmse(got, want[i]) <= 1e-9, (i, nmse(got, want[i]))
            assert nmse(got, O.mul_mat(t, raw, xs[i], 1)) <= 1e-9
            assert nmse(got, O.mul_mat(t, raw, xs[i], 0)) <= 5e-4
        s.mulmat_noalloc(w, xb[0], db[1], N, stream=st.s.value)      # and once more: same input, same bits
        st.sync()
        assert np.array_equal(s.to_host(db[1], (N, M), np.float32), s.to_host(db[0], (N, M), np.float32))
        st.close()
        w.free()


def test_noalloc_refuses_what_does_not_fit():
    rng = np.random.default_rng(7)
    M, K = 64, 4096
    per_col = K + K // 32 * 8
    N = SCRATCH // per_col + 1
    raw = O.random_blocks(O.Q4_0, M, K, rng)
    with B.Session(0) as s:
        w = s.weights(O.Q4_0, raw, K)
        xb, db = s.malloc(4 * N * K), s.malloc(4 * N * M)
        with pytest.raises(B.B200Error, match="static array holds"):
            s.mulmat_noalloc(w, xb, db, N)
        n_ok = SCRATCH // per_col - 1                         # just inside the limit: runs
        s.h2d(xb, rng.uniform(-1, 1, (N, K)).astype(np.float32))
        s.sync()                                                  # (the session's stream and the default stream are not ordered)
        s.mulmat_noalloc(w, xb, db, n_ok)
        assert C.CDLL("libcudart.so.12").cudaDeviceSynchronize() == 0
        got = s.to_host(db, (N, M), np.float32)[:n_ok]
        x = s.to_host(xb, (N, K), np.float32)[:n_ok]
        assert nmse(got, O.mul_mat(O.Q4_0, raw, x, 1)) <= 1e-9
        w.free()

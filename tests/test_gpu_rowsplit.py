"""Fused row split over peer memory (b200_graph_compute_rowsplit): needs 2 GPUs in one process.  Each "rank" is a session
on its own device holding its row slice of every matrix; the GEMVs store their slices into both ranks' copies of dst over
NVLink and a flag exchange per op replaces the all-gather.  Checked against the oracle on the full matrices."""
import numpy as np
import pytest

import pyoracle as O
import b200ops as B

pytestmark = pytest.mark.gpu


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.mark.parametrize("W", [2, 4, 8])
@pytest.mark.parametrize("t", [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K])
@pytest.mark.parametrize("use_graph", [0, 1])
def test_rowsplit_ranks(t, use_graph, W):
    """2, 4 and 8 ranks (every rank holds a DIFFERENT row slice: a gather that mixes ranks up cannot pass); five submissions in a row
    walk the two halves of the LL staging area several times"""
    if B.device_count() < W:
        pytest.skip("needs %d GPUs (gpurun --gpus %d)" % (W, W))
    rng = np.random.default_rng(11 + t)
    M = K = 2048
    raw1, raw2, raw3 = (O.random_blocks(t, M, K, rng, 0.05) for _ in range(3))
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    sess = [B.Session(r) for r in range(W)]
    heaps = [B.SymHeap(s, 1 << 20) for s in sess]
    for r in range(W):
        heaps[r].attach_ptrs([hp.base for hp in heaps], list(range(W)), r, W)
    bufs, nodes_per_rank, keep = [], [], []
    for r in range(W):
        s, hp = sess[r], heaps[r]
        xb, y1, y2, z, out = (hp.malloc(4 * M) for _ in range(5))
        s.h2d(xb, x)
        rows = M // W
        w1, w2, w3 = (s.weights(t, raw[r * rows:(r + 1) * rows], K) for raw in (raw1, raw2, raw3))
        keep.append((w1, w2, w3))

        def mm(w, src, dst):
            n = B.Node()
            n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(src.ptr, [K, 1]), B.f32_tensor(dst.ptr, [M, 1], op=B.OP_MUL_MAT)
            return n
        add = B.Node()
        add.src0, add.src1, add.dst = B.f32_tensor(y1.ptr, [M]), B.f32_tensor(xb.ptr, [M]), B.f32_tensor(z.ptr, [M], op=B.OP_ADD)
        # y1 and y2 share the activation vector (one grouped launch, one exchange); z = y1 + x replicated; out = W3 z
        nodes_per_rank.append([mm(w1, xb, y1), mm(w2, xb, y2), add, mm(w3, z, out)])
        bufs.append((y1, y2, z, out))
    for rep in range(5):
        for r in range(W):                       # asynchronous launches: all ranks' kernels must be in flight together
            sess[r].graph_compute_rowsplit(nodes_per_rank[r], use_cuda_graph=use_graph)
        for r in range(W):
            sess[r].sync()
    mmo = lambda raw, v: O.mul_mat(t, raw, v.reshape(1, -1), 1).reshape(-1)
    for r in range(W):
        y1, y2, z, out = (sess[r].to_host(b, (M,), np.float32) for b in bufs[r])
        assert nmse(y1, mmo(raw1, x)) <= 1e-9 and nmse(y2, mmo(raw2, x)) <= 1e-9, r
        assert np.array_equal(z, y1 + x.reshape(-1))
        assert nmse(out, mmo(raw3, z)) <= 1e-9, r
    # all ranks hold identical full vectors
    for r in range(1, W):
        assert np.array_equal(sess[0].to_host(bufs[0][3], (M,), np.float32), sess[r].to_host(bufs[r][3], (M,), np.float32))
    for s in sess:
        s.close()


@pytest.mark.parametrize("W", [2, 4, 8])
@pytest.mark.parametrize("t", [O.Q4_0, O.Q6_K])
@pytest.mark.parametrize("use_graph", [0, 1])
def test_rowsplit_prefill_ranks(monkeypatch, t, use_graph, W):
    """N = 96 activation columns under row split: every rank runs the dense tensor-core GEMM on its row slice, pushes the slice into
    the peer's copy of dst over NVLink, and the flag exchange publishes it; a second MUL_MAT consumes the gathered result"""
    if B.device_count() < W:
        pytest.skip("needs %d GPUs (gpurun --gpus %d)" % (W, W))
    monkeypatch.setenv("B200_GEMM_MIN_N", "16")
    N = 96
    rng = np.random.default_rng(23 + t)
    M = K = 2048
    raw1, raw3 = (O.random_blocks(t, M, K, rng, 0.05) for _ in range(2))
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    sess = [B.Session(r) for r in range(W)]
    heaps = [B.SymHeap(s, 8 << 20) for s in sess]
    for r in range(W):
        heaps[r].attach_ptrs([hp.base for hp in heaps], list(range(W)), r, W)
    bufs, nodes_per_rank, keep = [], [], []
    for r in range(W):
        s, hp = sess[r], heaps[r]
        xb, y1, z, out = (hp.malloc(4 * M * N) for _ in range(4))
        s.h2d(xb, x)
        rows = M // W
        w1, w3 = (s.weights(t, raw[r * rows:(r + 1) * rows], K) for raw in (raw1, raw3))
        keep.append((w1, w3))

        def mm(w, src, dst):
            n = B.Node()
            n.src0, n.src1, n.dst = w.tensor(), B.f32_tensor(src.ptr, [K, N]), B.f32_tensor(dst.ptr, [M, N], op=B.OP_MUL_MAT)
            return n
        add = B.Node()
        add.src0, add.src1, add.dst = B.f32_tensor(y1.ptr, [M, N]), B.f32_tensor(xb.ptr, [M, N]), B.f32_tensor(z.ptr, [M, N], op=B.OP_ADD)
        nodes_per_rank.append([mm(w1, xb, y1), add, mm(w3, z, out)])
        bufs.append((y1, z, out))
    for rep in range(3):
        for r in range(W):
            sess[r].graph_compute_rowsplit(nodes_per_rank[r], use_cuda_graph=use_graph)
        for r in range(W):
            sess[r].sync()
    for r in range(W):
        y1, z, out = (sess[r].to_host(b, (N, M), np.float32) for b in bufs[r])
        assert nmse(y1, O.mul_mat(t, raw1, x, 1)) <= 1e-6, r
        assert np.array_equal(z, y1 + x)
        assert nmse(out, O.mul_mat(t, raw3, z, 1)) <= 1e-6, r
    for r in range(1, W):
        assert np.array_equal(sess[0].to_host(bufs[0][2], (N, M), np.float32), sess[r].to_host(bufs[r][2], (N, M), np.float32))
    for s in sess:
        s.close()


def test_second_device_large_shared_memory_kernels(monkeypatch):
    """kernels that opt in to more than 48 KB of dynamic shared memory (gemv5 tiles of long rows, the tensor-core GEMM) carry that
    attribute per device: a session on device 1 of the same process must work after device 0 has used them"""
    if B.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    monkeypatch.setenv("B200_GEMV5_MIN_MB", "0")
    rng = np.random.default_rng(5)
    M, K = 64, 14336                                  # q4_0 row = 8064 B -> 8-row tile = 64.5 KB
    raw = O.random_blocks(O.Q4_0, M, K, rng)
    x1 = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    xn = rng.uniform(-1, 1, (64, K)).astype(np.float32)
    for dev in (0, 1):
        with B.Session(dev) as s:
            w = s.weights(O.Q4_0, raw, K)
            assert nmse(s.mulmat(w, x1), O.mul_mat(O.Q4_0, raw, x1, 1)) <= 1e-9
            assert nmse(s.mulmat(w, xn), O.mul_mat(O.Q4_0, raw, xn, 1)) <= 1e-6        # 64 columns: the dense fp16 GEMM (operands rounded to fp16)
            w.free()

"""ctypes binding of oracle/liboracle.so (the CPU restatement, oracle/ggml_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product never imports this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K = 0, 2, 8, 9, 12, 14, 15   # enum ggml_type values
TYPE_NAMES = {F32: "f32", Q4_0: "q4_0", Q8_0: "q8_0", Q8_1: "q8_1", Q4_K: "q4_K", Q6_K: "q6_K", Q8_K: "q8_K"}
BLOCK = {F32: (1, 4), Q4_0: (32, 18), Q8_0: (32, 34), Q8_1: (32, 36), Q4_K: (256, 144), Q6_K: (256, 210), Q8_K: (256, 292)}

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-f", os.path.join(HERE, "Makefile")])


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        u8p, f32p, i8p, i32p, i64p = (C.POINTER(t) for t in (C.c_uint8, C.c_float, C.c_int8, C.c_int32, C.c_int64))
        L.oracle_fp16_to_fp32.restype = C.c_float
        L.oracle_fp16_to_fp32.argtypes = [C.c_uint16]
        L.oracle_fp32_to_fp16.restype = C.c_uint16
        L.oracle_fp32_to_fp16.argtypes = [C.c_float]
        L.oracle_row_size.restype = C.c_int64
        L.oracle_row_size.argtypes = [C.c_int, C.c_int64]
        for n in ("q4_0", "q8_0", "q8_1", "q8_K"):
            f = getattr(L, "oracle_quantize_row_" + n)
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_dequantize_row.restype = None
        L.oracle_dequantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        for n in ("oracle_vec_dot_q8_0", "oracle_vec_dot_q8_K", "oracle_vec_dot_q8_1"):
            f = getattr(L, n)
            f.restype = C.c_float
            f.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.oracle_block_dots.restype = C.c_int64
        L.oracle_block_dots.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_mul_mat.restype = C.c_int
        L.oracle_mul_mat.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int]
        L.oracle_add_f32.restype = None
        L.oracle_add_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def row_size(t, k):
    bl, by = BLOCK[t]
    assert k % bl == 0
    return k // bl * by


def quantize(t, x):
    """x: float32 [rows, k] -> uint8 [rows, row_size].  Only q4_0/q8_0 (weights) and q8_0/q8_1/q8_K (activations)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows, k = x.shape
    out = np.zeros((rows, row_size(t, k)), dtype=np.uint8)
    f = getattr(lib(), "oracle_quantize_row_" + TYPE_NAMES[t])
    for r in range(rows):
        f(_p(x[r]), _p(out[r]), k)
    return out


def dequantize(t, w, k):
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rows = w.shape[0]
    out = np.empty((rows, k), dtype=np.float32)
    for r in range(rows):
        lib().oracle_dequantize_row(t, _p(w[r]), _p(out[r]), k)
    return out


def vec_dot(kind, t, k, wrow, arow):
    f = getattr(lib(), "oracle_vec_dot_" + kind)
    return f(t, k, _p(np.ascontiguousarray(wrow)), _p(np.ascontiguousarray(arow)))


def block_dots(t, k, wrow, a_qs):
    n = k // 32 if t != Q6_K else k // 16
    out = np.empty(n, dtype=np.int32)
    a_qs = np.ascontiguousarray(a_qs, dtype=np.int8)
    got = lib().oracle_block_dots(t, k, _p(np.ascontiguousarray(wrow)), _p(a_qs), _p(out))
    assert got == n
    return out


def mul_mat(t, w, x, mode=0):
    """w: uint8 [M, row_size] (or float32 [M,K] for F32), x: float32 [N, K] -> float32 [N, M] (ggml dst: column n contiguous)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w)
    N, K = x.shape
    M = w.shape[0]
    dst = np.empty((N, M), dtype=np.float32)
    rc = lib().oracle_mul_mat(t, _p(w), _p(x), _p(dst), M, N, K, mode)
    assert rc == 0
    return dst


def add_f32(a, b):
    """a, b: float32 arrays in numpy order [ne3, ne2, ne1, ne0] (b broadcast by repetition, ggml rules)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    a4 = a.reshape((1,) * (4 - a.ndim) + a.shape)
    b4 = b.reshape((1,) * (4 - b.ndim) + b.shape)
    ne0 = np.array(a4.shape[::-1], dtype=np.int64)
    ne1 = np.array(b4.shape[::-1], dtype=np.int64)
    out = np.empty_like(a4)
    lib().oracle_add_f32(_p(a4), _p(b4), _p(out), _p(ne0), _p(ne1))
    return out.reshape(a.shape)


def random_blocks(t, rows, k, rng, scale=1.0):
    """Random-init quantised weights made directly in the wire format (valid for every type):
    random quant bytes, fp16 scales log-uniform-ish in a range that keeps |w| ~ O(0.1)."""
    bl, by = BLOCK[t]
    nb = k // bl
    raw = rng.integers(0, 256, size=(rows, nb, by), dtype=np.uint8)

    def put_f16(off, lo, hi, signed=True):
        v = (rng.uniform(lo, hi, size=(rows, nb)) * scale).astype(np.float32)
        if signed:
            v *= rng.choice(np.array([-1.0, 1.0], dtype=np.float32), size=(rows, nb))
        h = v.astype(np.float16).view(np.uint16)
        raw[:, :, off] = (h & 0xFF).astype(np.uint8)
        raw[:, :, off + 1] = (h >> 8).astype(np.uint8)

    if t == Q4_0:
        put_f16(0, 0.002, 0.02)
    elif t == Q8_0:
        put_f16(0, 0.0002, 0.002)
    elif t == Q4_K:
        put_f16(0, 0.00005, 0.0005, signed=False)
        put_f16(2, 0.00005, 0.0005, signed=False)
    elif t == Q6_K:
        put_f16(208, 0.000005, 0.00005)
    else:
        raise ValueError(t)
    return raw.reshape(rows, nb * by)


def rms_norm(x, eps):
    """x: float32 [..., ne00] -> same shape (ggml_compute_forward_rms_norm_f32)"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x)
    lib().oracle_rms_norm_f32(_p(x), _p(y), C.c_int64(x.shape[-1]), C.c_int64(x.size // max(x.shape[-1], 1)), C.c_float(eps))
    return y


def soft_max(x, mask=None, scale=1.0, max_bias=0.0):
    """x: float32 numpy order [ne03, ne02, ne01, ne00]; mask: float32 [>= ne01, ne00] or None (ggml_compute_forward_soft_max_f32)"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    x4 = x.reshape((1,) * (4 - x.ndim) + x.shape)
    y = np.empty_like(x4)
    m = None if mask is None else np.ascontiguousarray(mask, dtype=np.float32)
    ne3, ne2, ne1, ne0 = x4.shape
    lib().oracle_soft_max_f32(_p(x4), _p(m) if m is not None else None, _p(y), C.c_int64(ne0), C.c_int64(ne1), C.c_int64(ne2), C.c_int64(ne3),
                              C.c_float(scale), C.c_float(max_bias))
    return y.reshape(x.shape)


def pool_2d(x, op, k0, k1, s0, s1, p0, p1):
    """x: float32 numpy order [ne3, ne2, ne1, ne0]; op 0 = max, 1 = avg -> [ne3, ne2, py, px] (ggml_compute_forward_pool_2d)"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    ne3, ne2, ne1, ne0 = x.shape
    px = (ne0 + 2 * p0 - k0) // s0 + 1
    py = (ne1 + 2 * p1 - k1) // s1 + 1
    y = np.empty((ne3, ne2, py, px), dtype=np.float32)
    ne = np.array([ne0, ne1, ne2, ne3], dtype=np.int64)
    lib().oracle_pool_2d_f32(_p(x), _p(y), _p(ne), C.c_int64(px), C.c_int64(py), int(op), int(k0), int(k1), int(s0), int(s1), int(p0), int(p1))
    return y

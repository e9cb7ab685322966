"""ctypes view of libb200ops.so (include/ggml_b200_ops.h) for the tests and bench.py.

This is harness plumbing, not the product: the product is the C-ABI library and the ggml backend plugin.
There is no fallback of any kind: if the library is missing or no CUDA device is visible, everything raises.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(HERE, "lib", "libb200ops.so")     # B200_LIB: development builds

F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K = 0, 1, 2, 8, 9, 12, 14
OP_NONE, OP_ADD, OP_RMS_NORM, OP_MUL_MAT, OP_SOFT_MAX, OP_POOL_2D = 0, 2, 23, 27, 43, 53
OP_DUP, OP_MUL, OP_CPY, OP_CONT, OP_GET_ROWS, OP_ROPE, OP_UNARY = 1, 6, 32, 33, 38, 45, 73      # graph_ops.cu (node lists only)
UNARY_SILU, TYPE_I32, TENSOR_DYNAMIC = 10, 26, 0x4
TENSOR_HOST, TENSOR_REPACKED = 1, 2
BLOCK = {F32: (1, 4), Q4_0: (32, 18), Q8_0: (32, 34), Q8_1: (32, 36), Q4_K: (256, 144), Q6_K: (256, 210)}
TYPE_NAMES = {F32: "f32", Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K"}

# every symbol include/ggml_b200_ops.h declares (tests check the library exports all of them)
SYMBOLS = [
    "b200_open", "b200_close", "b200_setclocks", "b200_last_error", "b200_device_count", "b200_device_info",
    "b200_add", "b200_mulmat", "b200_mulmat_noalloc", "b200_rms_norm", "b200_soft_max", "b200_pool_2d", "b200_graph_compute", "b200_graph_compute_keyed", "b200_synchronize", "b200_quantize_q8_1", "b200_block_dots",
    "b200_malloc", "b200_free", "b200_memcpy_h2d", "b200_memcpy_d2h", "b200_memcpy_d2d", "b200_memset",
    "b200_memcpy_h2d_async", "b200_memcpy_d2h_async",
    "b200_host_alloc", "b200_host_free", "b200_weights_upload", "b200_weights_download", "b200_block_elems",
    "b200_block_bytes", "b200_event_create", "b200_event_record", "b200_event_elapsed_ms", "b200_event_destroy", "b200_event_wait", "b200_event_synchronize",
    "b200_launch_count", "b200_flush_l2", "b200_rowsplit_unique_id", "b200_rowsplit_init", "b200_rowsplit_finalize",
    "b200_mulmat_rowsplit", "b200_rowsplit_heap_create", "b200_rowsplit_heap_attach", "b200_rowsplit_heap_attach_ptrs",
    "b200_rowsplit_heap_destroy", "b200_graph_compute_rowsplit", "b200_debug_trace_begin", "b200_debug_trace_read",
]


class Tensor(C.Structure):
    _fields_ = [("type", C.c_int32), ("op", C.c_int32), ("flags", C.c_int32), ("reserved", C.c_int32),
                ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4), ("op_params", C.c_int32 * 16),
                ("data", C.c_void_p), ("data_len", C.c_int64),
                ("w_base", C.c_void_p), ("w_blocks", C.c_int64), ("w_first", C.c_int64)]


class Node(C.Structure):
    _fields_ = [("src0", Tensor), ("src1", Tensor), ("dst", Tensor), ("src2", Tensor)]


class B200Error(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error("libb200ops.so not built (%s): run `make -f ggml-hexagon_b200/Makefile`; there is no fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.b200_last_error.restype = C.c_char_p
        L.b200_launch_count.restype = C.c_int64
        L.b200_launch_count.argtypes = [C.c_uint64]
        L.b200_block_elems.restype = C.c_int64
        L.b200_block_bytes.restype = C.c_int64
        H, VP, I64, I32 = C.c_uint64, C.c_void_p, C.c_int64, C.c_int32
        TP = C.POINTER(Tensor)
        L.b200_open.argtypes = [C.c_char_p, C.POINTER(H)]
        L.b200_close.argtypes = [H]
        L.b200_setclocks.argtypes = [H, I32, I32, I32]
        L.b200_device_info.argtypes = [C.c_int, C.c_char_p, C.c_int, C.POINTER(I64), C.POINTER(I64), C.POINTER(I32)]
        L.b200_add.argtypes = [H, TP, TP, TP]
        L.b200_mulmat.argtypes = [H, TP, TP, TP]
        L.b200_rms_norm.argtypes = [H, TP, TP]
        L.b200_soft_max.argtypes = [H, TP, TP, TP]
        L.b200_pool_2d.argtypes = [H, TP, TP]
        L.b200_mulmat_rowsplit.argtypes = [H, TP, TP, TP]
        L.b200_graph_compute.argtypes = [H, C.POINTER(Node), I32, I32]
        L.b200_graph_compute_keyed.argtypes = [H, C.POINTER(Node), I32, I32, C.c_uint64]
        L.b200_synchronize.argtypes = [H]
        L.b200_quantize_q8_1.argtypes = [H, VP, VP, I64, I64]
        L.b200_block_dots.argtypes = [H, TP, VP, VP]
        L.b200_malloc.argtypes = [H, C.c_size_t, C.POINTER(VP)]
        L.b200_free.argtypes = [H, VP]
        L.b200_memcpy_h2d.argtypes = [H, VP, VP, C.c_size_t]
        L.b200_memcpy_d2h.argtypes = [H, VP, VP, C.c_size_t]
        L.b200_memcpy_h2d_async.argtypes = [H, VP, VP, C.c_size_t]
        L.b200_memcpy_d2h_async.argtypes = [H, VP, VP, C.c_size_t]
        L.b200_memcpy_d2d.argtypes = [H, VP, VP, C.c_size_t]
        L.b200_memset.argtypes = [H, VP, C.c_int, C.c_size_t]
        L.b200_host_alloc.argtypes = [H, C.c_size_t, C.POINTER(VP)]
        L.b200_host_free.argtypes = [H, VP]
        L.b200_weights_upload.argtypes = [H, I32, VP, I64, I64, I64, VP]
        L.b200_weights_download.argtypes = [H, I32, VP, I64, I64, I64, VP]
        L.b200_event_create.argtypes = [H, C.POINTER(VP)]
        L.b200_event_record.argtypes = [H, VP]
        L.b200_event_elapsed_ms.argtypes = [H, VP, VP, C.POINTER(C.c_float)]
        L.b200_event_destroy.argtypes = [H, VP]
        L.b200_event_wait.argtypes = [H, VP]
        L.b200_event_synchronize.argtypes = [H, VP]
        L.b200_flush_l2.argtypes = [H]
        L.b200_rowsplit_unique_id.argtypes = [VP]
        L.b200_rowsplit_init.argtypes = [H, VP, I32, I32]
        L.b200_rowsplit_finalize.argtypes = [H]
        L.b200_rowsplit_heap_create.argtypes = [H, C.c_size_t, C.POINTER(VP), VP]
        L.b200_rowsplit_heap_attach.argtypes = [H, VP, I32, I32]
        L.b200_rowsplit_heap_attach_ptrs.argtypes = [H, C.POINTER(VP), C.POINTER(I32), I32, I32]
        L.b200_rowsplit_heap_destroy.argtypes = [H]
        L.b200_graph_compute_rowsplit.argtypes = [H, C.POINTER(Node), I32, I32]
        L.b200_debug_trace_begin.argtypes = [H, I32]
        L.b200_debug_trace_read.argtypes = [H, VP, I32]
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise B200Error("b200 error %d: %s" % (rc, lib().b200_last_error().decode()))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def row_bytes(t, k):
    bl, by = BLOCK[t]
    assert k % bl == 0
    return k // bl * by


class DevBuf:
    """A device allocation (b200_malloc)."""

    def __init__(self, sess, nbytes):
        self.sess, self.nbytes = sess, int(nbytes)
        p = C.c_void_p()
        _ck(lib().b200_malloc(sess.h, self.nbytes, C.byref(p)))
        self.ptr = p.value

    def free(self):
        if self.ptr:
            _ck(lib().b200_free(self.sess.h, self.ptr))
            self.ptr = None


class HeapBuf:
    """A slice of the symmetric row-split heap (same offset on every rank)."""

    def __init__(self, ptr, nbytes):
        self.ptr, self.nbytes = ptr, nbytes

    def free(self):
        pass


class SymHeap:
    """b200_rowsplit_heap_create + a bump allocator: every rank must allocate the same sizes in the same order."""

    def __init__(self, sess, nbytes):
        self.sess = sess
        base = C.c_void_p()
        self.handle = (C.c_uint8 * 64)()
        _ck(lib().b200_rowsplit_heap_create(sess.h, nbytes, C.byref(base), self.handle))
        self.base, self.size, self.off = base.value, nbytes, 0

    def malloc(self, nbytes):
        nbytes = (int(nbytes) + 255) & ~255
        assert self.off + nbytes <= self.size, "symmetric heap exhausted"
        b = HeapBuf(self.base + self.off, nbytes)
        self.off += nbytes
        return b

    def attach(self, all_handles_bytes, rank, world):
        buf = (C.c_uint8 * (64 * world)).from_buffer_copy(bytes(all_handles_bytes))
        _ck(lib().b200_rowsplit_heap_attach(self.sess.h, buf, rank, world))

    def attach_ptrs(self, bases, devices, rank, world):
        pb = (C.c_void_p * world)(*bases)
        pd = (C.c_int32 * world)(*devices)
        _ck(lib().b200_rowsplit_heap_attach_ptrs(self.sess.h, pb, pd, rank, world))


class Weights:
    """Quantised (REPACKED) or f32 weight matrix resident in HBM: M rows of K."""

    def __init__(self, sess, t, raw, K):
        raw = np.ascontiguousarray(raw)
        self.sess, self.type, self.K, self.M = sess, t, int(K), raw.shape[0]
        self.nbytes = raw.nbytes
        self.buf = DevBuf(sess, raw.nbytes)
        if t == F32:
            assert raw.dtype == np.float32
            _ck(lib().b200_memcpy_h2d(sess.h, self.buf.ptr, _p(raw), raw.nbytes))
            self.blocks = 0
        else:
            assert raw.dtype == np.uint8 and raw.shape[1] == row_bytes(t, K)
            self.blocks = raw.nbytes // BLOCK[t][1]
            _ck(lib().b200_weights_upload(sess.h, t, self.buf.ptr, self.blocks, 0, self.blocks, _p(raw)))

    def tensor(self, row0=0, rows=None):
        rows = self.M - row0 if rows is None else rows
        t = Tensor()
        t.type = self.type
        rb = row_bytes(self.type, self.K)
        t.ne[:] = [self.K, rows, 1, 1]
        t.nb[:] = [BLOCK[self.type][1], rb, rb * rows, rb * rows]
        if self.type == F32:
            t.data = self.buf.ptr + row0 * rb
            t.data_len = rb * rows
        else:
            t.flags = TENSOR_REPACKED
            t.data = self.buf.ptr
            t.data_len = self.nbytes
            t.w_base = self.buf.ptr
            t.w_blocks = self.blocks
            t.w_first = row0 * (self.K // BLOCK[self.type][0])
        return t

    def download(self):
        out = np.empty((self.M, row_bytes(self.type, self.K)), dtype=np.uint8)
        _ck(lib().b200_weights_download(self.sess.h, self.type, self.buf.ptr, self.blocks, 0, self.blocks, _p(out)))
        return out

    def free(self):
        self.buf.free()


def f32_tensor(ptr, ne, nb=None, op=OP_NONE, host=False, data_len=0):
    """ne in ggml order (ne0 fastest)."""
    ne = list(ne) + [1] * (4 - len(ne))
    t = Tensor()
    t.type = F32
    t.op = op
    t.flags = TENSOR_HOST if host else 0
    t.ne[:] = ne
    if nb is None:
        nb = [4, 4 * ne[0], 4 * ne[0] * ne[1], 4 * ne[0] * ne[1] * ne[2]]
    t.nb[:] = nb
    t.data = ptr
    t.data_len = data_len or (4 * ne[0] * ne[1] * ne[2] * ne[3])
    return t


class Session:
    """b200_open / b200_close pair (the seat of ggmlop_dsp_open/close)."""

    def __init__(self, device=0):
        h = C.c_uint64(0)
        _ck(lib().b200_open(("cuda:%d" % device).encode(), C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if self.h is not None:
            _ck(lib().b200_close(self.h))
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # memory
    def malloc(self, nbytes):
        return DevBuf(self, nbytes)

    def to_device(self, arr):
        arr = np.ascontiguousarray(arr)
        b = DevBuf(self, max(arr.nbytes, 16))
        _ck(lib().b200_memcpy_h2d(self.h, b.ptr, _p(arr), arr.nbytes))
        return b

    def h2d(self, buf, arr):
        arr = np.ascontiguousarray(arr)
        _ck(lib().b200_memcpy_h2d(self.h, buf.ptr, _p(arr), arr.nbytes))

    def to_host(self, buf, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        _ck(lib().b200_memcpy_d2h(self.h, _p(out), buf.ptr if hasattr(buf, "ptr") else buf, out.nbytes))
        return out

    def weights(self, t, raw, K):
        return Weights(self, t, raw, K)

    def sync(self):
        _ck(lib().b200_synchronize(self.h))

    # ops
    def mulmat(self, w, x, cols_per_call=None):
        """x: float32 [N, K] host array -> float32 [N, M] (device tensors inside)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        N, K = x.shape
        xb = self.to_device(x)
        db = self.malloc(4 * N * w.M)
        s0, s1, d = w.tensor(), f32_tensor(xb.ptr, [K, N]), f32_tensor(db.ptr, [w.M, N], op=OP_MUL_MAT)
        _ck(lib().b200_mulmat(self.h, C.byref(s0), C.byref(s1), C.byref(d)))
        out = self.to_host(db, (N, w.M), np.float32)
        xb.free()
        db.free()
        return out

    def mulmat_noalloc(self, w, xb, db, N, stream=None):
        """b200_mulmat_noalloc on device buffers the caller made beforehand (xb: f32 [N, K], db: f32 [N, M]); the call itself takes no
        session and allocates nothing; asynchronous on `stream` (a cudaStream_t value; None = the default stream)"""
        s0, s1, d = w.tensor(), f32_tensor(xb.ptr, [w.K, N]), f32_tensor(db.ptr, [w.M, N], op=OP_MUL_MAT)
        L = lib()
        L.b200_mulmat_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _ck(L.b200_mulmat_noalloc(C.byref(s0), C.byref(s1), C.byref(d), C.c_void_p(stream) if stream else None))

    def mulmat_host(self, t, w_raw, x, K):
        """Everything in host memory (FastRPC-style call): returns float32 [N, M]."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        w_raw = np.ascontiguousarray(w_raw)
        N, M = x.shape[0], w_raw.shape[0]
        out = np.empty((N, M), dtype=np.float32)
        s0 = Tensor()
        s0.type = t
        rb = row_bytes(t, K)
        s0.ne[:] = [K, M, 1, 1]
        s0.nb[:] = [BLOCK[t][1], rb, rb * M, rb * M]
        s0.flags = TENSOR_HOST
        s0.data = w_raw.ctypes.data
        s0.data_len = w_raw.nbytes
        s1 = f32_tensor(x.ctypes.data, [K, N], host=True)
        d = f32_tensor(out.ctypes.data, [M, N], op=OP_MUL_MAT, host=True)
        _ck(lib().b200_mulmat(self.h, C.byref(s0), C.byref(s1), C.byref(d)))
        return out

    def add(self, a, b):
        """a, b numpy float32 in numpy order (ggml ne reversed); b broadcast by repetition."""
        a = np.ascontiguousarray(a, dtype=np.float32)
        b = np.ascontiguousarray(b, dtype=np.float32)
        ab, bb = self.to_device(a), self.to_device(b)
        db = self.malloc(max(a.nbytes, 16))
        nea = list(a.shape[::-1])
        neb = list(b.shape[::-1])
        s0, s1, d = f32_tensor(ab.ptr, nea), f32_tensor(bb.ptr, neb), f32_tensor(db.ptr, nea, op=OP_ADD)
        _ck(lib().b200_add(self.h, C.byref(s0), C.byref(s1), C.byref(d)))
        out = self.to_host(db, a.shape, np.float32)
        for x in (ab, bb, db):
            x.free()
        return out

    def rms_norm(self, x, eps):
        """x: float32 numpy order [..., ne00]"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        xb, db = self.to_device(x), self.malloc(max(x.nbytes, 16))
        ne = list(x.shape[::-1])
        s, d = f32_tensor(xb.ptr, ne), f32_tensor(db.ptr, ne, op=OP_RMS_NORM)
        d.op_params[0] = int(np.float32(eps).view(np.int32))
        _ck(lib().b200_rms_norm(self.h, C.byref(s), C.byref(d)))
        out = self.to_host(db, x.shape, np.float32)
        xb.free(); db.free()
        return out

    def soft_max(self, x, mask=None, scale=1.0, max_bias=0.0):
        """x: float32 [ne03, ne02, ne01, ne00] (numpy order); mask: float32 or float16 [>= ne01, ne00] or None"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        xb, db = self.to_device(x), self.malloc(max(x.nbytes, 16))
        ne = list(x.shape[::-1])
        s, d = f32_tensor(xb.ptr, ne), f32_tensor(db.ptr, ne, op=OP_SOFT_MAX)
        d.op_params[0] = int(np.float32(scale).view(np.int32))
        d.op_params[1] = int(np.float32(max_bias).view(np.int32))
        mb, mt = None, None
        if mask is not None:
            mask = np.ascontiguousarray(mask)
            assert mask.dtype in (np.float32, np.float16)
            mb = self.to_device(mask)
            es = mask.dtype.itemsize
            mt = Tensor()
            mt.type = F32 if es == 4 else F16
            mt.ne[:] = [mask.shape[1], mask.shape[0], 1, 1]
            mt.nb[:] = [es, es * mask.shape[1], mask.nbytes, mask.nbytes]
            mt.data, mt.data_len = mb.ptr, mask.nbytes
        _ck(lib().b200_soft_max(self.h, C.byref(s), C.byref(mt) if mt is not None else None, C.byref(d)))
        out = self.to_host(db, x.shape, np.float32)
        for b in (xb, db, mb):
            if b is not None:
                b.free()
        return out

    def pool_2d(self, x, op, k0, k1, s0, s1, p0, p1):
        """x: float32 or float16 [ne3, ne2, ne1, ne0] -> float32 [ne3, ne2, py, px]; op 0 = max, 1 = avg"""
        x = np.ascontiguousarray(x)
        ne3, ne2, ne1, ne0 = x.shape
        px, py = (ne0 + 2 * p0 - k0) // s0 + 1, (ne1 + 2 * p1 - k1) // s1 + 1
        xb, db = self.to_device(x), self.malloc(4 * ne3 * ne2 * px * py + 16)
        es = x.dtype.itemsize
        s = Tensor()
        s.type = F32 if es == 4 else F16
        s.ne[:] = [ne0, ne1, ne2, ne3]
        s.nb[:] = [es, es * ne0, es * ne0 * ne1, es * ne0 * ne1 * ne2]
        s.data, s.data_len = xb.ptr, x.nbytes
        d = f32_tensor(db.ptr, [px, py, ne2, ne3], op=OP_POOL_2D)
        for i, v in enumerate((op, k0, k1, s0, s1, p0, p1)):
            d.op_params[i] = int(v)
        _ck(lib().b200_pool_2d(self.h, C.byref(s), C.byref(d)))
        out = self.to_host(db, (ne3, ne2, py, px), np.float32)
        xb.free(); db.free()
        return out

    def quantize_q8_1(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        xb = self.to_device(x)
        yb = self.malloc(rows * k // 32 * 36)
        _ck(lib().b200_quantize_q8_1(self.h, xb.ptr, yb.ptr, k, rows))
        out = self.to_host(yb, (rows, k // 32 * 36), np.uint8)
        xb.free()
        yb.free()
        return out

    def block_dots(self, w, row, a_qs):
        a_qs = np.ascontiguousarray(a_qs, dtype=np.int8)
        n = w.K // 32 if w.type != Q6_K else w.K // 16
        ab = self.to_device(a_qs)
        ob = self.malloc(4 * n)
        t = w.tensor(row, 1)
        _ck(lib().b200_block_dots(self.h, C.byref(t), ab.ptr, ob.ptr))
        out = self.to_host(ob, (n,), np.int32)
        ab.free()
        ob.free()
        return out

    def graph_compute_rowsplit(self, nodes, use_cuda_graph=True):
        arr = (Node * len(nodes))(*nodes) if not isinstance(nodes, C.Array) else nodes
        _ck(lib().b200_graph_compute_rowsplit(self.h, arr, len(arr), int(use_cuda_graph)))
        return arr

    def graph_compute(self, nodes, use_cuda_graph=True):
        arr = (Node * len(nodes))(*nodes) if not isinstance(nodes, C.Array) else nodes
        _ck(lib().b200_graph_compute(self.h, arr, len(arr), int(use_cuda_graph)))
        return arr

    def graph_compute_keyed(self, nodes, key, use_cuda_graph=True):
        """the caller vouches that `key` names this node list (up to the data pointers of TENSOR_DYNAMIC destinations)"""
        arr = (Node * len(nodes))(*nodes) if not isinstance(nodes, C.Array) else nodes
        _ck(lib().b200_graph_compute_keyed(self.h, arr, len(arr), int(use_cuda_graph), C.c_uint64(key)))
        return arr

    # timing
    def event(self):
        e = C.c_void_p()
        _ck(lib().b200_event_create(self.h, C.byref(e)))
        return e

    def record(self, e):
        _ck(lib().b200_event_record(self.h, e))

    def elapsed_ms(self, a, b):
        ms = C.c_float()
        _ck(lib().b200_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return ms.value

    def launch_count(self):
        return lib().b200_launch_count(self.h)

    def flush_l2(self):
        _ck(lib().b200_flush_l2(self.h))


def device_count():
    return lib().b200_device_count()

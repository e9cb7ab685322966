"""ctypes binding of the UNMODIFIED reference built by oracle/Makefile.ref into oracle/_ref/<variant>/.

TEST INFRASTRUCTURE ONLY (checker for the oracle, golden-vector generator, timed CPU baseline).
Nothing here reads /root/reference at run time; it only loads the prebuilt libraries.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K = 0, 2, 8, 9, 12, 14, 15
_NAMES = {Q4_0: "q4_0", Q8_0: "q8_0", Q8_1: "q8_1", Q4_K: "q4_K", Q6_K: "q6_K", Q8_K: "q8_K"}


def cpu_flags():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def best_variant():
    fl = cpu_flags()
    need512 = {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512_vnni", "avx512vbmi"}
    order = []
    if need512 <= fl:
        order.append("avx512")
    if {"avx2", "fma", "f16c", "bmi2"} <= fl:
        order.append("avx2")
    order.append("scalar")
    for v in order:
        if os.path.exists(os.path.join(HERE, "_ref", v, "libggml-cpu.so")):
            return v
    return None


def ref_dir(variant=None):
    v = variant or best_variant()
    return None if v is None else os.path.join(HERE, "_ref", v)


class Ref:
    def __init__(self, variant=None):
        d = ref_dir(variant)
        if d is None or not os.path.exists(os.path.join(d, "libggml-cpu.so")):
            raise FileNotFoundError("oracle/_ref not built (run: make -f oracle/Makefile.ref VARIANT=...)")
        self.dir = d
        self.variant = os.path.basename(d)
        self.base = C.CDLL(os.path.join(d, "libggml-base.so"), mode=C.RTLD_GLOBAL)
        self.cpu = C.CDLL(os.path.join(d, "libggml-cpu.so"), mode=C.RTLD_GLOBAL)
        self.base.ggml_quantize_chunk.restype = C.c_size_t
        self.base.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
        self.base.ggml_row_size.restype = C.c_size_t
        self.base.ggml_row_size.argtypes = [C.c_int, C.c_int64]
        self.cpu.ggml_cpu_init()

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(C.c_void_p)

    def row_size(self, t, k):
        return self.base.ggml_row_size(t, k)

    def quantize_chunk(self, t, x):
        """ggml_quantize_chunk (ggml.c) — how test-backend-ops makes quantised tensors."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, self.row_size(t, k)), dtype=np.uint8)
        self.base.ggml_quantize_chunk(t, self._p(x), self._p(out), 0, rows, k, None)
        return out

    def quantize_row_ref(self, t, x):
        """quantize_row_<t>_ref (ggml-quants.c)"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, self.row_size(t, k)), dtype=np.uint8)
        f = getattr(self.base, "quantize_row_%s_ref" % _NAMES[t])
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for r in range(rows):
            f(self._p(x[r]), self._p(out[r]), k)
        return out

    def quantize_row_cpu(self, t, x):
        """quantize_row_<t> (ggml-cpu-quants.c; SIMD from_float used by ggml-cpu MUL_MAT)"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, self.row_size(t, k)), dtype=np.uint8)
        f = getattr(self.cpu, "quantize_row_%s" % _NAMES[t])
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for r in range(rows):
            f(self._p(x[r]), self._p(out[r]), k)
        return out

    def dequantize(self, t, w, k):
        w = np.ascontiguousarray(w, dtype=np.uint8)
        out = np.empty((w.shape[0], k), dtype=np.float32)
        f = getattr(self.base, "dequantize_row_%s" % _NAMES[t])
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for r in range(w.shape[0]):
            f(self._p(w[r]), self._p(out[r]), k)
        return out

    def vec_dot(self, t, k, wrow, arow):
        """ggml_vec_dot_<t>_<q8_0|q8_K> (ggml-cpu-quants.c)"""
        at = "q8_K" if t in (Q4_K, Q6_K) else "q8_0"
        f = getattr(self.cpu, "ggml_vec_dot_%s_%s" % (_NAMES[t], at))
        f.restype = None
        f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int]
        s = C.c_float(0)
        f(k, C.byref(s), 0, self._p(np.ascontiguousarray(wrow)), 0, self._p(np.ascontiguousarray(arow)), 0, 1)
        return s.value

    # ---- full graph through ggml-cpu (ggml_compute_forward_mul_mat / add), the parity oracle of north_star
    class _InitParams(C.Structure):
        _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]

    def _ctx(self, nbytes):
        b = self.base
        b.ggml_init.restype = C.c_void_p
        b.ggml_init.argtypes = [Ref._InitParams]
        for n in ("ggml_new_tensor_2d",):
            getattr(b, n).restype = C.c_void_p
        b.ggml_new_tensor_2d.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64]
        b.ggml_new_tensor_4d.restype = C.c_void_p
        b.ggml_new_tensor_4d.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64]
        b.ggml_mul_mat.restype = C.c_void_p
        b.ggml_mul_mat.argtypes = [C.c_void_p] * 3
        b.ggml_add.restype = C.c_void_p
        b.ggml_add.argtypes = [C.c_void_p] * 3
        b.ggml_new_graph.restype = C.c_void_p
        b.ggml_new_graph.argtypes = [C.c_void_p]
        b.ggml_build_forward_expand.restype = None
        b.ggml_build_forward_expand.argtypes = [C.c_void_p, C.c_void_p]
        b.ggml_get_data.restype = C.c_void_p
        b.ggml_get_data.argtypes = [C.c_void_p]
        b.ggml_nbytes.restype = C.c_size_t
        b.ggml_nbytes.argtypes = [C.c_void_p]
        b.ggml_free.argtypes = [C.c_void_p]
        self.cpu.ggml_graph_compute_with_ctx.restype = C.c_int
        self.cpu.ggml_graph_compute_with_ctx.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        return b.ggml_init(Ref._InitParams(nbytes, None, False))

    def mul_mat(self, t, w, x, n_threads=1):
        """dst[N, M] = ggml_mul_mat(W[type t, M rows of K], X[f32, N rows of K]) on ggml-cpu."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        w = np.ascontiguousarray(w)
        N, K = x.shape
        M = w.shape[0]
        ctx = self._ctx(w.nbytes + x.nbytes + 4 * N * M + 2 * (x.nbytes + 4096) * max(1, n_threads) + (64 << 20))
        b = self.base
        tw = b.ggml_new_tensor_2d(ctx, t, K, M)
        tx = b.ggml_new_tensor_2d(ctx, F32, K, N)
        assert b.ggml_nbytes(tw) == w.nbytes, (b.ggml_nbytes(tw), w.nbytes)
        C.memmove(b.ggml_get_data(tw), self._p(w), w.nbytes)
        C.memmove(b.ggml_get_data(tx), self._p(x), x.nbytes)
        td = b.ggml_mul_mat(ctx, tw, tx)
        g = b.ggml_new_graph(ctx)
        b.ggml_build_forward_expand(g, td)
        rc = self.cpu.ggml_graph_compute_with_ctx(ctx, g, n_threads)
        assert rc == 0
        out = np.empty((N, M), dtype=np.float32)
        C.memmove(self._p(out), b.ggml_get_data(td), out.nbytes)
        b.ggml_free(ctx)
        return out

    def add(self, a, bb, n_threads=1):
        a = np.ascontiguousarray(a, dtype=np.float32)
        bb = np.ascontiguousarray(bb, dtype=np.float32)
        a4 = a.reshape((1,) * (4 - a.ndim) + a.shape)
        b4 = bb.reshape((1,) * (4 - bb.ndim) + bb.shape)
        ctx = self._ctx(2 * a.nbytes + bb.nbytes + (16 << 20))
        b = self.base
        ta = b.ggml_new_tensor_4d(ctx, F32, *a4.shape[::-1])
        tb = b.ggml_new_tensor_4d(ctx, F32, *b4.shape[::-1])
        C.memmove(b.ggml_get_data(ta), self._p(a4), a4.nbytes)
        C.memmove(b.ggml_get_data(tb), self._p(b4), b4.nbytes)
        td = b.ggml_add(ctx, ta, tb)
        g = b.ggml_new_graph(ctx)
        b.ggml_build_forward_expand(g, td)
        assert self.cpu.ggml_graph_compute_with_ctx(ctx, g, n_threads) == 0
        out = np.empty_like(a4)
        C.memmove(self._p(out), b.ggml_get_data(td), out.nbytes)
        b.ggml_free(ctx)
        return out.reshape(a.shape)


def _unary_graph(ref, x4, build):
    """run one op on ggml-cpu: x4 float32 [ne3, ne2, ne1, ne0] -> output array of the op's shape"""
    b = ref.base
    ctx = ref._ctx(8 * x4.nbytes + (64 << 20))
    ta = b.ggml_new_tensor_4d(ctx, F32, *x4.shape[::-1])
    C.memmove(b.ggml_get_data(ta), x4.ctypes.data, x4.nbytes)
    td = build(ctx, ta)
    g = b.ggml_new_graph(ctx)
    b.ggml_build_forward_expand(g, td)
    assert ref.cpu.ggml_graph_compute_with_ctx(ctx, g, 1) == 0
    n = b.ggml_nbytes(td) // 4
    out = np.empty(n, dtype=np.float32)
    C.memmove(out.ctypes.data, b.ggml_get_data(td), 4 * n)
    b.ggml_free(ctx)
    return out


def ref_rms_norm(ref, x, eps):
    x = np.ascontiguousarray(x, dtype=np.float32)
    x4 = x.reshape((1,) * (4 - x.ndim) + x.shape)
    b = ref.base
    b.ggml_rms_norm.restype = C.c_void_p
    b.ggml_rms_norm.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
    return _unary_graph(ref, x4, lambda ctx, ta: b.ggml_rms_norm(ctx, ta, eps)).reshape(x.shape)


def ref_soft_max(ref, x, mask=None, scale=1.0, max_bias=0.0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    x4 = x.reshape((1,) * (4 - x.ndim) + x.shape)
    b = ref.base
    b.ggml_soft_max_ext.restype = C.c_void_p
    b.ggml_soft_max_ext.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float]

    def build(ctx, ta):
        tm = None
        if mask is not None:
            m = np.ascontiguousarray(mask, dtype=np.float32)
            tm = b.ggml_new_tensor_2d(ctx, F32, m.shape[1], m.shape[0])
            C.memmove(b.ggml_get_data(tm), m.ctypes.data, m.nbytes)
        return b.ggml_soft_max_ext(ctx, ta, tm, scale, max_bias)
    return _unary_graph(ref, x4, build).reshape(x.shape)


def ref_pool_2d(ref, x, op, k0, k1, s0, s1, p0, p1):
    x = np.ascontiguousarray(x, dtype=np.float32)
    b = ref.base
    b.ggml_pool_2d.restype = C.c_void_p
    b.ggml_pool_2d.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float]
    ne3, ne2, ne1, ne0 = x.shape
    px, py = (ne0 + 2 * p0 - k0) // s0 + 1, (ne1 + 2 * p1 - k1) // s1 + 1
    return _unary_graph(ref, x, lambda ctx, ta: b.ggml_pool_2d(ctx, ta, op, k0, k1, s0, s1, float(p0), float(p1))).reshape(ne3, ne2, py, px)


def bench_decode(ref, layer_shapes, out_shape, n_layers_total, n_layers_sample, wtype, n_threads, budget_s=15.0, seed=7):
    """Timed CPU baseline: the reference's ggml-cpu runs the decode MUL_MAT/ADD node list (batch 1) of
    `n_layers_sample` layers (+ the output matrix), all `n_threads` host threads, repeated until `budget_s`.
    Returns (seconds per FULL token extrapolated to n_layers_total, sample description)."""
    import time
    rng = np.random.default_rng(seed)
    b, cpu = ref.base, ref.cpu
    tot = 0
    for _, M, K in layer_shapes:
        tot += ref.row_size(wtype, K) * M
    tot = tot * n_layers_sample + ref.row_size(wtype, out_shape[2]) * out_shape[1]
    ctx = ref._ctx(tot + (512 << 20))
    b.ggml_new_tensor_1d.restype = C.c_void_p
    b.ggml_new_tensor_1d.argtypes = [C.c_void_p, C.c_int, C.c_int64]

    def fill_w(t, M, K):
        n = ref.row_size(wtype, K) * M
        raw = rng.integers(0, 256, size=n, dtype=np.uint8)
        # keep scales finite: overwrite with a small quantised random matrix pattern on the first rows only is overkill;
        # force the fp16 scale bytes of q4_0/q8_0 blocks to ~0.002
        if wtype in (Q4_0, Q8_0):
            bs = 18 if wtype == Q4_0 else 34
            v = raw.reshape(-1, bs)
            v[:, 0] = 0x19
            v[:, 1] = 0x18          # 0x1819 = 0.00199 in fp16
        C.memmove(b.ggml_get_data(t), raw.ctypes.data, n)

    def vec(n):
        t = b.ggml_new_tensor_1d(ctx, F32, n)
        x = rng.uniform(-1, 1, n).astype(np.float32)
        C.memmove(b.ggml_get_data(t), x.ctypes.data, 4 * n)
        return t

    dims = dict((n, (M, K)) for n, M, K in layer_shapes)
    h, f = dims["wq"][1], dims["down"][1]
    x = vec(h)
    g = b.ggml_new_graph(ctx)
    last = None
    for _ in range(n_layers_sample):
        ws = {}
        for n, M, K in layer_shapes:
            ws[n] = b.ggml_new_tensor_2d(ctx, wtype, K, M)
            fill_w(ws[n], M, K)
        q = b.ggml_mul_mat(ctx, ws["wq"], x)
        k = b.ggml_mul_mat(ctx, ws["wk"], x)
        v = b.ggml_mul_mat(ctx, ws["wv"], x)
        o = b.ggml_mul_mat(ctx, ws["wo"], q)
        hres = b.ggml_add(ctx, x, o)
        gt = b.ggml_mul_mat(ctx, ws["gate"], hres)
        up = b.ggml_mul_mat(ctx, ws["up"], hres)
        dn = b.ggml_mul_mat(ctx, ws["down"], gt)
        x = b.ggml_add(ctx, hres, dn)
        for t in (k, v, up, x):
            b.ggml_build_forward_expand(g, t)
        last = x
    g_out = b.ggml_new_graph(ctx)
    wo = b.ggml_new_tensor_2d(ctx, wtype, out_shape[2], out_shape[1])
    fill_w(wo, out_shape[1], out_shape[2])
    logits = b.ggml_mul_mat(ctx, wo, vec(h))
    b.ggml_build_forward_expand(g_out, logits)
    # ggml_graph_compute_with_ctx allocates its work buffer inside ctx at every call -> plan once instead
    class CPlan(C.Structure):
        _fields_ = [("work_size", C.c_size_t), ("work_data", C.c_void_p), ("n_threads", C.c_int),
                    ("threadpool", C.c_void_p), ("abort_callback", C.c_void_p), ("abort_callback_data", C.c_void_p)]
    cpu.ggml_graph_plan.restype = CPlan
    cpu.ggml_graph_plan.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    cpu.ggml_graph_compute.restype = C.c_int
    cpu.ggml_graph_compute.argtypes = [C.c_void_p, C.POINTER(CPlan)]
    # ONE thread pool for the whole measurement (with threadpool = NULL ggml-cpu creates and joins n_threads threads inside every
    # ggml_graph_compute call), as llama.cpp keeps one per context
    class TPParams(C.Structure):
        _fields_ = [("cpumask", C.c_bool * 512), ("n_threads", C.c_int), ("prio", C.c_int), ("poll", C.c_uint32),
                    ("strict_cpu", C.c_bool), ("paused", C.c_bool)]
    b.ggml_threadpool_params_init.restype = None
    b.ggml_threadpool_params_init.argtypes = [C.POINTER(TPParams), C.c_int]
    cpu.ggml_threadpool_new.restype = C.c_void_p
    cpu.ggml_threadpool_new.argtypes = [C.POINTER(TPParams)]
    cpu.ggml_threadpool_free.restype = None
    cpu.ggml_threadpool_free.argtypes = [C.c_void_p]
    tpp = TPParams()
    b.ggml_threadpool_params_init(C.byref(tpp), n_threads)
    pool = cpu.ggml_threadpool_new(C.byref(tpp))
    assert pool
    plans = []
    for gr in (g, g_out):
        p = cpu.ggml_graph_plan(gr, n_threads, pool)
        buf = np.empty(max(p.work_size, 16), dtype=np.uint8)
        p.work_data = buf.ctypes.data
        plans.append((gr, p, buf))

    def run():
        for gr, p, _ in plans:
            assert cpu.ggml_graph_compute(gr, C.byref(p)) == 0
    run()  # warm-up
    t_layers = t_out = 0.0
    reps = 0
    t_start = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        assert cpu.ggml_graph_compute(plans[0][0], C.byref(plans[0][1])) == 0
        t1 = time.perf_counter()
        assert cpu.ggml_graph_compute(plans[1][0], C.byref(plans[1][1])) == 0
        t2 = time.perf_counter()
        t_layers += t1 - t0
        t_out += t2 - t1
        reps += 1
        if time.perf_counter() - t_start > budget_s or reps >= 200:
            break
    cpu.ggml_threadpool_free(pool)
    b.ggml_free(ctx)
    per_token = t_layers / reps * (n_layers_total / n_layers_sample) + t_out / reps
    desc = ("%d of %d layers (7 MUL_MAT + 2 ADD each) + output matrix, one persistent ggml threadpool, %d reps, extrapolated x%.1f on the layer part"
            % (n_layers_sample, n_layers_total, reps, n_layers_total / n_layers_sample))
    return per_token, desc

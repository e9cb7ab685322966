"""ctypes view of the REFERENCE's public ggml C API (ggml.h / ggml-backend.h / ggml-alloc.h), as built unmodified into
oracle/_ref/<variant>/libggml*.so by oracle/Makefile.ref.  Test infrastructure: it lets the tests drive libggml-b200.so exactly
as a ggml host does (ggml_backend_load -> registry -> device -> buffer -> graph_compute) and run the same graph on the
reference's own CPU backend for comparison.  Nothing under ggml-hexagon_b200/ imports this."""
import ctypes as C
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "ggml-hexagon_b200", "lib", "libggml-b200.so")

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K = 0, 1, 2, 8, 12, 14
DEV_CPU, DEV_GPU = 0, 1
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q6_K: (256, 210)}


def ref_dir(variant=None):
    for v in ([variant] if variant else []) + [os.environ.get("REF_VARIANT", ""), "avx2", "avx512", "scalar"]:
        d = os.path.join(ROOT, "oracle", "_ref", v) if v else None
        if d and os.path.exists(os.path.join(d, "libggml.so")):
            return d
    return None


class InitParams(C.Structure):
    _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]


class DevCaps(C.Structure):
    _fields_ = [("async_", C.c_bool), ("host_buffer", C.c_bool), ("buffer_from_host_ptr", C.c_bool), ("events", C.c_bool)]


class DevProps(C.Structure):
    _fields_ = [("name", C.c_char_p), ("description", C.c_char_p), ("memory_free", C.c_size_t), ("memory_total", C.c_size_t),
                ("type", C.c_int), ("caps", DevCaps)]


_G = None


def lib(variant=None):
    """libggml.so (registry + CPU backend) with libggml-base.so's symbols; raises FileNotFoundError when oracle/_ref is absent"""
    global _G
    if _G is not None:
        return _G
    d = ref_dir(variant)
    if d is None:
        raise FileNotFoundError("oracle/_ref/<variant>/libggml.so not built (make -f oracle/Makefile.ref)")
    C.CDLL(os.path.join(d, "libggml-base.so"), mode=C.RTLD_GLOBAL)
    C.CDLL(os.path.join(d, "libggml-cpu.so"), mode=C.RTLD_GLOBAL)
    G = C.CDLL(os.path.join(d, "libggml.so"), mode=C.RTLD_GLOBAL)
    VP, SZ, I64 = C.c_void_p, C.c_size_t, C.c_int64
    sig = {
        "ggml_init": (VP, [InitParams]), "ggml_free": (None, [VP]),
        "ggml_tensor_overhead": (SZ, []), "ggml_graph_overhead": (SZ, []), "ggml_nbytes": (SZ, [VP]),
        "ggml_new_tensor_1d": (VP, [VP, C.c_int, I64]), "ggml_new_tensor_2d": (VP, [VP, C.c_int, I64, I64]),
        "ggml_new_tensor_3d": (VP, [VP, C.c_int, I64, I64, I64]), "ggml_new_tensor_4d": (VP, [VP, C.c_int, I64, I64, I64, I64]),
        "ggml_mul_mat": (VP, [VP, VP, VP]), "ggml_add": (VP, [VP, VP, VP]),
        "ggml_rms_norm": (VP, [VP, VP, C.c_float]), "ggml_soft_max": (VP, [VP, VP]),
        "ggml_soft_max_ext": (VP, [VP, VP, VP, C.c_float, C.c_float]),
        "ggml_view_1d": (VP, [VP, VP, I64, SZ]), "ggml_view_2d": (VP, [VP, VP, I64, I64, SZ, SZ]),
        "ggml_view_3d": (VP, [VP, VP, I64, I64, I64, SZ, SZ, SZ]),
        "ggml_mul": (VP, [VP, VP, VP]), "ggml_silu": (VP, [VP, VP]), "ggml_cpy": (VP, [VP, VP, VP]), "ggml_cont": (VP, [VP, VP]),
        "ggml_cont_2d": (VP, [VP, VP, I64, I64]), "ggml_reshape_2d": (VP, [VP, VP, I64, I64]), "ggml_reshape_3d": (VP, [VP, VP, I64, I64, I64]),
        "ggml_permute": (VP, [VP, VP, C.c_int, C.c_int, C.c_int, C.c_int]), "ggml_transpose": (VP, [VP, VP]),
        "ggml_rope_ext": (VP, [VP, VP, VP, VP, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]),
        "ggml_get_rows": (VP, [VP, VP, VP]), "ggml_element_size": (SZ, [VP]),
        "ggml_new_graph": (VP, [VP]), "ggml_build_forward_expand": (None, [VP, VP]),
        "ggml_backend_load": (VP, [C.c_char_p]), "ggml_backend_reg_dev_count": (SZ, [VP]), "ggml_backend_reg_dev_get": (VP, [VP, SZ]),
        "ggml_backend_reg_name": (C.c_char_p, [VP]), "ggml_backend_reg_get_proc_address": (VP, [VP, C.c_char_p]),
        "ggml_backend_dev_name": (C.c_char_p, [VP]), "ggml_backend_dev_description": (C.c_char_p, [VP]),
        "ggml_backend_dev_get_props": (None, [VP, C.POINTER(DevProps)]),
        "ggml_backend_dev_init": (VP, [VP, C.c_char_p]), "ggml_backend_dev_supports_op": (C.c_bool, [VP, VP]),
        "ggml_backend_dev_offload_op": (C.c_bool, [VP, VP]),
        "ggml_backend_dev_buffer_type": (VP, [VP]), "ggml_backend_dev_supports_buft": (C.c_bool, [VP, VP]),
        "ggml_backend_init_by_type": (VP, [C.c_int, C.c_char_p]),
        "ggml_backend_alloc_ctx_tensors": (VP, [VP, VP]), "ggml_backend_buffer_free": (None, [VP]),
        "ggml_backend_buffer_clear": (None, [VP, C.c_uint8]),
        "ggml_backend_tensor_set": (None, [VP, VP, SZ, SZ]), "ggml_backend_tensor_get": (None, [VP, VP, SZ, SZ]),
        "ggml_backend_tensor_set_async": (None, [VP, VP, VP, SZ, SZ]), "ggml_backend_tensor_get_async": (None, [VP, VP, VP, SZ, SZ]),
        "ggml_backend_tensor_copy": (None, [VP, VP]), "ggml_backend_tensor_copy_async": (None, [VP, VP, VP, VP]),
        "ggml_backend_tensor_memset": (None, [VP, C.c_uint8, SZ, SZ]),
        "ggml_backend_graph_compute": (C.c_int, [VP, VP]), "ggml_backend_graph_compute_async": (C.c_int, [VP, VP]),
        "ggml_backend_synchronize": (None, [VP]), "ggml_backend_free": (None, [VP]), "ggml_backend_name": (C.c_char_p, [VP]),
        "ggml_backend_event_new": (VP, [VP]), "ggml_backend_event_free": (None, [VP]), "ggml_backend_event_record": (None, [VP, VP]),
        "ggml_backend_event_synchronize": (None, [VP]), "ggml_backend_event_wait": (None, [VP, VP]),
        "ggml_backend_get_device": (VP, [VP]),
        "ggml_backend_sched_new": (VP, [C.POINTER(VP), C.POINTER(VP), C.c_int, SZ, C.c_bool]), "ggml_backend_sched_free": (None, [VP]),
        "ggml_backend_sched_alloc_graph": (C.c_bool, [VP, VP]), "ggml_backend_sched_graph_compute": (C.c_int, [VP, VP]),
        "ggml_backend_sched_get_n_splits": (C.c_int, [VP]), "ggml_backend_sched_reset": (None, [VP]),
        "ggml_backend_sched_get_tensor_backend": (VP, [VP, VP]),
        "ggml_quantize_chunk": (SZ, [C.c_int, VP, VP, I64, I64, I64, VP]),
    }
    for name, (res, args) in sig.items():
        f = getattr(G, name)
        f.restype, f.argtypes = res, args
    _G = G
    return G


class Ctx:
    """a no_alloc ggml context big enough for n tensors and one graph"""

    def __init__(self, n_tensors=64):
        G = lib()
        p = InitParams(G.ggml_tensor_overhead() * (n_tensors + 8) + G.ggml_graph_overhead() + (1 << 16), None, True)
        self.G, self.ctx = G, G.ggml_init(p)
        assert self.ctx

    def tensor(self, t, *ne):
        G = self.G
        f = [None, G.ggml_new_tensor_1d, G.ggml_new_tensor_2d, G.ggml_new_tensor_3d, G.ggml_new_tensor_4d][len(ne)]
        r = f(self.ctx, t, *ne)
        assert r
        return r

    def free(self):
        if self.ctx:
            self.G.ggml_free(self.ctx)
            self.ctx = None


def load_plugin():
    """ggml_backend_load(libggml-b200.so) -> (reg, [devices]); the registry keeps it loaded for the life of the process"""
    G = lib()
    if not os.path.exists(PLUGIN):
        raise FileNotFoundError(PLUGIN)
    reg = G.ggml_backend_load(PLUGIN.encode())
    if not reg:
        raise RuntimeError("ggml_backend_load(%s) declined (no CUDA device?)" % PLUGIN)
    return reg, [G.ggml_backend_reg_dev_get(reg, i) for i in range(G.ggml_backend_reg_dev_count(reg))]

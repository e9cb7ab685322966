// ggml-b200.cpp — the ggml backend plugin (libggml-b200.so): drop-in for libggml-hexagon.so in cDSP mode.
//
// Mirrors section 8 of the reference backend (ggml/src/ggml-hexagon/ggml-hexagon.cpp:5065-6110): registry entry,
// device, buffer type, buffer, backend, supports_op, graph_compute.  Everything below the ggml structs goes
// through the C-ABI op layer (include/ggml_b200_ops.h), which takes the seat of the FastRPC calls
// ggmlop_dsp_open/close/add/mulmat (ggml-hexagon.cpp:4975-5060).  No CPU compute path: an op we do not
// accelerate is declined in supports_op, never emulated.
//
// Compiled against the reference's own headers (ggml.h, ggml-backend.h, ggml-backend-impl.h); nothing of the
// reference is copied into this repository.
#include "ggml-b200.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#include "ggml_b200_ops.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

// everything goes through ggml's logger (GGML_LOG_*, ggml-impl.h), so a host's ggml_log_set callback sees the backend
#define B200_LOG(...)  do { if (getenv("GGML_B200_VERBOSE")) { GGML_LOG_INFO("[ggml-b200] " __VA_ARGS__); GGML_LOG_INFO("\n"); } } while (0)
#define B200_ERR(...)  do { GGML_LOG_ERROR("[ggml-b200] " __VA_ARGS__); GGML_LOG_ERROR("\n"); } while (0)

// ------------------------------------------------------------------------------------------------ contexts
// ONE op-layer session (= one CUDA stream, one scratch set, one CUDA-graph cache) per device, shared by the device's buffers and
// by every backend instance on it: buffer copies, uploads and graph_compute are ordered on that stream, and anything the op layer
// caches per weight address (graphs, scratch) is invalidated by the same session that rewrites the weights.  `mu` serialises the
// entry points (callers are single-threaded per backend, but buffers of one device may be touched from several threads).
struct b200_device_ctx {
    int         ordinal = 0;
    std::string name;            // "B200_<ordinal>"   (ggml_backend_hexagon_get_devname)
    std::string desc;            // CUDA device name
    b200_handle h = 0;
    std::mutex  mu;
    ggml_backend_buffer_type buft;
};
// a node list already marshalled for one ggml_cgraph: what ggmlhexagon_compute rebuilds per op and per call (string op key + three
// dsptensor structs, ggml-hexagon.cpp:4975-5049) is built once per distinct graph here.  `fp` is the fingerprint of everything the
// marshalling read (tensor addresses, data addresses, types, shapes, strides, op parameters); a call whose graph has the same
// fingerprint reuses `nodes` and hands the op layer `key`, so neither side re-hashes 160 KB of descriptors per token.
struct b200_graph_entry {
    std::vector<uint64_t>  fp;
    std::vector<b200_node> nodes;
    uint64_t               key = 0;
    uint64_t               last_use = 0;
    std::vector<std::pair<int, int>> dyn;   // (index in nodes, index in the cgraph) of the nodes flagged B200_TENSOR_DYNAMIC
    bool                   trust = false;   // capture a new list of this cgraph on its FIRST sighting: large lists (a whole model's layer stack) start out trusted,
                                            // a list whose predecessor was replayed several times (a decode loop whose cache view just grew by one block) earns it,
    uint32_t               misses = 0;      // and four lists in a row that were never seen twice lose it (capture + instantiate costs several launches' worth)
    uint32_t               seen = 0;        // calls with this fingerprint: a CUDA graph is captured from the second one on, so node lists that
                                            // change every call (a decode step whose KV-cache views move with the position) are launched directly
};
struct b200_backend_ctx {
    b200_device_ctx * dev = nullptr;
    std::string       name;
    std::unordered_map<const ggml_cgraph *, b200_graph_entry> graphs;
    std::vector<uint64_t> fp;        // scratch
    uint64_t          tick = 0;
};
struct b200_buffer_ctx {
    b200_device_ctx * dev = nullptr;
    void *            base = nullptr;
    size_t            size = 0;
};

static b200_handle dev_session(b200_device_ctx * d) {          // call with d->mu held
    if (!d->h) {
        char uri[32]; snprintf(uri, sizeof(uri), "cuda:%d", d->ordinal);
        if (b200_open(uri, &d->h) != 0) { B200_ERR("%s", b200_last_error()); d->h = 0; }
    }
    return d->h;
}
#define DEV_LOCK(d) std::lock_guard<std::mutex> lk_((d)->mu)

static bool is_qweight(enum ggml_type t) {
    return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q6_K;
}
static const ggml_tensor * root_of(const ggml_tensor * t) { return t->view_src ? t->view_src : t; }

// ------------------------------------------------------------------------------------------------ tensor -> C-ABI descriptor
// (what ggmlhexagon_compute does when it fills struct dsptensor, ggml-hexagon.cpp:5003-5049; 64-bit ne/nb here)
static bool to_b200(const ggml_tensor * t, b200_tensor * o, int op) {
    memset(o, 0, sizeof(*o));
    o->type = (int32_t) t->type; o->op = op;
    for (int i = 0; i < 4; i++) { o->ne[i] = t->ne[i]; o->nb[i] = (int64_t) t->nb[i]; }
    memcpy(o->op_params, t->op_params, sizeof(o->op_params));
    o->data = t->data;
    o->data_len = (int64_t) ggml_nbytes(t);
    if (is_qweight(t->type)) {
        // quantised tensors live in HBM in the SoA layout over their ROOT tensor's region (see buffer set_tensor)
        const ggml_tensor * r = root_of(t);
        const int64_t bb = (int64_t) ggml_type_size(t->type);
        const int64_t off = (const char *) t->data - (const char *) r->data;
        if (off < 0 || off % bb) return false;
        o->flags = B200_TENSOR_REPACKED;
        o->w_base = r->data; o->w_blocks = (int64_t) ggml_nbytes(r) / bb; o->w_first = off / bb;
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ supports_op
// Reference predicate: ggmlhexagon_can_handle_op_through_cdsp (ggml-hexagon.cpp:5065-5115): NONE -> true;
// ADD f32; MUL_MAT with f32 src1/dst and f32 or quantised src0.  We widen ADD to ggml's repeat-broadcast
// (binary-ops.cpp:50-117) and the quantised set to q4_0/q8_0/q4_K/q6_K, and decline layouts the kernels
// do not take (the caller then keeps the node on another backend; test-backend-ops prints "not supported").
static bool b200_supports_mul_mat(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0]; const ggml_tensor * b = op->src[1];
    if (!a || !b) return false;
    if (b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
    if (a->type == GGML_TYPE_F16) {          // attention: K / V cache views (graph_ops.cu); K contiguous on both sides, other strides free
        if (getenv("B200_GRAPH_OPS") && atoi(getenv("B200_GRAPH_OPS")) == 0) return false;
        if (a->nb[0] != sizeof(ggml_fp16_t) || b->nb[0] != sizeof(float) || op->nb[0] != sizeof(float)) return false;
        if (a->ne[2] <= 0 || a->ne[3] <= 0 || b->ne[2] % a->ne[2] || b->ne[3] % a->ne[3]) return false;
        return b->ne[2] * b->ne[3] <= 65535;
    }
    if (!(a->type == GGML_TYPE_F32 || is_qweight(a->type))) return false;
    if (b->nb[0] != sizeof(float) || op->nb[0] != sizeof(float)) return false;
    if (a->ne[2] <= 0 || a->ne[3] <= 0 || b->ne[2] % a->ne[2] || b->ne[3] % a->ne[3]) return false;
    if (is_qweight(a->type)) {
        const size_t bb = ggml_type_size(a->type);
        if (a->ne[0] % 32) return false;
        if (a->nb[0] != bb || a->nb[1] != ggml_row_size(a->type, a->ne[0])) return false;      // contiguous rows
        if (a->nb[2] % bb || a->nb[3] % bb) return false;
        if ((b->nb[1] | b->nb[2] | b->nb[3]) & 15) return false;                                // 128-bit activation loads
        if (b->data && ((uintptr_t) b->data & 15)) return false;                                // (views at odd offsets, once allocated)
        const ggml_tensor * r = root_of(a);
        if (a->data && r->data && (((const char *) a->data - (const char *) r->data) % (ptrdiff_t) bb)) return false;
    } else {
        if (a->nb[0] != sizeof(float)) return false;
    }
    return true;
}
static bool b200_supports_add(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0]; const ggml_tensor * b = op->src[1];
    if (!a || !b) return false;
    if (a->type != GGML_TYPE_F32 || b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
    if (!ggml_are_same_shape(a, op) || !ggml_can_repeat(b, a)) return false;
    return true;
}
// The three ops the hexagon caps table lists next to ADD / MUL_MAT without a kernel behind them (ggml-hexagon.cpp:619, 639, 649;
// empty stubs ggmlop_dsp_rmsnorm / _softmax / _pool2d, kernels/ggml-dsp.c:1353-1372): admitted here in the layouts ggml-cpu itself
// asserts (ggml-cpu.c:6260-6262, 8268-8269, 9764)
static bool b200_supports_rms_norm(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    return a && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ggml_are_same_shape(a, op) && a->nb[0] == sizeof(float) && op->nb[0] == sizeof(float);
}
static bool b200_supports_soft_max(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0]; const ggml_tensor * m = op->src[1];
    if (!a || a->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || !ggml_are_same_shape(a, op) || !ggml_is_contiguous(a) || !ggml_is_contiguous(op)) return false;
    if (m && !((m->type == GGML_TYPE_F32 || m->type == GGML_TYPE_F16) && ggml_is_contiguous(m) && m->ne[0] == a->ne[0] && m->ne[1] >= a->ne[1])) return false;
    return true;
}
static bool b200_supports_pool_2d(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    if (!a || !(a->type == GGML_TYPE_F32 || a->type == GGML_TYPE_F16) || op->type != GGML_TYPE_F32 || !ggml_is_contiguous(op)) return false;
    const int32_t * o = op->op_params;
    if (o[0] != GGML_OP_POOL_MAX && o[0] != GGML_OP_POOL_AVG) return false;
    return a->nb[0] == ggml_type_size(a->type) && a->nb[3] == a->nb[2] * (size_t) a->ne[2];
}
// The ops between the MUL_MATs of a Llama layer (graph_ops.cu), admitted in the layouts llama.cpp's graphs use; everything else stays
// with ggml-cpu.  With B200_GRAPH_OPS=0 the admission is the cDSP table's and nothing more.
static bool graph_ops_enabled(void) { static const bool on = !(getenv("B200_GRAPH_OPS") && atoi(getenv("B200_GRAPH_OPS")) == 0); return on; }
static bool f32_or_f16(enum ggml_type t) { return t == GGML_TYPE_F32 || t == GGML_TYPE_F16; }
static bool b200_supports_graph_op(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0]; const ggml_tensor * b = op->src[1];
    if (!graph_ops_enabled() || !a) return false;
    switch (op->op) {
        case GGML_OP_MUL:
            return b && a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ggml_are_same_shape(a, op) && ggml_can_repeat(b, a);
        case GGML_OP_UNARY:
            return ggml_get_unary_op(op) == GGML_UNARY_OP_SILU && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 &&
                   ggml_is_contiguous(a) && ggml_is_contiguous(op) && ggml_are_same_shape(a, op);
        case GGML_OP_ROPE: {
            if (!b || !f32_or_f16(a->type) || op->type != a->type || b->type != GGML_TYPE_I32) return false;
            if (op->src[2] && (op->src[2]->type != GGML_TYPE_F32 || op->src[2]->nb[0] != sizeof(float) || op->src[2]->ne[0] < op->op_params[1] / 2)) return false;   // frequency factors
            const int mode = op->op_params[2], n_dims = op->op_params[1];
            if (mode != 0 && mode != GGML_ROPE_TYPE_NEOX) return false;
            if (n_dims <= 0 || n_dims % 2 || n_dims > a->ne[0] || a->ne[0] % 2) return false;
            return a->nb[0] == ggml_type_size(a->type) && op->nb[0] == ggml_type_size(a->type) && ggml_are_same_shape(a, op) && b->ne[0] >= a->ne[2] && b->nb[0] == sizeof(int32_t);
        }
        case GGML_OP_CPY: case GGML_OP_DUP: case GGML_OP_CONT:
            return f32_or_f16(a->type) && f32_or_f16(op->type) && ggml_nelements(a) == ggml_nelements(op);
        case GGML_OP_GET_ROWS:
            return b && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && b->type == GGML_TYPE_I32 && a->nb[0] == sizeof(float) &&
                   op->nb[0] == sizeof(float) && b->nb[0] == sizeof(int32_t) && b->ne[3] == 1 && a->ne[2] == b->ne[1] && a->ne[3] == b->ne[2];
        default:
            return false;
    }
}
static bool b200_device_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    switch (op->op) {
        case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_ADD:      return b200_supports_add(op);
        case GGML_OP_MUL_MAT:  return b200_supports_mul_mat(op);
        case GGML_OP_RMS_NORM: return b200_supports_rms_norm(op);
        case GGML_OP_SOFT_MAX: return b200_supports_soft_max(op);
        case GGML_OP_POOL_2D:  return b200_supports_pool_2d(op);
        default:               return b200_supports_graph_op(op);
    }
}

// ------------------------------------------------------------------------------------------------ buffer
// (ggml_backend_hexagon_buffer_interface, ggml-hexagon.cpp:5361-5427.)  Device-private memory (is_host = false):
// callers reach the data only through set/get, which lets quantised tensors be kept in the SoA layout.
static void b200_buffer_free(ggml_backend_buffer_t buffer) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) buffer->context;
    if (c->base) { DEV_LOCK(c->dev); b200_free(dev_session(c->dev), c->base); }
    delete c;
}
static void * b200_buffer_get_base(ggml_backend_buffer_t buffer) { return ((b200_buffer_ctx *) buffer->context)->base; }

static void qrange(const ggml_tensor * t, size_t offset, size_t size, void ** base, int64_t * T, int64_t * first, int64_t * n) {
    const ggml_tensor * r = root_of(t);
    const int64_t bb = (int64_t) ggml_type_size(t->type);
    const int64_t pos = ((const char *) t->data - (const char *) r->data) + (int64_t) offset;
    GGML_ASSERT(pos % bb == 0 && (int64_t) size % bb == 0 && "b200: quantised tensors are read/written in whole blocks");
    *base = r->data; *T = (int64_t) ggml_nbytes(r) / bb; *first = pos / bb; *n = (int64_t) size / bb;
}
static void b200_buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * t, const void * data, size_t offset, size_t size) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) buffer->context;
    DEV_LOCK(c->dev);
    b200_handle h = dev_session(c->dev);
    int rc;
    if (is_qweight(t->type)) {
        void * base; int64_t T, first, n; qrange(t, offset, size, &base, &T, &first, &n);
        rc = b200_weights_upload(h, (int32_t) t->type, base, T, first, n, data);
    } else {
        rc = b200_memcpy_h2d(h, (char *) t->data + offset, data, size);
    }
    if (rc) GGML_ABORT("b200 set_tensor: %s", b200_last_error());
}
static void b200_buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * t, void * data, size_t offset, size_t size) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) buffer->context;
    DEV_LOCK(c->dev);
    b200_handle h = dev_session(c->dev);
    int rc;
    if (is_qweight(t->type)) {
        void * base; int64_t T, first, n; qrange(t, offset, size, &base, &T, &first, &n);
        rc = b200_weights_download(h, (int32_t) t->type, base, T, first, n, data);
    } else {
        rc = b200_memcpy_d2h(h, data, (const char *) t->data + offset, size);
    }
    if (rc) GGML_ABORT("b200 get_tensor: %s", b200_last_error());
}
static void b200_buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * t, uint8_t value, size_t offset, size_t size) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) buffer->context;
    DEV_LOCK(c->dev);
    b200_handle h = dev_session(c->dev);
    // a constant byte pattern is the same in the native and in the SoA layout only over whole tensors; partial
    // fills of a quantised tensor are not a thing callers do (ggml-backend.cpp uses memset for f32/kv zeroing)
    GGML_ASSERT(!is_qweight(t->type) || (offset == 0 && size == ggml_nbytes(t)));
    if (b200_memset(h, (char *) t->data + offset, value, size) || b200_synchronize(h)) GGML_ABORT("b200 memset_tensor: %s", b200_last_error());
}
static void b200_buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) buffer->context;
    DEV_LOCK(c->dev);
    b200_handle h = dev_session(c->dev);
    if (b200_memset(h, c->base, value, c->size) || b200_synchronize(h)) GGML_ABORT("b200 clear: %s", b200_last_error());
}
static const ggml_backend_buffer_i b200_buffer_iface = {
    /* free_buffer   */ b200_buffer_free,
    /* get_base      */ b200_buffer_get_base,
    /* init_tensor   */ nullptr,
    /* memset_tensor */ b200_buffer_memset_tensor,
    /* set_tensor    */ b200_buffer_set_tensor,
    /* get_tensor    */ b200_buffer_get_tensor,
    /* cpy_tensor    */ nullptr,          // copies go through get/set (layout-aware)
    /* clear         */ b200_buffer_clear,
    /* reset         */ nullptr,
};

// ------------------------------------------------------------------------------------------------ buffer type
// (ggml_backend_hexagon_buffer_type_*, ggml-hexagon.cpp:5429-5510.)
static const char * b200_buft_name(ggml_backend_buffer_type_t buft) { return ((b200_device_ctx *) buft->context)->name.c_str(); }
static ggml_backend_buffer_t b200_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    b200_device_ctx * d = (b200_device_ctx *) buft->context;
    DEV_LOCK(d);
    b200_handle h = dev_session(d);
    if (!h) return nullptr;
    void * p = nullptr;
    if (b200_malloc(h, size + 256, &p) != 0) {       // NULL on OOM, as ggml-hexagon.cpp:5470-5472
        B200_LOG("alloc_buffer(%zu) failed: %s", size, b200_last_error());
        return nullptr;
    }
    b200_buffer_ctx * c = new b200_buffer_ctx{d, p, size};
    return ggml_backend_buffer_init(buft, b200_buffer_iface, c, size);
}
static size_t b200_buft_alignment(ggml_backend_buffer_type_t) { return 256; }
static size_t b200_buft_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * t) {
    // room for the 16-byte-granular bulk copies of the 2-byte scale arrays to over-read
    return ggml_nbytes(t) + (is_qweight(t->type) ? 64 : 0);
}
static bool b200_buft_is_host(ggml_backend_buffer_type_t) { return false; }
static const ggml_backend_buffer_type_i b200_buft_iface = {
    /* get_name       */ b200_buft_name,
    /* alloc_buffer   */ b200_buft_alloc,
    /* get_alignment  */ b200_buft_alignment,
    /* get_max_size   */ nullptr,
    /* get_alloc_size */ b200_buft_alloc_size,
    /* is_host        */ b200_buft_is_host,
};

// ------------------------------------------------------------------------------------------------ backend
// (ggml_backend_hexagon_* backend interface, ggml-hexagon.cpp:5517-5574, table :5836-5850.)
static ggml_guid_t b200_guid(void) {
    static ggml_guid guid = {0xb2, 0x00, 0x5a, 0x10, 0x0a, 0x67, 0x67, 0x6d, 0x6c, 0x2d, 0x62, 0x32, 0x30, 0x30, 0x00, 0x01};
    return &guid;
}
static const char * b200_backend_name(ggml_backend_t backend) { return ((b200_backend_ctx *) backend->context)->name.c_str(); }
// development (GGML_B200_PROFILE=1): host time spent inside graph_compute, split into marshalling and the op-layer call; printed at backend free
static double g_prof_fp = 0, g_prof_call = 0; static long g_prof_n = 0;
static double now_us(void) { return (double) ggml_time_us(); }
static void b200_backend_free(ggml_backend_t backend) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    if (getenv("GGML_B200_PROFILE") && g_prof_n)
        fprintf(stderr, "[ggml-b200] graph_compute: %ld calls, marshalling %.1f us/call, op layer %.1f us/call (host time)\n", g_prof_n, g_prof_fp / g_prof_n, g_prof_call / g_prof_n);
    {   // the session belongs to the device (its buffers outlive backends); only drain it
        DEV_LOCK(c->dev);
        if (c->dev->h) b200_synchronize(c->dev->h);
    }
    delete c;
    delete backend;
}
static void b200_backend_synchronize(ggml_backend_t backend) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    DEV_LOCK(c->dev);
    if (b200_synchronize(dev_session(c->dev))) GGML_ABORT("b200 synchronize: %s", b200_last_error());
}
static bool is_view_op(enum ggml_op op) {
    return op == GGML_OP_NONE || op == GGML_OP_RESHAPE || op == GGML_OP_VIEW || op == GGML_OP_PERMUTE || op == GGML_OP_TRANSPOSE;
}
static int b200_op_of(const ggml_tensor * node) {
    switch (node->op) {
        case GGML_OP_ADD:      return B200_OP_ADD;
        case GGML_OP_MUL_MAT:  return B200_OP_MUL_MAT;
        case GGML_OP_RMS_NORM: return B200_OP_RMS_NORM;
        case GGML_OP_SOFT_MAX: return B200_OP_SOFT_MAX;
        case GGML_OP_POOL_2D:  return B200_OP_POOL_2D;
        case GGML_OP_MUL:      return B200_OP_MUL;
        case GGML_OP_UNARY:    return B200_OP_UNARY;
        case GGML_OP_ROPE:     return B200_OP_ROPE;
        case GGML_OP_CPY:      return B200_OP_CPY;
        case GGML_OP_DUP:      return B200_OP_DUP;
        case GGML_OP_CONT:     return B200_OP_CONT;
        case GGML_OP_GET_ROWS: return B200_OP_GET_ROWS;
        default:               return -1;
    }
}
// node loop of ggmlhexagon_backend_graph_compute_general (ggml-hexagon.cpp:5555-5574): view ops are no-ops, every other
// node is dst = op(src0, src1).  The whole list goes down in ONE call; asynchronous (ggml_backend_graph_compute adds
// the synchronize, ggml-backend.cpp:327-335).
static std::atomic<uint64_t> g_graph_key{1};
// A KV-cache store -- ggml_cpy into a view of a larger tensor (llama-graph.cpp:1341-1363) -- lands at an offset that moves with the position
// while everything else about the graph stays the same from one token to the next.  Such a node is marked B200_TENSOR_DYNAMIC: its data
// pointer stays out of the fingerprint (the cache tensor's own address goes in instead), and the op layer reads the destination of each call
// from a device table, so that one captured CUDA graph serves every position.
static bool is_dynamic_store(const ggml_tensor * node) {
    return node->op == GGML_OP_CPY && node->view_src && node->src[1] && node->src[1]->view_src == node->view_src &&
           node->view_src->buffer && ggml_nbytes(node) < ggml_nbytes(node->view_src);
}
static void fp_tensor(std::vector<uint64_t> & fp, const ggml_tensor * t, bool masked = false) {
    if (!t) { fp.push_back(0); return; }
    fp.push_back((uint64_t) (uintptr_t) t); fp.push_back(masked ? 0x64796eull : (uint64_t) (uintptr_t) t->data); fp.push_back((uint64_t) (uintptr_t) t->view_src);
    fp.push_back((uint64_t) t->type | ((uint64_t) t->op << 32));
    for (int i = 0; i < 4; i++) { fp.push_back((uint64_t) t->ne[i]); fp.push_back((uint64_t) t->nb[i]); }
    fp.push_back(t->view_src ? (uint64_t) (uintptr_t) t->view_src->data : 0);
}
static enum ggml_status b200_backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    const double t_in = now_us();
    // ---- fingerprint of the graph as this function would read it
    std::vector<uint64_t> & fp = c->fp;
    fp.clear();
    for (int i = 0; i < cgraph->n_nodes; i++) {
        const ggml_tensor * node = cgraph->nodes[i];
        if (is_view_op(node->op) || ggml_is_empty(node)) continue;
        const bool dyn = is_dynamic_store(node);
        fp_tensor(fp, node, dyn); fp_tensor(fp, node->src[0]); fp_tensor(fp, node->src[1], dyn); fp_tensor(fp, node->src[2]);
        uint64_t prm[2]; memcpy(prm, node->op_params, sizeof(prm));
        fp.push_back(prm[0]); fp.push_back(prm[1]);
    }
    if (fp.empty()) return GGML_STATUS_SUCCESS;
    b200_graph_entry & e = c->graphs[cgraph];
    e.last_use = ++c->tick;
    if (e.fp != fp) {
        e.nodes.clear(); e.dyn.clear();
        for (int i = 0; i < cgraph->n_nodes; i++) {
            ggml_tensor * node = cgraph->nodes[i];
            if (is_view_op(node->op) || ggml_is_empty(node)) continue;
            const int op = b200_op_of(node);
            if (op < 0) {
                B200_ERR("graph_compute: op %s was never admitted by supports_op", ggml_op_name(node->op));
                c->graphs.erase(cgraph);
                return GGML_STATUS_FAILED;
            }
            b200_node n;
            memset(&n, 0, sizeof(n));
            if (!to_b200(node->src[0], &n.src0, B200_OP_NONE) || (node->src[1] && !to_b200(node->src[1], &n.src1, B200_OP_NONE)) ||
                (node->src[2] && !to_b200(node->src[2], &n.src2, B200_OP_NONE)) || !to_b200(node, &n.dst, op)) {
                B200_ERR("graph_compute: node %s has a layout supports_op declines", node->name);
                c->graphs.erase(cgraph);
                return GGML_STATUS_FAILED;
            }
            if (is_dynamic_store(node)) { n.dst.flags |= B200_TENSOR_DYNAMIC; e.dyn.push_back({(int) e.nodes.size(), i}); }
            e.nodes.push_back(n);
        }
        if (getenv("GGML_B200_DUMP") && cgraph->n_nodes > 64) {          // development: the order the scheduler hands the nodes over in
            static int dumped = 0;
            if (dumped++ < 2) for (int i = 0; i < cgraph->n_nodes && i < 80; i++) {
                const ggml_tensor * t = cgraph->nodes[i];
                fprintf(stderr, "[ggml-b200] node %3d %-10s %-24s [%lld,%lld,%lld,%lld] data=%p src0=%p src1=%p\n", i, ggml_op_desc(t), t->name, (long long) t->ne[0], (long long) t->ne[1],
                              (long long) t->ne[2], (long long) t->ne[3], t->data, t->src[0] ? t->src[0]->data : nullptr, t->src[1] ? t->src[1]->data : nullptr);
            }
        }
        static const int eager_min = getenv("GGML_B200_GRAPH_EAGER_MIN") ? atoi(getenv("GGML_B200_GRAPH_EAGER_MIN")) : 256;
        if (e.seen >= 2) e.misses = 0; else if (e.seen == 1) e.misses++;
        e.trust = e.misses < 4 && (e.seen >= 6 || (int) e.nodes.size() >= eager_min);
        e.fp = fp; e.seen = 0;
        e.key = g_graph_key.fetch_add(1);          // a new key per distinct list: the op layer's caches never see one key with two lists
        if (c->graphs.size() > 512) {              // (a scheduler keeps a few dozen split graphs alive; drop the oldest half beyond that)
            std::vector<std::pair<uint64_t, const ggml_cgraph *>> age;
            for (auto & kv : c->graphs) age.push_back({kv.second.last_use, kv.first});
            std::sort(age.begin(), age.end());
            for (size_t i = 0; i < age.size() / 2; i++) if (age[i].second != cgraph) c->graphs.erase(age[i].second);
        }
    }
    b200_graph_entry & g = c->graphs[cgraph];
    // long node lists (a decode step) are captured once as a CUDA graph and replayed; short ones are launched directly
    static const int graph_min = getenv("GGML_B200_GRAPH_MIN") ? atoi(getenv("GGML_B200_GRAPH_MIN")) : 8;
    DEV_LOCK(c->dev);
    for (auto & d : g.dyn) {                    // same graph, this call's positions
        g.nodes[d.first].dst.data  = cgraph->nodes[d.second]->data;
        g.nodes[d.first].src1.data = cgraph->nodes[d.second]->src[1]->data;
    }
    g.seen++;
    const double t_mid = now_us();
    const int rc = b200_graph_compute_keyed(dev_session(c->dev), g.nodes.data(), (int32_t) g.nodes.size(), (int) g.nodes.size() >= graph_min && (g.seen >= 2 || g.trust), g.key);
    { const double t_out = now_us(); g_prof_fp += t_mid - t_in; g_prof_call += t_out - t_mid; g_prof_n++; }
    if (rc == B200_ENOMEM) return GGML_STATUS_ALLOC_FAILED;
    if (rc) { B200_ERR("graph_compute: %s", b200_last_error()); return GGML_STATUS_FAILED; }
    return GGML_STATUS_SUCCESS;
}
// Stream-ordered tensor copies (the hexagon backend leaves these slots empty, ggml-hexagon.cpp:5836-5850; a GPU backend wants
// them so that the scheduler's input copies queue behind / in front of graph_compute without a host round trip).  The device has
// ONE stream, so the quantised path (repack kernel, synchronous) is ordered with them too.
static void b200_backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * t, const void * data, size_t offset, size_t size) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    ggml_backend_buffer_t buf = t->view_src ? t->view_src->buffer : t->buffer;
    if (is_qweight(t->type)) { buf->iface.set_tensor(buf, t, data, offset, size); return; }
    DEV_LOCK(c->dev);
    if (b200_memcpy_h2d_async(dev_session(c->dev), (char *) t->data + offset, data, size)) GGML_ABORT("b200 set_tensor_async: %s", b200_last_error());
}
static void b200_backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * t, void * data, size_t offset, size_t size) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    ggml_backend_buffer_t buf = t->view_src ? t->view_src->buffer : t->buffer;
    if (is_qweight(t->type)) { buf->iface.get_tensor(buf, t, data, offset, size); return; }
    DEV_LOCK(c->dev);
    if (b200_memcpy_d2h_async(dev_session(c->dev), data, (const char *) t->data + offset, size)) GGML_ABORT("b200 get_tensor_async: %s", b200_last_error());
}
static const char * b200_buft_name(ggml_backend_buffer_type_t buft);
// device-to-device copy between two b200 backends (same or different GPU) without a host round trip: the source stream records an
// event, the destination stream waits for it and copies (ggml_backend_i::cpy_tensor_async, ggml-backend-impl.h:97-98; the
// scheduler's fallback is synchronize + get + set through host memory, ggml-backend.cpp:1379-1388).  Quantised tensors keep their
// SoA layout relative to their root tensor, so only whole contiguous non-quantised tensors travel this way; false = declined.
static bool b200_backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const ggml_tensor * src, ggml_tensor * dst) {
    if (!ggml_backend_is_b200(backend_src) || !ggml_backend_is_b200(backend_dst)) return false;
    ggml_backend_buffer_t bs = src->view_src ? src->view_src->buffer : src->buffer;
    ggml_backend_buffer_t bd = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!bs || !bd || bs->buft->iface.get_name != b200_buft_name || bd->buft->iface.get_name != b200_buft_name) return false;
    if (is_qweight(src->type) || src->type != dst->type || !ggml_is_contiguous(src) || !ggml_is_contiguous(dst) || ggml_nbytes(src) != ggml_nbytes(dst)) return false;
    b200_device_ctx * ds = ((b200_backend_ctx *) backend_src->context)->dev;
    b200_device_ctx * dd = ((b200_backend_ctx *) backend_dst->context)->dev;
    if (ds == dd) {
        DEV_LOCK(dd);
        return b200_memcpy_d2d(dev_session(dd), dst->data, src->data, ggml_nbytes(src)) == 0;
    }
    void * ev = nullptr;
    {
        DEV_LOCK(ds);
        if (b200_event_create(dev_session(ds), &ev) || b200_event_record(dev_session(ds), ev)) return false;
    }
    bool ok;
    {
        DEV_LOCK(dd);
        ok = b200_event_wait(dev_session(dd), ev) == 0 && b200_memcpy_d2d(dev_session(dd), dst->data, src->data, ggml_nbytes(src)) == 0;
    }
    { DEV_LOCK(ds); b200_event_destroy(dev_session(ds), ev); }       // (destruction is deferred by CUDA until the event has completed)
    return ok;
}
// events (ggml_backend_i::event_record / event_wait, ggml-backend-impl.h:112-116): CUDA events on the device's stream
static void b200_backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    DEV_LOCK(c->dev);
    if (b200_event_record(dev_session(c->dev), event->context)) GGML_ABORT("b200 event_record: %s", b200_last_error());
}
static void b200_backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    b200_backend_ctx * c = (b200_backend_ctx *) backend->context;
    DEV_LOCK(c->dev);
    if (b200_event_wait(dev_session(c->dev), event->context)) GGML_ABORT("b200 event_wait: %s", b200_last_error());
}
static const ggml_backend_i b200_backend_iface = {
    /* get_name           */ b200_backend_name,
    /* free               */ b200_backend_free,
    /* set_tensor_async   */ b200_backend_set_tensor_async,
    /* get_tensor_async   */ b200_backend_get_tensor_async,
    /* cpy_tensor_async   */ b200_backend_cpy_tensor_async,
    /* synchronize        */ b200_backend_synchronize,
    /* graph_plan_create  */ nullptr,
    /* graph_plan_free    */ nullptr,
    /* graph_plan_update  */ nullptr,
    /* graph_plan_compute */ nullptr,
    /* graph_compute      */ b200_backend_graph_compute,
    /* event_record       */ b200_backend_event_record,
    /* event_wait         */ b200_backend_event_wait,
};

// ------------------------------------------------------------------------------------------------ device
// (ggml_backend_hexagon_device_*, ggml-hexagon.cpp:5630-5834.)
static const char * b200_device_name(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->name.c_str(); }
static const char * b200_device_desc(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->desc.c_str(); }
static void b200_device_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    int64_t t = 0, f = 0;
    b200_device_info(((b200_device_ctx *) dev->context)->ordinal, nullptr, 0, &t, &f, nullptr);
    *free = (size_t) f; *total = (size_t) t;
}
static enum ggml_backend_dev_type b200_device_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }   // as cDSP mode, :5640-5644
static void b200_device_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name = b200_device_name(dev);
    props->description = b200_device_desc(dev);
    props->type = b200_device_type(dev);
    b200_device_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = { /* async */ true, /* host_buffer */ false, /* buffer_from_host_ptr */ false, /* events */ true };
}
static ggml_backend_t b200_device_init_backend(ggml_backend_dev_t dev, const char *) {
    b200_device_ctx * d = (b200_device_ctx *) dev->context;
    {
        DEV_LOCK(d);
        if (!dev_session(d)) { B200_ERR("init_backend: %s", b200_last_error()); return nullptr; }
    }
    b200_backend_ctx * c = new b200_backend_ctx();
    c->dev = d; c->name = d->name;
    return new ggml_backend{ /* guid */ b200_guid(), /* iface */ b200_backend_iface, /* device */ dev, /* context */ c };
}
// events (ggml_backend_device_i::event_new / event_free / event_synchronize, ggml-backend-impl.h:170-178)
static ggml_backend_event_t b200_device_event_new(ggml_backend_dev_t dev) {
    b200_device_ctx * d = (b200_device_ctx *) dev->context;
    DEV_LOCK(d);
    void * ev = nullptr;
    if (!dev_session(d) || b200_event_create(d->h, &ev)) { B200_ERR("event_new: %s", b200_last_error()); return nullptr; }
    return new ggml_backend_event{ /* device */ dev, /* context */ ev };
}
static void b200_device_event_free(ggml_backend_dev_t dev, ggml_backend_event_t event) {
    b200_device_ctx * d = (b200_device_ctx *) dev->context;
    DEV_LOCK(d);
    if (d->h) b200_event_destroy(d->h, event->context);
    delete event;
}
static void b200_device_event_synchronize(ggml_backend_dev_t dev, ggml_backend_event_t event) {
    b200_device_ctx * d = (b200_device_ctx *) dev->context;
    b200_handle h;
    { DEV_LOCK(d); h = dev_session(d); }
    if (b200_event_synchronize(h, event->context)) GGML_ABORT("b200 event_synchronize: %s", b200_last_error());   // (blocks: not under the lock)
}
static ggml_backend_buffer_type_t b200_device_buffer_type(ggml_backend_dev_t dev) { return &((b200_device_ctx *) dev->context)->buft; }
static bool b200_device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    return buft->iface.get_name == b200_buft_name && buft->context == dev->context;           // own buffers only, :5806-5816
}
static bool b200_device_offload_op(ggml_backend_dev_t, const ggml_tensor * op) {
    // worth pulling a weight over for: batched MUL_MAT (same threshold idea as ggml-cuda's min batch size)
    return op->op == GGML_OP_MUL_MAT && op->ne[1] >= 32;
}
static const ggml_backend_device_i b200_device_iface = {
    /* get_name             */ b200_device_name,
    /* get_description      */ b200_device_desc,
    /* get_memory           */ b200_device_memory,
    /* get_type             */ b200_device_type,
    /* get_props            */ b200_device_props,
    /* init_backend         */ b200_device_init_backend,
    /* get_buffer_type      */ b200_device_buffer_type,
    /* get_host_buffer_type */ nullptr,
    /* buffer_from_host_ptr */ nullptr,
    /* supports_op          */ b200_device_supports_op,
    /* supports_buft        */ b200_device_supports_buft,
    /* offload_op           */ b200_device_offload_op,
    /* event_new            */ b200_device_event_new,
    /* event_free           */ b200_device_event_free,
    /* event_synchronize    */ b200_device_event_synchronize,
};

// ------------------------------------------------------------------------------------------------ registry
// (ggml_backend_hexagon_reg_*, ggml-hexagon.cpp:5885-5994.)
struct b200_reg_ctx {
    std::vector<ggml_backend_device *> devices;
    std::vector<b200_device_ctx *>     ctxs;
};
static const char * b200_reg_name(ggml_backend_reg_t) { return GGML_B200_BACKEND_NAME; }
static size_t b200_reg_device_count(ggml_backend_reg_t reg) { return ((b200_reg_ctx *) reg->context)->devices.size(); }
static ggml_backend_dev_t b200_reg_device(ggml_backend_reg_t reg, size_t i) {
    b200_reg_ctx * r = (b200_reg_ctx *) reg->context;
    GGML_ASSERT(i < r->devices.size());
    return r->devices[i];
}
static void b200_set_n_threads(ggml_backend_t, int) {}     // accepted and ignored, as ggml-hexagon.cpp:5921-5932
static void * b200_reg_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_set_n_threads") == 0) return (void *) b200_set_n_threads;
    // "ggml_backend_split_buffer_type": row split runs one process per GPU over NCCL (b200_mulmat_rowsplit), not as
    // a single-process split buffer, so the name is not exported
    return nullptr;
}
static const ggml_backend_reg_i b200_reg_iface = {
    /* get_name         */ b200_reg_name,
    /* get_device_count */ b200_reg_device_count,
    /* get_device       */ b200_reg_device,
    /* get_proc_address */ b200_reg_proc_address,
};

ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static std::mutex mu;
    static ggml_backend_reg reg;
    static bool done = false;
    std::lock_guard<std::mutex> lk(mu);                      // init guarded by a mutex, as ggml-hexagon.cpp:5953-5954
    if (!done) {
        b200_reg_ctx * r = new b200_reg_ctx();               // static lifetime: the registry never frees a reg
        int n = b200_device_count();
        if (n > GGML_B200_MAX_DEVICES) n = GGML_B200_MAX_DEVICES;
        reg = ggml_backend_reg{ /* api_version */ GGML_BACKEND_API_VERSION, /* iface */ b200_reg_iface, /* context */ r };
        for (int i = 0; i < n; i++) {
            b200_device_ctx * d = new b200_device_ctx();
            char nm[128] = "";
            b200_device_info(i, nm, sizeof(nm), nullptr, nullptr, nullptr);
            d->ordinal = i; d->name = std::string("B200_") + std::to_string(i); d->desc = nm;
            d->buft = ggml_backend_buffer_type{ /* iface */ b200_buft_iface, /* device */ nullptr, /* context */ d };
            ggml_backend_device * dev = new ggml_backend_device{ /* iface */ b200_device_iface, /* reg */ &reg, /* context */ d };
            d->buft.device = dev;
            r->devices.push_back(dev); r->ctxs.push_back(d);
        }
        done = true;
    }
    return &reg;
}

int ggml_backend_b200_get_device_count(void) { return (int) b200_reg_device_count(ggml_backend_b200_reg()); }
const char * ggml_backend_b200_get_devname(size_t dev_num) {
    ggml_backend_reg_t reg = ggml_backend_b200_reg();
    if (dev_num >= b200_reg_device_count(reg)) return "unknown";
    return b200_device_name(b200_reg_device(reg, dev_num));
}
ggml_backend_t ggml_backend_b200_init(size_t dev_num, const char * params) {
    ggml_backend_reg_t reg = ggml_backend_b200_reg();
    if (dev_num >= b200_reg_device_count(reg)) return nullptr;
    return b200_device_init_backend(b200_reg_device(reg, dev_num), params);
}
bool ggml_backend_is_b200(ggml_backend_t backend) { return backend != nullptr && ggml_guid_matches(backend->guid, b200_guid()); }

// dynamic-loading entry points (ggml-backend-impl.h:215-246): no CUDA device -> score 0, the loader skips us
static int b200_score(void) { return b200_device_count() > 0 ? 100 : 0; }
GGML_BACKEND_DL_IMPL(ggml_backend_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(b200_score)

// ops_api.cu — implementation of the C-ABI declared in include/ggml_b200_ops.h.
// Host-side runtime of the op layer: context (device, stream, scratch), argument validation, dispatch to the
// kernels, CUDA-graph cache for node lists, staging for host-pointer tensors, NCCL row-split.
// No CPU compute path exists here: every op is a kernel launch or an error.
#include "../../include/ggml_b200_ops.h"
#include "b200_internal.cuh"

#include <cuda.h>

#include <dlfcn.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

using namespace b200;

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
static int fail(int code, const char * fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    return code;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(e_ == cudaErrorMemoryAllocation ? B200_ENOMEM : B200_ECUDA, "%s -> %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

extern "C" const char * b200_last_error(void) { return g_err; }

// ------------------------------------------------------------------------------------------------ NCCL (dlopen'ed)
typedef struct { char internal[128]; } nccl_uid;
typedef void * nccl_comm;
struct NcclApi {
    void * lib = nullptr;
    int (*GetUniqueId)(nccl_uid *) = nullptr;
    int (*CommInitRank)(nccl_comm *, int, nccl_uid, int) = nullptr;
    int (*CommDestroy)(nccl_comm) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char * (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;
static int nccl_load() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib) return 0;
    const char * names[] = { "libnccl.so.2", "libnccl.so" };
    void * l = nullptr;
    for (const char * n : names) { l = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (l) break; }
    if (!l) return fail(B200_ENCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define SYM(f) *(void **) &g_nccl.f = dlsym(l, "nccl" #f); if (!g_nccl.f) return fail(B200_ENCCL, "nccl" #f " missing")
    SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(AllGather); SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
#undef SYM
    g_nccl.lib = l;
    return 0;
}
#define NC(call) do { int r_ = (call); if (r_ != 0) return fail(B200_ENCCL, "%s -> %s", #call, g_nccl.GetErrorString(r_)); } while (0)

// ------------------------------------------------------------------------------------------------ context
struct GraphEntry { cudaGraphExec_t exec = nullptr; int64_t n_kernels = 0; int32_t n_nodes = 0; uint64_t check = 0; /* second, independent hash of the node list */ };

constexpr int PIN_SLOTS = 8; constexpr size_t PIN_SLOT_BYTES = 256 * 1024;
struct Ctx {
    uint64_t magic = 0xB200B200B200B200ull;
    int dev = 0;
    cudaStream_t st = nullptr;
    // q8_1 activation scratch (two-plane SoA), capacity in 32-element blocks
    int8_t * act_qs = nullptr; ActDS * act_ds = nullptr; int64_t act_cap = 0;
    // quantise on exit: the GEMV writes the NEXT op's q8_1 activations into the second scratch, which then becomes act_qs / act_ds (pointer swap)
    // pinned bounce slots for small synchronous copies from / to pageable host memory (b200_memcpy_h2d / _d2h): a pageable cudaMemcpyAsync is
    // staged by the driver and then waited for; through our own pinned slot an upload returns at once (the slot owns the bytes) and a
    // download is one DMA + one memcpy
    uint8_t * pin = nullptr; cudaEvent_t pin_ev[PIN_SLOTS] = {}; int pin_next = 0;
    int8_t * act_qs2 = nullptr; ActDS * act_ds2 = nullptr; uint2 * qx_slots = nullptr;
    struct { bool on = false, done = false; const void * vec = nullptr; } qx_hint;
    // device staging (host tensors, repack)
    uint8_t * stage = nullptr; size_t stage_cap = 0;
    uint8_t * flush = nullptr; size_t flush_bytes = 0;
    int64_t launches = 0;
    bool in_graph = false;        // inside b200_graph_compute: src1 quantisation may be reused between nodes
    uint64_t scratch_gen = 0;     // bumped whenever a scratch buffer moves: captured graphs hold the old addresses and must not match any more
    const void * q_ptr = nullptr; int64_t q_K = 0, q_N = 0, q_stride = 0;
    int q_kind = 0;               // what the cached conversion of q_ptr is: 0 = q8_1 planes in act_qs / act_ds, 1 = fp16 d * q in gm_qb (dense GEMM)
    std::unordered_map<uint64_t, GraphEntry> graphs;
    // tensor-core GEMM scratch (gemm_i8.cu): int8 activations [N][K], their block scales, unpacked q4_0 weights [M][K]
    int8_t * gm_qb = nullptr; size_t gm_qb_cap = 0;
    float *  gm_db = nullptr; size_t gm_db_cap = 0;
    int8_t * gm_a  = nullptr; size_t gm_a_cap = 0;       // exact mode only: unpacked integer weights of one MUL_MAT
    float *  gm_ws = nullptr; size_t gm_ws_cap = 0;
    float *  gm_tail = nullptr; size_t gm_tail_cap = 0;  // stream-K tail partial tiles of the dense GEMM (gemm_dense_tail_ws_bytes)
    float *  gm_sp = nullptr; size_t gm_sp_cap = 0;      // split-K partial results      // K-quants: f32 block scales (and negated block minima) of the unpacked weights
    bool capturing = false;       // inside cudaStreamBeginCapture .. EndCapture: no allocation, no synchronisation
    // B200_TENSOR_DYNAMIC destinations (KV-cache stores): a device table of the current dst pointers, refreshed before every launch of a captured
    // graph, so that one CUDA graph serves every position of the cache
    char ** dyn_tab = nullptr; int dyn_i = 0; bool dyn_on = false;
    std::unordered_set<uint64_t> taken_over;
    uint64_t last_graph_key = 0;  // the CUDA graph captured last: donor of cudaGraphExecUpdate when the next capture has the same topology
    int sms = 0;
    // row split
    nccl_comm comm = nullptr; int rank = 0, world = 1;
    // fused row split over peer memory: symmetric heap [flags 4 KB | user area], peers' mappings of theirs
    char * rs_heap = nullptr; size_t rs_heap_bytes = 0;
    char * rs_peer[8] = {nullptr}; bool rs_ipc[8] = {false};
    bool rs_attached = false, rs_on = false;
    // LL protocol bookkeeping of the submission being built: exchanges so far, and what the gather after the current GEMV collects
    // look-ahead of run_nodes for the next row-split exchange: fold the ADD / q8_1 quantisation that follow into the gather kernel
    struct { bool on = false, done = false; const void * of = nullptr; int qmat = -1; const float * add_src = nullptr; float * add_dst = nullptr; } rs_hint;
    uint32_t rs_ops = 0; int ll_nmat = 0; float * ll_dst[3] = {}; int ll_M[3] = {}; int ll_off[3] = {}; uint32_t ll_op = 0;
};

static Ctx * ctx_of(b200_handle h) {
    Ctx * c = (Ctx *) (uintptr_t) h;
    if (!c || c->magic != 0xB200B200B200B200ull) return nullptr;
    return c;
}
#define CTX(h) Ctx * c = ctx_of(h); if (!c) return fail(B200_EINVAL, "bad handle"); \
    { cudaError_t e_ = cudaSetDevice(c->dev); if (e_ != cudaSuccess) return fail(B200_ECUDA, "cudaSetDevice: %s", cudaGetErrorString(e_)); }

constexpr int64_t QX_MAX_BLOCKS = 8192;            // 32-element blocks of a result vector that can be quantised on exit (262144 rows)
static int ensure_act(Ctx * c, int64_t blocks) {
    if (blocks <= c->act_cap) return 0;
    if (c->act_qs) {
        CU(cudaStreamSynchronize(c->st)); cudaFree(c->act_qs); cudaFree(c->act_ds); cudaFree(c->act_qs2); cudaFree(c->act_ds2);
        c->act_qs = c->act_qs2 = nullptr; c->act_ds = c->act_ds2 = nullptr; c->act_cap = 0;
    }
    int64_t cap = blocks < 4096 ? 4096 : blocks;
    CU(cudaMalloc((void **) &c->act_qs, (size_t) cap * 32));
    CU(cudaMalloc((void **) &c->act_ds, (size_t) cap * sizeof(ActDS)));
    CU(cudaMalloc((void **) &c->act_qs2, (size_t) cap * 32));
    CU(cudaMalloc((void **) &c->act_ds2, (size_t) cap * sizeof(ActDS)));
    if (!c->qx_slots) { CU(cudaMalloc((void **) &c->qx_slots, QX_MAX_BLOCKS * 32 * sizeof(uint2))); CU(cudaMemsetAsync(c->qx_slots, 0, QX_MAX_BLOCKS * 32 * sizeof(uint2), c->st)); }
    c->act_cap = cap;
    c->q_ptr = nullptr;
    c->scratch_gen++;
    return 0;
}
template <typename T>
static int ensure_buf(Ctx * c, T ** p, size_t * cap, size_t bytes) {
    if (bytes <= *cap) return 0;
    if (*p) { CU(cudaStreamSynchronize(c->st)); cudaFree(*p); *p = nullptr; *cap = 0; }
    CU(cudaMalloc((void **) p, bytes + 256));
    *cap = bytes;
    c->scratch_gen++;
    c->q_ptr = nullptr;
    return 0;
}
static int ensure_stage(Ctx * c, size_t bytes) {
    if (bytes <= c->stage_cap) return 0;
    if (c->stage) { CU(cudaStreamSynchronize(c->st)); cudaFree(c->stage); c->stage = nullptr; c->stage_cap = 0; }
    size_t cap = bytes < (size_t) (1 << 20) ? (size_t) (1 << 20) : bytes;
    CU(cudaMalloc((void **) &c->stage, cap));
    c->stage_cap = cap;
    return 0;
}

// B200_GEMM_MODE=exact keeps the per-block exact int8 path (gemm_i8.cu) for many-column MUL_MATs; default: dense (gemm_dense.cu),
// which dequantises the native blocks inside the kernel -- no converted copy of any weight exists in HBM
static bool gemm_dense_enabled() { const char * e = getenv("B200_GEMM_MODE"); return !(e && strcmp(e, "exact") == 0); }

extern "C" int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int b200_device_info(int device, char * name, int len, int64_t * total, int64_t * free_b, int32_t * sms) {
    cudaDeviceProp p;
    CU(cudaGetDeviceProperties(&p, device));
    if (name && len > 0) { strncpy(name, p.name, len - 1); name[len - 1] = 0; }
    if (sms) *sms = p.multiProcessorCount;
    if (total || free_b) {
        int cur = 0; cudaGetDevice(&cur);
        CU(cudaSetDevice(device));
        size_t f = 0, t = 0;
        CU(cudaMemGetInfo(&f, &t));
        if (total) *total = (int64_t) t;
        if (free_b) *free_b = (int64_t) f;
        cudaSetDevice(cur);
    }
    return 0;
}

extern "C" int b200_open(const char * uri, b200_handle * h) {
    if (!h) return fail(B200_EINVAL, "h is NULL");
    int dev = 0;
    if (uri && *uri) {
        if (strncmp(uri, "cuda:", 5) != 0) return fail(B200_EINVAL, "uri must be cuda:<ordinal>, got '%s'", uri);
        dev = atoi(uri + 5);
    }
    int n = b200_device_count();
    if (n <= 0) return fail(B200_ENODEV, "no CUDA device visible (the B200 op layer has no CPU fallback)");
    if (dev < 0 || dev >= n) return fail(B200_ENODEV, "device %d out of range (%d visible)", dev, n);
    CU(cudaSetDevice(dev));
    Ctx * c = new Ctx();
    c->dev = dev;
    CU(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
    CU(cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, dev));
    int rc = ensure_act(c, 4096);
    if (rc) { delete c; return rc; }
    *h = (b200_handle) (uintptr_t) c;
    return 0;
}

extern "C" int b200_close(b200_handle h) {
    CTX(h);
    cudaStreamSynchronize(c->st);
    for (auto & kv : c->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    for (int r = 0; r < 8; r++) if (c->rs_ipc[r] && c->rs_peer[r]) cudaIpcCloseMemHandle(c->rs_peer[r]);
    cudaFree(c->rs_heap);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    cudaFree(c->act_qs); cudaFree(c->act_ds); cudaFree(c->act_qs2); cudaFree(c->act_ds2); cudaFree(c->qx_slots); cudaFree(c->stage); cudaFree(c->flush);
    if (c->pin) { cudaFreeHost(c->pin); for (int i = 0; i < PIN_SLOTS; i++) if (c->pin_ev[i]) cudaEventDestroy(c->pin_ev[i]); }
    cudaFree(c->gm_qb); cudaFree(c->gm_db); cudaFree(c->gm_a); cudaFree(c->gm_ws); cudaFree(c->gm_sp); cudaFree(c->gm_tail); cudaFree(c->dyn_tab);
    cudaStreamDestroy(c->st);
    c->magic = 0;
    delete c;
    return 0;
}

extern "C" int b200_setclocks(b200_handle h, int32_t, int32_t, int32_t) { CTX(h); return 0; }
extern "C" int b200_synchronize(b200_handle h) { CTX(h); CU(cudaStreamSynchronize(c->st)); return 0; }
extern "C" int64_t b200_launch_count(b200_handle h) { Ctx * c = ctx_of(h); return c ? c->launches : -1; }

extern "C" int64_t b200_block_elems(int32_t type) {
    switch (type) { case T_F32: case T_Q4_0: case T_Q8_0: case T_Q8_1: case T_Q4_K: case T_Q6_K: return block_elems(type); }
    return 0;
}
extern "C" int64_t b200_block_bytes(int32_t type) {
    switch (type) { case T_F32: case T_Q4_0: case T_Q8_0: case T_Q8_1: case T_Q4_K: case T_Q6_K: return block_bytes(type); }
    return 0;
}

// ------------------------------------------------------------------------------------------------ memory
extern "C" int b200_malloc(b200_handle h, size_t size, void ** ptr) {
    CTX(h); if (!ptr) return fail(B200_EINVAL, "ptr is NULL");
    *ptr = nullptr;
    cudaError_t e = cudaMalloc(ptr, size + 256);      // slack: 16-B granular bulk copies of 2-byte scale arrays may over-read
    if (e != cudaSuccess) { cudaGetLastError(); return fail(B200_ENOMEM, "cudaMalloc(%zu): %s", size, cudaGetErrorString(e)); }
    return 0;
}
extern "C" int b200_free(b200_handle h, void * ptr) {
    CTX(h);
    CU(cudaStreamSynchronize(c->st)); CU(cudaFree(ptr)); return 0;
}
static int pin_slot(Ctx * c, uint8_t ** slot, cudaEvent_t * ev) {
    if (!c->pin) {
        CU(cudaMallocHost((void **) &c->pin, PIN_SLOTS * PIN_SLOT_BYTES));
        for (int i = 0; i < PIN_SLOTS; i++) CU(cudaEventCreateWithFlags(&c->pin_ev[i], cudaEventDisableTiming));
    }
    const int i = c->pin_next; c->pin_next = (i + 1) % PIN_SLOTS;
    CU(cudaEventSynchronize(c->pin_ev[i]));          // (the copy that last used the slot; long done unless 8 uploads are in flight)
    *slot = c->pin + (size_t) i * PIN_SLOT_BYTES; *ev = c->pin_ev[i];
    return 0;
}
static bool capturing(Ctx * c) { cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone; cudaStreamIsCapturing(c->st, &st); return st != cudaStreamCaptureStatusNone; }
extern "C" int b200_memcpy_h2d(b200_handle h, void * d, const void * s, size_t n) {
    CTX(h); c->q_ptr = nullptr;
    if (n > 0 && n <= PIN_SLOT_BYTES && !capturing(c)) {          // small: through a pinned slot, no wait (later work on the stream sees it in order)
        uint8_t * slot; cudaEvent_t ev;
        int rc = pin_slot(c, &slot, &ev); if (rc) return rc;
        memcpy(slot, s, n);
        CU(cudaMemcpyAsync(d, slot, n, cudaMemcpyHostToDevice, c->st));
        CU(cudaEventRecord(ev, c->st));
        return 0;
    }
    CU(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, c->st)); CU(cudaStreamSynchronize(c->st)); return 0;
}
extern "C" int b200_memcpy_d2h(b200_handle h, void * d, const void * s, size_t n) {
    CTX(h);
    if (n > 0 && n <= PIN_SLOT_BYTES && !capturing(c)) {
        uint8_t * slot; cudaEvent_t ev;
        int rc = pin_slot(c, &slot, &ev); if (rc) return rc;
        CU(cudaMemcpyAsync(slot, s, n, cudaMemcpyDeviceToHost, c->st));
        CU(cudaEventRecord(ev, c->st));
        CU(cudaEventSynchronize(ev));
        memcpy(d, slot, n);
        return 0;
    }
    CU(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, c->st)); CU(cudaStreamSynchronize(c->st)); return 0;
}
extern "C" int b200_memcpy_h2d_async(b200_handle h, void * d, const void * s, size_t n) {
    CTX(h); c->q_ptr = nullptr; CU(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, c->st)); return 0;
}
extern "C" int b200_memcpy_d2h_async(b200_handle h, void * d, const void * s, size_t n) {
    CTX(h); CU(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, c->st)); return 0;
}
extern "C" int b200_memcpy_d2d(b200_handle h, void * d, const void * s, size_t n) {
    CTX(h); c->q_ptr = nullptr; CU(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToDevice, c->st)); return 0;
}
extern "C" int b200_memset(b200_handle h, void * d, int v, size_t n) {
    CTX(h); c->q_ptr = nullptr; CU(cudaMemsetAsync(d, v, n, c->st)); return 0;
}
extern "C" int b200_host_alloc(b200_handle h, size_t size, void ** ptr) { CTX(h); CU(cudaMallocHost(ptr, size ? size : 1)); return 0; }
extern "C" int b200_host_free(b200_handle h, void * ptr) { CTX(h); CU(cudaFreeHost(ptr)); return 0; }

extern "C" int b200_flush_l2(b200_handle h) {
    CTX(h);
    if (!c->flush) { c->flush_bytes = (size_t) 256 << 20; CU(cudaMalloc((void **) &c->flush, c->flush_bytes)); }
    CU(launch_fill(c->flush, c->flush_bytes, 0x01010101u, c->st));
    return 0;
}

// ------------------------------------------------------------------------------------------------ weights
static bool is_qweight(int t) { return t == T_Q4_0 || t == T_Q8_0 || t == T_Q4_K || t == T_Q6_K; }

extern "C" int b200_weights_upload(b200_handle h, int32_t type, void * base, int64_t T, int64_t first, int64_t n, const void * host) {
    CTX(h);
    if (!is_qweight(type)) return fail(B200_EINVAL, "weights_upload: type %d is not a quantised weight type", type);
    if (first < 0 || n < 0 || first + n > T) return fail(B200_EINVAL, "weights_upload: block range [%lld,+%lld) outside %lld", (long long) first, (long long) n, (long long) T);
    const size_t bytes = (size_t) n * block_bytes(type);
    int rc = ensure_stage(c, bytes); if (rc) return rc;
    CU(cudaMemcpyAsync(c->stage, host, bytes, cudaMemcpyHostToDevice, c->st));
    CU(launch_repack(type, true, (uint8_t *) base, T, first, n, c->stage, c->st)); c->launches++;
    CU(cudaStreamSynchronize(c->st));
    return 0;
}
extern "C" int b200_weights_download(b200_handle h, int32_t type, const void * base, int64_t T, int64_t first, int64_t n, void * host) {
    CTX(h);
    if (!is_qweight(type)) return fail(B200_EINVAL, "weights_download: type %d is not a quantised weight type", type);
    if (first < 0 || n < 0 || first + n > T) return fail(B200_EINVAL, "weights_download: bad block range");
    const size_t bytes = (size_t) n * block_bytes(type);
    int rc = ensure_stage(c, bytes); if (rc) return rc;
    CU(launch_repack(type, false, (uint8_t *) base, T, first, n, c->stage, c->st)); c->launches++;
    CU(cudaMemcpyAsync(host, c->stage, bytes, cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return 0;
}

// ------------------------------------------------------------------------------------------------ quantize / block dots
extern "C" int b200_quantize_q8_1(b200_handle h, const float * x, void * y, int64_t k, int64_t nrows) {
    CTX(h);
    if (k <= 0 || k % 32 || nrows < 0) return fail(B200_EINVAL, "quantize_q8_1: k=%lld must be a positive multiple of 32", (long long) k);
    if (((uintptr_t) x & 15) || ((uintptr_t) y & 3)) return fail(B200_EINVAL, "quantize_q8_1: x must be 16-byte and y 4-byte aligned");
    if (nrows == 0) return 0;
    CU(launch_quantize_q8_1_aos(x, k, nrows, y, c->st)); c->launches++;
    return 0;
}

static int wptrs_of(const b200_tensor * t, WPtrs * out) {
    if (!(t->flags & B200_TENSOR_REPACKED)) return fail(B200_EINVAL, "quantised src0 must be REPACKED device weights (b200_weights_upload)");
    if (!t->w_base || t->w_blocks <= 0 || t->w_first < 0) return fail(B200_EINVAL, "REPACKED tensor without w_base/w_blocks");
    *out = wptrs(t->type, t->w_base, t->w_blocks, t->w_first);
    return 0;
}

extern "C" int b200_block_dots(b200_handle h, const b200_tensor * w, const int8_t * a_qs, int32_t * out) {
    CTX(h);
    if (!w || !is_qweight(w->type)) return fail(B200_EINVAL, "block_dots: quantised weight row expected");
    const int64_t K = w->ne[0];
    if (K <= 0 || K % block_elems(w->type)) return fail(B200_EINVAL, "block_dots: K=%lld not a multiple of the block size", (long long) K);
    WPtrs wp; int rc = wptrs_of(w, &wp); if (rc) return rc;
    rc = ensure_stage(c, (size_t) K + (size_t) (K / 32) * sizeof(ActDS) + 64); if (rc) return rc;
    int8_t * sq = (int8_t *) c->stage;
    ActDS * sd = (ActDS *) (c->stage + ((K + 15) / 16) * 16);
    CU(launch_block_dots_impl(w->type, wp, K, a_qs, out, sq, sd, c->st)); c->launches++;
    return 0;
}

// ------------------------------------------------------------------------------------------------ ADD
static int add_dev(Ctx * c, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    if (s0->type != T_F32 || s1->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "add: f32 only (ggml-hexagon.cpp:5079-5085)");
    AddArgs a{};
    a.a = (const char *) s0->data; a.b = (const char *) s1->data; a.d = (char *) d->data;
    for (int i = 0; i < 4; i++) {
        if (s0->ne[i] != d->ne[i]) return fail(B200_EINVAL, "add: src0 and dst shapes differ in dim %d", i);
        if (s1->ne[i] <= 0 || d->ne[i] % s1->ne[i]) return fail(B200_EINVAL, "add: src1 dim %d (%lld) does not repeat into %lld", i, (long long) s1->ne[i], (long long) d->ne[i]);
        a.ne[i] = d->ne[i]; a.neb[i] = s1->ne[i];
        a.nba[i] = s0->nb[i]; a.nbb[i] = s1->nb[i]; a.nbd[i] = d->nb[i];
    }
    CU(launch_add_f32(a, c->st)); c->launches++;
    // dst may alias a cached quantised activation
    c->q_ptr = nullptr;
    return 0;
}

// ------------------------------------------------------------------------------------------------ RMS_NORM / SOFT_MAX / POOL_2D
static void wrote(Ctx * c, const b200_tensor * d);
static float f32_param(const b200_tensor * t, int i) { float f; memcpy(&f, &t->op_params[i], 4); return f; }
static int rms_norm_dev(Ctx * c, const b200_tensor * s, b200_tensor * d) {
    if (s->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "rms_norm: f32 only");
    for (int i = 0; i < 4; i++) if (s->ne[i] != d->ne[i]) return fail(B200_EINVAL, "rms_norm: src and dst shapes differ");
    if (s->nb[0] != 4 || d->nb[0] != 4) return fail(B200_ENOTSUP, "rms_norm: rows must be contiguous");
    const float eps = f32_param(d, 0);
    if (!(eps >= 0.0f)) return fail(B200_EINVAL, "rms_norm: eps < 0");
    CU(launch_rms_norm_f32(s->data, d->data, s->ne, s->nb, d->nb, eps, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static bool contiguous_f32_t(const b200_tensor * t) {
    int64_t st = 4;
    for (int i = 0; i < 4; i++) { if (t->ne[i] != 1 && t->nb[i] != st) return false; st *= t->ne[i]; }
    return true;
}
static int soft_max_dev(Ctx * c, const b200_tensor * s, const b200_tensor * m, b200_tensor * d) {
    if (s->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "soft_max: f32 only");
    for (int i = 0; i < 4; i++) if (s->ne[i] != d->ne[i]) return fail(B200_EINVAL, "soft_max: src and dst shapes differ");
    if (!contiguous_f32_t(s) || !contiguous_f32_t(d)) return fail(B200_ENOTSUP, "soft_max: src and dst must be contiguous");
    const bool has_mask = m && m->data;
    if (has_mask) {
        if (m->type != T_F32 && m->type != T_F16) return fail(B200_ENOTSUP, "soft_max: mask must be f32 or f16");
        const int64_t es = m->type == T_F16 ? 2 : 4;
        if (m->ne[0] != s->ne[0] || m->ne[1] < s->ne[1] || m->nb[0] != es || m->nb[1] != es * m->ne[0]) return fail(B200_EINVAL, "soft_max: mask must be a contiguous [ne00, >= ne01] matrix");
    }
    CU(launch_soft_max_f32((const float *) s->data, has_mask ? m->data : nullptr, has_mask && m->type == T_F16, (float *) d->data, s->ne,
                           f32_param(d, 0), f32_param(d, 1), c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int pool_2d_dev(Ctx * c, const b200_tensor * s, b200_tensor * d) {
    if ((s->type != T_F32 && s->type != T_F16) || d->type != T_F32) return fail(B200_ENOTSUP, "pool_2d: f32 / f16 in, f32 out");
    const int32_t * o = d->op_params;
    if ((o[0] != 0 && o[0] != 1) || o[1] <= 0 || o[2] <= 0 || o[3] <= 0 || o[4] <= 0 || o[5] < 0 || o[6] < 0) return fail(B200_EINVAL, "pool_2d: bad parameters");
    const int64_t es = s->type == T_F16 ? 2 : 4;
    if (s->nb[0] != es || s->nb[3] != s->nb[2] * s->ne[2]) return fail(B200_ENOTSUP, "pool_2d: src planes must be evenly strided");
    if (!contiguous_f32_t(d) || d->ne[2] != s->ne[2] || d->ne[3] != s->ne[3]) return fail(B200_EINVAL, "pool_2d: dst must be contiguous [px, py, ne2, ne3]");
    CU(launch_pool_2d(s->data, s->type == T_F16, (float *) d->data, s->ne, s->nb, d->ne[0], d->ne[1], o, c->st)); c->launches++;
    c->q_ptr = nullptr;
    return 0;
}
// ------------------------------------------------------------------------------------------------ graph_ops.cu: MUL / SILU / ROPE / CPY / GET_ROWS / f16 MUL_MAT
constexpr int T_I32 = 26;                                   // GGML_TYPE_I32 (ggml.h:349-390)
constexpr int DYN_MAX = 1024;
// an op wrote d: a cached conversion of the activations at q_ptr stays valid unless d touches the memory it was taken from (the ops of
// this section sit BETWEEN MUL_MATs that share their src1 -- q / k / v with the ROPEs in between -- so dropping the cache blindly would
// quantise the same vector three times)
static void wrote(Ctx * c, const b200_tensor * d) {
    if (!c->q_ptr) return;
    int64_t span = d->type == T_F16 ? 2 : 4;
    for (int i = 0; i < 4; i++) span += (d->ne[i] - 1) * d->nb[i];
    const int64_t qspan = (c->q_N > 1 ? (c->q_N - 1) * c->q_stride : 0) + c->q_K * 4;
    const char * a = (const char *) d->data; const char * b = (const char *) c->q_ptr;
    if (a < b + qspan && b < a + span) c->q_ptr = nullptr;
}
static int mul_dev(Ctx * c, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    if (s0->type != T_F32 || s1->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "mul: f32 only");
    AddArgs a{};
    a.a = (const char *) s0->data; a.b = (const char *) s1->data; a.d = (char *) d->data; a.mul = 1;
    for (int i = 0; i < 4; i++) {
        if (s0->ne[i] != d->ne[i]) return fail(B200_EINVAL, "mul: src0 and dst shapes differ in dim %d", i);
        if (s1->ne[i] <= 0 || d->ne[i] % s1->ne[i]) return fail(B200_EINVAL, "mul: src1 dim %d does not repeat into dst", i);
        a.ne[i] = d->ne[i]; a.neb[i] = s1->ne[i];
        a.nba[i] = s0->nb[i]; a.nbb[i] = s1->nb[i]; a.nbd[i] = d->nb[i];
    }
    CU(launch_add_f32(a, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int64_t nelems(const b200_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
static int unary_dev(Ctx * c, const b200_tensor * s, b200_tensor * d) {
    if (d->op_params[0] != B200_UNARY_SILU) return fail(B200_ENOTSUP, "unary: only SILU (%d) is implemented, got %d", B200_UNARY_SILU, d->op_params[0]);
    if (s->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "silu: f32 only");
    for (int i = 0; i < 4; i++) if (s->ne[i] != d->ne[i]) return fail(B200_EINVAL, "silu: src and dst shapes differ");
    if (!contiguous_f32_t(s) || !contiguous_f32_t(d)) return fail(B200_ENOTSUP, "silu: src and dst must be contiguous");
    CU(launch_silu_f32((const float *) s->data, (float *) d->data, nelems(s), c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int rope_dev(Ctx * c, const b200_tensor * s, const b200_tensor * pos, const b200_tensor * ff, b200_tensor * d, char * st_y = nullptr, char * const * st_dyn = nullptr) {
    const bool f16 = s->type == T_F16;
    if ((s->type != T_F32 && s->type != T_F16) || d->type != s->type || pos->type != T_I32) return fail(B200_ENOTSUP, "rope: f32 or f16 data (same on both sides), i32 positions");
    const int64_t es = f16 ? 2 : 4;
    for (int i = 0; i < 4; i++) if (s->ne[i] != d->ne[i]) return fail(B200_EINVAL, "rope: src and dst shapes differ");
    if (s->nb[0] != es || d->nb[0] != es) return fail(B200_ENOTSUP, "rope: rows must be contiguous");
    const int32_t * o = d->op_params;
    const int n_dims = o[1], mode = o[2], n_ctx_orig = o[4];
    if (mode != 0 && mode != 2) return fail(B200_ENOTSUP, "rope: mode %d (only normal = 0 and NeoX = 2)", mode);
    if (n_dims <= 0 || n_dims % 2 || n_dims > s->ne[0] || s->ne[0] % 2) return fail(B200_EINVAL, "rope: n_dims = %d on rows of %lld", n_dims, (long long) s->ne[0]);
    if (pos->ne[0] < s->ne[2] || pos->nb[0] != 4) return fail(B200_EINVAL, "rope: one position per ne[2] index");
    RopeArgs a{};
    a.x = (const char *) s->data; a.y = (char *) d->data; a.pos = (const int32_t *) pos->data;
    for (int i = 0; i < 4; i++) { a.ne[i] = s->ne[i]; a.nbs[i] = s->nb[i]; a.nbd[i] = d->nb[i]; }
    a.n_dims = n_dims; a.neox = mode == 2;
    if (ff && ff->data) {
        if (ff->type != T_F32 || ff->nb[0] != 4 || ff->ne[0] < n_dims / 2) return fail(B200_EINVAL, "rope: frequency factors must be f32 [>= n_dims / 2]");
        a.ff = (const float *) ff->data;
    }
    const float freq_base = f32_param(d, 5), beta_fast = f32_param(d, 9), beta_slow = f32_param(d, 10);
    a.freq_scale = f32_param(d, 6); a.ext_factor = f32_param(d, 7); a.attn_factor = f32_param(d, 8);
    a.theta_scale = powf(freq_base, -2.0f / n_dims);
    // ggml_rope_yarn_corr_dims (ggml.c:3741-3753)
    auto corr = [&](float n_rot) { return n_dims * logf(n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(freq_base)); };
    a.corr0 = fmaxf(0.0f, floorf(corr(beta_fast))); a.corr1 = fminf((float) (n_dims - 1), ceilf(corr(beta_slow)));
    a.st_y = st_y; a.st_dyn = st_dyn;
    CU(launch_rope_f32(a, f16, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int cpy_dev(Ctx * c, const b200_tensor * s, b200_tensor * d) {
    if ((s->type != T_F32 && s->type != T_F16) || (d->type != T_F32 && d->type != T_F16)) return fail(B200_ENOTSUP, "cpy: f32 / f16 only");
    if (nelems(s) != nelems(d)) return fail(B200_EINVAL, "cpy: element counts differ");
    CpyArgs a{};
    a.x = (const char *) s->data; a.y = (char *) d->data;
    for (int i = 0; i < 4; i++) { a.nes[i] = s->ne[i]; a.nbs[i] = s->nb[i]; a.ned[i] = d->ne[i]; a.nbd[i] = d->nb[i]; }
    if (d->flags & B200_TENSOR_DYNAMIC) {
        if (c->dyn_on) { if (c->dyn_i >= DYN_MAX) return fail(B200_ENOTSUP, "more than %d dynamic destinations in one node list", DYN_MAX); a.dyn = c->dyn_tab + c->dyn_i; }
        c->dyn_i++;
        c->q_ptr = nullptr;         // (where it lands differs from launch to launch)
    }
    CU(launch_cpy(a, s->type == T_F16, d->type == T_F16, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int get_rows_dev(Ctx * c, const b200_tensor * s, const b200_tensor * ids, b200_tensor * d) {
    if (s->type != T_F32 || d->type != T_F32 || ids->type != T_I32) return fail(B200_ENOTSUP, "get_rows: f32 rows, i32 ids");
    if (s->nb[0] != 4 || d->nb[0] != 4 || ids->nb[0] != 4) return fail(B200_ENOTSUP, "get_rows: dim 0 must be contiguous");
    if (d->ne[0] != s->ne[0] || d->ne[1] != ids->ne[0] || d->ne[2] != ids->ne[1] || d->ne[3] != ids->ne[2] || ids->ne[3] != 1 ||
        s->ne[2] != ids->ne[1] || s->ne[3] != ids->ne[2]) return fail(B200_EINVAL, "get_rows: shapes do not match");
    GetRowsArgs a{};
    a.x = (const char *) s->data; a.ids = (const char *) ids->data; a.y = (char *) d->data;
    a.ne00 = s->ne[0]; a.ne01 = s->ne[1]; a.ne10 = ids->ne[0]; a.ne11 = ids->ne[1]; a.ne12 = ids->ne[2];
    for (int i = 0; i < 4; i++) { a.nbs[i] = s->nb[i]; a.nbd[i] = d->nb[i]; }
    for (int i = 0; i < 3; i++) a.nbi[i] = ids->nb[i];
    CU(launch_get_rows_f32(a, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static int mm_f16_dev(Ctx * c, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    if (s0->type != T_F16 || s1->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "mulmat f16: f16 x f32 -> f32");
    const int64_t K = s0->ne[0], M = s0->ne[1], N = s1->ne[1];
    if (s1->ne[0] != K) return fail(B200_EINVAL, "mulmat f16: K mismatch");
    if (d->ne[0] != M || d->ne[1] != N || d->ne[2] != s1->ne[2] || d->ne[3] != s1->ne[3]) return fail(B200_EINVAL, "mulmat f16: dst shape mismatch");
    if (s0->ne[2] <= 0 || s0->ne[3] <= 0 || s1->ne[2] % s0->ne[2] || s1->ne[3] % s0->ne[3]) return fail(B200_EINVAL, "mulmat f16: batch dims do not broadcast");
    if (s0->nb[0] != 2 || s1->nb[0] != 4 || d->nb[0] != 4) return fail(B200_ENOTSUP, "mulmat f16: K (src0, src1) and M (dst) must be contiguous");
    if (M == 0 || N == 0) return 0;
    MmF16Args a{};
    a.a = (const char *) s0->data; a.b = (const char *) s1->data; a.d = (char *) d->data;
    a.M = M; a.N = N; a.K = K; a.ne12 = s1->ne[2]; a.ne13 = s1->ne[3]; a.r2 = s1->ne[2] / s0->ne[2]; a.r3 = s1->ne[3] / s0->ne[3];
    for (int i = 0; i < 4; i++) { a.nba[i] = s0->nb[i]; a.nbb[i] = s1->nb[i]; a.nbd[i] = d->nb[i]; }
    CU(launch_mm_f16(a, c->st)); c->launches++;
    wrote(c, d);
    return 0;
}
static bool any_host(const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    return ((a ? a->flags : 0) | (b ? b->flags : 0) | (d ? d->flags : 0)) & B200_TENSOR_HOST;
}
extern "C" int b200_rms_norm(b200_handle h, const b200_tensor * s, b200_tensor * d) {
    CTX(h); if (!s || !d) return fail(B200_EINVAL, "NULL tensor");
    if (any_host(s, nullptr, d)) return fail(B200_ENOTSUP, "rms_norm takes device tensors only");
    return rms_norm_dev(c, s, d);
}
extern "C" int b200_soft_max(b200_handle h, const b200_tensor * s, const b200_tensor * m, b200_tensor * d) {
    CTX(h); if (!s || !d) return fail(B200_EINVAL, "NULL tensor");
    if (any_host(s, m, d)) return fail(B200_ENOTSUP, "soft_max takes device tensors only");
    return soft_max_dev(c, s, m, d);
}
extern "C" int b200_pool_2d(b200_handle h, const b200_tensor * s, b200_tensor * d) {
    CTX(h); if (!s || !d) return fail(B200_EINVAL, "NULL tensor");
    if (any_host(s, nullptr, d)) return fail(B200_ENOTSUP, "pool_2d takes device tensors only");
    return pool_2d_dev(c, s, d);
}

// ------------------------------------------------------------------------------------------------ MUL_MAT
static int mulmat_check(const b200_tensor * s0, const b200_tensor * s1, const b200_tensor * d) {
    if (s1->type != T_F32 || d->type != T_F32) return fail(B200_ENOTSUP, "mulmat: src1 and dst must be f32");
    if (!(is_qweight(s0->type) || s0->type == T_F32)) return fail(B200_ENOTSUP, "mulmat: src0 type %d not supported", s0->type);
    const int64_t K = s0->ne[0], M = s0->ne[1], N = s1->ne[1];
    if (s1->ne[0] != K) return fail(B200_EINVAL, "mulmat: K mismatch %lld vs %lld", (long long) K, (long long) s1->ne[0]);
    if (d->ne[0] != M || d->ne[1] != N || d->ne[2] != s1->ne[2] || d->ne[3] != s1->ne[3]) return fail(B200_EINVAL, "mulmat: dst shape mismatch");
    if (s0->ne[2] <= 0 || s0->ne[3] <= 0 || s1->ne[2] % s0->ne[2] || s1->ne[3] % s0->ne[3]) return fail(B200_EINVAL, "mulmat: batch dims do not broadcast");
    if (K % block_elems(s0->type)) return fail(B200_EINVAL, "mulmat: K=%lld not a multiple of the block size", (long long) K);
    if (s1->nb[0] != 4 || d->nb[0] != 4) return fail(B200_ENOTSUP, "mulmat: src1/dst must be contiguous in dim 0");
    if (is_qweight(s0->type)) {
        if (K % 32) return fail(B200_EINVAL, "mulmat: K must be a multiple of 32");
        const int64_t bb = block_bytes(s0->type);
        if (s0->nb[0] != bb || s0->nb[1] != K / block_elems(s0->type) * bb) return fail(B200_ENOTSUP, "mulmat: quantised src0 must have contiguous rows");
        if (s0->nb[2] % bb || s0->nb[3] % bb) return fail(B200_ENOTSUP, "mulmat: src0 batch strides not block aligned");
        if (((uintptr_t) s1->data & 15) || (s1->nb[1] & 15) || (s1->nb[2] & 15) || (s1->nb[3] & 15)) return fail(B200_ENOTSUP, "mulmat: src1 rows must be 16-byte aligned");
    } else {
        if (s0->nb[0] != 4) return fail(B200_ENOTSUP, "mulmat: f32 src0 must be contiguous in dim 0");
    }
    return 0;
}

struct FusedAdd { const float * src = nullptr; float * dst = nullptr; };

// symmetric heap: [4096 B control: flags +0, op counter +256, ctrl +260/+264, LL submission base count +320][2 LL staging halves][user]
constexpr size_t RS_LL_BYTES = (size_t) 2 * LL_HALF_ELEMS * sizeof(uint2);
constexpr size_t RS_HDR = 4096 + RS_LL_BYTES;
static bool rs_proto_ll() { static const bool ll = !(getenv("B200_RS_PROTO") && strcmp(getenv("B200_RS_PROTO"), "flag") == 0); return ll; }

// ---- fused row split: peer deltas for the GEMV stores, and the exchange that follows each row-split op
// dsts[i]: this rank's slice of matrix i's dst; rows[i]: rows per rank; vec: one activation column (LL protocol possible)
static int rs_fill(Ctx * c, GemvArgs & a, const void * const * dsts, const int64_t * bytes, int n, const int64_t * rows = nullptr, bool vec = false) {
    a.npeers = 0; a.ll_slots = nullptr; c->ll_nmat = 0;
    if (!c->rs_on || c->world == 1) return 0;
    for (int i = 0; i < n; i++) {
        const char * p = (const char *) dsts[i];
        if (p < c->rs_heap + RS_HDR || p + bytes[i] > c->rs_heap + c->rs_heap_bytes) return fail(B200_EINVAL, "rowsplit: dst is outside the symmetric heap (b200_rowsplit_heap_create)");
    }
    for (int r = 0; r < c->world; r++) if (r != c->rank) {
        a.peer_flag[a.npeers] = (long long) (uintptr_t) (c->rs_peer[r] + 4 * c->rank);
        a.peer_delta[a.npeers++] = (long long) (c->rs_peer[r] - c->rs_heap);
    }
    if (vec && rows && rs_proto_ll()) {
        // LL protocol: {value, seq} slots in the staging area, polled by ll_gather_kernel on every peer
        int64_t off = 0;
        for (int i = 0; i < n; i++) { c->ll_off[i] = (int) off; off += rows[i] * c->world; }
        if (off <= LL_HALF_ELEMS) {
            c->rs_ops++;
            a.ll_slots = (uint2 *) (c->rs_heap + 4096); a.ll_count = (const uint32_t *) (c->rs_heap + 320); a.ll_op = c->rs_ops; a.ll_rank = c->rank;
            c->ll_nmat = n; c->ll_op = c->rs_ops;
            for (int i = 0; i < n; i++) { a.ll_off[i] = c->ll_off[i]; c->ll_M[i] = (int) rows[i]; c->ll_dst[i] = (float *) dsts[i] - (int64_t) c->rank * rows[i]; }
            return 0;
        }
    }
    // flag protocol: a separate 32-thread exchange kernel after the op (signal + wait), launched by rs_exchange
    return 0;
}
// dense GEMM activations in gm_qb: fp16 [N][K], then (256-byte aligned) the column scales, N rounded up to 256 floats
static size_t dense_act_cs_off(int64_t N, int64_t K) { return ((size_t) N * K * 2 + 255) & ~(size_t) 255; }
static size_t dense_act_bytes(int64_t N, int64_t K) { return dense_act_cs_off(N, K) + (size_t) ((N + 255) & ~(int64_t) 255) * sizeof(float); }
static bool qx_enabled() { return !(getenv("B200_QX") && atoi(getenv("B200_QX")) == 0); }      // (read per submission: the tests switch it)
// quantise on exit: if run_nodes announced that the result vector `qx_hint.vec` feeds the next MUL_MAT, let the GEMV that produces it
// write its q8_1 form into the second scratch (one column, M % 32 == 0, no row split)
static void qx_arm(Ctx * c, GemvArgs & a) {
    a.qx_qs = nullptr;
    if (!c->qx_hint.on || c->rs_on || !c->act_qs2) return;
    for (int m = 0; m < a.nmat; m++) {
        const float * v = a.mat[m].add_dst ? a.mat[m].add_dst : a.mat[m].dst;
        if ((const void *) v == c->qx_hint.vec && a.mat[m].M % 32 == 0 && a.mat[m].M / 32 <= QX_MAX_BLOCKS && a.mat[m].M / 32 <= c->act_cap) {
            a.qx_qs = c->act_qs2; a.qx_ds = c->act_ds2; a.qx_slots = c->qx_slots; a.qx_mat = m;
            return;
        }
    }
}
static void qx_launched(Ctx * c, const GemvArgs & a) {      // the armed GEMV is in the stream: the second scratch is the current one from here on
    if (!a.qx_qs) return;
    std::swap(c->act_qs, c->act_qs2); std::swap(c->act_ds, c->act_ds2);
    c->qx_hint.done = true;
}
static bool rs_fuse_enabled() { static const bool on = !(getenv("B200_RS_FUSE") && atoi(getenv("B200_RS_FUSE")) == 0); return on; }
static bool act_fusion_enabled() { static const bool on = !(getenv("B200_GEMM_ACTFUSE") && atoi(getenv("B200_GEMM_ACTFUSE")) == 0); return on; }
static int rs_exchange(Ctx * c) {
    if (!c->rs_on || c->world == 1) return 0;
    if (c->ll_nmat > 0) {
        LlGatherArgs g{};
        g.ll_slots = (uint2 *) (c->rs_heap + 4096); g.ll_count = (const uint32_t *) (c->rs_heap + 320); g.ll_op = c->ll_op;
        g.nmat = c->ll_nmat; g.world = c->world; g.rank = c->rank;
        for (int i = 0; i < c->ll_nmat; i++) { g.dst[i] = c->ll_dst[i]; g.M[i] = c->ll_M[i]; g.ll_off[i] = c->ll_off[i]; }
        c->ll_nmat = 0;
        if (c->rs_hint.on) {
            int qm = -1;
            for (int i = 0; i < g.nmat; i++) if ((const void *) g.dst[i] == c->rs_hint.of && g.M[i] % 32 == 0) qm = i;
            bool ok = qm >= 0;
            for (int i = 0; i < g.nmat; i++) ok = ok && g.M[i] % 32 == 0;
            if (ok && (int64_t) g.M[qm] * g.world / 32 <= c->act_cap) {
                LlFuseArgs f{}; f.qmat = qm; f.add_src = c->rs_hint.add_src; f.add_dst = c->rs_hint.add_dst; f.qs = c->act_qs; f.ds = c->act_ds;
                CU(launch_ll_gather_fused(g, f, c->st)); c->launches++;
                c->rs_hint.done = true;
                return 0;
            }
        }
        CU(launch_ll_gather(g, c->st)); c->launches++;
        return 0;
    }
    long long addr[7]; int ranks[7]; int np = 0;
    for (int r = 0; r < c->world; r++) if (r != c->rank) { addr[np] = (long long) (uintptr_t) (c->rs_peer[r] + 4 * c->rank); ranks[np] = r; np++; }
    CU(launch_rowsplit_exchange((uint32_t *) c->rs_heap, (uint32_t *) (c->rs_heap + 256), np, addr, ranks, c->st));
    c->launches++;
    return 0;
}

static WPtrs wptrs_offset(const b200_tensor * s0, WPtrs wp, int64_t byte_off) {
    const int64_t blk_off = byte_off / block_bytes(s0->type);
    WFields f = wfields(s0->type);
    for (int i = 0; i < f.n; i++) wp.p[i] += blk_off * f.size[i];
    return wp;
}

// column count from which the tensor-core GEMM takes over from the GEMV (B200_GEMM_MIN_N overrides).  Measured crossover
// against the shared-memory-tile GEMV (gemv5n) on B200: q4_0 4096 x 4096 33 vs 37 us at 48 columns, 43 vs 37 us at 64;
// q4_K 44 vs 50 us at 32 columns, 77 vs 50 us at 48
static int gemm_min_cols(int type) {
    const int env = getenv("B200_GEMM_MIN_N") ? atoi(getenv("B200_GEMM_MIN_N")) : -1;
    return env >= 0 ? env : (type == T_Q4_K || type == T_Q6_K ? 36 : 52);
}
// Scratch of the GEMM path for one MUL_MAT, sized for either MMA kind, allocated BEFORE a stream capture starts
// (cudaMalloc / cudaStreamSynchronize are not allowed inside one); mulmat_dev's own ensure_buf calls then never reallocate.
static int presize_gemm(Ctx * c, const b200_tensor * s0, const b200_tensor * s1) {
    const int64_t K = s0->ne[0], M = s0->ne[1], N = s1->ne[1];
    const int mn = gemm_min_cols(s0->type);
    if (!is_qweight(s0->type) || mn <= 0 || N < mn || K % 32 != 0 || M >= 0x7fffffff || N >= 0x7fffffff || K >= 0x3fffffff) return 0;
    const size_t ldb = (size_t) gemm_i8_ldb((int) N);
    int rc;
    if (gemm_dense_enabled()) {
        rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, dense_act_bytes(N, K)); if (rc) return rc;
        if (M % 32 == 0) { rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, dense_act_bytes(N, M)); if (rc) return rc; }      // (room for the fused conversion of a residual sum)
        const int ksp = gemm_dense_ksplit((int) M, (int) N, (int) K, c->sms);
        if (ksp > 1) { rc = ensure_buf(c, &c->gm_sp, &c->gm_sp_cap, (size_t) ksp * M * N * sizeof(float)); if (rc) return rc; }
        rc = ensure_buf(c, &c->gm_tail, &c->gm_tail_cap, gemm_dense_tail_ws_bytes(c->sms)); if (rc) return rc;
        return 0;
    }
    rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, (size_t) N * K * 2); if (rc) return rc;
    rc = ensure_buf(c, &c->gm_db, &c->gm_db_cap, (size_t) 2 * (K / 16) * ldb * sizeof(float)); if (rc) return rc;
    rc = ensure_buf(c, &c->gm_a, &c->gm_a_cap, (size_t) M * K * 2); if (rc) return rc;
    rc = ensure_buf(c, &c->gm_ws, &c->gm_ws_cap, (size_t) 2 * M * (K / 16) * sizeof(float)); if (rc) return rc;
    int ksp = gemm_i8_ksplit((int) M, (int) N, (int) K, 1, c->sms);
    const int k0 = gemm_i8_ksplit((int) M, (int) N, (int) (2 * K), 0, c->sms), k2 = gemm_i8_ksplit((int) M, (int) N, (int) K, 2, c->sms);
    if (k0 > ksp) ksp = k0;
    if (k2 > ksp) ksp = k2;
    if (ksp > 1) { rc = ensure_buf(c, &c->gm_sp, &c->gm_sp_cap, (size_t) ksp * M * N * sizeof(float)); if (rc) return rc; }
    return 0;
}

struct FusedAdd;
static int mulmat_dev(Ctx * c, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d, int64_t row_off_dst /*rowsplit*/,
                      const FusedAdd * fa = nullptr) {
    int rc = mulmat_check(s0, s1, d); if (rc) return rc;
    const int64_t K = s0->ne[0], M = s0->ne[1], N = s1->ne[1];
    if (M == 0 || N == 0 || K == 0) return 0;
    const int64_t r2 = s1->ne[2] / s0->ne[2], r3 = s1->ne[3] / s0->ne[3];
    const bool q = is_qweight(s0->type);
    WPtrs wp0{};
    if (q) { rc = wptrs_of(s0, &wp0); if (rc) return rc; }
    const int64_t NB = K / 32;
    const int gemm_min_n = gemm_min_cols(s0->type);
    const bool use_gemm = q && N >= gemm_min_n && gemm_min_n > 0 && K % 32 == 0 &&
                          M < 0x7fffffff && N < 0x7fffffff && K < 0x7fffffff && (!fa || N > 1) && ((uintptr_t) wp0.p[0] & 15) == 0 &&
                          (!c->rs_on || gemm_dense_enabled());   // row split: dense path + slice push
    if (fa && N > 1 && !(use_gemm && gemm_dense_enabled())) return fail(B200_EINVAL, "fused residual ADD needs the dense GEMM path");
    if (q) { rc = ensure_act(c, NB * N); if (rc) return rc; }
    const int64_t dcs = d->nb[1] / 4;
    for (int64_t i13 = 0; i13 < s1->ne[3]; i13++)
    for (int64_t i12 = 0; i12 < s1->ne[2]; i12++) {
        const char * x = (const char *) s1->data + i12 * s1->nb[2] + i13 * s1->nb[3];
        float * y = (float *) ((char *) d->data + i12 * d->nb[2] + i13 * d->nb[3]) + row_off_dst;
        const int64_t i02 = i12 / r2, i03 = i13 / r3;
        if (q && use_gemm && gemm_dense_enabled()) {
            // ---- many columns: dense fp16 tensor-core GEMM, block scales folded into the operands (gemm_dense.cu)
            const WPtrs wp = wptrs_offset(s0, wp0, i02 * s0->nb[2] + i03 * s0->nb[3]);
            rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, dense_act_bytes(N, K)); if (rc) return rc;
            float * const colscale = (float *) (c->gm_qb + dense_act_cs_off(N, K));
            // (consecutive MUL_MATs of a node list over the same activations -- q / k / v, gate / up -- convert them once)
            const bool reuse = c->in_graph && c->q_kind == 1 && c->q_ptr == (const void *) x && c->q_K == K && c->q_N == N && c->q_stride == s1->nb[1];
            if (!reuse) {
                CU(launch_dequant_act_f16((const float *) x, s1->nb[1], (int) K, (int) N, (__half *) c->gm_qb, colscale, c->st)); c->launches++;
                c->q_ptr = x; c->q_K = K; c->q_N = N; c->q_stride = s1->nb[1]; c->q_kind = 1;
            }
            const int ksp = gemm_dense_ksplit((int) M, (int) N, (int) K, c->sms);
            if (ksp > 1) { rc = ensure_buf(c, &c->gm_sp, &c->gm_sp_cap, (size_t) ksp * M * N * sizeof(float)); if (rc) return rc; }
            if (fa && ksp <= 1) return fail(B200_EINVAL, "fused residual ADD on a dense MUL_MAT without split-K");      // (can_fuse_add_dense checks)
            // the fused ADD's result is usually the next MUL_MAT's src1 (K = M): the reduce converts it into gm_qb on the way (the GEMM of this
            // op has read gm_qb by then), and the next op finds its activations converted.  Only where gm_qb already has the room (no
            // reallocation under a capture) and inside a node list.
            const bool next_act = fa && c->in_graph && M % 32 == 0 && M <= 8192 && c->gm_qb_cap >= dense_act_bytes(N, M) && act_fusion_enabled();
            rc = ensure_buf(c, &c->gm_tail, &c->gm_tail_cap, gemm_dense_tail_ws_bytes(c->sms)); if (rc) return rc;
            int nl = 0;
            CU(launch_gemm_dense(s0->type, wp, (const __half *) c->gm_qb, colscale, y, dcs, (int) M, (int) N, (int) K, c->sms, ksp > 1 ? c->gm_sp : nullptr, c->gm_tail, &nl, c->st,
                                 fa ? fa->src : nullptr, fa ? fa->dst : nullptr, next_act ? c->gm_qb : nullptr,
                                 next_act ? (float *) (c->gm_qb + dense_act_cs_off(N, M)) : nullptr));
            c->launches += nl;
            if (next_act) { c->q_ptr = fa->dst; c->q_K = M; c->q_N = N; c->q_stride = M * 4; c->q_kind = 1; }
            if (c->rs_on && c->world > 1) {
                // row split: y is this rank's M rows of every dst column; the slice goes into every peer's copy over NVLink, and the
                // flag exchange at the end of the op (rs_exchange below) publishes it
                const char * p = (const char *) y; const int64_t span = ((N - 1) * dcs + M) * 4;
                if (p < c->rs_heap + RS_HDR || p + span > c->rs_heap + c->rs_heap_bytes)
                    return fail(B200_EINVAL, "rowsplit: dst is outside the symmetric heap (b200_rowsplit_heap_create)");
                long long deltas[7]; int np = 0;
                for (int r = 0; r < c->world; r++) if (r != c->rank) deltas[np++] = (long long) (c->rs_peer[r] - c->rs_heap);
                CU(launch_rowsplit_push(y, M, N, dcs, np, deltas, c->st)); c->launches++;
            }
        } else if (q && use_gemm) {
            // ---- many columns, B200_GEMM_MODE=exact: int8 tensor-core GEMM with exact per-block sums (gemm_i8.cu)
            const WPtrs wp = wptrs_offset(s0, wp0, i02 * s0->nb[2] + i03 * s0->nb[3]);
            const long long ldb = gemm_i8_ldb((int) N);
            const int t = s0->type;
            // int8 operands with s32 block sums (kind::i8), or the same integers as fp16 operands with exact f32 block sums
            // (kind::f16: twice the operand bytes, no int -> float conversion).  Measured default: i8, except q6_K whose 16-element
            // scale groups are native K16 fp16 MMAs but need zero padding as int8.  B200_GEMM_KIND=i8|f16 forces one (development).
            const int kind_env = !getenv("B200_GEMM_KIND") ? -1 : strcmp(getenv("B200_GEMM_KIND"), "i8") == 0 ? 0 : 1;
            const bool kind_i8 = kind_env < 0 ? t != T_Q6_K : kind_env == 0;
            const float * dA_f = nullptr, * mA = nullptr;
            const __half * dA_h = t == T_Q8_0 || t == T_Q4_0 ? (const __half *) wp.p[1] : nullptr;
            if (kind_i8) {
                // MMA-side shape: q6_K scales change every 16 elements, so each 16-group is padded to its own 32-byte block (K doubles)
                const long long Kg = t == T_Q6_K ? 2 * K : K, KBg = Kg / 32;
                const int qmode = t == T_Q4_K ? 1 : t == T_Q6_K ? 2 : 0;
                rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, (size_t) N * Kg); if (rc) return rc;
                rc = ensure_buf(c, &c->gm_db, &c->gm_db_cap, (size_t) (qmode == 1 ? 2 : 1) * KBg * ldb * sizeof(float)); if (rc) return rc;
                CU(launch_quantize_rows_i8((const float *) x, s1->nb[1], (int) K, (int) N, c->gm_qb, c->gm_db, ldb, qmode, c->st)); c->launches++;
                const int8_t * A = (const int8_t *) wp.p[0];
                if (t != T_Q8_0) { rc = ensure_buf(c, &c->gm_a, &c->gm_a_cap, (size_t) M * Kg); if (rc) return rc; A = c->gm_a; }
                if (t == T_Q4_0) {
                    CU(launch_unpack_q4_0_i8(wp.p[0], M * NB, c->gm_a, c->st)); c->launches++;
                } else if (t == T_Q4_K) {
                    rc = ensure_buf(c, &c->gm_ws, &c->gm_ws_cap, (size_t) 2 * M * KBg * sizeof(float)); if (rc) return rc;
                    dA_f = c->gm_ws; mA = c->gm_ws + M * KBg;
                    CU(launch_unpack_q4_K_i8(wp, M * (K / 256), c->gm_a, c->gm_ws, c->gm_ws + M * KBg, c->st)); c->launches++;
                } else if (t == T_Q6_K) {
                    rc = ensure_buf(c, &c->gm_ws, &c->gm_ws_cap, (size_t) M * KBg * sizeof(float)); if (rc) return rc;
                    dA_f = c->gm_ws;
                    CU(launch_unpack_q6_K_i8(wp, M * (K / 256), c->gm_a, c->gm_ws, c->st)); c->launches++;
                }
                const int ksp = gemm_i8_ksplit((int) M, (int) N, (int) Kg, 0, c->sms);
                if (ksp > 1) { rc = ensure_buf(c, &c->gm_sp, &c->gm_sp_cap, (size_t) ksp * M * N * sizeof(float)); if (rc) return rc; }
                CU(launch_gemm_i8(0, A, dA_h, dA_f, mA, c->gm_qb, c->gm_db, ldb, y, dcs, (int) M, (int) N, (int) Kg, c->sms, ksp > 1 ? c->gm_sp : nullptr, c->st));
                c->launches += 1 + (mA ? 1 : 0) + (ksp > 1 ? 1 : 0);
            } else {
                const int kind = t == T_Q6_K ? 2 : 1;
                const long long KBg = K / (kind == 2 ? 16 : 32);
                const int qmode = t == T_Q4_K ? 1 : t == T_Q6_K ? 2 : 0;
                rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, (size_t) N * K * 2); if (rc) return rc;
                rc = ensure_buf(c, &c->gm_db, &c->gm_db_cap, (size_t) (qmode == 1 ? 2 : 1) * KBg * ldb * sizeof(float)); if (rc) return rc;
                CU(launch_quantize_rows_f16((const float *) x, s1->nb[1], (int) K, (int) N, (__half *) c->gm_qb, c->gm_db, ldb, qmode, c->st)); c->launches++;
                rc = ensure_buf(c, &c->gm_a, &c->gm_a_cap, (size_t) M * K * 2); if (rc) return rc;
                if (t == T_Q4_K || t == T_Q6_K) {
                    rc = ensure_buf(c, &c->gm_ws, &c->gm_ws_cap, (size_t) 2 * M * KBg * sizeof(float)); if (rc) return rc;
                    dA_f = c->gm_ws; if (t == T_Q4_K) mA = c->gm_ws + M * KBg;
                }
                CU(launch_unpack_f16(t, wp, M * (K / block_elems(t)), (__half *) c->gm_a, c->gm_ws, c->gm_ws + M * KBg, c->st)); c->launches++;
                const int ksp = gemm_i8_ksplit((int) M, (int) N, (int) K, kind, c->sms);
                if (ksp > 1) { rc = ensure_buf(c, &c->gm_sp, &c->gm_sp_cap, (size_t) ksp * M * N * sizeof(float)); if (rc) return rc; }
                CU(launch_gemm_i8(kind, c->gm_a, dA_h, dA_f, mA, c->gm_qb, c->gm_db, ldb, y, dcs, (int) M, (int) N, (int) K, c->sms, ksp > 1 ? c->gm_sp : nullptr, c->st));
                c->launches += 1 + (mA ? 1 : 0) + (ksp > 1 ? 1 : 0);
            }
            c->q_ptr = nullptr;
        } else if (q) {
            GemvArgs a{};
            a.nmat = 1; a.K = K; a.dcs = dcs;
            a.mat[0].w = wptrs_offset(s0, wp0, i02 * s0->nb[2] + i03 * s0->nb[3]);
            a.mat[0].M = M;
            {
                if (fa && N == 1) { a.mat[0].add_src = fa->src; a.mat[0].add_dst = fa->dst; }
                { const void * dp[1] = {y}; const int64_t db_[1] = {(N - 1) * dcs * 4 + M * 4}; const int64_t rw[1] = {M};
                  rc = rs_fill(c, a, dp, db_, 1, rw, N == 1 && s1->ne[2] == 1 && s1->ne[3] == 1); if (rc) return rc; }
                bool done = false;
                if (N == 1) {                               // one column: persistent ring GEMV, quantises the activations itself (gemv6.cu)
                    a.mat[0].dst = y; a.x = (const float *) x; a.x_col_stride = s1->nb[1];
                    cudaError_t e6 = cudaSuccess;
                    done = launch_gemv6(s0->type, a, c->sms, c->st, &e6);
                    if (done) { CU(e6); c->launches++; continue; }
                }
                const bool reuse = c->in_graph && c->q_kind == 0 && c->q_ptr == (const void *) x && c->q_K == K && c->q_N == N && c->q_stride == s1->nb[1];
                if (!reuse) {
                    CU(launch_quantize_q8_1_soa((const float *) x, s1->nb[1], K, N, c->act_qs, c->act_ds, c->st)); c->launches++;
                    c->q_ptr = x; c->q_K = K; c->q_N = N; c->q_stride = s1->nb[1]; c->q_kind = 0;
                }
                if (N > 1 && N < 0x7fff) {                  // small batch: one launch over shared-memory tiles, weights read once
                    a.act = ActView{c->act_qs, c->act_ds, NB};
                    a.mat[0].dst = y;
                    cudaError_t e5 = cudaSuccess;
                    done = launch_gemv_cols(s0->type, a, (int) N, c->st, &e5);
                    if (done) { CU(e5); c->launches++; }
                }
                if (!done)
                for (int64_t c0 = 0; c0 < N; c0 += 8) {
                    const int nc = (int) (N - c0 < 8 ? N - c0 : 8);
                    a.act = ActView{c->act_qs + c0 * NB * 32, c->act_ds + c0 * NB, NB};
                    a.mat[0].dst = y + c0 * dcs;
                    if (N == 1) qx_arm(c, a);
                    CU(launch_gemv2(s0->type, a, nc, c->st)); c->launches++;
                    if (N == 1) qx_launched(c, a);
                }
            }
        } else {
            const char * w = (const char *) s0->data + i02 * s0->nb[2] + i03 * s0->nb[3];
            for (int64_t c0 = 0; c0 < N; c0 += 8) {
                const int nc = (int) (N - c0 < 8 ? N - c0 : 8);
                CU(launch_gemv_f32((const float *) w, s0->nb[1], M, K, (const float *) (x + c0 * s1->nb[1]), s1->nb[1], nc, y + c0 * dcs, dcs, c->st)); c->launches++;
            }
        }
    }
    if (q) { rc = rs_exchange(c); if (rc) return rc; }
    // dst may overwrite the memory the cached quantisation was taken from
    {
        const char * lo = (const char *) d->data; const char * hi = lo + d->data_len;
        const char * qp = (const char *) c->q_ptr;
        const char * qe = qp ? qp + (c->q_N > 1 ? (c->q_N - 1) * c->q_stride : 0) + c->q_K * 4 : nullptr;
        if (qp && (d->data_len <= 0 || (qp < hi && lo < qe))) c->q_ptr = nullptr;
    }
    return 0;
}

// ---- graph-level fusion helpers (b200_graph_compute only)
static thread_local int g_rs_world = 1;     // > 1 inside b200_graph_compute_rowsplit: MUL_MAT dst has world x the rows of src0
static bool is_vec_mulmat(const b200_node & n) {       // quantised W x one f32 column, plain 2-D, device tensors
    return n.dst.op == B200_OP_MUL_MAT && is_qweight(n.src0.type) && (n.src0.flags & B200_TENSOR_REPACKED) &&
           n.src1.ne[1] == 1 && n.src1.ne[2] == 1 && n.src1.ne[3] == 1 && n.src0.ne[2] == 1 && n.src0.ne[3] == 1;
}
static bool overlap(const void * a, int64_t na, const void * b, int64_t nb) {
    const char * x = (const char *) a; const char * y = (const char *) b;
    return x < y + nb && y < x + na;
}
static bool is_vec_f32(const b200_tensor & t, int64_t n) {
    return t.type == T_F32 && t.ne[0] == n && t.ne[1] == 1 && t.ne[2] == 1 && t.ne[3] == 1 && t.nb[0] == 4;
}
// MUL_MAT node i followed by ADD node i+1 that consumes its dst: one launch writes both results
static bool can_fuse_add(const b200_node & mm, const b200_node & ad, FusedAdd * fa) {
    if (g_rs_world > 1 || !is_vec_mulmat(mm) || ad.dst.op != B200_OP_ADD) return false;
    const int64_t M = mm.src0.ne[1];
    if (!is_vec_f32(ad.src0, M) || !is_vec_f32(ad.src1, M) || !is_vec_f32(ad.dst, M) || !is_vec_f32(mm.dst, M)) return false;
    const void * other;
    if (ad.src0.data == mm.dst.data) other = ad.src1.data; else if (ad.src1.data == mm.dst.data) other = ad.src0.data; else return false;
    const int64_t bytes = M * 4, K4 = mm.src0.ne[0] * 4;
    // the add result must not land on anything other CTAs of the same launch still read or write
    // (the ADD in place on the product -- ad.dst == mm.dst exactly, as ggml-alloc lays out `cur = ggml_add(cur, residual)` -- is fine: the
    //  thread that stores the product stores the sum to the same address right after it)
    if (overlap(ad.dst.data, bytes, mm.src1.data, K4) || (ad.dst.data != mm.dst.data && overlap(ad.dst.data, bytes, mm.dst.data, bytes))) return false;
    if (other != ad.dst.data && overlap(ad.dst.data, bytes, other, bytes)) return false;
    if (overlap(mm.dst.data, bytes, other, bytes) || overlap(mm.dst.data, bytes, mm.src1.data, K4)) return false;
    fa->src = (const float *) other; fa->dst = (float *) ad.dst.data;
    return true;
}
// consecutive vector MUL_MATs over the same activation vector and weight type: one grouped launch (q/k/v, gate/up)
static int group_len(const b200_node * nodes, int i, int n) {
    if (!is_vec_mulmat(nodes[i])) return 1;
    int len = 1;
    while (len < 3 && i + len < n) {
        const b200_node & a = nodes[i]; const b200_node & b = nodes[i + len];
        if (!is_vec_mulmat(b) || b.src0.type != a.src0.type || b.src1.data != a.src1.data || b.src0.ne[0] != a.src0.ne[0] ||
            b.src1.nb[0] != 4 || !is_vec_f32(b.dst, b.src0.ne[1] * g_rs_world)) break;
        // no member may write what another member (or the shared activation) occupies
        const int64_t W4 = 4 * g_rs_world;
        bool clash = overlap(b.dst.data, b.src0.ne[1] * W4, a.src1.data, a.src0.ne[0] * 4);
        for (int j = 0; j < len && !clash; j++) clash = overlap(b.dst.data, b.src0.ne[1] * W4, nodes[i + j].dst.data, nodes[i + j].src0.ne[1] * W4);
        if (clash) break;
        len++;
    }
    if (len > 1 && (!is_vec_f32(nodes[i].dst, nodes[i].src0.ne[1] * g_rs_world) ||
                    overlap(nodes[i].dst.data, nodes[i].src0.ne[1] * 4 * g_rs_world, nodes[i].src1.data, nodes[i].src0.ne[0] * 4))) return 1;
    return len;
}
// consecutive many-column MUL_MATs over the same activations on the dense GEMM (prefill q / k / v, gate / up): one grouped launch of the
// CTA-pair kernel.  One 4096-row matrix has 32 tiles for 74 CTA pairs; the group's tile list fills the waves and needs no split-K.
static bool is_dense_mulmat(Ctx * c, const b200_node & n) {
    if (n.dst.op != B200_OP_MUL_MAT || !is_qweight(n.src0.type) || !(n.src0.flags & B200_TENSOR_REPACKED) || c->rs_on || g_rs_world > 1) return false;
    const b200_tensor & s0 = n.src0; const b200_tensor & s1 = n.src1;
    if (s0.ne[2] != 1 || s0.ne[3] != 1 || s1.ne[2] != 1 || s1.ne[3] != 1 || s1.nb[0] != 4 || n.dst.nb[0] != 4) return false;
    const int64_t K = s0.ne[0], M = s0.ne[1], N = s1.ne[1];
    const int mn = gemm_min_cols(s0.type);
    return mn > 0 && N >= mn && K % 32 == 0 && M > 0 && M < 0x7fffffff && N < 0x7fffffff && K < 0x3fffffff;
}
// many-column MUL_MAT on the dense GEMM with split-K, followed by the ADD that consumes its dst (wo / down + residual in prefill): the
// split-K reduce writes both results, the ADD launch disappears
static bool can_fuse_add_dense(Ctx * c, const b200_node & mm, const b200_node & ad, FusedAdd * fa) {
    if (!gemm_dense_enabled() || !is_dense_mulmat(c, mm) || ad.dst.op != B200_OP_ADD) return false;
    const int64_t K = mm.src0.ne[0], M = mm.src0.ne[1], N = mm.src1.ne[1];
    if (gemm_dense_ksplit((int) M, (int) N, (int) K, c->sms) <= 1) return false;
    auto plain = [&](const b200_tensor & t) { return t.type == T_F32 && t.ne[0] == M && t.ne[1] == N && t.ne[2] == 1 && t.ne[3] == 1 && t.nb[0] == 4 && t.nb[1] == M * 4; };
    if (!plain(ad.src0) || !plain(ad.src1) || !plain(ad.dst) || !plain(mm.dst)) return false;
    const void * other;
    if (ad.src0.data == mm.dst.data) other = ad.src1.data; else if (ad.src1.data == mm.dst.data) other = ad.src0.data; else return false;
    const int64_t bytes = M * N * 4, xspan = (N - 1) * mm.src1.nb[1] + K * 4;
    if (overlap(ad.dst.data, bytes, mm.src1.data, xspan) || overlap(ad.dst.data, bytes, mm.dst.data, bytes)) return false;
    if (other != ad.dst.data && overlap(ad.dst.data, bytes, other, bytes)) return false;      // in place on the residual is fine (element-wise)
    if (overlap(mm.dst.data, bytes, other, bytes) || overlap(mm.dst.data, bytes, mm.src1.data, xspan)) return false;
    if (!reduce_splits_can_fuse_residual(M, M, mm.dst.data, other, ad.dst.data)) return false;
    fa->src = (const float *) other; fa->dst = (float *) ad.dst.data;
    return true;
}
static int dense_group_len(Ctx * c, const b200_node * nodes, int i, int n) {
    if (!gemm_dense_enabled() || !is_dense_mulmat(c, nodes[i])) return 1;
    const b200_node & a = nodes[i];
    const int64_t K = a.src0.ne[0], N = a.src1.ne[1];
    const int64_t xspan = (N - 1) * a.src1.nb[1] + K * 4;
    auto dspan = [](const b200_node & x) { return (int64_t) ((x.src1.ne[1] - 1) * x.dst.nb[1] + x.src0.ne[1] * 4); };
    if (overlap(a.dst.data, dspan(a), a.src1.data, xspan)) return 1;
    int len = 1;
    while (len < 3 && i + len < n) {
        const b200_node & b = nodes[i + len];
        if (!is_dense_mulmat(c, b) || b.src0.type != a.src0.type || b.src1.data != a.src1.data || b.src0.ne[0] != K || b.src1.ne[1] != N || b.src1.nb[1] != a.src1.nb[1]) break;
        // no member may write what another member (or the shared activations) occupies
        bool clash = overlap(b.dst.data, dspan(b), a.src1.data, xspan);
        for (int j = 0; j < len && !clash; j++) clash = overlap(b.dst.data, dspan(b), nodes[i + j].dst.data, dspan(nodes[i + j]));
        if (clash) break;
        len++;
    }
    return len;
}
static int mulmat_dev(Ctx * c, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d, int64_t row_off_dst, const FusedAdd * fa);
static int mulmat_dense_group_dev(Ctx * c, const b200_node * nodes, int len) {
    const int64_t K = nodes[0].src0.ne[0], N = nodes[0].src1.ne[1];
    WPtrs A[3]; float * out[3]; long long ocs[3]; int Ms[3];
    for (int j = 0; j < len; j++) {
        b200_tensor d = nodes[j].dst;
        int rc = mulmat_check(&nodes[j].src0, &nodes[j].src1, &d); if (rc) return rc;
        rc = wptrs_of(&nodes[j].src0, &A[j]); if (rc) return rc;
        out[j] = (float *) d.data; ocs[j] = d.nb[1] / 4; Ms[j] = (int) nodes[j].src0.ne[1];
    }
    int rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, dense_act_bytes(N, K)); if (rc) return rc;
    float * const colscale = (float *) (c->gm_qb + dense_act_cs_off(N, K));
    const b200_tensor & s1 = nodes[0].src1;
    const bool reuse = c->q_kind == 1 && c->q_ptr == (const void *) s1.data && c->q_K == K && c->q_N == N && c->q_stride == s1.nb[1];
    if (!reuse) {
        CU(launch_dequant_act_f16((const float *) s1.data, s1.nb[1], (int) K, (int) N, (__half *) c->gm_qb, colscale, c->st)); c->launches++;
        c->q_ptr = s1.data; c->q_K = K; c->q_N = N; c->q_stride = s1.nb[1]; c->q_kind = 1;
    }
    rc = ensure_buf(c, &c->gm_tail, &c->gm_tail_cap, gemm_dense_tail_ws_bytes(c->sms)); if (rc) return rc;
    int nl = 0;
    CU(launch_gemm_dense_group(nodes[0].src0.type, A, out, ocs, Ms, len, (const __half *) c->gm_qb, colscale, (int) N, (int) K, c->sms, c->gm_tail, &nl, c->st)); c->launches += nl;
    return 0;           // (dense_group_len made sure that no dst touches the activations the cached conversion came from)
}

static int mulmat_group_dev(Ctx * c, const b200_node * nodes, int len) {
    GemvArgs a{};
    const int64_t K = nodes[0].src0.ne[0], NB = K / 32;
    a.nmat = len; a.K = K; a.dcs = 0;
    const float * x = (const float *) nodes[0].src1.data;
    const void * dsts[3]; int64_t dbytes[3];
    for (int j = 0; j < len; j++) {
        b200_tensor dl = nodes[j].dst; dl.ne[0] = nodes[j].src0.ne[1];            // row split: this rank's slice of dst
        int rc = mulmat_check(&nodes[j].src0, &nodes[j].src1, &dl); if (rc) return rc;
        WPtrs wp; rc = wptrs_of(&nodes[j].src0, &wp); if (rc) return rc;
        a.mat[j].w = wp; a.mat[j].M = nodes[j].src0.ne[1];
        a.mat[j].dst = (float *) nodes[j].dst.data + (c->rs_on ? c->rank * nodes[j].src0.ne[1] : 0);
        dsts[j] = a.mat[j].dst; dbytes[j] = a.mat[j].M * 4;
    }
    { int64_t rw[3]; for (int j = 0; j < len; j++) rw[j] = a.mat[j].M;
      int rc = rs_fill(c, a, dsts, dbytes, len, rw, true); if (rc) return rc; }
    {   // persistent ring GEMV over the whole group, activations quantised inside the kernel (gemv6.cu)
        a.x = x; a.x_col_stride = nodes[0].src1.nb[1];
        cudaError_t e6 = cudaSuccess;
        if (launch_gemv6(nodes[0].src0.type, a, c->sms, c->st, &e6)) {
            CU(e6); c->launches++;
            for (int j = 0; j < len; j++)            // a dst may overwrite the memory an earlier cached quantisation was taken from
                if (c->q_ptr && overlap(nodes[j].dst.data, nodes[j].src0.ne[1] * 4 * g_rs_world, c->q_ptr, c->q_K * 4)) c->q_ptr = nullptr;
            return rs_exchange(c);
        }
    }
    int rc = ensure_act(c, NB); if (rc) return rc;
    const bool reuse = c->q_kind == 0 && c->q_ptr == (const void *) x && c->q_K == K && c->q_N == 1;
    if (!reuse) {
        CU(launch_quantize_q8_1_soa(x, K * 4, K, 1, c->act_qs, c->act_ds, c->st)); c->launches++;
        c->q_ptr = x; c->q_K = K; c->q_N = 1; c->q_stride = nodes[0].src1.nb[1]; c->q_kind = 0;
    }
    a.act = ActView{c->act_qs, c->act_ds, NB};
    qx_arm(c, a);
    CU(launch_gemv2(nodes[0].src0.type, a, 1, c->st)); c->launches++;
    qx_launched(c, a);
    { int rc2 = rs_exchange(c); if (rc2) return rc2; }
    for (int j = 0; j < len; j++)
        if (c->q_ptr && overlap(nodes[j].dst.data, nodes[j].src0.ne[1] * 4, c->q_ptr, K * 4)) c->q_ptr = nullptr;
    return 0;
}

// ---- host-pointer tensors: stage in, run, stage out, synchronise (FastRPC semantics: dsptensor.data is host memory)
static size_t tensor_span(const b200_tensor * t) {
    if (t->data_len > 0) return (size_t) t->data_len;
    size_t s = 0;
    if (t->type == T_F32) { s = 4; for (int i = 0; i < 4; i++) s += (size_t) (t->ne[i] - 1) * t->nb[i]; }
    else { s = (size_t) t->nb[1] * t->ne[1]; if (t->ne[2] > 1) s = (size_t) t->nb[2] * t->ne[2]; if (t->ne[3] > 1) s = (size_t) t->nb[3] * t->ne[3]; }
    return s;
}

static int run_with_staging(Ctx * c, int op, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    const bool h0 = s0->flags & B200_TENSOR_HOST, h1 = s1->flags & B200_TENSOR_HOST, hd = d->flags & B200_TENSOR_HOST;
    if (!(h0 || h1 || hd)) return op == B200_OP_ADD ? add_dev(c, s0, s1, d) : mulmat_dev(c, s0, s1, d, 0);
    b200_tensor t0 = *s0, t1 = *s1, td = *d;
    const size_t n0 = h0 ? tensor_span(s0) : 0, n1 = h1 ? tensor_span(s1) : 0, nd = hd ? tensor_span(d) : 0;
    auto al = [](size_t v) { return (v + 255) & ~(size_t) 255; };
    const bool q0 = h0 && is_qweight(s0->type);
    // layout of the staging buffer: [src1][dst][src0 SoA][src0 native]
    const size_t o1 = 0, od = o1 + al(n1), o0 = od + al(nd), o0n = o0 + al(n0);
    int rc = ensure_stage(c, o0n + (q0 ? al(n0) : 0) + 256); if (rc) return rc;
    if (h1) { CU(cudaMemcpyAsync(c->stage + o1, s1->data, n1, cudaMemcpyHostToDevice, c->st)); t1.data = c->stage + o1; t1.flags &= ~B200_TENSOR_HOST; }
    if (h0) {
        if (q0) {
            const int64_t bb = block_bytes(s0->type);
            if (n0 % bb) return fail(B200_EINVAL, "host src0 span is not a whole number of blocks");
            const int64_t T = (int64_t) n0 / bb;
            CU(cudaMemcpyAsync(c->stage + o0n, s0->data, n0, cudaMemcpyHostToDevice, c->st));
            CU(launch_repack(s0->type, true, c->stage + o0, T, 0, T, c->stage + o0n, c->st)); c->launches++;
            t0.data = c->stage + o0; t0.w_base = c->stage + o0; t0.w_blocks = T; t0.w_first = 0;
            t0.flags = (t0.flags & ~B200_TENSOR_HOST) | B200_TENSOR_REPACKED;
        } else {
            CU(cudaMemcpyAsync(c->stage + o0, s0->data, n0, cudaMemcpyHostToDevice, c->st));
            t0.data = c->stage + o0; t0.flags &= ~B200_TENSOR_HOST;
        }
    }
    if (hd) { td.data = c->stage + od; td.flags &= ~B200_TENSOR_HOST; td.data_len = (int64_t) nd; }
    c->q_ptr = nullptr;
    rc = op == B200_OP_ADD ? add_dev(c, &t0, &t1, &td) : mulmat_dev(c, &t0, &t1, &td, 0);
    c->q_ptr = nullptr;
    if (rc) return rc;
    if (hd) CU(cudaMemcpyAsync(d->data, c->stage + od, nd, cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return 0;
}

extern "C" int b200_add(b200_handle h, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    CTX(h); if (!s0 || !s1 || !d) return fail(B200_EINVAL, "NULL tensor");
    return run_with_staging(c, B200_OP_ADD, s0, s1, d);
}
extern "C" int b200_mulmat(b200_handle h, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    CTX(h); if (!s0 || !s1 || !d) return fail(B200_EINVAL, "NULL tensor");
    return run_with_staging(c, B200_OP_MUL_MAT, s0, s1, d);
}

// ---- producer fusions of a Llama layer (graph_ops.cu).  Each one replaces launches the node list asks for by ONE kernel with the same
// roundings, and only where the skipped intermediate is overwritten in place by the pair's second node (so nothing can observe it).
static bool graph_fuse_enabled() { const char * e = getenv("B200_GRAPH_FUSE"); return !(e && atoi(e) == 0); }      // (read per call: tests flip it)
static bool same_shape(const b200_tensor & a, const b200_tensor & b) { for (int i = 0; i < 4; i++) if (a.ne[i] != b.ne[i]) return false; return true; }
static bool al16(const void * p) { return ((uintptr_t) p & 15) == 0; }
// a quantised MUL_MAT that reads the single f32 column x[K] next: its q8_1 activations can be written by the producer of x
static bool reads_vec_q(const b200_node & nd, const void * x, int64_t K) {
    return nd.dst.op == B200_OP_MUL_MAT && is_qweight(nd.src0.type) && nd.src1.type == T_F32 && nd.src1.data == x && nd.src0.ne[0] == K &&
           nd.src1.ne[0] == K && nd.src1.ne[1] == 1 && nd.src1.ne[2] == 1 && nd.src1.ne[3] == 1 && nd.src1.nb[0] == 4;
}
static void set_qcache(Ctx * c, const b200_tensor & d) { c->q_ptr = d.data; c->q_K = d.ne[0]; c->q_N = 1; c->q_stride = d.nb[1]; c->q_kind = 0; }
// a dense-GEMM MUL_MAT that reads the N columns x[K] (rows K * 4 bytes apart) next: its fp16 activations + column scales can be written by the
// producer of x (gemm_dense.cu row_act_kernel).  Returns the scratch to write into (room made), or NULL.
static bool reads_cols_dense(Ctx * c, const b200_node & nd, const void * x, int64_t K, int64_t N) {
    return gemm_dense_enabled() && is_dense_mulmat(c, nd) && nd.src1.type == T_F32 && nd.src1.data == x && nd.src0.ne[0] == K && nd.src1.ne[0] == K &&
           nd.src1.ne[1] == N && nd.src1.nb[1] == K * 4 && K % 32 == 0 && K <= ROW_ACT_MAX_K;
}
static int dense_act_room(Ctx * c, int64_t N, int64_t K, void ** act, float ** cs) {
    int rc = ensure_buf(c, &c->gm_qb, &c->gm_qb_cap, dense_act_bytes(N, K)); if (rc) return rc;
    *act = c->gm_qb; *cs = (float *) (c->gm_qb + dense_act_cs_off(N, K));
    return 0;
}
static void set_qcache_dense(Ctx * c, const void * x, int64_t K, int64_t N) { c->q_ptr = x; c->q_K = K; c->q_N = N; c->q_stride = K * 4; c->q_kind = 1; }
// RMS_NORM x -> X ; MUL X, w -> X (in place, w one row): one kernel; [+ q8_1 of X when a vector MUL_MAT reads it next]
static int try_rms_norm_mul(Ctx * c, const b200_node * nodes, int i, int n, int * used) {
    if (!graph_fuse_enabled() || c->rs_on || i + 1 >= n) return 0;
    const b200_node & r = nodes[i]; const b200_node & m = nodes[i + 1];
    if (m.dst.op != B200_OP_MUL || m.src0.data != r.dst.data || m.dst.data != r.dst.data) return 0;
    if (r.src0.type != T_F32 || r.dst.type != T_F32 || m.src1.type != T_F32 || m.dst.type != T_F32) return 0;
    if (!same_shape(r.src0, r.dst) || !same_shape(m.dst, r.dst) || !same_shape(m.src0, r.dst)) return 0;
    const int64_t ne00 = r.dst.ne[0], rows = r.dst.ne[1];
    if (r.dst.ne[2] != 1 || r.dst.ne[3] != 1 || ne00 % 4 || r.src0.nb[0] != 4 || r.dst.nb[0] != 4 || m.dst.nb[0] != 4 || m.dst.nb[1] != r.dst.nb[1]) return 0;
    if (m.src1.ne[0] != ne00 || m.src1.ne[1] != 1 || m.src1.ne[2] != 1 || m.src1.ne[3] != 1 || m.src1.nb[0] != 4) return 0;
    if (!al16(r.src0.data) || !al16(r.dst.data) || !al16(m.src1.data) || (r.src0.nb[1] & 15) || (r.dst.nb[1] & 15)) return 0;
    const float eps = f32_param(&r.dst, 0);
    if (!(eps >= 0.0f)) return 0;
    if (rows > 1 && i + 2 < n && r.dst.nb[1] == ne00 * 4 && reads_cols_dense(c, nodes[i + 2], r.dst.data, ne00, rows)) {
        void * act; float * cs; int rc = dense_act_room(c, rows, ne00, &act, &cs); if (rc) return rc;
        RowActArgs a{};
        a.a = (const char *) r.src0.data; a.as = r.src0.nb[1]; a.b = (const char *) m.src1.data; a.y = (char *) r.dst.data; a.ys = r.dst.nb[1];
        a.act = act; a.colscale = cs; a.K = (int) ne00; a.eps = eps;
        CU(launch_row_act(0, a, (int) rows, c->st)); c->launches++;
        wrote(c, &r.dst);
        set_qcache_dense(c, r.dst.data, ne00, rows);
        *used = 2;
        return 0;
    }
    const bool quant = rows == 1 && ne00 % 128 == 0 && i + 2 < n && reads_vec_q(nodes[i + 2], r.dst.data, ne00);
    if (quant) { int rc = ensure_act(c, ne00 / 32); if (rc) return rc; }
    CU(launch_rms_norm_mul((const float *) r.src0.data, (const float *) m.src1.data, (float *) r.dst.data, ne00, rows, r.src0.nb[1] / 4, r.dst.nb[1] / 4, eps,
                           quant ? c->act_qs : nullptr, quant ? c->act_ds : nullptr, c->st)); c->launches++;
    wrote(c, &r.dst);
    if (quant) set_qcache(c, m.dst);
    *used = 2;
    return 0;
}
// SILU(G) * U -> X for N > 1 columns of K that a dense GEMM reads next: one kernel that also converts X; false = not that case
static int silu_mul_cols(Ctx * c, const b200_node * nodes, int next, int n, const b200_tensor & x, const void * g, const void * u, bool * done) {
    *done = false;
    const int64_t K = x.ne[0], N = x.ne[1];
    if (N <= 1 || x.ne[2] != 1 || x.ne[3] != 1 || next >= n || !reads_cols_dense(c, nodes[next], x.data, K, N)) return 0;
    void * act; float * cs; int rc = dense_act_room(c, N, K, &act, &cs); if (rc) return rc;
    RowActArgs a{};
    a.a = (const char *) g; a.as = K * 4; a.b = (const char *) u; a.bs = K * 4; a.y = (char *) x.data; a.ys = K * 4; a.act = act; a.colscale = cs; a.K = (int) K;
    CU(launch_row_act(1, a, (int) N, c->st)); c->launches++;
    wrote(c, &x);
    set_qcache_dense(c, x.data, K, N);
    *done = true;
    return 0;
}
// SILU G -> X ; [MUL_MAT ... -> U] ; MUL X, U -> X (in place): the SILU is folded into the MUL; [+ q8_1 of X for the vector MUL_MAT that follows]
static int try_silu_mul(Ctx * c, const b200_node * nodes, int i, int n, int * used, int * deferred_mul) {
    if (!graph_fuse_enabled() || c->rs_on) return 0;
    const b200_node & sn = nodes[i];
    if (sn.dst.op_params[0] != B200_UNARY_SILU || sn.src0.type != T_F32 || sn.dst.type != T_F32) return 0;
    int j = i + 1;
    if (j < n && nodes[j].dst.op == B200_OP_MUL_MAT) {          // the up projection in between must not touch G or X
        const b200_node & u = nodes[j];
        const int64_t xs = nelems(&sn.dst) * 4, us = nelems(&u.dst) * 4;
        if (u.src1.data == sn.dst.data || overlap(u.dst.data, us, sn.dst.data, xs) || overlap(u.dst.data, us, sn.src0.data, xs)) return 0;
        j++;
    }
    if (j >= n) return 0;
    const b200_node & m = nodes[j];
    if (m.dst.op != B200_OP_MUL || m.src0.data != sn.dst.data || m.dst.data != sn.dst.data) return 0;
    if (m.src1.type != T_F32 || m.dst.type != T_F32 || !same_shape(m.src0, sn.dst) || !same_shape(m.src1, sn.dst) || !same_shape(m.dst, sn.dst) || !same_shape(sn.src0, sn.dst)) return 0;
    if (!contiguous_f32_t(&sn.src0) || !contiguous_f32_t(&sn.dst) || !contiguous_f32_t(&m.src1)) return 0;
    const int64_t tot = nelems(&sn.dst);
    if (tot % 4 || !al16(sn.src0.data) || !al16(sn.dst.data) || !al16(m.src1.data)) return 0;
    if (overlap(m.src1.data, tot * 4, sn.dst.data, tot * 4)) return 0;
    if (j == i + 1) {           // adjacent: run the fused kernel now
        bool done = false; int rc0 = silu_mul_cols(c, nodes, j + 1, n, m.dst, sn.src0.data, m.src1.data, &done); if (rc0) return rc0;
        if (done) { *used = 2; return 0; }
        const int64_t K = sn.dst.ne[0];
        const bool quant = tot == K && K % 128 == 0 && j + 1 < n && reads_vec_q(nodes[j + 1], m.dst.data, K);
        if (quant) { int rc = ensure_act(c, K / 32); if (rc) return rc; }
        CU(launch_silu_mul_f32((const float *) sn.src0.data, (const float *) m.src1.data, (float *) m.dst.data, tot, K, quant ? c->act_qs : nullptr, quant ? c->act_ds : nullptr, c->st)); c->launches++;
        wrote(c, &m.dst);
        if (quant) set_qcache(c, m.dst);
        *used = 2;
    } else {                    // one MUL_MAT in between: skip the SILU now, the MUL node (index j) runs the fused kernel
        *deferred_mul = j; *used = 1;
    }
    return 0;
}
// MUL_MAT(K view f16, q) -> SOFT_MAX in place (mask, scale) -> MUL_MAT(V^T view f16, p), one query column: one launch per layer instead of three
// (graph_ops.cu attn_decode_kernel; bit-equal to the three kernels, whose vectorised paths it requires)
static int64_t attn_decode_max_kv() { const char * e = getenv("B200_ATTN_FUSE_MAX_KV"); const int64_t v = e ? atoll(e) : 1024; return v < AD_MAX_KV ? v : AD_MAX_KV; }
static int try_attn_decode(Ctx * c, const b200_node * nodes, int i, int n, int * used) {
    if (!graph_fuse_enabled() || c->rs_on || i + 2 >= n) return 0;
    const b200_node & kq = nodes[i]; const b200_node & sm = nodes[i + 1]; const b200_node & pv = nodes[i + 2];
    if (sm.dst.op != B200_OP_SOFT_MAX || pv.dst.op != B200_OP_MUL_MAT || pv.src0.type != T_F16 || kq.src0.type != T_F16) return 0;
    if (kq.src1.type != T_F32 || kq.dst.type != T_F32 || pv.src1.type != T_F32 || pv.dst.type != T_F32 || sm.src0.type != T_F32 || sm.dst.type != T_F32) return 0;
    const int64_t hd = kq.src0.ne[0], n_kv = kq.src0.ne[1], nhk = kq.src0.ne[2], nh = kq.src1.ne[2];
    if (kq.src0.ne[3] != 1 || kq.src1.ne[3] != 1 || kq.src1.ne[1] != 1 || kq.src1.ne[0] != hd || nhk <= 0 || nh % nhk) return 0;
    if (hd % 2 || n_kv % 2 || n_kv <= 0 || n_kv > attn_decode_max_kv() || nh >= 512) return 0;       // (longer contexts: the three kernels spread over more CTAs than one per head)
    if (kq.src0.nb[0] != 2 || kq.src1.nb[0] != 4) return 0;
    const b200_tensor & s = kq.dst;            // scores / probabilities [n_kv, 1, nh], contiguous
    if (s.ne[0] != n_kv || s.ne[1] != 1 || s.ne[2] != nh || s.ne[3] != 1 || s.nb[0] != 4 || s.nb[2] != n_kv * 4) return 0;
    if (sm.src0.data != s.data || sm.dst.data != s.data || !same_shape(sm.src0, s) || !same_shape(sm.dst, s) || !contiguous_f32_t(&sm.src0) || !contiguous_f32_t(&sm.dst)) return 0;
    const b200_tensor & m = sm.src1;
    if (!m.data || m.type != T_F32 || m.ne[0] != n_kv || m.ne[1] < 1 || m.nb[0] != 4 || m.nb[1] != n_kv * 4) return 0;
    if (f32_param(&sm.dst, 1) != 0.0f) return 0;                                         // ALiBi slopes: not in a Llama graph
    if (pv.src1.data != s.data || !same_shape(pv.src1, s) || pv.src1.nb[0] != 4 || pv.src1.nb[2] != n_kv * 4) return 0;
    if (pv.src0.ne[0] != n_kv || pv.src0.ne[1] != hd || pv.src0.ne[2] != nhk || pv.src0.ne[3] != 1 || pv.src0.nb[0] != 2) return 0;
    const b200_tensor & o = pv.dst;
    if (o.ne[0] != hd || o.ne[1] != 1 || o.ne[2] != nh || o.ne[3] != 1 || o.nb[0] != 4) return 0;
    // the alignment under which the separate kernels take their 2-element paths
    auto a4 = [](const void * p, int64_t x, int64_t y) { return (((uintptr_t) p | (uintptr_t) x | (uintptr_t) y) & 3) == 0; };
    auto a8 = [](const void * p, int64_t x, int64_t y) { return (((uintptr_t) p | (uintptr_t) x | (uintptr_t) y) & 7) == 0; };
    if (!a4(kq.src0.data, kq.src0.nb[1], kq.src0.nb[2]) || !a4(pv.src0.data, pv.src0.nb[1], pv.src0.nb[2])) return 0;
    if (!a8(kq.src1.data, kq.src1.nb[1], kq.src1.nb[2]) || !a8(s.data, s.nb[1], s.nb[2])) return 0;
    AttnDecodeArgs a{};
    a.k = (const char *) kq.src0.data; a.v = (const char *) pv.src0.data; a.q = (const char *) kq.src1.data; a.mask = (const float *) m.data;
    a.prob = (char *) s.data; a.out = (char *) o.data;
    a.hd = hd; a.n_kv = n_kv; a.r2 = nh / nhk; a.nbk1 = kq.src0.nb[1]; a.nbk2 = kq.src0.nb[2]; a.nbv1 = pv.src0.nb[1]; a.nbv2 = pv.src0.nb[2];
    a.nbq2 = kq.src1.nb[2]; a.nbp2 = s.nb[2]; a.nbo2 = o.nb[2]; a.scale = f32_param(&sm.dst, 0);
    CU(launch_attn_decode(a, (int) nh, c->st)); c->launches++;
    wrote(c, &s); wrote(c, &o);
    *used = 3;
    return 0;
}
// ROPE -> X (f32, contiguous) ... CPY X -> f16 cache rows, flat (llama-graph.cpp:1341-1345, with the V projection ordered in between): the
// ROPE kernel stores the f16 row as well and the CPY node is skipped when the loop reaches it.  `*cpy_at` = index of that node, `*slot` =
// its position among the list's dynamic destinations.
static bool rope_store_target(Ctx * c, const b200_node * nodes, int i, int n, int * cpy_at, int * slot) {
    if (!graph_fuse_enabled() || c->rs_on) return false;
    const b200_node & r = nodes[i];
    if (r.dst.type != T_F32 || !contiguous_f32_t(&r.dst)) return false;
    const int64_t E = nelems(&r.dst);
    int ndyn = 0;
    for (int j = i + 1; j < n && j <= i + 4; j++) {
        const b200_node & x = nodes[j];
        if (x.dst.op == B200_OP_CPY && x.src0.data == r.dst.data) {
            if (x.src0.type != T_F32 || x.dst.type != T_F16 || nelems(&x.src0) != E || nelems(&x.dst) != E || !contiguous_f32_t(&x.src0)) return false;
            int64_t st = 2;                                   // dst: f16, contiguous in logical order
            for (int k = 0; k < 4; k++) { if (x.dst.ne[k] != 1 && x.dst.nb[k] != st) return false; st *= x.dst.ne[k]; }
            if (overlap(x.dst.data, E * 2, r.dst.data, E * 4) || overlap(x.dst.data, E * 2, r.src0.data, E * 4)) return false;
            *cpy_at = j; *slot = ndyn;
            return true;
        }
        // what may sit in between: a vector / matrix MUL_MAT over quantised weights that touches neither X nor the cache rows
        if (x.dst.op != B200_OP_MUL_MAT || !is_qweight(x.src0.type) || x.src1.data == r.dst.data) return false;
        int64_t span = 4; for (int k = 0; k < 4; k++) span += (x.dst.ne[k] - 1) * x.dst.nb[k];
        if (overlap(x.dst.data, span, r.dst.data, E * 4)) return false;
        if (x.dst.flags & B200_TENSOR_DYNAMIC) ndyn++;
    }
    return false;
}
// CONT -> contiguous X[K] read next by a vector MUL_MAT: copy + q8_1 in one kernel
static int try_cont_quant(Ctx * c, const b200_node * nodes, int i, int n, int * used) {
    if (!graph_fuse_enabled() || c->rs_on || i + 1 >= n) return 0;
    const b200_node & cn = nodes[i];
    if (cn.src0.type != T_F32 || cn.dst.type != T_F32 || (cn.dst.flags & B200_TENSOR_DYNAMIC)) return 0;
    const int64_t K = cn.dst.ne[0];
    if (cn.dst.ne[1] > 1 && cn.dst.ne[2] == 1 && cn.dst.ne[3] == 1 && contiguous_f32_t(&cn.dst) && nelems(&cn.src0) == nelems(&cn.dst) &&
        cn.src0.nb[0] == 4 && cn.src0.ne[0] % 4 == 0 && al16(cn.src0.data) && al16(cn.dst.data) && !((cn.src0.nb[1] | cn.src0.nb[2] | cn.src0.nb[3]) & 15) &&
        reads_cols_dense(c, nodes[i + 1], cn.dst.data, K, cn.dst.ne[1])) {
        const int64_t N = cn.dst.ne[1];
        void * act; float * cs; int rc = dense_act_room(c, N, K, &act, &cs); if (rc) return rc;
        RowActArgs a{};
        a.a = (const char *) cn.src0.data; a.y = (char *) cn.dst.data; a.ys = K * 4; a.act = act; a.colscale = cs; a.K = (int) K;
        for (int k = 0; k < 4; k++) { a.nes[k] = cn.src0.ne[k]; a.nbs[k] = cn.src0.nb[k]; }
        CU(launch_row_act(2, a, (int) N, c->st)); c->launches++;
        wrote(c, &cn.dst);
        set_qcache_dense(c, cn.dst.data, K, N);
        *used = 1;
        return 0;
    }
    if (nelems(&cn.dst) != K || nelems(&cn.src0) != K || K % 128 || !contiguous_f32_t(&cn.dst) || !reads_vec_q(nodes[i + 1], cn.dst.data, K)) return 0;
    if (cn.src0.nb[0] != 4 || cn.src0.ne[0] % 4 || !al16(cn.src0.data) || !al16(cn.dst.data) || ((cn.src0.nb[1] | cn.src0.nb[2] | cn.src0.nb[3]) & 15)) return 0;
    int rc = ensure_act(c, K / 32); if (rc) return rc;
    CpyArgs a{};
    a.x = (const char *) cn.src0.data; a.y = (char *) cn.dst.data;
    for (int k = 0; k < 4; k++) { a.nes[k] = cn.src0.ne[k]; a.nbs[k] = cn.src0.nb[k]; a.ned[k] = cn.dst.ne[k]; a.nbd[k] = cn.dst.nb[k]; }
    CU(launch_cont_quant(a, K, c->act_qs, c->act_ds, c->st)); c->launches++;
    wrote(c, &cn.dst);
    set_qcache(c, cn.dst);
    *used = 1;
    return 0;
}

// ------------------------------------------------------------------------------------------------ graph compute
static int run_nodes(Ctx * c, const b200_node * nodes, int n) {
    static const bool fuse = !(getenv("B200_FUSE") && atoi(getenv("B200_FUSE")) == 0);
    c->in_graph = true; c->q_ptr = nullptr; c->rs_ops = 0; c->dyn_i = 0;
    int rc = 0;
    int silu_mul_at = -1;            // index of a MUL node that also applies the SILU skipped two nodes earlier (try_silu_mul)
    int cpy_done_at = -1;            // index of a CPY node whose store the ROPE before it has already done (rope_store_target)
    for (int i = 0; i < n && !rc; ) {
        b200_tensor d = nodes[i].dst;
        int used = 1;
        if (silu_mul_at >= 0 && i > silu_mul_at) { rc = fail(B200_EINVAL, "internal: node %d that carries a deferred SILU was skipped", silu_mul_at); break; }
        if (cpy_done_at >= 0 && i > cpy_done_at) { rc = fail(B200_EINVAL, "internal: the CPY node %d whose store was done early was never reached", cpy_done_at); break; }
        switch (d.op) {
            case B200_OP_NONE: break;
            case B200_OP_ADD: rc = add_dev(c, &nodes[i].src0, &nodes[i].src1, &d); break;
            case B200_OP_RMS_NORM: {
                int u = 0; rc = try_rms_norm_mul(c, nodes, i, n, &u);
                if (!rc && u) used = u; else if (!rc) rc = rms_norm_dev(c, &nodes[i].src0, &d);
            } break;
            case B200_OP_SOFT_MAX: rc = soft_max_dev(c, &nodes[i].src0, &nodes[i].src1, &d); break;
            case B200_OP_POOL_2D:  rc = pool_2d_dev(c, &nodes[i].src0, &d); break;
            case B200_OP_MUL:
                if (i == silu_mul_at) {         // SILU(G) * U -> X; G is the skipped SILU's input, X its output = this node's src0 = dst
                    const b200_node & sn = nodes[i - 2];
                    silu_mul_at = -1;
                    bool done = false; rc = silu_mul_cols(c, nodes, i + 1, n, d, sn.src0.data, nodes[i].src1.data, &done);
                    if (rc || done) break;
                    const int64_t K = d.ne[0], tot = nelems(&d);
                    const bool quant = tot == K && K % 128 == 0 && i + 1 < n && reads_vec_q(nodes[i + 1], d.data, K);
                    if (quant) rc = ensure_act(c, K / 32);
                    if (!rc) { CU(launch_silu_mul_f32((const float *) sn.src0.data, (const float *) nodes[i].src1.data, (float *) d.data, tot, K,
                                                      quant ? c->act_qs : nullptr, quant ? c->act_ds : nullptr, c->st)); c->launches++; }
                    wrote(c, &d);
                    if (!rc && quant) set_qcache(c, d);
                    silu_mul_at = -1;
                } else rc = mul_dev(c, &nodes[i].src0, &nodes[i].src1, &d);
                break;
            case B200_OP_UNARY: {
                int u = 0, dm = -1; rc = try_silu_mul(c, nodes, i, n, &u, &dm);
                if (!rc && u) { used = u; if (dm >= 0) silu_mul_at = dm; } else if (!rc) rc = unary_dev(c, &nodes[i].src0, &d);
            } break;
            case B200_OP_ROPE: {
                int at = -1, slot = 0;
                if (cpy_done_at < 0 && rope_store_target(c, nodes, i, n, &at, &slot)) {
                    const b200_tensor & cd = nodes[at].dst;
                    const bool dyn = (cd.flags & B200_TENSOR_DYNAMIC) && c->dyn_on;
                    if (dyn && c->dyn_i + slot >= DYN_MAX) { rc = fail(B200_ENOTSUP, "more than %d dynamic destinations in one node list", DYN_MAX); break; }
                    rc = rope_dev(c, &nodes[i].src0, &nodes[i].src1, &nodes[i].src2, &d, (char *) cd.data, dyn ? c->dyn_tab + c->dyn_i + slot : nullptr);
                    // (the store lands in the cache tensor whatever the position: the quantised activations q / k / v share stay valid across it)
                    if (!rc) { cpy_done_at = at; wrote(c, &cd); }
                } else rc = rope_dev(c, &nodes[i].src0, &nodes[i].src1, &nodes[i].src2, &d);
            } break;
            case B200_OP_DUP: case B200_OP_CONT: case B200_OP_CPY: {    // (ggml_cpy(a, b): the node is a view of b, so dst carries b's layout)
                if (i == cpy_done_at) {          // stored by the ROPE that produced its source; only its slot in the dynamic table is consumed
                    if (d.flags & B200_TENSOR_DYNAMIC) c->dyn_i++;
                    cpy_done_at = -1;
                    break;
                }
                int u = 0; if (d.op == B200_OP_CONT) rc = try_cont_quant(c, nodes, i, n, &u);
                if (!rc && !u) rc = cpy_dev(c, &nodes[i].src0, &d);
            } break;
            case B200_OP_GET_ROWS: rc = get_rows_dev(c, &nodes[i].src0, &nodes[i].src1, &d); break;
            case B200_OP_MUL_MAT: {
                if (nodes[i].src0.type == T_F16) {
                    int u = 0; rc = try_attn_decode(c, nodes, i, n, &u);
                    if (!rc && u) used = u; else if (!rc) rc = mm_f16_dev(c, &nodes[i].src0, &nodes[i].src1, &d);
                    break;
                }
                FusedAdd fa;
                const int gl = fuse ? group_len(nodes, i, n) : 1;
                // row split: what consumes this op's gathered vectors next?  [ADD with a plain vector ->] vector MUL_MAT: the gather kernel
                // does the ADD and quantises for that MUL_MAT (rs_exchange), the ADD node is skipped and the MUL_MAT finds its activations cached
                const b200_node * nx = nullptr; bool skip_add = false;
                c->rs_hint.on = false; c->rs_hint.done = false;
                if (fuse && rs_fuse_enabled() && c->rs_on && c->world > 1 && is_vec_mulmat(nodes[i])) {
                    const int j = i + gl;
                    for (int mbr = 0; mbr < gl && !c->rs_hint.on && j < n; mbr++) {
                        const b200_node & me = nodes[i + mbr];
                        const int64_t full = me.src0.ne[1] * c->world;
                        if (!is_vec_f32(me.dst, full)) continue;
                        if (is_vec_mulmat(nodes[j]) && nodes[j].src1.data == me.dst.data && nodes[j].src0.ne[0] == full && nodes[j].src1.nb[0] == 4) {
                            c->rs_hint.on = true; c->rs_hint.of = me.dst.data; c->rs_hint.add_src = nullptr; c->rs_hint.add_dst = nullptr; nx = &nodes[j];
                        } else if (nodes[j].dst.op == B200_OP_ADD && j + 1 < n && is_vec_f32(nodes[j].src0, full) && is_vec_f32(nodes[j].src1, full) && is_vec_f32(nodes[j].dst, full) &&
                                   (nodes[j].src0.data == me.dst.data) != (nodes[j].src1.data == me.dst.data) &&
                                   is_vec_mulmat(nodes[j + 1]) && nodes[j + 1].src1.data == nodes[j].dst.data && nodes[j + 1].src0.ne[0] == full && nodes[j + 1].src1.nb[0] == 4) {
                            const void * other = nodes[j].src0.data == me.dst.data ? nodes[j].src1.data : nodes[j].src0.data;
                            bool clash = false;                  // the sum must not land on another member's gathered vector
                            for (int o = 0; o < gl; o++) if (o != mbr) clash = clash || overlap(nodes[j].dst.data, full * 4, nodes[i + o].dst.data, nodes[i + o].src0.ne[1] * c->world * 4);
                            if (!clash && (((uintptr_t) other | (uintptr_t) nodes[j].dst.data | (uintptr_t) me.dst.data) & 15) == 0) {
                                c->rs_hint.on = true; c->rs_hint.of = me.dst.data; c->rs_hint.add_src = (const float *) other; c->rs_hint.add_dst = (float *) nodes[j].dst.data;
                                nx = &nodes[j + 1]; skip_add = true;
                            }
                        }
                    }
                    if (c->rs_hint.on) { rc = ensure_act(c, nx->src0.ne[0] / 32); if (rc) break; }
                }
                const int dgl = fuse && gl == 1 ? dense_group_len(c, nodes, i, n) : 1;
                // one GPU: which result vector of this op (a member's dst, or the sum of the ADD fused into it) is the next vector MUL_MAT's src1?
                // The GEMV then quantises it on exit (gemv.cu, gemv_quant_on_exit) and that MUL_MAT finds its activations cached.
                c->qx_hint.on = false; c->qx_hint.done = false;
                const b200_node * qx_next = nullptr;
                if (fuse && qx_enabled() && !c->rs_on && is_vec_mulmat(nodes[i])) {
                    FusedAdd fq;
                    const bool add_fused = gl == 1 && i + 1 < n && can_fuse_add(nodes[i], nodes[i + 1], &fq);
                    const int j = i + (gl > 1 ? gl : (add_fused ? 2 : 1));
                    if (j < n && is_vec_mulmat(nodes[j]) && nodes[j].src1.nb[0] == 4) {
                        const void * want = nodes[j].src1.data;
                        const int64_t Kn = nodes[j].src0.ne[0];
                        if (add_fused) { if ((const void *) fq.dst == want && nodes[i].src0.ne[1] == Kn) c->qx_hint.on = true; }
                        else for (int mbr = 0; mbr < gl; mbr++) if (nodes[i + mbr].dst.data == want && nodes[i + mbr].src0.ne[1] == Kn && is_vec_f32(nodes[i + mbr].dst, Kn)) c->qx_hint.on = true;
                        if (c->qx_hint.on && Kn % 32 == 0) { c->qx_hint.vec = want; qx_next = &nodes[j]; rc = ensure_act(c, Kn / 32); if (rc) break; }
                        else c->qx_hint.on = false;
                    }
                }
                if (gl > 1) { rc = mulmat_group_dev(c, nodes + i, gl); used = gl; }
                else if (dgl > 1) { rc = mulmat_dense_group_dev(c, nodes + i, dgl); used = dgl; }
                else if (fuse && i + 1 < n && can_fuse_add_dense(c, nodes[i], nodes[i + 1], &fa)) { rc = mulmat_dev(c, &nodes[i].src0, &nodes[i].src1, &d, 0, &fa); used = 2; }
                else if (fuse && i + 1 < n && can_fuse_add(nodes[i], nodes[i + 1], &fa)) { rc = mulmat_dev(c, &nodes[i].src0, &nodes[i].src1, &d, 0, &fa); used = 2; }
                else if (c->rs_on) {
                    const int64_t rows = nodes[i].src0.ne[1];
                    if (d.ne[0] != rows * c->world) { rc = fail(B200_EINVAL, "rowsplit: dst has %lld rows, expected %lld x %d", (long long) d.ne[0], (long long) rows, c->world); break; }
                    b200_tensor dl = d; dl.ne[0] = rows;
                    rc = mulmat_dev(c, &nodes[i].src0, &nodes[i].src1, &dl, c->rank * rows);
                }
                else rc = mulmat_dev(c, &nodes[i].src0, &nodes[i].src1, &d, 0);
                if (!rc && c->rs_hint.on && c->rs_hint.done) {       // the gather quantised for nx (and did the ADD in front of it)
                    c->q_ptr = nx->src1.data; c->q_K = nx->src0.ne[0]; c->q_N = 1; c->q_stride = nx->src1.nb[1]; c->q_kind = 0;
                    if (skip_add) used++;
                }
                c->rs_hint.on = false; c->rs_hint.done = false;
                if (!rc && c->qx_hint.on && c->qx_hint.done) {       // the GEMV left the next op's q8_1 activations in (what is now) act_qs / act_ds
                    c->q_ptr = qx_next->src1.data; c->q_K = qx_next->src0.ne[0]; c->q_N = 1; c->q_stride = qx_next->src1.nb[1]; c->q_kind = 0;
                }
                c->qx_hint.on = false; c->qx_hint.done = false;
                break;
            }
            default: rc = fail(B200_ENOTSUP, "graph node %d: op %d not offloaded", i, d.op);
        }
        i += used;
    }
    if (!rc && c->rs_on && c->rs_ops > 0) {      // LL protocol: the sequence numbers of the next submission continue after this one's
        cudaError_t e = launch_ll_bump((uint32_t *) (c->rs_heap + 320), c->rs_ops, c->st);
        if (e != cudaSuccess) rc = fail(B200_ECUDA, "ll_bump: %s", cudaGetErrorString(e));
        c->launches++;
    }
    c->in_graph = false; c->q_ptr = nullptr;
    return rc;
}

// Key of a node list in the CUDA-graph cache.  Word-wise with four independent lanes: a byte-wise FNV over the 160 KB of a
// Llama decode node list cost ~200 us of host time per submission (seen as the gap between device time and e2e time).
static uint64_t fnv1a(const void * p, size_t n, uint64_t hsh = 1469598103934665603ull) {
    const uint8_t * b = (const uint8_t *) p;
    uint64_t h0 = hsh, h1 = hsh ^ 0x9e3779b97f4a7c15ull, h2 = hsh ^ 0xc2b2ae3d27d4eb4full, h3 = hsh ^ 0x165667b19e3779f9ull;
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t w[4];
        memcpy(w, b + i, 32);
        h0 = (h0 ^ w[0]) * 1099511628211ull; h0 ^= h0 >> 29;
        h1 = (h1 ^ w[1]) * 1099511628211ull; h1 ^= h1 >> 29;
        h2 = (h2 ^ w[2]) * 1099511628211ull; h2 ^= h2 >> 29;
        h3 = (h3 ^ w[3]) * 1099511628211ull; h3 ^= h3 >> 29;
    }
    uint64_t hh = h0 ^ (h1 * 0xff51afd7ed558ccdull) ^ (h2 * 0xc4ceb9fe1a85ec53ull) ^ (h3 * 0x9fb21c651e98df25ull);
    for (; i < n; i++) { hh ^= b[i]; hh *= 1099511628211ull; }
    hh ^= hh >> 33; hh *= 0xff51afd7ed558ccdull; hh ^= hh >> 33;
    return hh;
}

static thread_local uint64_t g_caller_key = 0;     // b200_graph_compute_keyed: the caller vouches that this key names exactly this node list
extern "C" int b200_graph_compute_keyed(b200_handle h, const b200_node * nodes, int32_t n, int32_t use_cuda_graph, uint64_t key) {
    g_caller_key = key;
    const int rc = b200_graph_compute(h, nodes, n, use_cuda_graph);
    g_caller_key = 0;
    return rc;
}
extern "C" int b200_graph_compute(b200_handle h, const b200_node * nodes, int32_t n, int32_t use_cuda_graph) {
    CTX(h);
    if (n < 0 || (n > 0 && !nodes)) return fail(B200_EINVAL, "bad node list");
    if (n == 0) return 0;
    bool any_host = false; int64_t need_act = 0;
    for (int i = 0; i < n; i++) {
        const b200_node & nd = nodes[i];
        any_host |= (nd.src0.flags | nd.src1.flags | nd.dst.flags) & B200_TENSOR_HOST;
        if (nd.dst.op == B200_OP_MUL_MAT && is_qweight(nd.src0.type)) {
            b200_tensor dl = nd.dst; if (c->rs_on) dl.ne[0] = nd.src0.ne[1];
            int rc = mulmat_check(&nd.src0, &nd.src1, &dl); if (rc) return rc;
            const int64_t b = nd.src0.ne[0] / 32 * nd.src1.ne[1];
            if (b > need_act) need_act = b;
        }
    }
    if (any_host) return fail(B200_ENOTSUP, "graph_compute takes device tensors only");
    int rc = ensure_act(c, need_act); if (rc) return rc;
    if (use_cuda_graph && (!c->rs_on || gemm_dense_enabled()))      // (row split: src0 is the rank's slice, so the sizes are the slice's)
        for (int i = 0; i < n; i++)
            if (nodes[i].dst.op == B200_OP_MUL_MAT) { rc = presize_gemm(c, &nodes[i].src0, &nodes[i].src1); if (rc) return rc; }
    int n_dyn = 0;
    for (int i = 0; i < n; i++) if (nodes[i].dst.flags & B200_TENSOR_DYNAMIC) {
        const int op = nodes[i].dst.op;
        if (op != B200_OP_CPY && op != B200_OP_DUP && op != B200_OP_CONT) return fail(B200_ENOTSUP, "B200_TENSOR_DYNAMIC is for the destination of CPY / DUP / CONT nodes");
        n_dyn++;
    }
    if (!use_cuda_graph) return run_nodes(c, nodes, n);
    if (n_dyn > DYN_MAX) return fail(B200_ENOTSUP, "more than %d dynamic destinations in one node list", DYN_MAX);
    if (n_dyn) {        // this call's destinations -> device table, stream-ordered in front of the graph launch (and outside the capture)
        if (!c->dyn_tab) CU(cudaMalloc((void **) &c->dyn_tab, DYN_MAX * sizeof(char *)));
        uint8_t * slot; cudaEvent_t ev;
        rc = pin_slot(c, &slot, &ev); if (rc) return rc;
        char ** hp = (char **) slot; int j = 0;
        for (int i = 0; i < n; i++) if (nodes[i].dst.flags & B200_TENSOR_DYNAMIC) hp[j++] = (char *) nodes[i].dst.data;
        CU(cudaMemcpyAsync(c->dyn_tab, hp, (size_t) n_dyn * sizeof(char *), cudaMemcpyHostToDevice, c->st));
        CU(cudaEventRecord(ev, c->st));
    }

    // cache key: the caller's (b200_graph_compute_keyed) or a hash of the whole list; a hit on a hashed key is confirmed by the node
    // count and a second, independent hash before anything captured for another list could be replayed
    const uint64_t salt = fnv1a(&c->scratch_gen, sizeof(c->scratch_gen)) ^ (c->rs_on ? 0x9e3779b97f4a7c15ull : 0);
    const bool keyed = g_caller_key != 0;
    const uint64_t key = keyed ? (g_caller_key * 0x9e3779b97f4a7c15ull) ^ salt ^ 0x5bd1e995 : fnv1a(nodes, sizeof(b200_node) * (size_t) n, salt);
    const uint64_t check = keyed ? g_caller_key : fnv1a(nodes, sizeof(b200_node) * (size_t) n, 0x243f6a8885a308d3ull);
    auto it = c->graphs.find(key);
    if (it != c->graphs.end() && (it->second.n_nodes != n || it->second.check != check)) {        // key collision: never replay another list's graph
        cudaGraphExecDestroy(it->second.exec); c->graphs.erase(it); it = c->graphs.end();
    }
    if (it == c->graphs.end()) {
        const int64_t before = c->launches;
        CU(cudaStreamBeginCapture(c->st, cudaStreamCaptureModeThreadLocal));
        c->capturing = true; c->dyn_on = n_dyn > 0;
        rc = run_nodes(c, nodes, n);
        c->capturing = false; c->dyn_on = false;
        cudaGraph_t g = nullptr;
        cudaError_t e = cudaStreamEndCapture(c->st, &g);
        const int64_t captured = c->launches - before;
        c->launches = before;
        if (rc) { if (g) cudaGraphDestroy(g); return rc; }
        if (e != cudaSuccess) return fail(B200_ECUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
        GraphEntry ge; ge.n_kernels = captured; ge.n_nodes = n; ge.check = check;
        // a decode loop re-captures the same topology with other shapes every time its cache view grows by a block: updating the executable
        // graph captured last in place is several times cheaper than instantiating a new one (the old list is captured again if it comes back)
        static const bool upd = !(getenv("B200_GRAPH_UPDATE") && atoi(getenv("B200_GRAPH_UPDATE")) == 0);
        // (a key whose graph was taken over before and that now comes back -- two lists alternating, or a loop that starts over -- gets an
        //  executable graph of its own, so that lists which do recur end up replayed without any capture)
        const bool came_back = c->taken_over.count(key) != 0;
        auto donor = upd && !came_back && c->last_graph_key ? c->graphs.find(c->last_graph_key) : c->graphs.end();
        if (donor != c->graphs.end() && donor->second.n_nodes == n && donor->second.n_kernels == captured) {
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(donor->second.exec, g, &info) == cudaSuccess) { ge.exec = donor->second.exec; c->taken_over.insert(donor->first); }
            else { cudaGetLastError(); cudaGraphExecDestroy(donor->second.exec); }      // (another kernel somewhere: not the same topology after all;
            c->graphs.erase(donor);                                                      //  a failed update may leave the donor half-updated: drop it)
        }
        e = ge.exec ? cudaSuccess : cudaGraphInstantiate(&ge.exec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) return fail(B200_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
        c->last_graph_key = key;
        if (c->graphs.size() > 64) { for (auto & kv : c->graphs) cudaGraphExecDestroy(kv.second.exec); c->graphs.clear(); c->taken_over.clear(); }
        if (c->taken_over.size() > 4096) c->taken_over.clear();
        it = c->graphs.emplace(key, ge).first;
    }
    static const bool time_it = getenv("B200_GRAPH_TIME") != nullptr;          // development: device time of each replay, to stderr
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    if (time_it) { cudaEventCreate(&t0); cudaEventCreate(&t1); cudaEventRecord(t0, c->st); }
    CU(cudaGraphLaunch(it->second.exec, c->st));
    if (time_it) {
        cudaEventRecord(t1, c->st); cudaEventSynchronize(t1);
        float ms = 0; cudaEventElapsedTime(&ms, t0, t1);
        fprintf(stderr, "[b200] graph replay: %d nodes, %lld kernels, %.3f ms\n", n, (long long) it->second.n_kernels, ms);
        cudaEventDestroy(t0); cudaEventDestroy(t1);
    }
    c->launches += it->second.n_kernels;      // a replay runs exactly the kernels captured
    return 0;
}

// ------------------------------------------------------------------------------------------------ development: gemv6 timeline
// 4 x %globaltimer per CTA (start, dependency resolved, activations quantised, done) for the next max_launches gemv6 launches
static long long * g_trace_buf = nullptr; static int g_trace_cap = 0;
extern "C" int b200_debug_trace_begin(b200_handle h, int32_t max_launches) {
    CTX(h);
    if (g_trace_buf) { cudaFree(g_trace_buf); g_trace_buf = nullptr; }
    if (max_launches <= 0) { gemv6_trace(nullptr, 0); return 0; }
    const size_t bytes = (size_t) max_launches * 4096 * sizeof(long long);
    CU(cudaMalloc((void **) &g_trace_buf, bytes));
    CU(cudaMemsetAsync(g_trace_buf, 0, bytes, c->st));
    g_trace_cap = max_launches;
    gemv6_trace(g_trace_buf, max_launches);
    return 0;
}
extern "C" int b200_debug_trace_read(b200_handle h, int64_t * out, int32_t max_launches) {
    CTX(h);
    CU(cudaStreamSynchronize(c->st));
    int n = gemv6_trace_count(); if (n > max_launches) n = max_launches;
    if (n > 0) CU(cudaMemcpy(out, g_trace_buf, (size_t) n * 4096 * sizeof(long long), cudaMemcpyDeviceToHost));
    gemv6_trace(nullptr, 0);
    return n;
}

// ------------------------------------------------------------------------------------------------ events
extern "C" int b200_event_create(b200_handle h, void ** ev) { CTX(h); cudaEvent_t e; CU(cudaEventCreate(&e)); *ev = e; return 0; }
extern "C" int b200_event_record(b200_handle h, void * ev) { CTX(h); CU(cudaEventRecord((cudaEvent_t) ev, c->st)); return 0; }
extern "C" int b200_event_elapsed_ms(b200_handle h, void * a, void * b, float * ms) {
    CTX(h); CU(cudaEventSynchronize((cudaEvent_t) b)); CU(cudaEventElapsedTime(ms, (cudaEvent_t) a, (cudaEvent_t) b)); return 0;
}
extern "C" int b200_event_wait(b200_handle h, void * ev) { CTX(h); CU(cudaStreamWaitEvent(c->st, (cudaEvent_t) ev, 0)); return 0; }
extern "C" int b200_event_synchronize(b200_handle h, void * ev) { CTX(h); CU(cudaEventSynchronize((cudaEvent_t) ev)); return 0; }
extern "C" int b200_event_destroy(b200_handle h, void * ev) { CTX(h); CU(cudaEventDestroy((cudaEvent_t) ev)); return 0; }

// ------------------------------------------------------------------------------------------------ row split (NCCL)
extern "C" int b200_rowsplit_unique_id(void * id128) {
    int rc = nccl_load(); if (rc) return rc;
    nccl_uid id; NC(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, 128);
    return 0;
}
extern "C" int b200_rowsplit_init(b200_handle h, const void * id128, int32_t rank, int32_t world) {
    CTX(h);
    if (world < 1 || rank < 0 || rank >= world) return fail(B200_EINVAL, "rowsplit_init: rank %d of %d", rank, world);
    int rc = nccl_load(); if (rc) return rc;
    nccl_uid id; memcpy(&id, id128, 128);
    NC(g_nccl.CommInitRank(&c->comm, world, id, rank));
    c->rank = rank; c->world = world;
    return 0;
}
extern "C" int b200_rowsplit_finalize(b200_handle h) {
    CTX(h);
    if (c->comm) { CU(cudaStreamSynchronize(c->st)); NC(g_nccl.CommDestroy(c->comm)); c->comm = nullptr; }
    c->world = 1; c->rank = 0;
    return 0;
}
extern "C" int b200_mulmat_rowsplit(b200_handle h, const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d) {
    CTX(h); if (!s0 || !s1 || !d) return fail(B200_EINVAL, "NULL tensor");
    if (!c->comm && c->world != 1) return fail(B200_EINVAL, "rowsplit not initialised");
    const int64_t rows = s0->ne[1], Mtot = d->ne[0], N = s1->ne[1];
    if (rows * c->world != Mtot) return fail(B200_EINVAL, "rowsplit: %lld rows x %d ranks != M %lld", (long long) rows, c->world, (long long) Mtot);
    if (s1->ne[2] != 1 || s1->ne[3] != 1) return fail(B200_ENOTSUP, "rowsplit: 2-D only");
    if (d->nb[0] != 4 || d->nb[1] != Mtot * 4) return fail(B200_ENOTSUP, "rowsplit: dst must be contiguous");
    // local rows straight into this rank's slice of the full dst
    b200_tensor dl = *d; dl.ne[0] = rows;
    int rc = mulmat_check(s0, s1, &dl); if (rc) return rc;
    c->q_ptr = nullptr;
    rc = mulmat_dev(c, s0, s1, &dl, c->rank * rows); if (rc) return rc;
    c->q_ptr = nullptr;
    if (c->world == 1) return 0;
    float * base = (float *) d->data;
    NC(g_nccl.GroupStart());
    for (int64_t n = 0; n < N; n++) {
        float * col = base + n * Mtot;
        int r = g_nccl.AllGather(col + c->rank * rows, col, (size_t) rows, /*ncclFloat32*/ 7, c->comm, c->st);
        if (r) { g_nccl.GroupEnd(); return fail(B200_ENCCL, "ncclAllGather: %s", g_nccl.GetErrorString(r)); }
    }
    NC(g_nccl.GroupEnd());
    return 0;
}

// ------------------------------------------------------------------------------------------------ fused row split over peer memory
// One process per GPU.  Every rank allocates the same symmetric heap, publishes its IPC handle, maps the peers' heaps
// (NVLink P2P), and puts every activation vector of the decode step at the same offset.  A row-split MUL_MAT then stores
// its slice straight into every rank's copy of dst from inside the GEMV (gemv.cu), and one 32-thread exchange kernel per
// op replaces the ncclAllGather: release-store of the op counter into each peer, acquire-spin on theirs.
extern "C" int b200_rowsplit_heap_create(b200_handle h, size_t bytes, void ** user_base, void * ipc_handle64) {
    CTX(h);
    if (c->rs_heap) return fail(B200_EINVAL, "rowsplit heap already created");
    if (!user_base) return fail(B200_EINVAL, "user_base is NULL");
    const size_t total = ((bytes + 255) & ~(size_t) 255) + RS_HDR;
    CU(cudaMalloc((void **) &c->rs_heap, total));
    CU(cudaMemsetAsync(c->rs_heap, 0, total, c->st));
    CU(cudaStreamSynchronize(c->st));
    c->rs_heap_bytes = total;
    if (ipc_handle64) {
        cudaIpcMemHandle_t hd;
        static_assert(sizeof(hd) == 64, "ipc handle size");
        CU(cudaIpcGetMemHandle(&hd, c->rs_heap));
        memcpy(ipc_handle64, &hd, 64);
    }
    *user_base = c->rs_heap + RS_HDR;
    return 0;
}
static int rs_check_attach(Ctx * c, int rank, int world) {
    if (!c->rs_heap) return fail(B200_EINVAL, "rowsplit: create the heap first");
    if (world < 1 || world > 8 || rank < 0 || rank >= world) return fail(B200_EINVAL, "rowsplit: rank %d of %d (at most 8)", rank, world);
    return 0;
}
extern "C" int b200_rowsplit_heap_attach(b200_handle h, const void * handles, int32_t rank, int32_t world) {
    CTX(h);
    int rc = rs_check_attach(c, rank, world); if (rc) return rc;
    for (int r = 0; r < world; r++) {
        if (r == rank) { c->rs_peer[r] = c->rs_heap; continue; }
        cudaIpcMemHandle_t hd; memcpy(&hd, (const char *) handles + 64 * r, 64);
        void * p = nullptr;
        CU(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
        c->rs_peer[r] = (char *) p; c->rs_ipc[r] = true;
    }
    c->rank = rank; c->world = world; c->rs_attached = true;
    return 0;
}
// same-process variant (tests: two sessions on two devices of one process): peers' heap bases are plain device pointers
extern "C" int b200_rowsplit_heap_attach_ptrs(b200_handle h, void * const * peer_user_bases, const int32_t * peer_devices, int32_t rank, int32_t world) {
    CTX(h);
    int rc = rs_check_attach(c, rank, world); if (rc) return rc;
    for (int r = 0; r < world; r++) {
        if (r == rank) { c->rs_peer[r] = c->rs_heap; continue; }
        if (peer_devices[r] != c->dev) {
            cudaError_t e = cudaDeviceEnablePeerAccess(peer_devices[r], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(B200_ECUDA, "cudaDeviceEnablePeerAccess(%d): %s", peer_devices[r], cudaGetErrorString(e));
            cudaGetLastError();
        }
        c->rs_peer[r] = (char *) peer_user_bases[r] - RS_HDR;
    }
    c->rank = rank; c->world = world; c->rs_attached = true;
    return 0;
}
extern "C" int b200_rowsplit_heap_destroy(b200_handle h) {
    CTX(h);
    CU(cudaStreamSynchronize(c->st));
    for (int r = 0; r < 8; r++) { if (c->rs_ipc[r] && c->rs_peer[r]) cudaIpcCloseMemHandle(c->rs_peer[r]); c->rs_peer[r] = nullptr; c->rs_ipc[r] = false; }
    if (c->rs_heap) cudaFree(c->rs_heap);
    c->rs_heap = nullptr; c->rs_heap_bytes = 0; c->rs_attached = false;
    return 0;
}
// node list of a row-split decode step: MUL_MAT src0 = THIS rank's row slice [rows x K] (REPACKED), dst = the full vector
// [rows * world] inside the symmetric heap (complete on every rank when the op's exchange has run); ADD nodes are replicated.
extern "C" int b200_graph_compute_rowsplit(b200_handle h, const b200_node * nodes, int32_t n, int32_t use_cuda_graph) {
    CTX(h);
    if (c->world > 1 && !c->rs_attached) return fail(B200_EINVAL, "rowsplit: attach the peers' heaps first");
    c->rs_on = true; g_rs_world = c->world;
    const int rc = b200_graph_compute(h, nodes, n, use_cuda_graph ? 2 : 0);
    c->rs_on = false; g_rs_world = 1;
    return rc;
}

// ------------------------------------------------------------------------------------------------ MUL_MAT without allocations
// b200_mulmat_noalloc: the few-column (decode) MUL_MAT for callers that cannot allocate device memory -- code under a stream capture,
// serving loops, harnesses that forbid cudaMalloc.  Same two kernels as b200_mulmat's GEMV path (q8_1 quantiser, then gemv4 / gemv5),
// but the quantised activations go into a file-scope device array instead of the session's cudaMalloc'ed scratch, and no session,
// no stream and no event is created: the caller passes device pointers and its own stream.
// Every GPU that loads this library holds the B200_NOALLOC_SCRATCH_BYTES of g_noalloc_scratch until the process exits.  All calls on a
// device share it: calls that overlap on the GPU (different streams, no dependency between them) race on it.  The array keeps the last
// call's contents; nothing has to start at zero -- the quantiser writes every block of every column before the GEMV reads it.
// Layout: int8 quants [N][K] at offset 0, ActDS {d, block sums} [N][K / 32] behind them at the next multiple of 256 bytes.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is the host-side stand-in of the symbol, not a device address: the kernels get what
// cudaGetSymbolAddress returns, asked once per device with that device current.
static int noalloc_scratch_of_current_device(unsigned char ** out) {
    static std::mutex mu;
    static unsigned char * ptr[64] = {nullptr};
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(B200_EINVAL, "b200_mulmat_noalloc: device %d out of range", dev);
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[dev]) {
        void * p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(B200_ECUDA, "b200_mulmat_noalloc: cudaGetSymbolAddress(g_noalloc_scratch) on device %d -> %s", dev, cudaGetErrorString(e));
        ptr[dev] = (unsigned char *) p;
    }
    *out = ptr[dev];
    return 0;
}

extern "C" int b200_mulmat_noalloc(const b200_tensor * s0, const b200_tensor * s1, b200_tensor * d, void * stream) {
    if (!s0 || !s1 || !d) return fail(B200_EINVAL, "b200_mulmat_noalloc: null tensor");
    int rc = mulmat_check(s0, s1, d); if (rc) return rc;
    if (!is_qweight(s0->type)) return fail(B200_ENOTSUP, "b200_mulmat_noalloc: src0 must be q4_0 / q8_0 / q4_K / q6_K REPACKED device weights");
    if ((s0->flags | s1->flags | d->flags) & B200_TENSOR_HOST) return fail(B200_ENOTSUP, "b200_mulmat_noalloc: device tensors only (staging host tensors would allocate)");
    if (s0->ne[2] != 1 || s0->ne[3] != 1 || s1->ne[2] != 1 || s1->ne[3] != 1) return fail(B200_ENOTSUP, "b200_mulmat_noalloc: 2-D operands only");
    const int64_t K = s0->ne[0], M = s0->ne[1], N = s1->ne[1];
    if (M == 0 || N == 0 || K == 0) return 0;
    const int64_t NB = K / 32;
    const size_t qs_bytes = ((size_t) N * NB * 32 + 255) & ~(size_t) 255, need = qs_bytes + (size_t) N * NB * sizeof(ActDS);
    if (need > (size_t) B200_NOALLOC_SCRATCH_BYTES)
        return fail(B200_EINVAL, "b200_mulmat_noalloc: K = %lld x N = %lld needs %zu bytes of activation scratch, the static array holds %d (use b200_mulmat)",
                    (long long) K, (long long) N, need, (int) B200_NOALLOC_SCRATCH_BYTES);
    WPtrs wp{};
    rc = wptrs_of(s0, &wp); if (rc) return rc;
    unsigned char * scratch = nullptr;
    rc = noalloc_scratch_of_current_device(&scratch); if (rc) return rc;
    int8_t * qs = (int8_t *) scratch;
    ActDS * ds = (ActDS *) (scratch + qs_bytes);
    cudaStream_t st = (cudaStream_t) stream;
    CU(launch_quantize_q8_1_soa((const float *) s1->data, s1->nb[1], K, N, qs, ds, st));
    GemvArgs a{};
    a.nmat = 1; a.K = K; a.dcs = d->nb[1] / 4;
    a.mat[0].w = wp; a.mat[0].M = M;
    for (int64_t c0 = 0; c0 < N; c0 += 8) {
        const int nc = (int) (N - c0 < 8 ? N - c0 : 8);
        a.act = ActView{qs + c0 * NB * 32, ds + c0 * NB, NB};
        a.mat[0].dst = (float *) d->data + c0 * a.dcs;
        CU(launch_gemv2(s0->type, a, nc, st));
    }
    return 0;
}

/*
 * ggml_b200_ops.h — C-ABI of the B200 op layer (libb200ops.so).
 *
 * This is the seat of the reference's FastRPC interface between the host backend and the
 * Hexagon cDSP kernels: `ggmlop_dsp_open/close/setclocks/add/mulmat` taking `struct dsptensor`
 * (ggml/src/ggml-hexagon/kernels/ggmlop_ap_skel.h:235-244 and :266-280), called from
 * ggmlhexagon_compute (ggml/src/ggml-hexagon/ggml-hexagon.cpp:4975-5060).
 * Plain C: pointers, sizes, ints.  No ggml, torch or CUDA types in any signature.
 *
 * Every function returns 0 on success or a negative B200_E* code (the FastRPC functions return 0 or
 * an AEE error the same way); b200_last_error() gives the text.  There is no CPU fallback: on a
 * machine without a CUDA device b200_open fails with B200_ENODEV and nothing else is usable.
 */
#ifndef GGML_B200_OPS_H
#define GGML_B200_OPS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

/* tensor element types: numeric values of enum ggml_type (ggml/include/ggml.h:349-390) */
enum b200_type {
    B200_TYPE_F32  = 0,
    B200_TYPE_F16  = 1,   /* SOFT_MAX mask, POOL_2D input, CPY either side, MUL_MAT src0 (KV cache) */
    B200_TYPE_Q4_0 = 2,
    B200_TYPE_Q8_0 = 8,
    B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q4_K = 12,
    B200_TYPE_Q6_K = 14,
};

/* ops: numeric values of enum ggml_op for the two ops the cDSP path offloads
 * (ggml-hexagon.cpp:5079-5098; ggml/include/ggml.h:426-, GGML_OP_ADD = 2, GGML_OP_MUL_MAT = 27) */
enum b200_op {
    B200_OP_NONE     = 0,
    B200_OP_ADD      = 2,
    B200_OP_RMS_NORM = 23,   /* the three ops the cDSP table lists and leaves as stubs (ggml-hexagon.cpp:619,639,649; ggml-dsp.c:1353-1372) */
    B200_OP_MUL_MAT  = 27,
    B200_OP_SOFT_MAX = 43,
    B200_OP_POOL_2D  = 53,
    /* the ops between the MUL_MATs of a Llama layer, taken only in node lists (b200_graph_compute): with them on the device the
     * scheduler keeps a whole layer stack in one split instead of a host round trip per MUL_MAT (graph_ops.cu) */
    B200_OP_DUP      = 1,
    B200_OP_MUL      = 6,    /* f32, src1 broadcast as in ADD */
    B200_OP_CPY      = 32,   /* src0 -> src1's layout (dst is a view of src1); f32 / f16 either side */
    B200_OP_CONT     = 33,
    B200_OP_GET_ROWS = 38,   /* f32 rows, i32 ids */
    B200_OP_ROPE     = 45,   /* f32 or f16 (same on both sides); mode 0 (normal) or 2 (NeoX); src1 = i32 positions; src2 = optional f32 frequency factors */
    B200_OP_UNARY    = 73,   /* op_params[0] = B200_UNARY_SILU */
};
#define B200_UNARY_SILU 10   /* GGML_UNARY_OP_SILU */
#define B200_TYPE_I32   26   /* GGML_TYPE_I32: ROPE positions, GET_ROWS ids */

enum b200_err {
    B200_OK       = 0,
    B200_ENODEV   = -1,  /* no CUDA device / driver */
    B200_EINVAL   = -2,  /* bad handle / argument / unsupported shape or type */
    B200_ENOMEM   = -3,  /* device allocation failed */
    B200_ECUDA    = -4,  /* CUDA runtime error (text in b200_last_error) */
    B200_ENCCL    = -5,  /* NCCL error */
    B200_ENOTSUP  = -6,  /* valid request the kernels do not implement (caller must not route it here) */
};

/* flags of b200_tensor */
#define B200_TENSOR_HOST      0x1  /* data is a HOST pointer: the call copies it in/out (as FastRPC does with dsptensor.data) */
#define B200_TENSOR_REPACKED  0x2  /* quantised weight already in the device SoA layout written by b200_weights_upload */
#define B200_TENSOR_DYNAMIC   0x4  /* dst of a CPY / DUP / CONT node: `data` may differ between calls made under one b200_graph_compute_keyed key
                                    * (a KV-cache store moves with the position); everything else about the node list is the same */

/*
 * Replaces `struct dsptensor` (ggmlop_ap_skel.h:235-244).  Same fields, but ne/nb are 64-bit
 * (the reference narrows ggml's int64/size_t to int32 at ggml-hexagon.cpp:5006-5049, which cannot
 * describe a 128256x8192 matrix in bytes).  ne/nb have ggml_tensor meaning (ggml.h:578-610):
 * ne[0] is the contiguous dimension in ELEMENTS, nb[i] the stride in BYTES of the native ggml block
 * layout (nb[0] = block size; nb[1] = row size ...), also for REPACKED weights.
 */
typedef struct b200_tensor {
    int32_t type;          /* enum b200_type */
    int32_t op;            /* enum b200_op (dst only) */
    int32_t flags;
    int32_t reserved;
    int64_t ne[4];
    int64_t nb[4];
    int32_t op_params[16];
    void *  data;          /* device pointer (or host pointer with B200_TENSOR_HOST) */
    int64_t data_len;      /* bytes addressable from data */
    /* REPACKED weights only: the SoA region is laid out over the whole allocation it was uploaded into */
    void *  w_base;        /* base of that allocation (device) */
    int64_t w_blocks;      /* total blocks (32- or 256-element units) in that allocation */
    int64_t w_first;       /* index of this tensor's first block inside it */
} b200_tensor;

/* one node of a compute graph: dst = op(src0, src1)   (ggml_cgraph node, ggml-hexagon.cpp:5555-5574) */
typedef struct b200_node {
    b200_tensor src0, src1, dst;
    b200_tensor src2;      /* third operand where the op has one (ROPE: f32 frequency factors); data = NULL otherwise */
} b200_node;

typedef uint64_t b200_handle;   /* replaces remote_handle64 */

/* ---- session (ggmlop_dsp_open / _close / _setclocks, ggmlop_ap_skel.h:266-275; ggml-hexagon.cpp:4945-4951) */
/* uri: "cuda:<ordinal>" (NULL or "" = "cuda:0").  Creates the context: device, stream, q8_1 scratch. */
B200_API int b200_open(const char * uri, b200_handle * h);
B200_API int b200_close(b200_handle h);
/* kept for interface symmetry; GPU clocks are never touched, returns 0 */
B200_API int b200_setclocks(b200_handle h, int32_t power_level, int32_t latency, int32_t dcvs_enable);
B200_API const char * b200_last_error(void);
B200_API int b200_device_count(void);
/* name[<=len], total/free HBM bytes, SM count */
B200_API int b200_device_info(int device, char * name, int len, int64_t * total, int64_t * free_b, int32_t * sms);

/* ---- the two ops (ggmlop_dsp_add / ggmlop_dsp_mulmat, ggmlop_ap_skel.h:276-277; ggml-dsp.c:1069,1192) */
/* dst = src0 + src1, f32; src1 broadcast by repetition in every dim (ggml-dsp.c:991-1066, binary-ops.cpp:50-117) */
B200_API int b200_add(b200_handle h, const b200_tensor * src0, const b200_tensor * src1, b200_tensor * dst);
/* dst[M,N,ne12,ne13] = src0[K,M,ne02,ne03] x src1[K,N,ne12,ne13] (ggml-dsp.c:1192-1351, ggml-cpu.c:6745-6937).
 * src0: q4_0/q8_0/q4_K/q6_K REPACKED device weights, or f32; src1, dst: f32. Asynchronous on the handle's stream
 * unless a tensor carries B200_TENSOR_HOST (then the call returns after dst is back in host memory). */
B200_API int b200_mulmat(b200_handle h, const b200_tensor * src0, const b200_tensor * src1, b200_tensor * dst);
/* MUL_MAT that allocates nothing.  No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under a CUDA graph capture, tight serving loops, harnesses that forbid it).  No session handle: src0 = 2-D
 * q4_0/q8_0/q4_K/q6_K REPACKED device weights, src1 / dst = 2-D f32 device tensors of the CURRENT device, `stream` = the caller's
 * cudaStream_t (NULL: the default stream); asynchronous on it.  The q8_1 activations live in a static device array of
 * B200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit: (K * N rounded up to 256) + K / 32 * N * 8 bytes must fit
 * (e.g. K = 4096: N <= 204; K = 28672: N <= 29), else B200_EINVAL with the sizes in b200_last_error.  Calls on one device share the
 * array: calls that overlap on the GPU race.  Results are bit-identical to b200_mulmat's GEMV path. */
#define B200_NOALLOC_SCRATCH_BYTES (1 << 20)
B200_API int b200_mulmat_noalloc(const b200_tensor * src0, const b200_tensor * src1, b200_tensor * dst, void * stream);
/* ---- the ops ggml-hexagon lists beside them without an implementation (ggmlop_dsp_rmsnorm / _softmax / _pool2d, ggml-dsp.c:1353-1372);
 *      device tensors only; parameters travel in dst->op_params exactly as in ggml_tensor.op_params
 * RMS_NORM: y = x / sqrt(mean(x^2) + eps) per row, f32, src nb[0] = 4 (ggml-cpu.c:6254-6300); eps = op_params[0] (float bits) */
B200_API int b200_rms_norm(b200_handle h, const b200_tensor * src, b200_tensor * dst);
/* SOFT_MAX(x * scale + slope(head) * mask) per row; x, dst contiguous f32; mask NULL, f32 or f16 [ne00, >= ne01] (ggml-cpu.c:8261-8357);
 * scale = op_params[0], max_bias = op_params[1] (float bits) */
B200_API int b200_soft_max(b200_handle h, const b200_tensor * src, const b200_tensor * mask, b200_tensor * dst);
/* POOL_2D max / average over [ne0, ne1] planes, f32 or f16 in, f32 out (ggml-cpu.c:9758-9830); op_params[0..6] = op, k0, k1, s0, s1, p0, p1 */
B200_API int b200_pool_2d(b200_handle h, const b200_tensor * src, b200_tensor * dst);

/* a whole node list in one call (graph_compute, ggml-hexagon.cpp:5555-5574); asynchronous.
 * use_cuda_graph = 0: one launch per node (fused where adjacent nodes allow).
 * use_cuda_graph != 0: replayable fast path, cached per distinct node list: the per-node launches are captured once as a CUDA graph
 *                     (programmatic dependent launch edges) and replayed. */
B200_API int b200_graph_compute(b200_handle h, const b200_node * nodes, int32_t n_nodes, int32_t use_cuda_graph);
/* the same, for a caller that keeps its own cache of marshalled node lists (the ggml plugin): `key` != 0 names this exact list --
 * the caller guarantees one key never stands for two different lists -- so the library does not hash the descriptors again */
B200_API int b200_graph_compute_keyed(b200_handle h, const b200_node * nodes, int32_t n_nodes, int32_t use_cuda_graph, uint64_t key);
B200_API int b200_synchronize(b200_handle h);

/* ---- activation quantiser, exposed for parity tests (= quantize_row_q8_1_ref, ggml-quants.c:220-253)
 * x: device f32 [nrows][k] contiguous, y: device block_q8_1 AoS [nrows][k/32] (36 B blocks). */
B200_API int b200_quantize_q8_1(b200_handle h, const float * x, void * y, int64_t k, int64_t nrows);
/* per-32-group (q6_K: per-16-group) exact integer dot products of one weight row with int8 activations
 * (the bit-exact contract, oracle_block_dots).  a_qs: device int8[k]; out: device int32[k/32 or k/16]. */
B200_API int b200_block_dots(b200_handle h, const b200_tensor * w_row, const int8_t * a_qs, int32_t * out);

/* ---- device memory (buffer_type.alloc_buffer / buffer.set_tensor / get_tensor / memset / clear,
 *      ggml-backend-impl.h:17-58; ggml-hexagon.cpp:5361-5478) */
B200_API int b200_malloc(b200_handle h, size_t size, void ** ptr);
B200_API int b200_free(b200_handle h, void * ptr);
/* h2d returns when src_host may be reused; the bytes are in place for everything submitted to the session afterwards (up to 256 KB travel
 * through a pinned slot of the session and are NOT waited for: call b200_synchronize before touching dst_dev from another stream).
 * d2h returns when dst_host holds the bytes. */
B200_API int b200_memcpy_h2d(b200_handle h, void * dst_dev, const void * src_host, size_t n);
B200_API int b200_memcpy_d2h(b200_handle h, void * dst_host, const void * src_dev, size_t n);
B200_API int b200_memcpy_d2d(b200_handle h, void * dst_dev, const void * src_dev, size_t n);
/* stream-ordered variants (backend.set_tensor_async / get_tensor_async, ggml-backend-impl.h:93-94; the hexagon backend leaves
 * these slots empty): they return at once, the copy is ordered with the ops of the session and is complete after
 * b200_synchronize().  Host memory should come from b200_host_alloc for the copy to overlap. */
B200_API int b200_memcpy_h2d_async(b200_handle h, void * dst_dev, const void * src_host, size_t n);
B200_API int b200_memcpy_d2h_async(b200_handle h, void * dst_host, const void * src_dev, size_t n);
B200_API int b200_memset(b200_handle h, void * dst_dev, int value, size_t n);
/* pinned host staging memory for the host-buffer (e2e) path */
B200_API int b200_host_alloc(b200_handle h, size_t size, void ** ptr);
B200_API int b200_host_free(b200_handle h, void * ptr);

/* ---- quantised weights: native ggml blocks (host) <-> device SoA layout (DESIGN.md "HBM layout").
 * The region [base, base + total_blocks*block_bytes) holds total_blocks blocks of `type`;
 * upload/download blocks [first, first+n).  Byte count is identical to the native layout. */
B200_API int b200_weights_upload(b200_handle h, int32_t type, void * base, int64_t total_blocks,
                                 int64_t first, int64_t n, const void * host_blocks);
B200_API int b200_weights_download(b200_handle h, int32_t type, const void * base, int64_t total_blocks,
                                   int64_t first, int64_t n, void * host_blocks);
B200_API int64_t b200_block_elems(int32_t type);   /* 32 / 256 / 1 */
B200_API int64_t b200_block_bytes(int32_t type);   /* 18 / 34 / 144 / 210 / 4 */

/* ---- device-side timing on the handle's stream (CUDA events) */
B200_API int b200_event_create(b200_handle h, void ** ev);
B200_API int b200_event_record(b200_handle h, void * ev);
B200_API int b200_event_elapsed_ms(b200_handle h, void * start, void * stop, float * ms); /* syncs on stop */
B200_API int b200_event_destroy(b200_handle h, void * ev);
/* ordering between sessions / devices (ggml_backend_i::event_wait, ggml_backend_device_i::event_synchronize,
 * ggml-backend-impl.h:112-116, 170-178): the session's stream waits for an event recorded on any session; the host waits for it */
B200_API int b200_event_wait(b200_handle h, void * ev);
B200_API int b200_event_synchronize(b200_handle h, void * ev);
/* kernels launched through this handle since open (the `gpu_launches` evidence of bench.py) */
B200_API int64_t b200_launch_count(b200_handle h);
/* overwrite a >L2-sized scratch buffer so the next kernel starts from a cold L2 */
B200_API int b200_flush_l2(b200_handle h);

/* ---- development instrumentation: %globaltimer stamps of the decode GEMV's CTAs (tools/g6_trace.py).  begin(n): record the next n
 *      launches (captured CUDA graphs keep recording into the same buffer on every replay); read: 4096 int64 per launch
 *      = 4 stamps {start, dependency resolved, activations quantised, done} per CTA; returns the number of launches recorded */
B200_API int b200_debug_trace_begin(b200_handle h, int32_t max_launches);
B200_API int b200_debug_trace_read(b200_handle h, int64_t * out, int32_t max_launches);

/* ---- row-split MUL_MAT across the GPUs of one box: one process per GPU, NCCL over NVLink
 *      (takes the seat of ggml-cuda's split buffers, ggml-cuda.cu:727-753,1365-1673) */
B200_API int b200_rowsplit_unique_id(void * id128);                       /* 128 bytes, rank 0 creates, all ranks get */
B200_API int b200_rowsplit_init(b200_handle h, const void * id128, int32_t rank, int32_t world);
B200_API int b200_rowsplit_finalize(b200_handle h);
/* src0 = THIS rank's row slice [rows_local x K] (REPACKED); src1 replicated; dst_full [M_total x N] on every rank.
 * rows_local must be equal on all ranks (M_total = world * rows_local).  GEMV/GEMM into the local slice of
 * dst_full, then all-gather of the slices. */
B200_API int b200_mulmat_rowsplit(b200_handle h, const b200_tensor * src0_slice, const b200_tensor * src1,
                                  b200_tensor * dst_full);


/* ---- fused row split over peer memory, instead of the ncclAllGather: the GEMV stores its results into every peer's heap over
 *      NVLink as {value, sequence number} pairs which a gather kernel on the peer polls ("LL" protocol; B200_RS_PROTO=flag and
 *      ops with several columns: plain stores + a flag exchange per op).  Every rank: heap_create (same size everywhere),
 *      exchange the 64-byte IPC handles by any means, heap_attach, carve its activation vectors out of the heap at the same
 *      offsets on every rank.  All ranks must submit the same node lists in the same order. */
B200_API int b200_rowsplit_heap_create(b200_handle h, size_t bytes, void ** user_base, void * ipc_handle64);
B200_API int b200_rowsplit_heap_attach(b200_handle h, const void * all_handles /* world x 64 bytes, by rank */, int32_t rank, int32_t world);
/* same-process variant: the peers' user_base pointers and CUDA ordinals */
B200_API int b200_rowsplit_heap_attach_ptrs(b200_handle h, void * const * peer_user_bases, const int32_t * peer_devices, int32_t rank, int32_t world);
B200_API int b200_rowsplit_heap_destroy(b200_handle h);
/* MUL_MAT nodes: src0 = this rank's row slice, dst = full vector in the heap; ADD nodes replicated; N = 1..8 columns */
B200_API int b200_graph_compute_rowsplit(b200_handle h, const b200_node * nodes, int32_t n_nodes, int32_t use_cuda_graph);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_OPS_H */

// b200_internal.cuh — shared declarations of the B200 op layer (device layouts + launchers).
// Not part of the public C-ABI (include/ggml_b200_ops.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

enum : int { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q8_0 = 8, T_Q8_1 = 9, T_Q4_K = 12, T_Q6_K = 14 };

// ---------------------------------------------------------------- HBM layout of quantised weights
// A weight allocation of T blocks keeps the native byte count (T * block_bytes) but stores the
// fields of all blocks as separate arrays ("SoA"), so that every hot load is a naturally aligned
// 128-bit load and consecutive lanes read consecutive 16 B:
//   q4_0 : [qs  T x 16 B][d  T x f16]
//   q8_0 : [qs  T x 32 B][d  T x f16]
//   q4_K : [qs  T x 128 B][hdr T x 16 B = f16 d, f16 dmin, 12 B packed scales/mins]
//   q6_K : [ql  T x 128 B][qh T x 64 B][sc T x 16 x i8][d T x f16]
// Native layouts: ggml/src/ggml-common.h:167-172 (q4_0), :211-215 (q8_0), :285-296 (q4_K), :320-326 (q6_K).
struct WFields {            // field table used by the repack kernels (native offset, size), SoA order
    int n;
    int native_off[4];
    int size[4];
    int block_bytes;
};

__host__ __device__ inline WFields wfields(int type) {
    WFields f{};
    switch (type) {
        case T_Q4_0: f.n = 2; f.native_off[0] = 2;  f.size[0] = 16;  f.native_off[1] = 0;  f.size[1] = 2;  f.block_bytes = 18;  break;
        case T_Q8_0: f.n = 2; f.native_off[0] = 2;  f.size[0] = 32;  f.native_off[1] = 0;  f.size[1] = 2;  f.block_bytes = 34;  break;
        case T_Q4_K: f.n = 2; f.native_off[0] = 16; f.size[0] = 128; f.native_off[1] = 0;  f.size[1] = 16; f.block_bytes = 144; break;
        case T_Q6_K: f.n = 4; f.native_off[0] = 0;  f.size[0] = 128; f.native_off[1] = 128; f.size[1] = 64;
                              f.native_off[2] = 192; f.size[2] = 16; f.native_off[3] = 208; f.size[3] = 2; f.block_bytes = 210; break;
        default: f.n = 0; f.block_bytes = 0;
    }
    return f;
}

inline int block_elems(int type) { return type == T_F32 ? 1 : (type == T_Q4_K || type == T_Q6_K) ? 256 : 32; }
inline int block_bytes(int type) { return type == T_F32 ? 4 : type == T_Q8_1 ? 36 : wfields(type).block_bytes; }

// pointers to the field arrays, already offset to the first block of the tensor at hand
struct WPtrs { const uint8_t * p[4]; };

inline WPtrs wptrs(int type, const void * base, int64_t T, int64_t first) {
    WFields f = wfields(type);
    WPtrs w{};
    const uint8_t * b = (const uint8_t *) base;
    int64_t off = 0;
    for (int i = 0; i < f.n; i++) { w.p[i] = b + off + first * f.size[i]; off += T * f.size[i]; }
    return w;
}

// ---------------------------------------------------------------- activation scratch (q8_1, SoA, two planes)
// For a column of K f32 activations (NB = K/32 blocks):
//   qs : int8 [2][NB][16]  plane 0 = elements 0..15 of each block, plane 1 = elements 16..31
//   ds : {float d (the fp16-rounded scale of quantize_row_q8_1_ref), int16 isum0, int16 isum1} [NB]
//        isum_p = exact integer sum of the 16 quants of plane p
struct ActDS { float d; short isum0; short isum1; };
static_assert(sizeof(ActDS) == 8, "ActDS");

struct ActView {            // N columns
    const int8_t * qs; const ActDS * ds;
    int64_t NB;             // blocks per column
};

// ---------------------------------------------------------------- launchers (each returns cudaGetLastError())
// x: f32, row r at (const char*)x + r*row_stride_bytes.
cudaError_t launch_quantize_q8_1_soa(const float * x, int64_t row_stride_bytes, int64_t K, int64_t nrows,
                                     int8_t * qs, ActDS * ds, cudaStream_t st);
cudaError_t launch_quantize_q8_1_aos(const float * x, int64_t K, int64_t nrows, void * y, cudaStream_t st);

// ---- GEMV (gemv.cu).  Up to 3 matrices sharing one activation vector run as one launch.
struct GemvMat {
    WPtrs w; int64_t M;
    float * dst;                 // dst[c*dcs + row]
    const float * add_src;       // optional fused GGML_OP_ADD: add_dst[c*dcs + row] = add_src[c*dcs + row] + dst[...]
    float * add_dst;             // (nullptr = none)
};
struct GemvArgs {
    GemvMat mat[3]; int nmat; int cta_end[3];
    int steps; int64_t K; int64_t dcs;
    const float * x; int64_t x_col_stride;     // gemv6: the f32 activation column (quantised inside the kernel)
    ActView act;                                // gemv4 / gemv5 / gemv5n: pre-quantised scratch
    // row split over the GPUs of one box: every result is also stored into the peers' copies of dst over NVLink
    // (peer address = local address + peer_delta[p]; dst lies in the symmetric heap every rank allocated alike)
    int npeers; long long peer_delta[7];
    long long peer_flag[7];                  // (flag protocol: this rank's slot in peer p's flag array; used by the exchange kernel)
    // "LL" protocol (default for vector ops): instead of plain stores + a flag exchange, every result goes to the peers as ONE
    // 8-byte store {value, sequence number} into a staging area of the symmetric heap; the receivers poll the slots themselves
    // (ll_gather_kernel), so no fence, no flag and no second NVLink round trip.  seq = *ll_count + ll_op (ops of this submission
    // so far + 1); slot = ll_slots + (seq & 1) * LL_HALF_ELEMS + ll_off[mat] + ll_rank * M + row  (two halves: a rank can be at
    // most one exchange ahead of the slowest reader)
    uint2 * ll_slots; const uint32_t * ll_count; uint32_t ll_op; int ll_off[3]; int ll_rank;
    // quantise on exit (one column, qx_qs != null): the result vector of matrix qx_mat (the fused ADD's sum if it has one, else dst; M % 32 == 0)
    // is the next MUL_MAT's src1: the last CTA of every 32-element block collects the block through qx_slots ({value, flag} per row, all
    // zero between launches) and writes its q8_1 form into the OTHER activation scratch qx_qs / qx_ds -- no quantise launch in front of the next GEMV
    int8_t * qx_qs; ActDS * qx_ds; uint2 * qx_slots; int qx_mat;
};
constexpr int LL_HALF_ELEMS = 131072;                     // {value, seq} slots per staging half
struct LlGatherArgs { uint2 * ll_slots; const uint32_t * ll_count; uint32_t ll_op; int nmat; float * dst[3]; int M[3]; int ll_off[3]; int world, rank; };
// polls the peers' slots of one exchange and writes the values into the local dst vectors; starts while the GEMV still runs (PDL)
cudaError_t launch_ll_gather(const LlGatherArgs & a, cudaStream_t st);
// the gather + what follows it in the node list: the residual ADD on vector qmat (add_dst = gathered + add_src, or null) and the q8_1
// quantisation of the result into qs / ds for the MUL_MAT that consumes it (every M % 32 == 0)
struct LlFuseArgs { int qmat; const float * add_src; float * add_dst; int8_t * qs; ActDS * ds; };
cudaError_t launch_ll_gather_fused(const LlGatherArgs & a, const LlFuseArgs & f, cudaStream_t st);
// end of a submission: *ll_count += nops
cudaError_t launch_ll_bump(uint32_t * ll_count, uint32_t nops, cudaStream_t st);
// row-split prefill: copy this rank's rows x N slice of dst (local, column stride dcs floats) into every peer's dst (heap offset + delta)
cudaError_t launch_rowsplit_push(const float * slice, long long rows, long long N, long long dcs, int npeers, const long long * peer_delta, cudaStream_t st);
// one warp: tell every peer "my slice of the op is in your memory" (release, system scope), then wait for theirs
cudaError_t launch_rowsplit_exchange(uint32_t * local_flags, uint32_t * counter, int npeers, const long long * peer_flag_addr /* [npeers] */,
                                     const int * peer_rank /* [npeers] */, cudaStream_t st);
// 1 .. 8 columns of a.act (gemv5 tiles for large one-column launches, else the register path gemv4).  PDL unless B200_PDL=0.
cudaError_t launch_gemv2(int type, GemvArgs & a, int ncols, cudaStream_t st);
// all ncols (2 .. 31) columns of a.act in ONE launch over shared-memory tiles (gemv5n); false: not eligible, *err untouched
bool launch_gemv_cols(int type, GemvArgs & a, int ncols, cudaStream_t st, cudaError_t * err);
bool gemv_pdl_enabled();
// gemv6.cu: ONE f32 activation column a.x, quantised inside the kernel; persistent CTAs (one per SM), per-warp cp.async.bulk rings.
// false: not eligible (*err untouched) -- the caller quantises into a.act and runs gemv4 / gemv5
bool launch_gemv6(int type, GemvArgs & a, int sms, cudaStream_t st, cudaError_t * err);
void gemv6_trace(long long * buf, int cap_launches);      // development: 4 x %globaltimer per CTA of the next launches (4096 values per launch)
int  gemv6_trace_count();
// f32 weights (row stride in bytes), f32 activations (col stride bytes)
cudaError_t launch_gemv_f32(const float * w, int64_t w_row_stride, int64_t M, int64_t K,
                            const float * x, int64_t x_col_stride, int ncols,
                            float * dst, int64_t dst_col_stride, cudaStream_t st);
// scratch: sq = int8[K], sd = ActDS[K/32] (two-plane copy of a_qs built inside the kernel)
cudaError_t launch_block_dots_impl(int type, WPtrs w, int64_t K, const int8_t * a_qs, int32_t * out,
                                   int8_t * sq, ActDS * sd, cudaStream_t st);

// ---- prefill GEMM on the tensor cores (gemm_i8.cu): int8 tcgen05.mma per 32-element block, fp32 scale-accumulate
// mode 0: q [N][K], dB [K/32][ldb];  1: + d * isum in [K/32][ldb] behind dB (q4_K);  2: 16-element groups spread to 32-byte
// blocks, q [N][2K], dB [K/16][ldb] (q6_K)
cudaError_t launch_quantize_rows_i8(const float * x, long long row_stride_bytes, int K, int N, int8_t * q, float * dB, long long ldb, int mode, cudaStream_t st);
cudaError_t launch_unpack_q4_0_i8(const void * qs, long long nblocks, int8_t * out, cudaStream_t st);
cudaError_t launch_unpack_q4_K_i8(const WPtrs & w, long long nsuper, int8_t * out, float * wS, float * wM, cudaStream_t st);
cudaError_t launch_unpack_q6_K_i8(const WPtrs & w, long long nsuper, int8_t * out, float * wS, cudaStream_t st);
int gemm_i8_ldb(int N);
// kind 0: int8 operands (kind::i8 MMA); 1: the same integers as fp16 (kind::f16, f32 sums); 2: fp16 with 16-element blocks (q6_K)
cudaError_t launch_gemm_i8(int kind, const void * A, const __half * dA_h, const float * dA_f, const float * mA, const void * B, const float * dB,
                           long long ldb, float * out, long long out_cs, int M, int N, int K, int sms, float * split_ws, cudaStream_t st);
int gemm_i8_ksplit(int M, int N, int K, int kind, int sms);     // > 1: launch_gemm_i8 wants a workspace of that many x N x M floats
// fp16 operand preparation for kind 1 / 2 (modes as launch_quantize_rows_i8, without the q6_K padding)
cudaError_t launch_quantize_rows_f16(const float * x, long long row_stride_bytes, int K, int N, __half * q, float * dB, long long ldb, int mode, cudaStream_t st);
cudaError_t launch_unpack_f16(int type, const WPtrs & w, long long nblocks, __half * out, float * wS, float * wM, cudaStream_t st);
// res_dst != null: also res_dst[i] = res_src[i] + out[i] over [N][M] contiguous (the residual ADD node behind the MUL_MAT, fused)
cudaError_t launch_reduce_splits(const float * ws, int S, long long MN, int M, float * out, long long out_cs, int add_to_out, cudaStream_t st,
                                 const float * res_src = nullptr, float * res_dst = nullptr, void * act_out = nullptr);
// a producer of a dense GEMM's src1 that also converts it (gemm_dense.cu row_act_kernel): mode 0 RMS_NORM * w, 1 SILU(a) * b, 2 gather (CONT)
constexpr int ROW_ACT_MAX_K = 14 * 1024;
struct RowActArgs {
    const char * a; const char * b; char * y; void * act; float * colscale;
    long long as, bs, ys;         // byte strides between rows of a, b (mode 1), y
    int K; float eps;
    long long nes[4], nbs[4];     // mode 2: logical shape and byte strides of the source
};
cudaError_t launch_row_act(int mode, const RowActArgs & p, int N, cudaStream_t st);
bool reduce_splits_can_fuse_residual(long long M, long long out_cs, const void * out, const void * res_src, const void * res_dst);
// ---- dense prefill GEMM (gemm_dense.cu): block scales folded into fp16 operands, the whole K accumulated in TMEM
cudaError_t launch_dequant_act_f16(const float * x, long long row_stride_bytes, int K, int N, __half * out, float * colscale, cudaStream_t st);
int gemm_dense_ksplit(int M, int N, int K, int sms);            // > 1: launch_gemm_dense wants a workspace of that many x N x M floats
// A: quantised weights of `type` in the SoA layout (dequantised inside the kernel); B: fp16 activations [N][K] scaled per column, colscale: the
// power of two each dst column is multiplied by (both from launch_dequant_act_f16; colscale holds N rounded up to 256 floats, 16-byte aligned)
cudaError_t launch_gemm_dense(int type, const WPtrs & A, const __half * B, const float * colscale, float * out, long long out_cs, int M, int N, int K, int sms,
                              float * split_ws,
                              float * tail_ws,              // gemm_dense_tail_ws_bytes(sms) of scratch for the stream-K tail of the last wave, or nullptr (none)
                              int * nlaunch, cudaStream_t st,      // *nlaunch: kernels launched (GEMM, + tail fix-up, + split reduce)
                              const float * res_src = nullptr, float * res_dst = nullptr,   // residual ADD fused into the split-K reduce (needs split_ws and ksplit > 1)
                              void * act_out = nullptr, float * act_cs_out = nullptr);   // + the next dense GEMM's activations / column scales from res_dst (M % 32 == 0, M <= 8192)
size_t gemm_dense_tail_ws_bytes(int sms);
// 2..3 weight matrices of one type over the same activations in one launch (q / k / v, gate / up)
cudaError_t launch_gemm_dense_group(int type, const WPtrs * A, float * const * out, const long long * out_cs, const int * M, int nmat,
                                    const __half * B, const float * colscale, int N, int K, int sms, float * tail_ws, int * nlaunch, cudaStream_t st);

struct AddArgs {
    const char * a; const char * b; char * d;
    int64_t ne[4];      // dst / src0 shape
    int64_t neb[4];     // src1 shape
    int64_t nba[4], nbb[4], nbd[4];   // byte strides
    int mul;            // 0: a + b   1: a * b (GGML_OP_MUL, same broadcast rules)
};
cudaError_t launch_add_f32(const AddArgs & a, cudaStream_t st);

// ---- graph_ops.cu: the ops between the MUL_MATs of a Llama layer (all strides in bytes, ggml order)
struct RopeArgs {
    const char * x; char * y; const int32_t * pos; const float * ff /* frequency factors [n_dims / 2] or NULL */;
    int64_t ne[4], nbs[4], nbd[4];
    int n_dims, neox;
    float theta_scale, freq_scale, ext_factor, attn_factor, corr0, corr1;
    // optional second copy of the result as f16, flat in logical order (the K-cache store that follows ROPE(K)): at st_y, or where *st_dyn points
    char * st_y; char * const * st_dyn;
};
cudaError_t launch_rope_f32(const RopeArgs & p, bool f16, cudaStream_t st);       // f16: both sides __half (rotated in f32)
struct CpyArgs { const char * x; char * y; char * const * dyn /* device: where this launch's dst starts, or NULL = y */; int64_t nes[4], nbs[4], ned[4], nbd[4]; };
cudaError_t launch_cpy(const CpyArgs & p, bool src_f16, bool dst_f16, cudaStream_t st);
struct GetRowsArgs { const char * x; const char * ids; char * y; int64_t ne00, ne01, ne10, ne11, ne12; int64_t nbs[4], nbi[3], nbd[4]; };
cudaError_t launch_get_rows_f32(const GetRowsArgs & p, cudaStream_t st);
struct MmF16Args { const char * a; const char * b; char * d; int64_t M, N, K, ne12, ne13, r2, r3; int64_t nba[4], nbb[4], nbd[4]; };
cudaError_t launch_mm_f16(const MmF16Args & p, cudaStream_t st);
// one query column: K^T q, soft_max(scale * . + mask), V p in one launch (n_kv <= AD_MAX_KV; every operand row contiguous and aligned)
constexpr int AD_MAX_PER_THREAD = 16; constexpr int AD_MAX_KV = 256 * AD_MAX_PER_THREAD;
struct AttnDecodeArgs {
    const char * k; const char * v; const char * q; const float * mask; char * prob; char * out;
    int64_t hd, n_kv, r2, nbk1, nbk2, nbv1, nbv2, nbq2, nbp2, nbo2; float scale;
};
cudaError_t launch_attn_decode(const AttnDecodeArgs & p, int n_head, cudaStream_t st);
cudaError_t launch_silu_f32(const float * x, float * y, int64_t n, cudaStream_t st);
cudaError_t launch_silu_mul_f32(const float * g, const float * u, float * y, int64_t n, int64_t K, int8_t * qs, ActDS * ds, cudaStream_t st);
cudaError_t launch_rms_norm_mul(const float * x, const float * w, float * y, int64_t ne00, int64_t rows, int64_t xs, int64_t ys, float eps,
                                int8_t * qs, ActDS * ds, cudaStream_t st);
cudaError_t launch_cont_quant(const CpyArgs & p, int64_t K, int8_t * qs, ActDS * ds, cudaStream_t st);

// ---- norm_ops.cu: RMS_NORM / SOFT_MAX / POOL_2D (ne, nb: ggml order, bytes)
cudaError_t launch_rms_norm_f32(const void * x, void * y, const int64_t * ne, const int64_t * nb_src, const int64_t * nb_dst, float eps, cudaStream_t st);
cudaError_t launch_soft_max_f32(const float * x, const void * mask, bool mask_f16, float * y, const int64_t * ne, float scale, float max_bias, cudaStream_t st);
cudaError_t launch_pool_2d(const void * x, bool src_f16, float * y, const int64_t * ne, const int64_t * nb, int64_t px, int64_t py, const int32_t * opts, cudaStream_t st);

cudaError_t launch_repack(int type, bool to_soa, uint8_t * soa_base, int64_t T, int64_t first, int64_t n,
                          uint8_t * native, cudaStream_t st);
cudaError_t launch_fill(void * p, size_t n_bytes, uint32_t v, cudaStream_t st);

}  // namespace b200

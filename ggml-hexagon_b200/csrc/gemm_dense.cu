// gemm_dense.cu — prefill MUL_MAT as a DENSE fp16 tensor-core contraction: tcgen05.mma kind::f16 accumulating the whole K in
// TMEM, one TMEM read per output tile, the quantised weights dequantised INSIDE the kernel.
//
// Same seat as gemm_i8.cu (ggml_compute_forward_mul_mat_one_chunk for src1 with many columns, ggml-dsp.c:1091-1189 /
// ggml-cpu.c:6655-6743), different contract.  gemm_i8.cu keeps every per-block integer sum exact and therefore has to pull each
// 32-element block's 128 x 128 sums out of the tensor core and scale them on the CUDA cores: it is bound by that epilogue, not
// by the tensor pipe.  Here the block scales are folded into the OPERANDS instead:
//     A[m, k] = fp16( d_w[m, b(k)] * q_w[m, k] )      = dequantize_row_<type> (ggml-quants.c:307-325, 347-363, 1352-1374,
//                                                        1681-1710) rounded to fp16
//     B[n, k] = fp16( d_a[n, b(k)] * q_a[n, k] )      = the q8_1 quantisation the reference applies to src1
//                                                        (quantize_row_q8_1_ref, ggml-quants.c:220-253), dequantised again
//     dst[m, n] = sum_k A[m, k] * B[n, k]             f32 accumulation in TMEM
// so the contraction is dense and the accumulator runs over the whole K.  What changes against the reference is only the
// rounding of the two dequantised operands to fp16 (relative 2^-12 each; NMSE against ggml-cpu ~1e-7, the bar is 5e-4) and the
// f32 summation order.  The q8_1 step is kept so that the result follows the reference's own quantisation of the activations.
// Operand ranges (the reference's f32 block dots have no such limits beyond the fp16 block scales themselves):
//   * activations: every column n of B is scaled by a power of two 2^g(n) that brings its largest |d_a q_a| into [2^8, 2^9) (so that
//     |x| up to the reference's own limit, d_a = amax / 127 <= 65504, does not saturate), the epilogue multiplies dst column n by 2^-g(n):
//     exact, the product is the same as without the scale wherever fp16 could hold the unscaled operand;
//   * q4_0 / q8_0 weights: d_w is an fp16 number already and |q| <= 128, so d_w q_w is exact in the subnormal range (an integer times
//     2^-24) and rounds once above it; it overflows only for |d_w| > 511, which no quantiser produces for weights that fit fp16;
//   * K-quant weights: d sc q - dmin m is computed in f32 and would lose bits below 2^-14: the converters scale a row whose super-block
//     scale max(|d|, |dmin|) is below 2^-14 up by a power of two (taken from the first super-block of the work unit), the epilogue
//     multiplies the row by its inverse.
//
// A never exists in HBM as fp16 (round 1 kept 2-byte "shadows" of every weight matrix: 3.6x the q4_0 bytes in HBM and in L2 -> SM
// traffic): eight CONVERTER warps per CTA read the native SoA blocks (16 - 64 bytes per row and 64-element K slab, 128-bit loads,
// two slabs ahead in registers), dequantise with packed fp16 / fp32 arithmetic and write the K-major, 128-byte-swizzled fp16 stage
// the MMA descriptor expects -- the layout a TMA tile load would have produced.
//
// CTA pair (cluster of 2, one TPC), 256 x BN tile per pair, 18 warps per CTA:
//   warp 0        TMA producer: B = fp16 activations [N][K], boxes of 64 k (128 bytes) x BN / 2 rows, 128-B swizzle
//   warp 1        (leader CTA) MMA issue: tcgen05.mma.cta_group::2 kind::f16, M = 256, 4 K16 MMAs per stage
//   warps 2..9    epilogue: tcgen05.ld 32 lanes x 32 columns, coalesced f32 stores; the other accumulator buffer is in use meanwhile
//   warps 10..17  converters: thread = (row, half slab): native blocks -> fp16 -> swizzled stage; fence.proxy.async; arrive
#include "tc_common.cuh"
#include "wtypes.cuh"

#include <stdio.h>
#include <vector>

namespace b200 {

constexpr int GD_KS = 64;                                    // fp16 elements of K per stage = one 128-B swizzle row
constexpr int GD_EPI_WARPS = 8;                              // 4 TMEM lane quarters x 2 column halves
constexpr int GD_CVT_WARPS = 8;                              // converters: 128 rows x 2 half slabs of 32 elements
constexpr int GD_CVT_WARP0 = 2 + GD_EPI_WARPS;
constexpr int GD_THREADS = 64 + GD_EPI_WARPS * 32 + GD_CVT_WARPS * 32;

struct GemmDenseArgs {
    float * out; long long out_cs;              // dst[n * out_cs + m]
    int M, N, K;
    int mtiles, ntiles;
    int ksplit; long long split_stride;        // split-K: work item = (tile, split); split sp writes its partial sums to out + sp * split_stride
    unsigned long long * trace;                // B200_GEMM_TRACE=1: 8 globaltimer stamps per CTA (pair kernel), else nullptr
    // pair kernel: up to GD_MAXMAT weight matrices over the same activations in one launch (q / k / v, gate / up): the tile list
    // runs through all of them, which fills the waves that one 4096-row matrix cannot (32 tiles for 74 CTA pairs)
    int nmat;
    const uint8_t * w[3][4];                   // SoA field arrays of each weight matrix (b200_internal.cuh), at the matrix's first block
    float * outs[3]; long long out_css[3]; int Ms[3];
    int tile_end[3];                           // tiles (m-tile x n-tile) up to and including matrix j; INT_MAX past nmat
    // the last, partial wave as a stream-K tail (tail_T > 0; ksplit == 1): its `tail_T / nks` tiles are cut into one chunk of tail_L
    // consecutive (tile, slab) steps per CTA pair, so that a wave with few tiles left costs tiles / pairs of a tile time instead of a whole
    // one.  A chunk is at most two segments (the end of one tile, the beginning of the next); segment s of pair c leaves its 256 x BN
    // partial sums in tail_ws slot 2 c + s, and gd_tail_fixup_kernel adds the slots of every tile in chunk order (deterministic).
    int full_units, tail_T, tail_L;
    float * tail_ws;
    const float * colscale;                    // [N rounded up to 256]: dst column n = accumulator x colscale[n] (launch_dequant_act_f16 wrote B x 1 / colscale)
};
struct GdUnit { int tile, sp, ks0, ks1, slot; };    // slabs [ks0, ks1) of `tile`; slot >= 0: partial sums to that tail_ws slot
__device__ __forceinline__ GdUnit gd2_unit(const GemmDenseArgs & p, const int work, const int nks, const int pair, const int npairs) {
    GdUnit u;
    if (work < p.full_units) {
        u.tile = work / p.ksplit; u.sp = work - u.tile * p.ksplit; u.slot = -1;
        u.ks0 = (int) ((long long) nks * u.sp / p.ksplit); u.ks1 = (int) ((long long) nks * (u.sp + 1) / p.ksplit);
    } else {
        const int seg = work - p.full_units >= npairs ? 1 : 0;
        const int a = pair * p.tail_L, b = min(a + p.tail_L, p.tail_T);
        const int t0 = a / nks, e0 = (t0 + 1) * nks;
        u.sp = 0; u.slot = 2 * pair + seg;
        if (seg == 0) { u.tile = p.full_units + t0; u.ks0 = a - t0 * nks; u.ks1 = min(b, e0) - t0 * nks; }      // (a >= b: empty)
        else          { u.tile = p.full_units + t0 + 1; u.ks0 = 0; u.ks1 = b - e0; }                             // (b <= e0: empty)
    }
    return u;
}
constexpr int GD_MAXMAT = 3;
__device__ __forceinline__ void gd_stamp(const GemmDenseArgs & p, int slot) {
    if (p.trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(size_t) blockIdx.x * 8 + slot] = t;
    }
}

// Epilogue of one warp: its 32 TMEM lanes (= 32 consecutive rows m, one per lane) x NCH chunks of 32 columns, written to
// colp[i * cs] (a column of dst is contiguous in m, so every store instruction is one 128-byte line).  The pointer walks by cs:
// recomputing (n0 + i) * cs + m with its bounds test per element cost ~18 instructions and a branch per store, 7 us per
// 128 x 256 tile (B200_GEMM_TRACE timeline); the next chunk's tcgen05.ld is in flight while a chunk is stored.
__device__ __forceinline__ uint4 gd_lds16(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
// csc: shared-memory address of the chunk's 32 column scales; rs: the row's scale; both powers of two
__device__ __forceinline__ void gd_store_chunk(float * colp, const long long cs, const int (&v)[32], const int ncols, const uint32_t csc, const float rs) {
    #pragma unroll
    for (int i4 = 0; i4 < 8; i4++) {
        const uint4 t = gd_lds16(csc + i4 * 16);                   // (warp-uniform address: a broadcast)
        const float f[4] = {__uint_as_float(t.x) * rs, __uint_as_float(t.y) * rs, __uint_as_float(t.z) * rs, __uint_as_float(t.w) * rs};
        #pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = 4 * i4 + k;
            if (ncols >= 32 || i < ncols) *colp = __int_as_float(v[i]) * f[k];
            colp += cs;
        }
    }
}
template <int NCH>
__device__ __forceinline__ void gd_epilogue_warp(const uint32_t taddr, float * colp, const long long cs, const int ncols, const bool row_ok,
                                                 const uint32_t csc, const float rs) {
    static_assert(NCH % 2 == 0, "chunks are processed in pairs");
    int va[32], vb[32];
    gm_tmem_ld32(taddr, va);
    gm_tmem_ld_wait();
    #pragma unroll
    for (int c = 0; c < NCH; c += 2) {
        gm_tmem_ld32(taddr + (c + 1) * 32, vb);
        if (row_ok && ncols > c * 32) gd_store_chunk(colp + (long long) c * 32 * cs, cs, va, ncols - c * 32, csc + c * 128, rs);
        gm_tmem_ld_wait();
        if (c + 2 < NCH) gm_tmem_ld32(taddr + (c + 2) * 32, va);
        if (row_ok && ncols > (c + 1) * 32) gd_store_chunk(colp + (long long) (c + 1) * 32 * cs, cs, vb, ncols - (c + 1) * 32, csc + (c + 1) * 128, rs);
        if (c + 2 < NCH) gm_tmem_ld_wait();
    }
}
// K-quants: the power of two 2^f (f >= 0) that brings a row whose super-block scale dmax = max(|d|, |dmin|) lies below 2^-14 into [2^-14, 2^-13):
// the smallest non-zero element d sc q (sc, q integers) is a normal fp16 number then, and the other super-blocks of the unit may have scales
// up to 2^17 times the first one's before 127 x 32 x d' passes 65504 (the conversion saturates there)
__device__ __forceinline__ float gd_row_up(const float dmax) {
    if (!(dmax > 0.0f)) return 1.0f;
    const int E = (int) ((__float_as_uint(dmax) >> 23) & 0xFFu) - 126;        // dmax = m 2^E, m in [0.5, 1)   (fp16 subnormals are normal f32 numbers)
    const int f = -13 - E;
    return f > 0 ? __uint_as_float((uint32_t) (f + 127) << 23) : 1.0f;
}
__device__ __forceinline__ float gd_hdr_dmax(const int type, const uint32_t w) {      // q4_K: w = {d, dmin} halfs; q6_K: w = d in the low half
    const float d = fabsf(__half2float(__ushort_as_half((unsigned short) (w & 0xFFFFu))));
    return type == T_Q4_K ? fmaxf(d, fabsf(__half2float(__ushort_as_half((unsigned short) (w >> 16))))) : d;
}

// ------------------------------------------------------------------------------------------------ the CTA-pair kernel
// Two CTAs of one TPC (cluster of 2) share one 256 x BN tile through tcgen05.mma.cta_group::2: every CTA converts its own 128 weight
// rows and stages HALF of the activation rows, the MMA of the leader reads B from both shared memories, and each CTA's TMEM receives
// its 128 rows x BN columns.  (A 128 x 128 tile per SM wants 128 B/clk/SM of operands out of L2, which delivers ~43: round 1.)
//   full[s]   leader only; both CTAs' TMA loads of B (.cta_group::2) complete_tx on it, the leader posts the expected bytes of both
//   afull[s]  leader only; the 8 converter warps of BOTH CTAs arrive on it when their part of the stage's A tile is written and fenced
//   empty[s]  per CTA; tcgen05.commit.multicast arrives on both when the stage's MMAs have read shared memory
//   tfull[b]  per CTA; multicast commit after a tile's last MMA
//   tempty[b] leader only; the epilogue warps of BOTH CTAs arrive on it (remote arrive from the peer)
// Native blocks reach the converters through per-thread rings of cp.async copies (LDGSTS): DEPTH slabs in flight per thread cost no
// registers and -- unlike plain loads -- are not waited for by the MEMBAR that fence.proxy.async carries (with register-resident
// prefetches every stage's fence waited for the loads of the slab DEPTH stages ahead: 0.63 us per stage against 0.3 for the MMAs).
// Ring layout: N16 arrays [DEPTH][256 threads][16 B] and N4 arrays [DEPTH][256][4 B]: consecutive lanes, consecutive 16 B.
__host__ __device__ constexpr int gd2_n16(int type) { return type == T_Q4_0 ? 1 : type == T_Q8_0 ? 2 : type == T_Q4_K ? 3 : 4; }
__host__ __device__ constexpr int gd2_n4(int type)  { return type == T_Q4_K ? 0 : type == T_Q6_K ? 2 : 1; }
__host__ __device__ constexpr int gd2_depth(int type) { return type == T_Q4_0 ? 6 : type == T_Q8_0 ? 6 : type == T_Q4_K ? 5 : 3; }
__host__ __device__ constexpr int gd2_raw_bytes(int type) { return gd2_depth(type) * GD_CVT_WARPS * 32 * (16 * gd2_n16(type) + 4 * gd2_n4(type)); }
__host__ __device__ constexpr int gd2_stage_bytes(int bn) { return GM_BM * 128 + (bn / 2) * 128; }
__host__ __device__ constexpr int gd2_stages(int bn, int type) { return (232448 - 1280 - 2048 - gd2_raw_bytes(type)) / gd2_stage_bytes(bn); }
__host__ __device__ constexpr int gd2_smem(int bn, int type) { return gd2_stages(bn, type) * gd2_stage_bytes(bn) + gd2_raw_bytes(type) + 256 + 2048 + 1024; }
static_assert(gd2_stages(256, T_Q4_0) == 6 && gd2_stages(256, T_Q8_0) >= 5 && gd2_stages(256, T_Q4_K) >= 5 && gd2_stages(256, T_Q6_K) >= 5, "stage count");
// WTMA variant (the default): the native blocks arrive by TMA instead -- per-thread LDGSTS gathers of 16 + 4 bytes from 32 different rows cost
// 32 L1 tag cycles per warp instruction, ~512 per stage against the 512 clocks the stage's MMAs take (without the 4-byte scale gather a
// 4096^2 x 512 q4_0 GEMM went from 32.4 to 27.8 us).  One more warp per CTA loads RAW STAGES of gd3_rs slabs: for every SoA field a box
// of 128 rows x the field's bytes of those slabs (128-byte rows 128-B swizzled, so that the converters' 16-byte reads of consecutive rows
// spread over all banks), completing on rfull[rs]; the converter warps release a raw stage through rempty[rs] (8 arrivals).
// Needs 16-byte aligned field rows (q4_0 / q8_0 scale rows: K % 256 == 0), else the ring variant above runs.
// q6_K's d (2 bytes per row and super-block: no 16-byte box) stays a per-thread cp.async ring of GD3_DD super-blocks.
constexpr int GD_W_WARP = GD_CVT_WARP0 + GD_CVT_WARPS;       // the weight producer warp of the WTMA variant
constexpr int GD3_DD = 3;
__host__ __device__ constexpr int gd3_rs(int type) { return type == T_Q8_0 ? 2 : 4; }          // slabs per raw stage (K-quants: one super-block)
__host__ __device__ constexpr int gd3_rd(int type) { return type == T_Q6_K ? 2 : 3; }          // raw stages in flight
__host__ __device__ constexpr int gd3_nf(int type) { return type == T_Q6_K ? 3 : 2; }          // fields that come by TMA (= SoA field index)
__host__ __device__ constexpr int gd3_box(int type, int f) { return f == 0 ? 128 : (type == T_Q6_K && f == 1 ? 64 : 16); }      // box bytes along the row
__host__ __device__ constexpr int gd3_adv(int type, int f) { return f == 0 ? 128 : (type == T_Q6_K && f == 1 ? 64 : (type == T_Q8_0 ? 8 : 16)); }      // row bytes per raw stage
__host__ __device__ constexpr int gd3_foff(int f) { return f == 0 ? 0 : (f == 1 ? 16384 : 16384 + 8192); }
__host__ __device__ constexpr int gd3_raw_stage(int type) { return type == T_Q6_K ? 16384 + 8192 + 2048 : 16384 + 2048; }
__host__ __device__ constexpr int gd3_small_off(int type) { return type == T_Q6_K ? 16384 + 8192 : 16384; }      // the [128][16 B] scale / header tile
__host__ __device__ constexpr int gd3_raw_bytes(int type) { return gd3_rd(type) * gd3_raw_stage(type) + (type == T_Q6_K ? GD3_DD * GD_CVT_WARPS * 32 * 4 : 0); }
__host__ __device__ constexpr int gd3_stages(int bn, int type) { return (232448 - 1024 - 512 - 2048 - gd3_raw_bytes(type)) / gd2_stage_bytes(bn); }
__host__ __device__ constexpr int gd3_smem(int bn, int type) { return gd3_stages(bn, type) * gd2_stage_bytes(bn) + gd3_raw_bytes(type) + 512 + 2048 + 1024; }
static_assert(gd3_stages(256, T_Q4_0) >= 5 && gd3_stages(256, T_Q8_0) >= 5 && gd3_stages(256, T_Q4_K) >= 5 && gd3_stages(256, T_Q6_K) >= 5, "stage count");
struct GdMaps { CUtensorMap b[1]; CUtensorMap w[3][3]; };      // activations; weights [matrix][field]

__device__ __forceinline__ uint32_t gd2_mapa(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void gd2_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void gd2_tma_2d(uint32_t dst, const CUtensorMap * map, uint32_t bar_cluster, int c0, int c1) {
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
                 :: "r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
// whole-warp forms (one elected lane issues), see tc_common.cuh
__device__ __forceinline__ void gd2_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|e, 0xffffffff;\n\t@e tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void gd2_commit_both(uint32_t bar) {          // same barrier offset in both CTAs of the pair
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}"
                 :: "r"(bar), "h"((uint16_t) 3) : "memory");
}
__device__ __forceinline__ void gd2_arrive_cluster(uint32_t bar_cluster) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" :: "r"(bar_cluster) : "memory");
}
// remote arrive for the converters, once per stage and warp: default (.release.cta) semantics, as CUTLASS' ClusterBarrier::arrive(cta_id).
// The .release.cluster form above compiles to MEMBAR.ALL.GPU + ERRBAR + CGAERRBAR (seen in SASS) -- about a microsecond per stage; what the
// arrive has to order is this CTA's own shared-memory stores, already made visible to the async proxy by fence.proxy.async.
__device__ __forceinline__ void gd2_arrive_remote(uint32_t bar_cluster) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" :: "r"(bar_cluster) : "memory");
}
// predicated forms (no branch around the single-lane arrive: BSSY / BRA / BSYNC per use otherwise)
__device__ __forceinline__ void gd2_arrive_remote_if(uint32_t bar_cluster, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %1, 0;\n\t@p mbarrier.arrive.shared::cluster.b64 _, [%0];\n\t}" :: "r"(bar_cluster), "r"((uint32_t) pred) : "memory");
}
__device__ __forceinline__ void gd2_arrive_local_if(uint32_t bar, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 st;\n\tsetp.ne.b32 p, %1, 0;\n\t@p mbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(bar), "r"((uint32_t) pred) : "memory");
}
__device__ __forceinline__ void gd2_mbar_wait_cluster(uint32_t bar, uint32_t parity) {          // acquire at cluster scope: the arrivals are remote
    uint32_t ok = 0;
    const long long t0 = clock64();
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && clock64() - t0 > GM_TIMEOUT_CYCLES) __trap();
    }
}

// ---- converters: native SoA blocks -> fp16(d * q), 32 consecutive elements of one row per thread and stage.
// Small unsigned integers become fp16 by OR-ing them into the mantissa of 1024.0 (0x6400): exact, one PRMT per two values; the
// bias is removed by an exact HSUB2, the scale applied by ONE rounding HMUL2 -- the same single rounding as fp16(float(d) * q).
__device__ __forceinline__ uint32_t gd_prmt(uint32_t a, uint32_t b, uint32_t sel) { uint32_t r; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel)); return r; }
__device__ __forceinline__ uint32_t gd_h2sub(uint32_t a, uint32_t b) { uint32_t r; asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t gd_h2mul(uint32_t a, uint32_t b) { uint32_t r; asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
// 4 bytes of w (values 0 .. 255) -> 4 halfs (v - bias) * d2; bias2 = {1024 + bias, 1024 + bias}
__device__ __forceinline__ uint2 gd_u8x4_h(uint32_t w, uint32_t bias2, uint32_t d2) {
    const uint32_t lo = gd_prmt(w, 0x64646464u, 0x5140u), hi = gd_prmt(w, 0x64646464u, 0x5342u);
    return make_uint2(gd_h2mul(gd_h2sub(lo, bias2), d2), gd_h2mul(gd_h2sub(hi, bias2), d2));
}
__device__ __forceinline__ uint32_t gd_pack_h2f(float a, float b) { uint32_t r; asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }
// 4 bytes of w (values 0 .. 255) -> 4 halfs of fp16(ds * v - ms), the product and difference in fp32 (one FFMA: a single rounding)
__device__ __forceinline__ uint2 gd_u8x4_f(uint32_t w, float ds, float ms) {
    const float f0 = __fmaf_rn(ds, (float) (w & 0xFFu), -ms), f1 = __fmaf_rn(ds, (float) ((w >> 8) & 0xFFu), -ms);
    const float f2 = __fmaf_rn(ds, (float) ((w >> 16) & 0xFFu), -ms), f3 = __fmaf_rn(ds, (float) (w >> 24), -ms);
    return make_uint2(gd_pack_h2f(f0, f1), gd_pack_h2f(f2, f3));
}
__device__ __forceinline__ uint32_t gd_h2fma(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
// 4 bytes of w (values 0 .. 255) -> 4 halfs v * ds2 + nms2 (v exact in fp16, one rounding in the FMA; ds2 / nms2 = the f32 scale and negated
// minimum rounded to fp16 once per sub-block): the K-quants' packed path, ~2.3x fewer instructions than I2F + FFMA + cvt per element
__device__ __forceinline__ uint2 gd_u8x4_hf(uint32_t w, uint32_t ds2, uint32_t nms2) {
    const uint32_t lo = gd_prmt(w, 0x64646464u, 0x5140u), hi = gd_prmt(w, 0x64646464u, 0x5342u);
    return make_uint2(gd_h2fma(gd_h2sub(lo, 0x64006400u), ds2, nms2), gd_h2fma(gd_h2sub(hi, 0x64006400u), ds2, nms2));
}
struct GdRaw { uint4 a, b, c, e; uint32_t s0, s1; float fs; };      // fs: K-quant row scale (power of two)      // one slab's native bytes of a thread (unused members cost nothing)
__device__ __forceinline__ void gd_cpa16(uint32_t dst, const void * src) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory"); }
__device__ __forceinline__ void gd_cpa4(uint32_t dst, const void * src)  { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst), "l"(src) : "memory"); }
__device__ __forceinline__ uint32_t gd_lds4(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
// slot addressing: q16[i] / q4[i] = this thread's 16-byte / 4-byte cell of field i in ring slot `slot`
struct GdSlot { uint32_t q16[4], q4[2]; };
template <int TYPE> struct GdCvt;
// q4_0: one 16-byte block = the thread's 32 elements (low nibbles = elements 0..15, high = 16..31); y = d * (q - 8)
template <> struct GdCvt<T_Q4_0> {
    static __device__ __forceinline__ void issue(const GdSlot & sl, const uint8_t * const * w, long long row, int K, int ks, int hk) {
        const long long blk = row * (K >> 5) + ks * 2 + hk;
        gd_cpa16(sl.q16[0], w[0] + blk * 16);
        gd_cpa4(sl.q4[0], w[1] + ((blk * 2) & ~3ll));              // the aligned word that holds the block's 2-byte scale
    }
    static __device__ __forceinline__ void fetch(GdRaw & r, const GdSlot & sl, const uint8_t * const *, long long row, int K, int ks, int hk) {
        r.a = gd_lds16(sl.q16[0]);
        const uint32_t wd = gd_lds4(sl.q4[0]);
        r.s0 = ((row * (K >> 5) + ks * 2 + hk) & 1) ? wd >> 16 : wd & 0xFFFFu;
    }
    static __device__ __forceinline__ void convert(const GdRaw & r, int, int, uint4 (&o)[4]) {
        const uint32_t d2 = r.s0 | (r.s0 << 16), bias = 0x64086408u, m = 0x0F0F0F0Fu;          // 1032 = 1024 + 8
        const uint2 a = gd_u8x4_h(r.a.x & m, bias, d2), b = gd_u8x4_h(r.a.y & m, bias, d2), c = gd_u8x4_h(r.a.z & m, bias, d2), d = gd_u8x4_h(r.a.w & m, bias, d2);
        const uint2 e = gd_u8x4_h((r.a.x >> 4) & m, bias, d2), f = gd_u8x4_h((r.a.y >> 4) & m, bias, d2), g = gd_u8x4_h((r.a.z >> 4) & m, bias, d2), h = gd_u8x4_h((r.a.w >> 4) & m, bias, d2);
        o[0] = make_uint4(a.x, a.y, b.x, b.y); o[1] = make_uint4(c.x, c.y, d.x, d.y);
        o[2] = make_uint4(e.x, e.y, f.x, f.y); o[3] = make_uint4(g.x, g.y, h.x, h.y);
    }
};
// q8_0: one 32-byte block; y = d * q, q signed: q + 128 is the unsigned byte q ^ 0x80
template <> struct GdCvt<T_Q8_0> {
    static __device__ __forceinline__ void issue(const GdSlot & sl, const uint8_t * const * w, long long row, int K, int ks, int hk) {
        const long long blk = row * (K >> 5) + ks * 2 + hk;
        gd_cpa16(sl.q16[0], w[0] + blk * 32); gd_cpa16(sl.q16[1], w[0] + blk * 32 + 16);
        gd_cpa4(sl.q4[0], w[1] + ((blk * 2) & ~3ll));
    }
    static __device__ __forceinline__ void fetch(GdRaw & r, const GdSlot & sl, const uint8_t * const *, long long row, int K, int ks, int hk) {
        r.a = gd_lds16(sl.q16[0]); r.b = gd_lds16(sl.q16[1]);
        const uint32_t wd = gd_lds4(sl.q4[0]);
        r.s0 = ((row * (K >> 5) + ks * 2 + hk) & 1) ? wd >> 16 : wd & 0xFFFFu;
    }
    static __device__ __forceinline__ void convert(const GdRaw & r, int, int, uint4 (&o)[4]) {
        const uint32_t d2 = r.s0 | (r.s0 << 16), bias = 0x64806480u, x = 0x80808080u;          // 1152 = 1024 + 128
        const uint2 a = gd_u8x4_h(r.a.x ^ x, bias, d2), b = gd_u8x4_h(r.a.y ^ x, bias, d2), c = gd_u8x4_h(r.a.z ^ x, bias, d2), d = gd_u8x4_h(r.a.w ^ x, bias, d2);
        const uint2 e = gd_u8x4_h(r.b.x ^ x, bias, d2), f = gd_u8x4_h(r.b.y ^ x, bias, d2), g = gd_u8x4_h(r.b.z ^ x, bias, d2), h = gd_u8x4_h(r.b.w ^ x, bias, d2);
        o[0] = make_uint4(a.x, a.y, b.x, b.y); o[1] = make_uint4(c.x, c.y, d.x, d.y);
        o[2] = make_uint4(e.x, e.y, f.x, f.y); o[3] = make_uint4(g.x, g.y, h.x, h.y);
    }
};
// q4_K: the slab is 32 bytes of super-block ks / 4, sub-blocks 2j (low nibbles, hk = 0) and 2j + 1 (high nibbles), j = ks % 4;
// y = (d * sc_g) * q - dmin * m_g, scales / mins 6-bit packed in the 16-byte header (get_scale_min_k4, ggml-quants.c:631-638)
template <> struct GdCvt<T_Q4_K> {
    static __device__ __forceinline__ void issue(const GdSlot & sl, const uint8_t * const * w, long long row, int K, int ks, int) {
        const long long sb = row * (K >> 8) + (ks >> 2);
        const uint8_t * q = w[0] + sb * 128 + (ks & 3) * 32;
        gd_cpa16(sl.q16[0], q); gd_cpa16(sl.q16[1], q + 16); gd_cpa16(sl.q16[2], w[1] + sb * 16);
    }
    static __device__ __forceinline__ void fetch(GdRaw & r, const GdSlot & sl, const uint8_t * const *, long long, int, int, int) {
        r.a = gd_lds16(sl.q16[0]); r.b = gd_lds16(sl.q16[1]); r.c = gd_lds16(sl.q16[2]);
    }
    static __device__ __forceinline__ void convert(const GdRaw & r, int ks, int hk, uint4 (&o)[4]) {
        const int g = 2 * (ks & 3) + hk;
        const uint32_t wsc[3] = {r.c.y, r.c.z, r.c.w};
        auto byte = [&](int k) -> uint32_t { return (wsc[k >> 2] >> (8 * (k & 3))) & 0xFFu; };
        uint32_t sc, mn;
        if (g < 4) { sc = byte(g) & 63u; mn = byte(g + 4) & 63u; }
        else       { sc = (byte(g + 4) & 0xFu) | ((byte(g - 4) >> 6) << 4); mn = (byte(g + 4) >> 4) | ((byte(g) >> 6) << 4); }
        const __half2 dd = *(const __half2 *) &r.c.x;
        const float dsf = __fmul_rn(__low2float(dd), (float) sc) * r.fs, msf = __fmul_rn(__high2float(dd), (float) mn) * r.fs;
        const uint32_t ds = gd_pack_h2f(dsf, dsf), ms = gd_pack_h2f(-msf, -msf);
        const uint32_t m = 0x0F0F0F0Fu; const int sh = hk * 4;
        const uint2 a = gd_u8x4_hf((r.a.x >> sh) & m, ds, ms), b = gd_u8x4_hf((r.a.y >> sh) & m, ds, ms), c = gd_u8x4_hf((r.a.z >> sh) & m, ds, ms), d = gd_u8x4_hf((r.a.w >> sh) & m, ds, ms);
        const uint2 e = gd_u8x4_hf((r.b.x >> sh) & m, ds, ms), f = gd_u8x4_hf((r.b.y >> sh) & m, ds, ms), gg = gd_u8x4_hf((r.b.z >> sh) & m, ds, ms), h = gd_u8x4_hf((r.b.w >> sh) & m, ds, ms);
        o[0] = make_uint4(a.x, a.y, b.x, b.y); o[1] = make_uint4(c.x, c.y, d.x, d.y);
        o[2] = make_uint4(e.x, e.y, f.x, f.y); o[3] = make_uint4(gg.x, gg.y, h.x, h.y);
    }
};
// q6_K: slab ks = 64 elements 128 h + 64 c + e of super-block ks / 4 (h = (ks / 2) % 2, c = ks % 2); the thread's e = 32 hk .. + 31 are
// l = 0 .. 31 of ql[64 h + 32 hk + l] (low nibble for c = 0, high for c = 1) and bits 2 (2 c + hk) .. + 1 of qh[32 h + l];
// y = d * sc[8 h + 4 c + 2 hk + l / 16] * (q - 32)   (dequantize_row_q6_K, ggml-quants.c:1681-1710)
template <> struct GdCvt<T_Q6_K> {
    static __device__ __forceinline__ void issue(const GdSlot & sl, const uint8_t * const * w, long long row, int K, int ks, int hk) {
        const long long sb = row * (K >> 8) + (ks >> 2);
        const int h = (ks >> 1) & 1;
        const uint8_t * ql = w[0] + sb * 128 + 64 * h + 32 * hk;
        const uint8_t * qh = w[1] + sb * 64 + 32 * h;
        gd_cpa16(sl.q16[0], ql); gd_cpa16(sl.q16[1], ql + 16); gd_cpa16(sl.q16[2], qh); gd_cpa16(sl.q16[3], qh + 16);
        gd_cpa4(sl.q4[0], w[2] + sb * 16 + 8 * h + 4 * (ks & 1));  // the 4 scales of this slab (the thread uses 2 hk, 2 hk + 1)
        gd_cpa4(sl.q4[1], w[3] + ((sb * 2) & ~3ll));
    }
    static __device__ __forceinline__ void fetch(GdRaw & r, const GdSlot & sl, const uint8_t * const *, long long row, int K, int ks, int hk) {
        r.a = gd_lds16(sl.q16[0]); r.b = gd_lds16(sl.q16[1]); r.c = gd_lds16(sl.q16[2]); r.e = gd_lds16(sl.q16[3]);
        const uint32_t sc4 = gd_lds4(sl.q4[0]), dw = gd_lds4(sl.q4[1]);
        r.s0 = hk ? sc4 >> 16 : sc4 & 0xFFFFu;
        r.s1 = ((row * (K >> 8) + (ks >> 2)) & 1) ? dw >> 16 : dw & 0xFFFFu;
    }
    static __device__ __forceinline__ void convert(const GdRaw & r, int ks, int hk, uint4 (&o)[4]) {
        const int c = ks & 1, sh = 2 * (2 * c + hk), ns = 4 * c;
        const uint32_t m = 0x0F0F0F0Fu, m3 = 0x03030303u;
        auto q = [&](uint32_t l, uint32_t hh) { return ((l >> ns) & m) | (((hh >> sh) & m3) << 4); };
        const float d = __half2float(__ushort_as_half((unsigned short) r.s1)) * r.fs;
        const float d0f = __fmul_rn(d, (float) (int8_t) (r.s0 & 0xFF)), d1f = __fmul_rn(d, (float) (int8_t) ((r.s0 >> 8) & 0xFF));
        const uint32_t d0 = gd_pack_h2f(d0f, d0f), d1 = gd_pack_h2f(d1f, d1f), bias = 0x64206420u;      // 1056 = 1024 + 32: (q - 32) exact, then one rounding multiply
        const uint2 a = gd_u8x4_h(q(r.a.x, r.c.x), bias, d0), b = gd_u8x4_h(q(r.a.y, r.c.y), bias, d0), cc = gd_u8x4_h(q(r.a.z, r.c.z), bias, d0), dd = gd_u8x4_h(q(r.a.w, r.c.w), bias, d0);
        const uint2 e = gd_u8x4_h(q(r.b.x, r.e.x), bias, d1), f = gd_u8x4_h(q(r.b.y, r.e.y), bias, d1), g = gd_u8x4_h(q(r.b.z, r.e.z), bias, d1), h = gd_u8x4_h(q(r.b.w, r.e.w), bias, d1);
        o[0] = make_uint4(a.x, a.y, b.x, b.y); o[1] = make_uint4(cc.x, cc.y, dd.x, dd.y);
        o[2] = make_uint4(e.x, e.y, f.x, f.y); o[3] = make_uint4(g.x, g.y, h.x, h.y);
    }
};

template <int BN, int TYPE, bool WTMA>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GD_THREADS + 32, 1)
gemm_dense2_kernel(const __grid_constant__ GdMaps maps, const GemmDenseArgs p) {
    constexpr int STAGES = WTMA ? gd3_stages(BN, TYPE) : gd2_stages(BN, TYPE), STAGE_A = GM_BM * 128, STAGE_B = (BN / 2) * 128;
    constexpr int OFF_B = STAGES * STAGE_A, OFF_RAW = OFF_B + STAGES * STAGE_B, OFF_BAR = OFF_RAW + (WTMA ? gd3_raw_bytes(TYPE) : gd2_raw_bytes(TYPE));
    constexpr int RD = WTMA ? gd3_rd(TYPE) : 0;
    extern __shared__ uint8_t gd_smem_raw[];
    const uint32_t base = (gm_smem_u32(gd_smem_raw) + 1023u) & ~1023u;
    uint8_t * smem = gd_smem_raw + (base - gm_smem_u32(gd_smem_raw));
    const uint32_t sA = base, sB = base + OFF_B, sRaw = base + OFF_RAW, bars = base + OFF_BAR;
    const uint32_t full0 = bars, empty0 = bars + STAGES * 8, afull0 = bars + 2 * STAGES * 8, tfull0 = bars + 3 * STAGES * 8, tempty0 = tfull0 + 2 * 8;
    const uint32_t rfull0 = tempty0 + 2 * 8, rempty0 = rfull0 + RD * 8;                  // (WTMA only)
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (smem + OFF_BAR + 3 * STAGES * 8 + 4 * 8 + 2 * RD * 8);
    static_assert(3 * STAGES * 8 + 4 * 8 + 2 * RD * 8 + 4 <= (WTMA ? 512 : 256), "barrier area");
    const uint32_t sCS = bars + (WTMA ? 512 : 256);                                    // column scales of the tile being drained: [2 TMEM buffers][BN] floats
    const CUtensorMap & tmB = maps.b[0];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    if (threadIdx.x == 0) gd_stamp(p, 0);
    // programmatic dependent launch: the next kernel of the stream may be scheduled as soon as SMs free up (its prologue runs beside our
    // tail); we ourselves set up barriers and TMEM beside the previous kernel's tail and wait for its results only before the first load
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int nks = (p.K + GD_KS - 1) / GD_KS;
    const int nwork = p.full_units + (p.tail_T > 0 ? 2 * npairs : 0);

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) { gm_mbar_init(full0 + i * 8, 1); gm_mbar_init(empty0 + i * 8, 1); gm_mbar_init(afull0 + i * 8, 2 * GD_CVT_WARPS); }
        for (int i = 0; i < 2; i++) { gm_mbar_init(tfull0 + i * 8, 1); gm_mbar_init(tempty0 + i * 8, 2 * GD_EPI_WARPS); }
        if (WTMA) for (int i = 0; i < RD; i++) { gm_mbar_init(rfull0 + i * 8, 1); gm_mbar_init(rempty0 + i * 8, GD_CVT_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {            // one warp of each CTA of the pair, collectively
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(gm_smem_u32((const void *) tmem_slot)), "n"(2 * BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    gm_fence_before();
    gd2_cluster_sync();
    gm_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (warp < GD_CVT_WARP0) asm volatile("griddepcontrol.wait;" ::: "memory");      // loads of B, stores of dst / the split workspace (weights depend on nothing)
    if (threadIdx.x == 0) gd_stamp(p, 1);

    if (warp == 0) {
        // ===== TMA producer of the activations (both CTAs).  Work items are numbered n-tile fastest: the pairs that share a weight tile
        // run at the same time, so the weights come from HBM once (the activations stay in L2 anyway)
        if (lane == 0) asm volatile("prefetch.tensormap [%0];" :: "l"(&tmB) : "memory");
        __syncwarp();
        {   // whole warp, one elected lane issues
            const uint32_t lfull0 = gd2_mapa(full0, 0);
            uint32_t it = 0;
            for (int work = pair; work < nwork; work += npairs) {
                const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
                const int tile = u.tile;
                const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
                const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
                const int mt = local / p.ntiles, nt = local - mt * p.ntiles;
                constexpr uint32_t tx = 2u * ((uint32_t) (BN / 2) * 128u);
                const int brow = nt * BN + (int) rank * (BN / 2);
                for (int ks = u.ks0; ks < u.ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(empty0 + s * 8, ph ^ 1);
                    if (rank == 0) gm_mbar_expect_tx_elect(full0 + s * 8, tx);
                    gd2_tma_2d(sB + s * STAGE_B, &tmB, lfull0 + s * 8, ks * GD_KS, brow);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: warp 1 of the leader CTA (whole warp, one elected lane issues), M = 256 over the pair
        if (rank == 0) {
            uint32_t it = 0, tc = 0;
            constexpr uint32_t idesc = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) ((2 * GM_BM) >> 4) << 24);
            for (int work = pair; work < nwork; work += npairs) {
                const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
                const int ks0 = u.ks0, ks1 = u.ks1;
                if (ks0 >= ks1) continue;                            // (an empty tail segment: every role skips it)
                const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
                gm_mbar_wait(tempty0 + buf * 8, bph ^ 1);
                gm_fence_after();
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(afull0 + s * 8, ph);                // both CTAs' converted weight rows are in shared memory (async-proxy visible)
                    gm_mbar_wait(full0 + s * 8, ph);                 // both halves of the activation tile have landed
                    gm_fence_after();
                    if (it == 0 && lane == 0) gd_stamp(p, 2);
                    const uint64_t ad = gm_smem_desc(sA + s * STAGE_A), bd = gm_smem_desc(sB + s * STAGE_B);
                    #pragma unroll
                    for (int j = 0; j < 4; j++)
                        gd2_mma_f16(tmem + buf * BN, ad + 2u * j, bd + 2u * j, idesc, (ks > ks0 || j > 0) ? 1u : 0u);
                    gd2_commit_both(empty0 + s * 8);
                }
                gd2_commit_both(tfull0 + buf * 8);
                tc++;
            }
            if (lane == 0) gd_stamp(p, 3);
        }
    } else if (warp < GD_CVT_WARP0) {
        // ===== epilogue (both CTAs): this CTA's TMEM holds rows rank * 128 .. +127 of the pair tile, all BN columns
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int m_local = q * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t) (q * 32) << 16);
        const uint32_t ltempty0 = gd2_mapa(tempty0, 0);
        uint32_t tc = 0;
        for (int work = pair; work < nwork; work += npairs) {
            const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
            if (u.ks0 >= u.ks1) continue;
            const int tile = u.tile;
            const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
            const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
            const int mt = local / p.ntiles, nt = local - mt * p.ntiles;
            const int m = mt * 2 * GM_BM + (int) rank * GM_BM + m_local;
            const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
            // this warp: TMEM lanes q * 32 .. +31, columns half * (BN / 2) .. + BN / 2 of the tile
            const int c0 = half * (BN / 2), n0 = nt * BN + c0;
            const uint32_t ta = tbase + buf * BN + c0;
            const bool row_ok = m < p.Ms[j];
            // while the tile is still being accumulated: this warp's BN / 2 column scales into shared memory (the four warps of a half write the
            // same values to the same places; each reads back only after its own stores)
            const uint32_t csc = sCS + (uint32_t) ((buf * BN + c0) * 4);
            if (lane < BN / 8) {
                const float4 t = __ldg((const float4 *) (p.colscale + n0) + lane);
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(csc + lane * 16), "f"(t.x), "f"(t.y), "f"(t.z), "f"(t.w) : "memory");
            }
            __syncwarp();
            float rs = 1.0f;
            if constexpr (TYPE == T_Q4_K || TYPE == T_Q6_K) {      // undo the converters' row scale: same function of the same header
                if (row_ok) {
                    const uint8_t * const wh = TYPE == T_Q4_K ? (j == 0 ? p.w[0][1] : (j == 1 ? p.w[1][1] : p.w[2][1])) : (j == 0 ? p.w[0][3] : (j == 1 ? p.w[1][3] : p.w[2][3]));
                    const long long sb = (long long) m * (p.K >> 8) + (u.ks0 >> 2);
                    const uint32_t hw = TYPE == T_Q4_K ? __ldg((const uint32_t *) (wh + sb * 16)) : (uint32_t) __ldg((const unsigned short *) (wh + sb * 2));
                    rs = __frcp_rn(gd_row_up(gd_hdr_dmax(TYPE, hw)));
                }
            }
            gm_mbar_wait(tfull0 + buf * 8, bph);
            gm_fence_after();
            if (warp == 2 && lane == 0) gd_stamp(p, tc == 0 ? 4 : 5);
            if (u.slot < 0) {
                float * const outp = p.outs[j] + (long long) u.sp * p.split_stride;
                const long long ocs = p.out_css[j];
                gd_epilogue_warp<BN / 64>(ta, outp + (long long) n0 * ocs + m, ocs, p.N - n0, row_ok, csc, rs);
            } else {                                                // tail segment: the whole 256 x BN partial tile, column-major, into its slot
                float * const slotp = p.tail_ws + (size_t) u.slot * (2 * GM_BM * BN);
                gd_epilogue_warp<BN / 64>(ta, slotp + (size_t) c0 * (2 * GM_BM) + (int) rank * GM_BM + m_local, 2 * GM_BM, BN, true, csc, rs);
            }
            gm_fence_before();
            __syncwarp();
            if (lane == 0) gd2_arrive_cluster(ltempty0 + buf * 8);
            tc++;
        }
        if (warp == 2 && lane == 0) gd_stamp(p, 6);
    } else if (warp < GD_W_WARP) {
        // ===== converters (both CTAs): thread = row r of this CTA's 128, half slab hk; native blocks -> swizzled fp16 stage.
        // 128-byte swizzle of a K-major tile: row r lies at (r / 8) * 1024 + (r % 8) * 128, its 16-byte chunk c at ((c ^ (r % 8)) * 16):
        // the 8 lanes of a store phase (8 consecutive rows, same c) hit 8 different chunk positions = all 32 banks once.
        const int cw = warp - GD_CVT_WARP0, r = (cw & 3) * 32 + lane, hk = cw >> 2, ct = cw * 32 + lane;
        const uint32_t rowoff = (uint32_t) ((r >> 3) * 1024 + (r & 7) * 128);
        uint32_t choff[4];
        #pragma unroll
        for (int c = 0; c < 4; c++) choff[c] = rowoff + (uint32_t) (((hk * 4 + c) ^ (r & 7)) << 4);
        const uint32_t lafull0 = gd2_mapa(afull0, 0);
        uint32_t it = 0;
        if constexpr (WTMA) {
            constexpr int RS = gd3_rs(TYPE), RAWB = gd3_raw_stage(TYPE), SOFF = gd3_small_off(TYPE);
            const uint32_t r7 = (uint32_t) (r & 7), r64 = (uint32_t) ((r >> 1) & 3);
            const uint32_t dcell0 = sRaw + RD * RAWB + (uint32_t) ct * 4;                 // q6_K: this thread's d cells, GD_CVT_WARPS * 32 * 4 bytes apart
            uint32_t rs = 0, rph = 0, st = 0, sph = 1;          // raw-stage ring position / parity; fp16 stage ring position / parity of its `empty` wait
            const bool lane0 = lane == 0;
            for (int work = pair; work < nwork; work += npairs) {
                const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
                const int ks0 = u.ks0, ks1 = u.ks1;
                if (ks0 >= ks1) continue;
                const int tile = u.tile;
                const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
                const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
                const int mt = local / p.ntiles;
                const long long m = (long long) mt * 2 * GM_BM + (int) rank * GM_BM + r;
                const bool row_ok = m < p.Ms[j];
                const int q0 = ks0 / RS, q1 = (ks1 + RS - 1) / RS;                         // raw stages are aligned in K (K-quants: super-blocks)
                const uint8_t * const wd = j == 0 ? p.w[0][3] : (j == 1 ? p.w[1][3] : p.w[2][3]);
                const long long sb0 = m * (p.K >> 8);
                float fs = 1.0f;
                if constexpr (TYPE == T_Q6_K) {
                    #pragma unroll 1
                    for (int d = 0; d < GD3_DD; d++) {
                        if (q0 + d < q1 && row_ok) gd_cpa4(dcell0 + (uint32_t) (d * GD_CVT_WARPS * 32 * 4), wd + (((sb0 + q0 + d) * 2) & ~3ll));
                        asm volatile("cp.async.commit_group;" ::: "memory");
                    }
                }
                #pragma unroll 1
                for (int q = q0; q < q1; q++) {
                    gm_mbar_wait(rfull0 + rs * 8, rph);                                  // this raw stage's boxes have landed
                    const uint32_t rb = sRaw + rs * RAWB, rowq = rb + (uint32_t) r * 128;
                    const uint4 sc = gd_lds16(rb + SOFF + (uint32_t) r * 16);           // the row's 8 block scales / K-quant header / 16 q6_K scales
                    const bool odd = TYPE == T_Q8_0 && (q & 1);                          // q8_0: the raw stage's 4 scales are the first or the second half of the 16 bytes
                    const uint32_t scw[4] = {odd ? sc.z : sc.x, odd ? sc.w : sc.y, sc.z, sc.w};
                    uint32_t dsel = 0;
                    if constexpr (TYPE == T_Q6_K) {
                        asm volatile("cp.async.wait_group %0;" :: "n"(GD3_DD - 1) : "memory");
                        const uint32_t dw = gd_lds4(dcell0 + (uint32_t) (((q - q0) % GD3_DD) * GD_CVT_WARPS * 32 * 4));
                        dsel = row_ok ? (((sb0 + q) & 1) ? dw >> 16 : dw & 0xFFFFu) : 0u;
                    }
                    if constexpr (TYPE == T_Q4_K || TYPE == T_Q6_K)      // the unit's row scale, from its first super-block (the epilogue derives the same)
                        if (q == q0) fs = gd_row_up(gd_hdr_dmax(TYPE, TYPE == T_Q4_K ? sc.x : dsel));
                    #pragma unroll
                    for (int jp = 0; jp < RS; jp += 2) {
                        uint4 o[2][4];
                        #pragma unroll
                        for (int u2 = 0; u2 < 2; u2++) {
                            const int jj = jp + u2, ks = q * RS + jj;
                            GdRaw raw;
                            raw.fs = fs;
                            if constexpr (TYPE == T_Q4_0) {
                                raw.a = gd_lds16(rowq + (((uint32_t) (2 * jj + hk) ^ r7) << 4));
                                raw.s0 = hk ? scw[jj] >> 16 : scw[jj] & 0xFFFFu;
                            } else if constexpr (TYPE == T_Q8_0) {
                                const uint32_t c = (uint32_t) (2 * (2 * jj + hk));
                                raw.a = gd_lds16(rowq + ((c ^ r7) << 4)); raw.b = gd_lds16(rowq + (((c + 1) ^ r7) << 4));
                                raw.s0 = hk ? scw[jj] >> 16 : scw[jj] & 0xFFFFu;
                            } else if constexpr (TYPE == T_Q4_K) {
                                raw.a = gd_lds16(rowq + (((uint32_t) (2 * jj) ^ r7) << 4)); raw.b = gd_lds16(rowq + (((uint32_t) (2 * jj + 1) ^ r7) << 4));
                                raw.c = sc;
                            } else {
                                const uint32_t h = (uint32_t) (jj >> 1), c = 4 * h + 2 * (uint32_t) hk, rowh = rb + 16384 + (uint32_t) r * 64;
                                raw.a = gd_lds16(rowq + ((c ^ r7) << 4)); raw.b = gd_lds16(rowq + (((c + 1) ^ r7) << 4));
                                raw.c = gd_lds16(rowh + (((2 * h) ^ r64) << 4)); raw.e = gd_lds16(rowh + (((2 * h + 1) ^ r64) << 4));      // 64-byte swizzle
                                raw.s0 = hk ? scw[jj] >> 16 : scw[jj] & 0xFFFFu;
                                raw.s1 = dsel;
                            }
                            GdCvt<TYPE>::convert(raw, ks, hk, o[u2]);
                        }
                        uint32_t sidx[2];
                        bool live[2];
                        #pragma unroll
                        for (int u2 = 0; u2 < 2; u2++) {
                            const int ks = q * RS + jp + u2;
                            live[u2] = ks >= ks0 && ks < ks1;                            // (a K split may begin or end inside a raw stage)
                            sidx[u2] = st;
                            if (live[u2]) {
                                gm_mbar_wait(empty0 + st * 8, sph);                      // the MMAs that read this slot are done
                                const uint32_t dst = sA + st * STAGE_A;
                                #pragma unroll
                                for (int c = 0; c < 4; c++)
                                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dst + choff[c]), "r"(o[u2][c].x), "r"(o[u2][c].y), "r"(o[u2][c].z), "r"(o[u2][c].w) : "memory");
                                if (++st == STAGES) { st = 0; sph ^= 1; }
                            }
                        }
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy stores -> visible to the tensor core
                        __syncwarp();
                        #pragma unroll
                        for (int u2 = 0; u2 < 2; u2++) gd2_arrive_remote_if(lafull0 + sidx[u2] * 8, lane0 && live[u2]);
                    }
                    // (every lane's reads of the raw stage went into the conversions above, which the __syncwarp()s have seen through)
                    gd2_arrive_local_if(rempty0 + rs * 8, lane0);
                    if (++rs == RD) { rs = 0; rph ^= 1; }
                    if constexpr (TYPE == T_Q6_K) {
                        if (q + GD3_DD < q1 && row_ok) gd_cpa4(dcell0 + (uint32_t) (((q - q0) % GD3_DD) * GD_CVT_WARPS * 32 * 4), wd + (((sb0 + q + GD3_DD) * 2) & ~3ll));
                        asm volatile("cp.async.commit_group;" ::: "memory");
                    }
                }
                if constexpr (TYPE == T_Q6_K) asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
        } else
        for (int work = pair; work < nwork; work += npairs) {
            const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
            const int ks0 = u.ks0, ks1 = u.ks1;
            if (ks0 >= ks1) continue;
            const int tile = u.tile;
            const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
            const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
            const int mt = local / p.ntiles;
            const long long m = (long long) mt * 2 * GM_BM + (int) rank * GM_BM + r;
            const bool row_ok = m < p.Ms[j];
            const uint8_t * w[4];
            #pragma unroll
            for (int f = 0; f < 4; f++) w[f] = j == 0 ? p.w[0][f] : (j == 1 ? p.w[1][f] : p.w[2][f]);
            // k range of a slab half that lies inside the row (q4_0 / q8_0 rows may end after the first half of the last slab)
            auto live = [&](int ks) { return row_ok && ks * GD_KS + hk * 32 < p.K; };
            float fs = 1.0f;                      // K-quants: the unit's row scale (see the WTMA branch)
            // D slabs in flight per thread: cp.async copies into the thread's ring cells, one commit group per slab (empty groups keep
            // the count uniform), wait_group D - 1 before a slab is read
            constexpr int D = gd2_depth(TYPE);
            auto slot_of = [&](int i) {
                GdSlot sl;
                const uint32_t cell = (uint32_t) ((i % D) * (GD_CVT_WARPS * 32) + ct);
                #pragma unroll
                for (int f = 0; f < 4; f++) sl.q16[f] = sRaw + (uint32_t) (f * D * GD_CVT_WARPS * 32 * 16) + cell * 16;
                #pragma unroll
                for (int f = 0; f < 2; f++) sl.q4[f] = sRaw + (uint32_t) (gd2_n16(TYPE) * D * GD_CVT_WARPS * 32 * 16 + f * D * GD_CVT_WARPS * 32 * 4) + cell * 4;
                return sl;
            };
            #pragma unroll 1
            for (int d = 0; d < D; d++) {
                if (ks0 + d < ks1 && live(ks0 + d)) GdCvt<TYPE>::issue(slot_of(d), w, m, p.K, ks0 + d, hk);
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
            // U slabs per iteration: their conversions interleave (two independent dependency chains per thread), and the fence --
            // a MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC of ~250 clocks (ncu source page) -- is paid once per U stages
            constexpr int U = 2;
            #pragma unroll 1
            for (int kb = ks0; kb < ks1; kb += U) {
                asm volatile("cp.async.wait_group %0;" :: "n"(D - U) : "memory");
                uint4 o[U][4];
                GdSlot sl[U];
                #pragma unroll
                for (int u2 = 0; u2 < U; u2++) {
                    const int ks = kb + u2;
                    sl[u2] = slot_of(ks - ks0);
                    if (ks < ks1 && live(ks)) {
                        GdRaw raw; GdCvt<TYPE>::fetch(raw, sl[u2], w, m, p.K, ks, hk);
                        if constexpr (TYPE == T_Q4_K || TYPE == T_Q6_K) if (ks == ks0) fs = gd_row_up(gd_hdr_dmax(TYPE, TYPE == T_Q4_K ? raw.c.x : raw.s1));
                        raw.fs = fs;
                        GdCvt<TYPE>::convert(raw, ks, hk, o[u2]);
                    }
                    else { o[u2][0] = o[u2][1] = o[u2][2] = o[u2][3] = make_uint4(0, 0, 0, 0); }
                }
                uint32_t sidx[U];
                #pragma unroll
                for (int u2 = 0; u2 < U; u2++) {
                    if (kb + u2 < ks1) {
                        const uint32_t s = (it + u2) % STAGES, ph = ((it + u2) / STAGES) & 1;
                        sidx[u2] = s;
                        gm_mbar_wait(empty0 + s * 8, ph ^ 1);                                       // the MMAs that read this slot are done
                        const uint32_t dst = sA + s * STAGE_A;
                        #pragma unroll
                        for (int c = 0; c < 4; c++)
                            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dst + choff[c]), "r"(o[u2][c].x), "r"(o[u2][c].y), "r"(o[u2][c].z), "r"(o[u2][c].w) : "memory");
                    }
                }
                // (the ring cells were read into registers and used above: the copies of the slabs D ahead may overwrite them now)
                #pragma unroll
                for (int u2 = 0; u2 < U; u2++) {
                    const int ks = kb + u2;
                    if (ks + D < ks1 && live(ks + D)) GdCvt<TYPE>::issue(sl[u2], w, m, p.K, ks + D, hk);
                    asm volatile("cp.async.commit_group;" ::: "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");                        // generic-proxy stores -> visible to the tensor core
                __syncwarp();
                #pragma unroll
                for (int u2 = 0; u2 < U; u2++)
                    if (kb + u2 < ks1) { if (lane == 0) gd2_arrive_remote(lafull0 + sidx[u2] * 8); it++; }
            }
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
    } else if (WTMA) {
        // ===== weight producer (both CTAs, WTMA): this CTA's 128 rows of every SoA field, one raw stage (gd3_rs slabs) per barrier
        constexpr int RS = gd3_rs(TYPE), RAWB = gd3_raw_stage(TYPE), NF = gd3_nf(TYPE);
        constexpr uint32_t TX = 128u * (uint32_t) (gd3_box(TYPE, 0) + gd3_box(TYPE, 1) + (NF > 2 ? gd3_box(TYPE, 2) : 0));
        uint32_t rit = 0;
        for (int work = pair; work < nwork; work += npairs) {
            const GdUnit u = gd2_unit(p, work, nks, pair, npairs);
            if (u.ks0 >= u.ks1) continue;
            const int tile = u.tile;
            const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
            const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
            const int mt = local / p.ntiles;
            const int row0 = mt * 2 * GM_BM + (int) rank * GM_BM;
            const int q0 = u.ks0 / RS, q1 = (u.ks1 + RS - 1) / RS;
            for (int q = q0; q < q1; q++, rit++) {
                const uint32_t rs = rit % RD, ph = (rit / RD) & 1;
                gm_mbar_wait(rempty0 + rs * 8, ph ^ 1);
                gm_mbar_expect_tx_elect(rfull0 + rs * 8, TX);
                #pragma unroll
                for (int f = 0; f < NF; f++)
                    gm_tma_2d_elect(sRaw + rs * RAWB + gd3_foff(f), &maps.w[j][f], rfull0 + rs * 8, (q * gd3_adv(TYPE, f)) & ~15, row0);      // (box starts are 16-byte aligned: q8_0's 8 bytes of scales per raw stage come with their neighbours)
            }
        }
    }
    __syncwarp();                // the cluster barrier is warp-aligned: the single-lane roles rejoin their warps first
    gm_fence_before();
    gd2_cluster_sync();          // the peer's shared memory and barriers stay alive until the leader's MMAs and commits are done
    if (threadIdx.x == 0) gd_stamp(p, 7);
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(2 * BN) : "memory");
}

// ------------------------------------------------------------------------------------------------ operand preparation
// activations: f32 [N][K] -> q8_1 (d = amax / 127 rounded to fp16, q = round(x / d); quantize_row_q8_1_ref ggml-quants.c:220-253)
// -> fp16(d * q * 2^g(n)), g(n) = the power of two that brings the column's largest |x| into [2^8, 2^9); colscale[n] = 2^-g(n).
// One CTA per column: pass 1 finds the column's amax, pass 2 (the row comes from L1 / L2 now) converts, 8 lanes per 32-element block.
__device__ __forceinline__ float gd_col_scale(const float amax, float * inv) {
    if (!(amax > 0.0f) || !(amax <= 3.0e38f)) { *inv = 1.0f; return 1.0f; }
    const int E = (int) ((__float_as_uint(amax) >> 23) & 0xFFu) - 126;        // amax = m 2^E, m in [0.5, 1)
    int g = 9 - E;
    g = g > 100 ? 100 : (g < -100 ? -100 : g);
    *inv = __uint_as_float((uint32_t) (127 - g) << 23);
    return __uint_as_float((uint32_t) (127 + g) << 23);
}
template <int NW>
__device__ __forceinline__ float gd_block_max(float v, float * sh) {      // max over the NW warps of the CTA, the same value in every thread
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sh[0];
    #pragma unroll
    for (int i = 1; i < NW; i++) r = fmaxf(r, sh[i]);
    return r;
}
// one float4 of a 32-element block (8 consecutive lanes hold the block) -> 4 halfs of fp16(d q sc)
__device__ __forceinline__ uint2 gd_q8_1_f16x4(const float4 v, const float sc) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const float dh = __half2float(__float2half_rn(d)) * sc;
    const float r0 = dh * roundf(__fmul_rn(v.x, id)), r1 = dh * roundf(__fmul_rn(v.y, id));
    const float r2 = dh * roundf(__fmul_rn(v.z, id)), r3 = dh * roundf(__fmul_rn(v.w, id));
    uint2 h;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.x) : "f"(r1), "f"(r0));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.y) : "f"(r3), "f"(r2));
    return h;
}
// One CTA of 256 threads per column.  The column (K floats, up to GD_ACT_SMEM_MAX bytes) arrives in shared memory by ONE bulk copy
// (cp.async.bulk, mbarrier completion): HBM is read once, at full burst, and both passes -- the column's amax, then the conversion --
// run out of shared memory.  Longer columns take two passes over global memory (STAGED = false).
constexpr int GD_ACT_SMEM_MAX = 112 * 1024;
template <bool STAGED>
__global__ void __launch_bounds__(256)
dequant_act_f16_kernel(const char * __restrict__ x, const long long row_stride, const int K, __half * __restrict__ out, float * __restrict__ colscale) {
    __shared__ float sh[8];
    __shared__ __align__(8) unsigned long long bar;
    extern __shared__ __align__(16) uint8_t gd_act_smem[];
    const int n = blockIdx.x, K4 = K >> 2;
    const float4 * grow = (const float4 *) (x + (long long) n * row_stride);
    const float4 * row = STAGED ? (const float4 *) gd_act_smem : grow;
    if (STAGED && threadIdx.x == 0) { gm_mbar_init(gm_smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");      // x comes from the kernels before us; out may still be read by the GEMM before us
    if (STAGED) {
        __syncthreads();
        if (threadIdx.x == 0) {
            gm_mbar_expect_tx(gm_smem_u32(&bar), (uint32_t) K * 4u);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(gm_smem_u32(gd_act_smem)), "l"(grow), "r"((uint32_t) K * 4u), "r"(gm_smem_u32(&bar)) : "memory");
        }
        gm_mbar_wait(gm_smem_u32(&bar), 0);
    }
    uint2 * o = (uint2 *) (out + (long long) n * K);
    float amax = 0.0f;
    for (int base = 0; base < K4; base += 1024) {
        float4 v[4];
        #pragma unroll
        for (int u = 0; u < 4; u++) { const int i = base + u * 256 + threadIdx.x; v[u] = i < K4 ? row[i] : make_float4(0.f, 0.f, 0.f, 0.f); }
        #pragma unroll
        for (int u = 0; u < 4; u++) amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[u].x), fabsf(v[u].y)), fmaxf(fabsf(v[u].z), fabsf(v[u].w))));
    }
    float inv;
    const float sc = gd_col_scale(gd_block_max<8>(amax, sh), &inv);
    if (threadIdx.x == 0) colscale[n] = inv;
    for (int base = 0; base < K4; base += 1024) {            // (uniform trip counts: the shuffles inside want whole warps)
        float4 v[4];
        #pragma unroll
        for (int u = 0; u < 4; u++) { const int i = base + u * 256 + threadIdx.x; v[u] = i < K4 ? row[i] : make_float4(0.f, 0.f, 0.f, 0.f); }
        #pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = base + u * 256 + threadIdx.x;
            const uint2 h = gd_q8_1_f16x4(v[u], sc);
            if (i < K4) o[i] = h;
        }
    }
}

// out: fp16 [N][K]; colscale: [N rounded up to 256] floats (the entries past N are never written and only multiply results nobody keeps)
cudaError_t launch_dequant_act_f16(const float * x, long long row_stride_bytes, int K, int N, __half * out, float * colscale, cudaStream_t st) {
    const bool staged = (long long) K * 4 <= GD_ACT_SMEM_MAX && ((uintptr_t) x & 15) == 0 && row_stride_bytes % 16 == 0;
    if (staged) {
        static bool attr_done[64] = {false};
        int dev = 0; cudaGetDevice(&dev); dev &= 63;
        if (!attr_done[dev]) {
            cudaError_t e = cudaFuncSetAttribute(dequant_act_f16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GD_ACT_SMEM_MAX);
            if (e != cudaSuccess) return e;
            attr_done[dev] = true;
        }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) N); cfg.blockDim = dim3(256); cfg.stream = st; cfg.dynamicSmemBytes = staged ? (size_t) K * 4 : 0;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return staged ? cudaLaunchKernelEx(&cfg, dequant_act_f16_kernel<true>, (const char *) x, (long long) row_stride_bytes, K, out, colscale)
                  : cudaLaunchKernelEx(&cfg, dequant_act_f16_kernel<false>, (const char *) x, (long long) row_stride_bytes, K, out, colscale);
}

// split-K reduce with the residual ADD fused, one CTA per dst column, whose result is the next dense GEMM's src1: out = sum of the splits,
// res_dst = res_src + out, act_out = fp16(q8_1(res_dst) 2^g) and colscale_out as dequant_act_f16_kernel would produce them (M <= 8192, M % 32 == 0)
__global__ void __launch_bounds__(256)
reduce_splits4_act_kernel(const float4 * __restrict__ ws, const int S, const long long MN4, const int M4, float4 * __restrict__ out, const long long out_cs4,
                          const float4 * __restrict__ res_src, float4 * __restrict__ res_dst, uint2 * __restrict__ act_out, float * __restrict__ colscale_out) {
    __shared__ float sh[8];
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long n = blockIdx.x;
    float4 z[8];
    float amax = 0.0f;
    #pragma unroll
    for (int it = 0; it < 8; it++) {
        const int m = it * 256 + threadIdx.x;
        z[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < M4) {
            const long long i = n * M4 + m;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int sp = 0; sp < S; sp++) {               // split order: deterministic
                const float4 w = ws[(long long) sp * MN4 + i];
                v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
            }
            out[n * out_cs4 + m] = v;
            const float4 r = res_src[i];
            z[it] = make_float4(r.x + v.x, r.y + v.y, r.z + v.z, r.w + v.w);      // the GGML_OP_ADD node that consumes this MUL_MAT
            res_dst[i] = z[it];
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(z[it].x), fabsf(z[it].y)), fmaxf(fabsf(z[it].z), fabsf(z[it].w))));
        }
    }
    float inv;
    const float sc = gd_col_scale(gd_block_max<8>(amax, sh), &inv);
    if (threadIdx.x == 0) colscale_out[n] = inv;
    #pragma unroll
    for (int it = 0; it < 8; it++) {
        if (it * 256 < M4) {                                // (uniform)
            const int m = it * 256 + threadIdx.x;
            const uint2 h = gd_q8_1_f16x4(z[it], sc);
            if (m < M4) act_out[n * M4 + m] = h;
        }
    }
}
cudaError_t launch_reduce_splits_act(const float * ws, int S, int M, int N, float * out, long long out_cs, cudaStream_t st,
                                     const float * res_src, float * res_dst, void * act_out, float * colscale_out) {
    if (M % 32 != 0 || M > 8192 || out_cs % 4 != 0 || !res_dst || !act_out || !colscale_out ||
        (((uintptr_t) ws | (uintptr_t) out | (uintptr_t) res_src | (uintptr_t) res_dst) & 15) != 0 || ((uintptr_t) act_out & 7) != 0) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) N); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, reduce_splits4_act_kernel, (const float4 *) ws, S, (long long) M * N / 4, M / 4, (float4 *) out, out_cs / 4,
                              (const float4 *) res_src, (float4 *) res_dst, (uint2 *) act_out, colscale_out);
}
// ---- producers of a dense GEMM's src1 that also write its converted form (ops_api.cu, the Llama-layer fusions at many columns):
// one CTA per column n; the column's K floats are produced, written, kept in registers, then converted exactly as dequant_act_f16_kernel
// would convert them after reading them back (column amax -> power-of-two scale -> fp16(d q 2^g) per 32-element block).
//   MODE 0  y = fl(fl(x * rsqrt(mean(x^2) + eps)) * w)      RMS_NORM followed by the in-place MUL with the norm weight
//   MODE 1  y = fl(silu(g) * u)                              SwiGLU gate
//   MODE 2  y = gather of 16-byte pieces                     CONT (head merge): source strides in `nes / nbs`
// NTHR threads per column: 256 for MODE 0 (the sum of squares keeps the order of rms_norm_mul_kernel), 512 for the long SwiGLU columns, whose
// only reduction is a max
template <int MODE, int NV, int NTHR>
__global__ void __launch_bounds__(NTHR)
row_act_kernel(const RowActArgs p) {
    __shared__ float sh[NTHR / 32];
    __shared__ double shd[8];
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long n = blockIdx.x;
    const int K4 = p.K >> 2;
    float4 * yr = (float4 *) (p.y + n * p.ys);
    float scale = 0.f;
    if (MODE == 0) {
        const float4 * xr = (const float4 *) (p.a + n * p.as);
        double sum = 0.0;
        for (int i = threadIdx.x; i < K4; i += NTHR) {
            const float4 v = xr[i];
            sum += (double) __fmul_rn(v.x, v.x) + (double) __fmul_rn(v.y, v.y) + (double) __fmul_rn(v.z, v.z) + (double) __fmul_rn(v.w, v.w);
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if ((threadIdx.x & 31) == 0) shd[threadIdx.x >> 5] = sum;
        __syncthreads();
        sum = shd[0];
        for (int i = 1; i < 8; i++) sum += shd[i];
        const float mean = (float) (sum / (double) p.K);
        scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, p.eps)));
    }
    float4 z[NV];
    float amax = 0.0f;
    #pragma unroll
    for (int it = 0; it < NV; it++) {
        const int i = it * NTHR + threadIdx.x;
        z[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < K4) {
            float4 r;
            if (MODE == 0) {
                const float4 v = ((const float4 *) (p.a + n * p.as))[i], g = ((const float4 *) p.b)[i];
                r = make_float4(__fmul_rn(__fmul_rn(v.x, scale), g.x), __fmul_rn(__fmul_rn(v.y, scale), g.y), __fmul_rn(__fmul_rn(v.z, scale), g.z), __fmul_rn(__fmul_rn(v.w, scale), g.w));
            } else if (MODE == 1) {
                const float4 a = ((const float4 *) (p.a + n * p.as))[i], b = ((const float4 *) (p.b + n * p.bs))[i];
                r.x = __fmul_rn(__fdiv_rn(a.x, __fadd_rn(1.0f, expf(-a.x))), b.x); r.y = __fmul_rn(__fdiv_rn(a.y, __fadd_rn(1.0f, expf(-a.y))), b.y);
                r.z = __fmul_rn(__fdiv_rn(a.z, __fadd_rn(1.0f, expf(-a.z))), b.z); r.w = __fmul_rn(__fdiv_rn(a.w, __fadd_rn(1.0f, expf(-a.w))), b.w);
            } else {
                long long t = (n * p.K + (long long) i * 4);                      // logical element index of the source
                const long long s0 = t % p.nes[0]; t /= p.nes[0];
                const long long s1 = t % p.nes[1]; t /= p.nes[1];
                const long long s2 = t % p.nes[2]; const long long s3 = t / p.nes[2];
                r = *(const float4 *) ((const char *) p.a + s0 * 4 + s1 * p.nbs[1] + s2 * p.nbs[2] + s3 * p.nbs[3]);
            }
            z[it] = r; yr[i] = r;
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(r.x), fabsf(r.y)), fmaxf(fabsf(r.z), fabsf(r.w))));
        }
    }
    float inv;
    const float sc = gd_col_scale(gd_block_max<NTHR / 32>(amax, sh), &inv);
    if (threadIdx.x == 0) p.colscale[n] = inv;
    uint2 * o = (uint2 *) p.act + n * K4;
    #pragma unroll
    for (int it = 0; it < NV; it++) {
        if (it * NTHR < K4) {                                // (uniform: the shuffles inside want whole warps)
            const int i = it * NTHR + threadIdx.x;
            const uint2 h = gd_q8_1_f16x4(z[it], sc);
            if (i < K4) o[i] = h;
        }
    }
}
template <int MODE>
static cudaError_t launch_row_act_m(const RowActArgs & p, int N, cudaStream_t st) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) N); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    const int nv = (p.K / 4 + 255) / 256;
    if (nv <= 4)  return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 4, 256>, p);
    if (MODE != 0) {                       // long columns: twice the threads, half the serial work per thread
        cfg.blockDim = dim3(512);
        if (nv <= 8)  return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 4, 512>, p);
        if (nv <= 12) return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 6, 512>, p);
        return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 7, 512>, p);
    }
    if (nv <= 8)  return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 8, 256>, p);
    if (nv <= 11) return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 11, 256>, p);
    return cudaLaunchKernelEx(&cfg, row_act_kernel<MODE, 14, 256>, p);
}
// K % 32 == 0, K <= ROW_ACT_MAX_K, every row 16-byte aligned; act: fp16 [N][K], colscale: [N] (the layout launch_dequant_act_f16 writes)
cudaError_t launch_row_act(int mode, const RowActArgs & p, int N, cudaStream_t st) {
    if (N <= 0 || p.K <= 0) return cudaSuccess;
    if (p.K % 32 != 0 || p.K > ROW_ACT_MAX_K || ((uintptr_t) p.act & 7) != 0) return cudaErrorInvalidValue;
    if (mode == 0) return launch_row_act_m<0>(p, N, st);
    if (mode == 1) return launch_row_act_m<1>(p, N, st);
    return launch_row_act_m<2>(p, N, st);
}
// ---- stream-K tail: dst tile = sum of the tail_ws slots of the chunks that cover it, in chunk order
struct GdFixArgs {
    const float * ws; int tail_T, tail_L, nks, full_tiles, ntiles, bn, N;
    float * outs[3]; long long out_css[3]; int Ms[3]; int tile_end[3];
};
__global__ void __launch_bounds__(256)
gd_tail_fixup_kernel(const GdFixArgs a) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int per_tile = a.bn * (2 * GM_BM / 4) / 256;                 // blocks per tile: one float4 (4 rows of a column) per thread
    const int tt = blockIdx.x / per_tile, e = (blockIdx.x - tt * per_tile) * 256 + threadIdx.x;
    const int n_local = e / (2 * GM_BM / 4), m4 = e - n_local * (2 * GM_BM / 4);
    const int tile = a.full_tiles + tt;
    const int j = (tile >= a.tile_end[0]) + (tile >= a.tile_end[1]);
    const int local = tile - (j == 0 ? 0 : a.tile_end[j - 1]);
    const int mt = local / a.ntiles, nt = local - mt * a.ntiles;
    const int n = nt * a.bn + n_local, m = mt * 2 * GM_BM + m4 * 4;
    if (n >= a.N || m >= a.Ms[j]) return;
    const int u0 = tt * a.nks, u1 = min(u0 + a.nks, a.tail_T);
    const int c0 = u0 / a.tail_L, c1 = (u1 - 1) / a.tail_L;
    const size_t tile_elems = (size_t) 2 * GM_BM * a.bn;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = c0; c <= c1; c++) {
        const int seg = (c * a.tail_L) / a.nks == tt ? 0 : 1;          // the chunk began in this tile, or in the one before
        const float4 v = *(const float4 *) (a.ws + (size_t) (2 * c + seg) * tile_elems + (size_t) n_local * (2 * GM_BM) + m4 * 4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float * o = a.outs[j] + (long long) n * a.out_css[j] + m;
    if (m + 3 < a.Ms[j] && (((uintptr_t) o) & 15) == 0) *(float4 *) o = acc;
    else { o[0] = acc.x; if (m + 1 < a.Ms[j]) o[1] = acc.y; if (m + 2 < a.Ms[j]) o[2] = acc.z; if (m + 3 < a.Ms[j]) o[3] = acc.w; }
}
size_t gemm_dense_tail_ws_bytes(int sms) { return (size_t) 2 * (sms / 2) * 2 * GM_BM * 256 * sizeof(float); }      // 2 slots per CTA pair, 256 x 256 floats

// ------------------------------------------------------------------------------------------------ host: launch
// N tile: 256 columns halve the number of times the weights travel from L2 to the SMs (few tiles are spread with split-K instead)
int gemm_dense_bn(int M, int N, int sms) {
    (void) M; (void) sms;
    if (getenv("B200_GEMM_BN")) return atoi(getenv("B200_GEMM_BN")) == 256 ? 256 : 128;
    return N > 128 ? 256 : 128;
}
int gemm_dense_ksplit(int M, int N, int K, int sms) {
    const int bn = gemm_dense_bn(M, N, sms);
    const int bm = 2 * GM_BM, units = sms / 2;
    const int tiles = ((M + bm - 1) / bm) * ((N + bn - 1) / bn);
    const int nks = (K + GD_KS - 1) / GD_KS;
    int s = units / (tiles > 0 ? tiles : 1);
    if (s > nks / 8) s = nks / 8;
    if (s > 8) s = 8;
    return s < 2 ? 1 : s;
}

template <int BN, int TYPE, bool WTMA>
static cudaError_t gemm_dense2_launch_t(const GdMaps & maps, const GemmDenseArgs & p, int grid, cudaStream_t st) {
    static bool attr_done[64] = {false};          // per device
    constexpr int smem = WTMA ? gd3_smem(BN, TYPE) : gd2_smem(BN, TYPE);
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    if (!attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_dense2_kernel<BN, TYPE, WTMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3(GD_THREADS + (WTMA ? 32 : 0)); cfg.dynamicSmemBytes = smem; cfg.stream = st;      // __cluster_dims__(2): grid is even
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, gemm_dense2_kernel<BN, TYPE, WTMA>, maps, p);
}
template <int BN, bool WTMA>
static cudaError_t gemm_dense2_launch(int type, const GdMaps & maps, const GemmDenseArgs & p, int grid, cudaStream_t st) {
    switch (type) {
        case T_Q4_0: return gemm_dense2_launch_t<BN, T_Q4_0, WTMA>(maps, p, grid, st);
        case T_Q8_0: return gemm_dense2_launch_t<BN, T_Q8_0, WTMA>(maps, p, grid, st);
        case T_Q4_K: return gemm_dense2_launch_t<BN, T_Q4_K, WTMA>(maps, p, grid, st);
        case T_Q6_K: return gemm_dense2_launch_t<BN, T_Q6_K, WTMA>(maps, p, grid, st);
    }
    return cudaErrorInvalidValue;
}
// one SoA field of a weight matrix as a byte matrix [rows][rowbytes]: boxes of 128 rows x `box` bytes
static bool make_map_field(CUtensorMap * m, const void * base, long long rows, long long rowbytes, int box) {
    encode_tiled_t enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t) rowbytes, (cuuint64_t) rows};
    const cuuint64_t strides[1] = {(cuuint64_t) rowbytes};
    const cuuint32_t bx[2] = {(cuuint32_t) box, (cuuint32_t) GM_BM};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapSwizzle sw = box == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (box == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE);
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, bx, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static int gd3_nf_rt(int type) { return type == T_Q6_K ? 3 : 2; }
static int gd3_box_rt(int type, int f) { return gd3_box(type, f); }

// nmat weight matrices A[j] (type `type`, SoA fields, M[j] rows of K) over the same activations B (fp16 [N][K], 16-B aligned rows);
// out[j][n * out_cs[j] + m]; split_ws: workspace of ksplit * N * M floats (nmat == 1), or nullptr (no split)
static cudaError_t launch_pair(int type, const WPtrs * A, float * const * out, const long long * out_cs, const int * M, int nmat,
                               const __half * B, const float * colscale, int N, int K, int sms, int ksplit, float * split_ws, float * tail_ws, int * nlaunch, cudaStream_t st) {
    const int bn = gemm_dense_bn(M[0], N, sms);
    GdMaps maps;
    CUtensorMap * const tmB = maps.b;
    GemmDenseArgs p{};
    p.N = N; p.K = K; p.ntiles = (N + bn - 1) / bn; p.nmat = nmat; p.ksplit = ksplit; p.colscale = colscale;
    if (!colscale || ((uintptr_t) colscale & 15) != 0) return cudaErrorInvalidValue;
    int tiles = 0;
    for (int j = 0; j < GD_MAXMAT; j++) {
        const int jj = j < nmat ? j : 0;
        for (int f = 0; f < 4; f++) {
            p.w[j][f] = A[jj].p[f];
            if (((uintptr_t) p.w[j][f] & 15) != 0 && f < wfields(type).n && wfields(type).size[f] >= 16) return cudaErrorInvalidValue;      // 128-bit loads
        }
        if (j < nmat) { tiles += ((M[j] + 2 * GM_BM - 1) / (2 * GM_BM)) * p.ntiles; p.outs[j] = out[j]; p.out_css[j] = out_cs[j]; p.Ms[j] = M[j]; p.tile_end[j] = tiles; }
        else { p.outs[j] = nullptr; p.out_css[j] = 0; p.Ms[j] = 0; p.tile_end[j] = 0x7fffffff; }
    }
    if (!make_map_operand(&tmB[0], B, N, K, 2, bn / 2)) return cudaErrorInvalidValue;
    // weights by TMA wherever every field row starts 16-byte aligned (B200_GEMM_WTMA=0: the cp.async ring variant, read per launch for the tests)
    bool wtma = !(getenv("B200_GEMM_WTMA") && atoi(getenv("B200_GEMM_WTMA")) == 0);
    const WFields wf = wfields(type);
    for (int j = 0; j < nmat && wtma; j++)
        for (int f = 0; f < gd3_nf_rt(type) && wtma; f++) {
            const long long rowbytes = (long long) (K / (type == T_Q4_K || type == T_Q6_K ? 256 : 32)) * wf.size[f];
            wtma = rowbytes % 16 == 0 && ((uintptr_t) A[j].p[f] & 15) == 0 && make_map_field(&maps.w[j][f], A[j].p[f], M[j], rowbytes, gd3_box_rt(type, f));
        }
    for (int j = nmat; j < GD_MAXMAT && wtma; j++) for (int f = 0; f < 3; f++) maps.w[j][f] = maps.w[0][f];
    if (wtma && gd3_nf_rt(type) == 2) for (int j = 0; j < GD_MAXMAT; j++) maps.w[j][2] = maps.w[j][0];
    if (ksplit > 1) { p.outs[0] = split_ws; p.out_css[0] = M[0]; p.split_stride = (long long) M[0] * N; }      // nmat == 1
    const int P = sms / 2;
    // stream-K tail over the last, partial wave (B200_GEMM_TAIL=0: off; =<percent>: the fill of the last wave up to which it is cut up).
    // Measured (q4_0, 512 columns): 11008 x 4096 (12 tiles left of 74) 55.9 -> 50.8 us; 14336 x 4096 (38 left) 58.1 -> 65.9 us: every partial
    // tile is 256 KB written and read again, which eats the gain from about a third of a wave on -> default 25
    p.full_units = tiles * ksplit; p.tail_T = 0; p.tail_L = 1; p.tail_ws = tail_ws;
    const int nks = (K + GD_KS - 1) / GD_KS;
    const int tail_pct = getenv("B200_GEMM_TAIL") ? atoi(getenv("B200_GEMM_TAIL")) : 25;
    const int rem = tiles % P;
    if (tail_ws && tail_pct > 0 && ksplit == 1 && rem > 0 && rem * 100 <= tail_pct * P && nks >= 16 && (tiles > P || 2 * rem > P)) {
        p.full_units = tiles - rem; p.tail_T = rem * nks; p.tail_L = (p.tail_T + P - 1) / P;
    }
    const int pairs = p.tail_T > 0 ? P : (p.full_units < P ? p.full_units : P);
    const int work = p.full_units + (p.tail_T > 0 ? 2 * pairs : 0);
    static const bool trace = getenv("B200_GEMM_TRACE") && atoi(getenv("B200_GEMM_TRACE")) > 0;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (trace) cudaStreamIsCapturing(st, &cap);                 // the trace allocates and synchronises: not inside a stream capture
    if (trace && cap == cudaStreamCaptureStatusNone && cudaMalloc((void **) &p.trace, (size_t) 2 * pairs * 8 * 8) == cudaSuccess)
        cudaMemsetAsync(p.trace, 0, (size_t) 2 * pairs * 8 * 8, st);
    const cudaError_t e = wtma ? (bn == 256 ? gemm_dense2_launch<256, true>(type, maps, p, 2 * pairs, st) : gemm_dense2_launch<128, true>(type, maps, p, 2 * pairs, st))
                               : (bn == 256 ? gemm_dense2_launch<256, false>(type, maps, p, 2 * pairs, st) : gemm_dense2_launch<128, false>(type, maps, p, 2 * pairs, st));
    if (nlaunch) *nlaunch = 1;
    if (e == cudaSuccess && p.tail_T > 0) {
        GdFixArgs fa{};
        fa.ws = tail_ws; fa.tail_T = p.tail_T; fa.tail_L = p.tail_L; fa.nks = nks; fa.full_tiles = p.full_units; fa.ntiles = p.ntiles; fa.bn = bn; fa.N = N;
        for (int j = 0; j < GD_MAXMAT; j++) { fa.outs[j] = p.outs[j]; fa.out_css[j] = p.out_css[j]; fa.Ms[j] = p.Ms[j]; fa.tile_end[j] = p.tile_end[j]; }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned) (rem * (bn * (2 * GM_BM / 4) / 256))); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
        const cudaError_t e2 = cudaLaunchKernelEx(&cfg, gd_tail_fixup_kernel, fa);
        if (e2 != cudaSuccess) return e2;
        if (nlaunch) *nlaunch = 2;
    }
    if (p.trace) {      // debugging aid (tools/gemm_bench.py under B200_GEMM_TRACE=1): per-CTA phase times in us since the first CTA started
        std::vector<unsigned long long> h((size_t) 2 * pairs * 8);
        cudaStreamSynchronize(st);
        cudaMemcpy(h.data(), p.trace, h.size() * 8, cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        unsigned long long t0 = ~0ull;
        for (int c = 0; c < 2 * pairs; c++) if (h[c * 8] && h[c * 8] < t0) t0 = h[c * 8];
        static const char * names[8] = {"start", "setup done", "first stage landed", "last MMA issued", "tile 0 complete", "last tile complete", "epilogue done", "end"};
        fprintf(stderr, "[gemm_dense2 trace] nmat=%d M=%d N=%d K=%d bn=%d ksplit=%d pairs=%d work=%d (full %d, tail %d steps in chunks of %d)\n", nmat, M[0], N, K, bn, ksplit, pairs, work, p.full_units, p.tail_T, p.tail_L);
        for (int sl = 0; sl < 8; sl++) {
            double mn = 1e30, mx = 0, sum = 0; int n = 0;
            for (int c = 0; c < 2 * pairs; c++) { const unsigned long long v = h[c * 8 + sl]; if (!v) continue; const double u = (double) (v - t0) * 1e-3; mn = u < mn ? u : mn; mx = u > mx ? u : mx; sum += u; n++; }
            if (n) fprintf(stderr, "  %-20s min %7.2f  avg %7.2f  max %7.2f us  (%d CTAs)\n", names[sl], mn, sum / n, mx, n);
        }
    }
    return e;
}

// A: quantised weights (type, SoA fields) [M][K]; B: fp16 [N][K] (K % 8 == 0) and its column scales as launch_dequant_act_f16 wrote them;
// out[n * out_cs + m].  split_ws: workspace of gemm_dense_ksplit(...) * N * M floats, or nullptr (no split); tail_ws:
// gemm_dense_tail_ws_bytes(sms), or nullptr (no stream-K tail); *nlaunch: kernels launched.
// act_out / act_cs_out: with res_dst, the next dense GEMM's activations (fp16 [N][M]) and column scales, converted by the reduce
cudaError_t launch_gemm_dense(int type, const WPtrs & A, const __half * B, const float * colscale, float * out, long long out_cs, int M, int N, int K, int sms,
                              float * split_ws, float * tail_ws, int * nlaunch, cudaStream_t st, const float * res_src, float * res_dst, void * act_out, float * act_cs_out) {
    const int ksplit = split_ws ? gemm_dense_ksplit(M, N, K, sms) : 1;
    if (res_dst && ksplit == 1) return cudaErrorInvalidValue;
    const cudaError_t e = launch_pair(type, &A, &out, &out_cs, &M, 1, B, colscale, N, K, sms, ksplit, split_ws, tail_ws, nlaunch, st);
    if (e != cudaSuccess || ksplit == 1) return e;
    if (nlaunch) (*nlaunch)++;
    if (act_out) return launch_reduce_splits_act(split_ws, ksplit, M, N, out, out_cs, st, res_src, res_dst, act_out, act_cs_out);
    return launch_reduce_splits(split_ws, ksplit, (long long) M * N, M, out, out_cs, 0, st, res_src, res_dst, nullptr);
}

// Several weight matrices of one type over the same activations in ONE launch (no split-K): 2 <= nmat <= 3
cudaError_t launch_gemm_dense_group(int type, const WPtrs * A, float * const * out, const long long * out_cs, const int * M, int nmat,
                                    const __half * B, const float * colscale, int N, int K, int sms, float * tail_ws, int * nlaunch, cudaStream_t st) {
    if (nmat < 1 || nmat > GD_MAXMAT) return cudaErrorNotSupported;
    return launch_pair(type, A, out, out_cs, M, nmat, B, colscale, N, K, sms, 1, nullptr, tail_ws, nlaunch, st);
}

}  // namespace b200

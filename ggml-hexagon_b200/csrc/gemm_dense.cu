// gemm_dense.cu — prefill MUL_MAT as a DENSE fp16 tensor-core contraction: tcgen05.mma kind::f16 accumulating the whole K in
// TMEM, operands staged by TMA, one TMEM read per output tile.
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
// |B| saturates at the fp16 maximum (activations beyond 65504 do not occur in the model families of the task).
//
// CTA (1 per SM, persistent over 128 x BN output tiles), 10 warps:
//   warp 0      TMA producer: A = fp16 weights [M][K], B = fp16 activations [N][K]; boxes of 64 k (128 bytes) x 128 / BN rows,
//               128-B swizzle, ring of GD_STAGES stages
//   warp 1      TMEM allocation + MMA issue (one elected lane): 4 K16 MMAs per stage into one of two accumulator buffers,
//               tcgen05.commit per stage (frees the ring slot) and per tile (hands the buffer to the epilogue)
//   warps 2..9  epilogue: tcgen05.ld 32 lanes x 32 columns, coalesced f32 stores (32 consecutive m per column); the other
//               buffer is being accumulated meanwhile
#include "tc_common.cuh"

#include <stdio.h>
#include <vector>

namespace b200 {

constexpr int GD_KS = 64;                                    // fp16 elements of K per stage = one 128-B swizzle row
constexpr int GD_EPI_WARPS = 8;                              // 4 TMEM lane quarters x 2 column halves
constexpr int GD_THREADS = 64 + GD_EPI_WARPS * 32;
__host__ __device__ constexpr int gd_stages(int bn) { return bn == 128 ? 6 : 4; }
__host__ __device__ constexpr int gd_smem(int bn) { return gd_stages(bn) * (GM_BM * 128 + bn * 128) + 256 + 1024; }
static_assert(gd_smem(128) <= 232448 && gd_smem(256) <= 232448, "shared memory");

struct GemmDenseArgs {
    float * out; long long out_cs;              // dst[n * out_cs + m]
    int M, N, K;
    int mtiles, ntiles;
    int ksplit; long long split_stride;        // split-K: work item = (tile, split); split sp writes its partial sums to out + sp * split_stride
    unsigned long long * trace;                // B200_GEMM_TRACE=1: 8 globaltimer stamps per CTA (pair kernel), else nullptr
    // pair kernel: up to GD_MAXMAT weight matrices over the same activations in one launch (q / k / v, gate / up): the tile list
    // runs through all of them, which fills the waves that one 4096-row matrix cannot (32 tiles for 74 CTA pairs)
    int nmat;
    float * outs[3]; long long out_css[3]; int Ms[3];
    int tile_end[3];                           // tiles (m-tile x n-tile) up to and including matrix j; INT_MAX past nmat
    // the last, partial wave in narrow units: work items >= full_units are 256 x (BN / narrow_f) column slices of the remaining tiles,
    // so that a wave with few tiles left takes 1/2 or 1/4 of a tile time instead of a whole one (narrow_f = 1: none; ksplit == 1)
    int full_units, narrow_f;
};
struct GdUnit { int tile, sp, n_off, n_w; };    // n_off: first column inside the tile; n_w: columns = N of the MMA
__device__ __forceinline__ GdUnit gd2_unit(const GemmDenseArgs & p, const int work, const int BN) {
    GdUnit u;
    if (work < p.full_units) { u.tile = work / p.ksplit; u.sp = work - u.tile * p.ksplit; u.n_off = 0; u.n_w = BN; }
    else {
        const int r = work - p.full_units, t = r / p.narrow_f;
        u.tile = p.full_units + t; u.sp = 0; u.n_w = BN / p.narrow_f; u.n_off = (r - t * p.narrow_f) * u.n_w;
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
__device__ __forceinline__ void gd_store_chunk(float * colp, const long long cs, const int (&v)[32], const int ncols) {
    if (ncols >= 32) {
        #pragma unroll
        for (int i = 0; i < 32; i++) { *colp = __int_as_float(v[i]); colp += cs; }
    } else {
        #pragma unroll
        for (int i = 0; i < 32; i++) { if (i < ncols) *colp = __int_as_float(v[i]); colp += cs; }
    }
}
template <int NCH>
__device__ __forceinline__ void gd_epilogue_warp(const uint32_t taddr, float * colp, const long long cs, const int ncols, const bool row_ok) {
    static_assert(NCH % 2 == 0, "chunks are processed in pairs");
    int va[32], vb[32];
    gm_tmem_ld32(taddr, va);
    gm_tmem_ld_wait();
    #pragma unroll
    for (int c = 0; c < NCH; c += 2) {
        gm_tmem_ld32(taddr + (c + 1) * 32, vb);
        if (row_ok && ncols > c * 32) gd_store_chunk(colp + (long long) c * 32 * cs, cs, va, ncols - c * 32);
        gm_tmem_ld_wait();
        if (c + 2 < NCH) gm_tmem_ld32(taddr + (c + 2) * 32, va);
        if (row_ok && ncols > (c + 1) * 32) gd_store_chunk(colp + (long long) (c + 1) * 32 * cs, cs, vb, ncols - (c + 1) * 32);
        if (c + 2 < NCH) gm_tmem_ld_wait();
    }
}

__device__ __forceinline__ void gd_epilogue_warp1(const uint32_t taddr, float * colp, const long long cs, const int ncols, const bool row_ok) {
    int va[32];
    gm_tmem_ld32(taddr, va);
    gm_tmem_ld_wait();
    if (row_ok && ncols > 0) gd_store_chunk(colp, cs, va, ncols);
}

template <int BN>
__global__ void __launch_bounds__(GD_THREADS, 1)
gemm_dense_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmDenseArgs p) {
    constexpr int STAGES = gd_stages(BN), STAGE_A = GM_BM * 128, STAGE_B = BN * 128;
    constexpr int OFF_B = STAGES * STAGE_A, OFF_BAR = OFF_B + STAGES * STAGE_B;
    extern __shared__ uint8_t gd_smem_raw[];
    const uint32_t base = (gm_smem_u32(gd_smem_raw) + 1023u) & ~1023u;           // 128-B swizzle atoms want 1024-B alignment
    uint8_t * smem = gd_smem_raw + (base - gm_smem_u32(gd_smem_raw));
    const uint32_t sA = base, sB = base + OFF_B, bars = base + OFF_BAR;
    // full: TMA landed; empty: the stage's MMAs have read it; tfull: a tile's accumulator is complete; tempty: the epilogue has read it
    const uint32_t full0 = bars, empty0 = bars + STAGES * 8, tfull0 = bars + 2 * STAGES * 8, tempty0 = tfull0 + 2 * 8;
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (smem + OFF_BAR + 2 * STAGES * 8 + 4 * 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nks = (p.K + GD_KS - 1) / GD_KS, nwork = p.mtiles * p.ntiles * p.ksplit;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) { gm_mbar_init(full0 + i * 8, 1); gm_mbar_init(empty0 + i * 8, 1); }
        for (int i = 0; i < 2; i++) { gm_mbar_init(tfull0 + i * 8, 1); gm_mbar_init(tempty0 + i * 8, GD_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {            // two f32 accumulator buffers of BN columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(gm_smem_u32((const void *) tmem_slot)), "n"(2 * BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    gm_fence_before();
    __syncthreads();
    gm_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer.  Work items are numbered m-tile fastest: the CTAs running at the same time share their B tile
        // (activations, small) in L2 and stream different weight rows
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmA) : "memory");
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmB) : "memory");
        }
        __syncwarp();
        {   // whole warp, one elected lane issues (as the MMA warp)
            uint32_t it = 0;
            for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
                const int tile = work / p.ksplit, sp = work - tile * p.ksplit;
                const int nt = tile / p.mtiles, mt = tile - nt * p.mtiles;
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(empty0 + s * 8, ph ^ 1);
                    gm_mbar_expect_tx_elect(full0 + s * 8, STAGE_A + STAGE_B);
                    gm_tma_2d_elect(sA + s * STAGE_A, &tmA, full0 + s * 8, ks * GD_KS, mt * GM_BM);
                    gm_tma_2d_elect(sB + s * STAGE_B, &tmB, full0 + s * 8, ks * GD_KS, nt * BN);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp runs the loop (uniform control flow), one elected lane issues (tc_common.cuh)
        {
            // instruction descriptor: D = f32 (bits 4-5 = 1), A / B = fp16 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
            const uint32_t idesc = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (GM_BM >> 4) << 24);
            uint32_t it = 0, tc = 0;
            for (int work = blockIdx.x; work < nwork; work += gridDim.x, tc++) {
                const int sp = work % p.ksplit;
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
                gm_mbar_wait(tempty0 + buf * 8, bph ^ 1);
                gm_fence_after();
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(full0 + s * 8, ph);
                    gm_fence_after();
                    const uint64_t ad = gm_smem_desc(sA + s * STAGE_A), bd = gm_smem_desc(sB + s * STAGE_B);
                    #pragma unroll
                    for (int j = 0; j < 4; j++)                   // every MMA consumes 16 fp16 = 32 bytes of K per row: start address field + 2
                        gm_mma_f16_elect(tmem + buf * BN, ad + 2u * j, bd + 2u * j, idesc, (ks > ks0 || j > 0) ? 1u : 0u);
                    gm_commit_elect(empty0 + s * 8);              // arrives when the stage's MMAs have read shared memory
                }
                gm_commit_elect(tfull0 + buf * 8);                // arrives when the tile's sums are in TMEM
            }
        }
    } else {
        // ===== epilogue: warp w reads TMEM lanes 32 * (w % 4) .. +31 (hardware rule), column half (w - 2) / 4
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int m_local = q * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t) (q * 32) << 16) + half * (BN / 2);
        uint32_t tc = 0;
        for (int work = blockIdx.x; work < nwork; work += gridDim.x, tc++) {
            const int tile = work / p.ksplit, sp = work - tile * p.ksplit;
            const int nt = tile / p.mtiles, mt = tile - nt * p.mtiles;
            float * const outp = p.out + (long long) sp * p.split_stride;
            const int m = mt * GM_BM + m_local;
            const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
            gm_mbar_wait(tfull0 + buf * 8, bph);            // every work item has k-steps: ksplit <= nks / 8
            gm_fence_after();
            const int n0 = nt * BN + half * (BN / 2);
            gd_epilogue_warp<BN / 64>(tbase + buf * BN, outp + (long long) n0 * p.out_cs + m, p.out_cs, p.N - n0, m < p.M);
            gm_fence_before();
            __syncwarp();
            if (lane == 0) gm_mbar_arrive(tempty0 + buf * 8);      // buffer free for the MMA issuer
        }
    }
    gm_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(2 * BN) : "memory");
}

// ------------------------------------------------------------------------------------------------ CTA-pair variant
// The 1-CTA kernel above is bound by the L2 -> shared-memory operand feed, not by the tensor pipe: a 128 x 128 tile wants
// 8 KB of operands per 64-clock MMA (128 B/clk/SM) while all of L2 delivers ~6300 B/clk (~43 B/clk/SM); ncu: tensor pipe 27 %
// busy (profiles/r01_l_ncu_gemm_dense_1cta_*).  Here two CTAs of one TPC (cluster of 2) share one 256 x BN tile through
// tcgen05.mma.cta_group::2: every CTA stages its own 128 weight rows and HALF of the activation rows, the MMA of the leader
// reads B from both shared memories, and each CTA's TMEM receives its 128 rows x BN columns.  With BN = 256 that is 8 KB per
// 128-clock MMA and SM = 64 B/clk/SM, and the total L2 traffic M N K 2 (1/256 + 1/256) is half of the 128 x 128 kernel's.
//   full[s]   leader only; both CTAs' TMA loads (.cta_group::2) complete_tx on it, the leader posts the expected bytes of both
//   empty[s]  per CTA; tcgen05.commit.multicast arrives on both when the stage's MMAs have read shared memory
//   tfull[b]  per CTA; multicast commit after a tile's last MMA
//   tempty[b] leader only; the epilogue warps of BOTH CTAs arrive on it (remote arrive from the peer)
__host__ __device__ constexpr int gd2_stage_bytes(int bn) { return GM_BM * 128 + (bn / 2) * 128; }
__host__ __device__ constexpr int gd2_stages(int bn) { return (232448 - 1280) / gd2_stage_bytes(bn); }
__host__ __device__ constexpr int gd2_smem(int bn) { return gd2_stages(bn) * gd2_stage_bytes(bn) + 256 + 1024; }
static_assert(gd2_stages(256) == 7 && gd2_stages(128) == 9, "stage count");

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

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GD_THREADS, 1)
gemm_dense2_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
                   const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBq,
                   const GemmDenseArgs p) {
    constexpr int STAGES = gd2_stages(BN), STAGE_A = GM_BM * 128, STAGE_B = (BN / 2) * 128;
    constexpr int OFF_B = STAGES * STAGE_A, OFF_BAR = OFF_B + STAGES * STAGE_B;
    extern __shared__ uint8_t gd_smem_raw[];
    const uint32_t base = (gm_smem_u32(gd_smem_raw) + 1023u) & ~1023u;
    uint8_t * smem = gd_smem_raw + (base - gm_smem_u32(gd_smem_raw));
    const uint32_t sA = base, sB = base + OFF_B, bars = base + OFF_BAR;
    const uint32_t full0 = bars, empty0 = bars + STAGES * 8, tfull0 = bars + 2 * STAGES * 8, tempty0 = tfull0 + 2 * 8;
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (smem + OFF_BAR + 2 * STAGES * 8 + 4 * 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    if (threadIdx.x == 0) gd_stamp(p, 0);
    // programmatic dependent launch: the next kernel of the stream may be scheduled as soon as SMs free up (its prologue runs beside our
    // tail); we ourselves set up barriers and TMEM beside the previous kernel's tail and wait for its results only before the first load
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int nks = (p.K + GD_KS - 1) / GD_KS;
    const int nwork = p.full_units + (p.tile_end[p.nmat - 1] * p.ksplit - p.full_units) * p.narrow_f;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) { gm_mbar_init(full0 + i * 8, 1); gm_mbar_init(empty0 + i * 8, 1); }
        for (int i = 0; i < 2; i++) { gm_mbar_init(tfull0 + i * 8, 1); gm_mbar_init(tempty0 + i * 8, 2 * GD_EPI_WARPS); }
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
    asm volatile("griddepcontrol.wait;" ::: "memory");      // every role: loads of B, stores of dst / the split workspace
    if (threadIdx.x == 0) gd_stamp(p, 1);

    if (warp == 0) {
        // ===== TMA producer (both CTAs).  Work items are numbered n-tile fastest: the pairs that share a weight tile run at the
        // same time, so the fp16 weights come from HBM once (the activations stay in L2 anyway)
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmA0) : "memory");
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmB) : "memory");
        }
        __syncwarp();
        {   // whole warp, one elected lane issues
            const uint32_t lfull0 = gd2_mapa(full0, 0);
            uint32_t it = 0;
            for (int work = pair; work < nwork; work += npairs) {
                const GdUnit u = gd2_unit(p, work, BN);
                const int tile = u.tile, sp = u.sp;
                const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
                const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
                const int mt = local / p.ntiles, nt = local - mt * p.ntiles;
                const CUtensorMap * const tmA = j == 0 ? &tmA0 : (j == 1 ? &tmA1 : &tmA2);
                const CUtensorMap * const tmBu = u.n_w == BN ? &tmB : (u.n_w == BN / 2 ? &tmBh : &tmBq);      // box rows = n_w / 2
                const uint32_t tx = 2u * (STAGE_A + (uint32_t) (u.n_w / 2) * 128u);
                const int brow = nt * BN + u.n_off + (int) rank * (u.n_w / 2);
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(empty0 + s * 8, ph ^ 1);
                    if (rank == 0) gm_mbar_expect_tx_elect(full0 + s * 8, tx);
                    gd2_tma_2d(sA + s * STAGE_A, tmA, lfull0 + s * 8, ks * GD_KS, mt * 2 * GM_BM + (int) rank * GM_BM);
                    gd2_tma_2d(sB + s * STAGE_B, tmBu, lfull0 + s * 8, ks * GD_KS, brow);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: warp 1 of the leader CTA (whole warp, one elected lane issues), M = 256 over the pair
        if (rank == 0) {
            uint32_t it = 0, tc = 0;
            for (int work = pair; work < nwork; work += npairs, tc++) {
                const GdUnit u = gd2_unit(p, work, BN);
                const int sp = u.sp;
                const uint32_t idesc = (1u << 4) | ((uint32_t) (u.n_w >> 3) << 17) | ((uint32_t) ((2 * GM_BM) >> 4) << 24);
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
                gm_mbar_wait(tempty0 + buf * 8, bph ^ 1);
                gm_fence_after();
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    gm_mbar_wait(full0 + s * 8, ph);
                    gm_fence_after();
                    if (it == 0 && lane == 0) gd_stamp(p, 2);
                    const uint64_t ad = gm_smem_desc(sA + s * STAGE_A), bd = gm_smem_desc(sB + s * STAGE_B);
                    #pragma unroll
                    for (int j = 0; j < 4; j++)
                        gd2_mma_f16(tmem + buf * BN, ad + 2u * j, bd + 2u * j, idesc, (ks > ks0 || j > 0) ? 1u : 0u);
                    gd2_commit_both(empty0 + s * 8);
                }
                gd2_commit_both(tfull0 + buf * 8);
            }
            if (lane == 0) gd_stamp(p, 3);
        }
    } else {
        // ===== epilogue (both CTAs): this CTA's TMEM holds rows rank * 128 .. +127 of the pair tile, all BN columns
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int m_local = q * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t) (q * 32) << 16);
        const uint32_t ltempty0 = gd2_mapa(tempty0, 0);
        uint32_t tc = 0;
        for (int work = pair; work < nwork; work += npairs, tc++) {
            const GdUnit u = gd2_unit(p, work, BN);
            const int tile = u.tile, sp = u.sp;
            const int j = (tile >= p.tile_end[0]) + (tile >= p.tile_end[1]);
            const int local = tile - (j == 0 ? 0 : p.tile_end[j - 1]);
            const int mt = local / p.ntiles, nt = local - mt * p.ntiles;
            float * const outp = p.outs[j] + (long long) sp * p.split_stride;
            const long long ocs = p.out_css[j];
            const int m = mt * 2 * GM_BM + (int) rank * GM_BM + m_local;
            const uint32_t buf = tc & 1, bph = (tc >> 1) & 1;
            gm_mbar_wait(tfull0 + buf * 8, bph);
            gm_fence_after();
            if (warp == 2 && lane == 0) gd_stamp(p, tc == 0 ? 4 : 5);
            // this warp: TMEM lanes q * 32 .. +31, columns half * (n_w / 2) .. + n_w / 2 of the unit
            const int c0 = half * (u.n_w / 2), n0 = nt * BN + u.n_off + c0;
            const uint32_t ta = tbase + buf * BN + c0;
            float * const colp = outp + (long long) n0 * ocs + m;
            const bool row_ok = m < p.Ms[j];
            if (u.n_w == BN)               gd_epilogue_warp<BN / 64>(ta, colp, ocs, p.N - n0, row_ok);
            else if (u.n_w == BN / 2) {
                if constexpr (BN == 256)   gd_epilogue_warp<BN / 128>(ta, colp, ocs, p.N - n0, row_ok);
                else                       gd_epilogue_warp1(ta, colp, ocs, p.N - n0, row_ok);
            } else                         gd_epilogue_warp1(ta, colp, ocs, p.N - n0, row_ok);
            gm_fence_before();
            __syncwarp();
            if (lane == 0) gd2_arrive_cluster(ltempty0 + buf * 8);
        }
        if (warp == 2 && lane == 0) gd_stamp(p, 6);
    }
    __syncwarp();                // the cluster barrier is warp-aligned: the single-lane roles rejoin their warps first
    gm_fence_before();
    gd2_cluster_sync();          // the peer's shared memory and barriers stay alive until the leader's MMAs and commits are done
    if (threadIdx.x == 0) gd_stamp(p, 7);
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(2 * BN) : "memory");
}

// ------------------------------------------------------------------------------------------------ operand preparation
__device__ __forceinline__ uint32_t gd_pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *(const uint32_t *) &h;
}
// 4 signed bytes of w times d -> 4 halfs
__device__ __forceinline__ uint2 gd_s8x4(uint32_t w, float d) {
    return make_uint2(gd_pack_h2(d * (float) (int8_t) (w & 0xFF), d * (float) (int8_t) ((w >> 8) & 0xFF)),
                      gd_pack_h2(d * (float) (int8_t) ((w >> 16) & 0xFF), d * (float) (int8_t) (w >> 24)));
}
// 4 unsigned bytes: d * (q - bias) - mn
__device__ __forceinline__ uint2 gd_u8x4(uint32_t w, float d, float bias, float mn) {
    return make_uint2(gd_pack_h2(d * ((float) (w & 0xFF) - bias) - mn, d * ((float) ((w >> 8) & 0xFF) - bias) - mn),
                      gd_pack_h2(d * ((float) ((w >> 16) & 0xFF) - bias) - mn, d * ((float) (w >> 24) - bias) - mn));
}
__device__ __forceinline__ void gd_store16(__half * out, const uint2 a, const uint2 b, const uint2 c, const uint2 d) {
    ((uint4 *) out)[0] = make_uint4(a.x, a.y, b.x, b.y);
    ((uint4 *) out)[1] = make_uint4(c.x, c.y, d.x, d.y);
}

// q8_0 (SoA: qs 32 B, d f16 per block; y = d * q, dequantize_row_q8_0 ggml-quants.c:347-363): one thread per 16 elements
__global__ void __launch_bounds__(256)
dequant_q8_0_f16_kernel(const uint4 * __restrict__ qs, const __half * __restrict__ dd, const long long n16, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (long long) gridDim.x * blockDim.x) {
        const uint4 v = qs[i];
        const float d = __half2float(dd[i >> 1]);
        gd_store16(out + i * 16, gd_s8x4(v.x, d), gd_s8x4(v.y, d), gd_s8x4(v.z, d), gd_s8x4(v.w, d));
    }
}
// q4_0 (qs 16 B: low nibbles = elements 0..15, high = 16..31; y = d * (q - 8), dequantize_row_q4_0 ggml-quants.c:307-325): one thread per block
__global__ void __launch_bounds__(256)
dequant_q4_0_f16_kernel(const uint4 * __restrict__ qs, const __half * __restrict__ dd, const long long nblocks, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nblocks; i += (long long) gridDim.x * blockDim.x) {
        const uint4 v = qs[i];
        const float d = __half2float(dd[i]);
        const uint32_t m = 0x0F0F0F0Fu;
        gd_store16(out + i * 32,      gd_u8x4(v.x & m, d, 8.f, 0.f), gd_u8x4(v.y & m, d, 8.f, 0.f), gd_u8x4(v.z & m, d, 8.f, 0.f), gd_u8x4(v.w & m, d, 8.f, 0.f));
        gd_store16(out + i * 32 + 16, gd_u8x4((v.x >> 4) & m, d, 8.f, 0.f), gd_u8x4((v.y >> 4) & m, d, 8.f, 0.f),
                                      gd_u8x4((v.z >> 4) & m, d, 8.f, 0.f), gd_u8x4((v.w >> 4) & m, d, 8.f, 0.f));
    }
}
// q4_K (qs 128 B + header {f16 d, f16 dmin, 12 B packed 6-bit scales / mins}; y = (d * sc_g) * q - dmin * m_g,
// dequantize_row_q4_K ggml-quants.c:1352-1374): one thread per 64 elements (low nibbles = sub-block 2j, high = 2j + 1)
__global__ void __launch_bounds__(256)
dequant_q4_K_f16_kernel(const uint4 * __restrict__ qs, const uint4 * __restrict__ hdr, const long long nsuper, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 4; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 2; const int j = (int) (i & 3);
        const uint4 v0 = qs[sb * 8 + 2 * j], v1 = qs[sb * 8 + 2 * j + 1];
        const uint4 h = hdr[sb];
        const __half2 dd = *(const __half2 *) &h.x;
        const float d = __low2float(dd), dmin = __high2float(dd);
        const uint32_t w[3] = {h.y, h.z, h.w};
        auto byte = [&](int k) -> uint32_t { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; };
        float ds[2], ms[2];
        #pragma unroll
        for (int e = 0; e < 2; e++) {
            const int g = 2 * j + e;
            uint32_t sc, mn;
            if (g < 4) { sc = byte(g) & 63u; mn = byte(g + 4) & 63u; }
            else       { sc = (byte(g + 4) & 0xFu) | ((byte(g - 4) >> 6) << 4); mn = (byte(g + 4) >> 4) | ((byte(g) >> 6) << 4); }
            ds[e] = __fmul_rn(d, (float) sc); ms[e] = __fmul_rn(dmin, (float) mn);
        }
        const uint32_t m = 0x0F0F0F0Fu;
        __half * o = out + sb * 256 + j * 64;
        gd_store16(o,      gd_u8x4(v0.x & m, ds[0], 0.f, ms[0]), gd_u8x4(v0.y & m, ds[0], 0.f, ms[0]), gd_u8x4(v0.z & m, ds[0], 0.f, ms[0]), gd_u8x4(v0.w & m, ds[0], 0.f, ms[0]));
        gd_store16(o + 16, gd_u8x4(v1.x & m, ds[0], 0.f, ms[0]), gd_u8x4(v1.y & m, ds[0], 0.f, ms[0]), gd_u8x4(v1.z & m, ds[0], 0.f, ms[0]), gd_u8x4(v1.w & m, ds[0], 0.f, ms[0]));
        gd_store16(o + 32, gd_u8x4((v0.x >> 4) & m, ds[1], 0.f, ms[1]), gd_u8x4((v0.y >> 4) & m, ds[1], 0.f, ms[1]),
                           gd_u8x4((v0.z >> 4) & m, ds[1], 0.f, ms[1]), gd_u8x4((v0.w >> 4) & m, ds[1], 0.f, ms[1]));
        gd_store16(o + 48, gd_u8x4((v1.x >> 4) & m, ds[1], 0.f, ms[1]), gd_u8x4((v1.y >> 4) & m, ds[1], 0.f, ms[1]),
                           gd_u8x4((v1.z >> 4) & m, ds[1], 0.f, ms[1]), gd_u8x4((v1.w >> 4) & m, ds[1], 0.f, ms[1]));
    }
}
// q6_K (ql 128 B, qh 64 B, 16 int8 scales, f16 d; y = d * sc_g * (q - 32), bit layout as dequantize_row_q6_K
// ggml-quants.c:1681-1710): 16 threads per super-block, thread t handles l = 4 (t & 7) .. +3 of half (t >> 3)
__global__ void __launch_bounds__(256)
dequant_q6_K_f16_kernel(const uint32_t * __restrict__ ql, const uint32_t * __restrict__ qh, const int8_t * __restrict__ sc,
                        const __half * __restrict__ dd, const long long nsuper, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 16; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 4; const int t = (int) (i & 15), half = t >> 3, l4 = t & 7;
        const uint32_t a = ql[sb * 32 + half * 16 + l4], b = ql[sb * 32 + half * 16 + 8 + l4], h = qh[sb * 16 + half * 8 + l4];
        const uint32_t v[4] = {
            (a & 0x0F0F0F0Fu)        | ((h << 4) & 0x30303030u),
            (b & 0x0F0F0F0Fu)        | ((h << 2) & 0x30303030u),
            ((a >> 4) & 0x0F0F0F0Fu) | ( h       & 0x30303030u),
            ((b >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u)};
        const float d = __half2float(dd[sb]);
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            const int e = half * 128 + r * 32 + l4 * 4;          // element index inside the super-block
            const float dg = __fmul_rn(d, (float) sc[sb * 16 + (e >> 4)]);
            *(uint2 *) (out + sb * 256 + e) = gd_u8x4(v[r], dg, 32.f, 0.f);
        }
    }
}

// activations: f32 [N][K] -> q8_1 (d = amax / 127 rounded to fp16, q = round(x / d); quantize_row_q8_1_ref ggml-quants.c:220-253)
// -> fp16(d * q), saturating.  4 lanes per 32-element block, two float4 loads and one 16-byte store each.
__global__ void __launch_bounds__(256)
dequant_act_f16_kernel(const char * __restrict__ x, const long long row_stride, const int K, const int N, __half * __restrict__ out) {
    const int NB = K / 32;
    const long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    const long long blk = gid >> 2;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");      // x comes from the kernels before us; out may still be read by the GEMM before us
    if (blk >= (long long) NB * N) return;                   // (whole groups of 4 lanes leave together: NB * N blocks, 4 lanes each)
    const int row = (int) (blk / NB), b = (int) (blk - (long long) row * NB), sub = (int) (gid & 3);
    const float4 * src = (const float4 *) (x + row * row_stride + ((long long) b * 32 + sub * 8) * 4);
    const float4 v = src[0], w = src[1];
    float amax = fmaxf(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))),
                       fmaxf(fmaxf(fabsf(w.x), fabsf(w.y)), fmaxf(fabsf(w.z), fabsf(w.w))));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const float dh = __half2float(__float2half_rn(d));
    const float r0 = dh * roundf(__fmul_rn(v.x, id)), r1 = dh * roundf(__fmul_rn(v.y, id));
    const float r2 = dh * roundf(__fmul_rn(v.z, id)), r3 = dh * roundf(__fmul_rn(v.w, id));
    const float r4 = dh * roundf(__fmul_rn(w.x, id)), r5 = dh * roundf(__fmul_rn(w.y, id));
    const float r6 = dh * roundf(__fmul_rn(w.z, id)), r7 = dh * roundf(__fmul_rn(w.w, id));
    uint4 h;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.x) : "f"(r1), "f"(r0));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.y) : "f"(r3), "f"(r2));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.z) : "f"(r5), "f"(r4));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.w) : "f"(r7), "f"(r6));
    *(uint4 *) (out + (long long) row * K + b * 32 + sub * 8) = h;
}

cudaError_t launch_dequant_act_f16(const float * x, long long row_stride_bytes, int K, int N, __half * out, cudaStream_t st) {
    const long long threads = (long long) (K / 8) * N;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) ((threads + 255) / 256)); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, dequant_act_f16_kernel, (const char *) x, row_stride_bytes, K, N, out);
}
// weights of `type` (SoA field pointers w, nblocks native blocks) -> dequantised fp16 [M][K]
cudaError_t launch_dequant_weights_f16(int type, const WPtrs & w, long long nblocks, __half * out, cudaStream_t st) {
    auto grid_for = [](long long threads) { long long b = (threads + 255) / 256; return (unsigned) (b > 148 * 32 ? 148 * 32 : (b < 1 ? 1 : b)); };
    switch (type) {
        case T_Q8_0: dequant_q8_0_f16_kernel<<<grid_for(nblocks * 2), 256, 0, st>>>((const uint4 *) w.p[0], (const __half *) w.p[1], nblocks * 2, out); break;
        case T_Q4_0: dequant_q4_0_f16_kernel<<<grid_for(nblocks), 256, 0, st>>>((const uint4 *) w.p[0], (const __half *) w.p[1], nblocks, out); break;
        case T_Q4_K: dequant_q4_K_f16_kernel<<<grid_for(nblocks * 4), 256, 0, st>>>((const uint4 *) w.p[0], (const uint4 *) w.p[1], nblocks, out); break;
        case T_Q6_K: dequant_q6_K_f16_kernel<<<grid_for(nblocks * 16), 256, 0, st>>>((const uint32_t *) w.p[0], (const uint32_t *) w.p[1], (const int8_t *) w.p[2],
                                                                                    (const __half *) w.p[3], nblocks, out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ host: launch
// B200_GEMM_CTA=1 keeps the 1-CTA kernel (128 x BN tiles); default: the CTA-pair kernel (256 x BN tiles)
bool gemm_dense_pair() { const char * e = getenv("B200_GEMM_CTA"); return !(e && atoi(e) == 1); }
// N tile: 256 columns halve the number of times the weights travel from L2 to the SMs.  1-CTA kernel: only where the 128 x 256
// tiles still fill the machine; pair kernel: always (few tiles are spread with split-K instead, the L2 feed is what binds)
int gemm_dense_bn(int M, int N, int sms) {
    if (getenv("B200_GEMM_BN")) return atoi(getenv("B200_GEMM_BN")) == 256 ? 256 : 128;
    if (gemm_dense_pair()) return N > 128 ? 256 : 128;
    const long long t256 = (long long) ((M + GM_BM - 1) / GM_BM) * ((N + 255) / 256);
    return N > 128 && t256 >= sms ? 256 : 128;
}
int gemm_dense_ksplit(int M, int N, int K, int sms) {
    const int bn = gemm_dense_bn(M, N, sms);
    const int bm = gemm_dense_pair() ? 2 * GM_BM : GM_BM, units = gemm_dense_pair() ? sms / 2 : sms;
    const int tiles = ((M + bm - 1) / bm) * ((N + bn - 1) / bn);
    const int nks = (K + GD_KS - 1) / GD_KS;
    int s = units / (tiles > 0 ? tiles : 1);
    if (s > nks / 8) s = nks / 8;
    if (s > 8) s = 8;
    return s < 2 ? 1 : s;
}

template <int BN>
static cudaError_t gemm_dense_launch_t(const CUtensorMap & tmA, const CUtensorMap & tmB, const GemmDenseArgs & p, int grid, cudaStream_t st) {
    static bool attr_done[64] = {false};          // per device
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    if (!attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_dense_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, gd_smem(BN));
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    gemm_dense_kernel<BN><<<(unsigned) grid, GD_THREADS, gd_smem(BN), st>>>(tmA, tmB, p);
    return cudaGetLastError();
}

template <int BN>
static cudaError_t gemm_dense2_launch_t(const CUtensorMap * tmA, const CUtensorMap * tmB, const GemmDenseArgs & p, int grid, cudaStream_t st) {
    static bool attr_done[64] = {false};          // per device
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    if (!attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_dense2_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, gd2_smem(BN));
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3(GD_THREADS); cfg.dynamicSmemBytes = gd2_smem(BN); cfg.stream = st;      // __cluster_dims__(2): grid is even
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, gemm_dense2_kernel<BN>, tmA[0], tmA[1], tmA[2], tmB[0], tmB[1], tmB[2], p);
}

// A: fp16 [M][K], B: fp16 [N][K] (16-B aligned rows: K % 8 == 0); out[n * out_cs + m]
// split_ws: workspace of gemm_dense_ksplit(...) * N * M floats, or nullptr (no split)
// pair kernel: nmat weight matrices A[j] (fp16 [M[j]][K]) over the same activations B (fp16 [N][K]); out[j][n * out_cs[j] + m]
static cudaError_t launch_pair(const __half * const * A, float * const * out, const long long * out_cs, const int * M, int nmat,
                               const __half * B, int N, int K, int sms, int ksplit, float * split_ws, cudaStream_t st) {
    const int bn = gemm_dense_bn(M[0], N, sms);
    CUtensorMap tmA[GD_MAXMAT], tmB[3];
    GemmDenseArgs p{};
    p.N = N; p.K = K; p.ntiles = (N + bn - 1) / bn; p.nmat = nmat; p.ksplit = ksplit;
    int tiles = 0;
    for (int j = 0; j < GD_MAXMAT; j++) {
        const int jj = j < nmat ? j : 0;
        if (!make_map_operand(&tmA[j], A[jj], M[jj], K, 2, GM_BM)) return cudaErrorInvalidValue;
        if (j < nmat) { tiles += ((M[j] + 2 * GM_BM - 1) / (2 * GM_BM)) * p.ntiles; p.outs[j] = out[j]; p.out_css[j] = out_cs[j]; p.Ms[j] = M[j]; p.tile_end[j] = tiles; }
        else { p.outs[j] = nullptr; p.out_css[j] = 0; p.Ms[j] = 0; p.tile_end[j] = 0x7fffffff; }
    }
    for (int f = 0; f < 3; f++)                 // activation boxes for full units and for the 1/2 and 1/4 wide units of the last wave
        if (!make_map_operand(&tmB[f], B, N, K, 2, (bn / 2) >> f)) return cudaErrorInvalidValue;
    if (ksplit > 1) { p.outs[0] = split_ws; p.out_css[0] = M[0]; p.split_stride = (long long) M[0] * N; }      // nmat == 1
    const int P = sms / 2;
    p.full_units = tiles * ksplit; p.narrow_f = 1;
    static const bool narrow_on = !(getenv("B200_GEMM_NARROW") && atoi(getenv("B200_GEMM_NARROW")) == 0);
    if (narrow_on && ksplit == 1 && bn == 256 && tiles > P) {
        const int rem = tiles % P;
        const int f = rem == 0 ? 1 : (4 * rem <= P ? 4 : (2 * rem <= P ? 2 : 1));
        if (f > 1) { p.full_units = tiles - rem; p.narrow_f = f; }
    }
    const int work = p.full_units + (tiles * ksplit - p.full_units) * p.narrow_f;
    const int pairs = work < P ? work : P;
    static const bool trace = getenv("B200_GEMM_TRACE") && atoi(getenv("B200_GEMM_TRACE")) > 0;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (trace) cudaStreamIsCapturing(st, &cap);                 // the trace allocates and synchronises: not inside a stream capture
    if (trace && cap == cudaStreamCaptureStatusNone && cudaMalloc((void **) &p.trace, (size_t) 2 * pairs * 8 * 8) == cudaSuccess)
        cudaMemsetAsync(p.trace, 0, (size_t) 2 * pairs * 8 * 8, st);
    const cudaError_t e = bn == 256 ? gemm_dense2_launch_t<256>(tmA, tmB, p, 2 * pairs, st) : gemm_dense2_launch_t<128>(tmA, tmB, p, 2 * pairs, st);
    if (p.trace) {      // debugging aid (tools/gemm_bench.py under B200_GEMM_TRACE=1): per-CTA phase times in us since the first CTA started
        std::vector<unsigned long long> h((size_t) 2 * pairs * 8);
        cudaStreamSynchronize(st);
        cudaMemcpy(h.data(), p.trace, h.size() * 8, cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        unsigned long long t0 = ~0ull;
        for (int c = 0; c < 2 * pairs; c++) if (h[c * 8] && h[c * 8] < t0) t0 = h[c * 8];
        static const char * names[8] = {"start", "setup done", "first stage landed", "last MMA issued", "tile 0 complete", "last tile complete", "epilogue done", "end"};
        fprintf(stderr, "[gemm_dense2 trace] nmat=%d M=%d N=%d K=%d bn=%d ksplit=%d pairs=%d work=%d (full %d, narrow x%d)\n", nmat, M[0], N, K, bn, ksplit, pairs, work, p.full_units, p.narrow_f);
        for (int sl = 0; sl < 8; sl++) {
            double mn = 1e30, mx = 0, sum = 0; int n = 0;
            for (int c = 0; c < 2 * pairs; c++) { const unsigned long long v = h[c * 8 + sl]; if (!v) continue; const double u = (double) (v - t0) * 1e-3; mn = u < mn ? u : mn; mx = u > mx ? u : mx; sum += u; n++; }
            if (n) fprintf(stderr, "  %-20s min %7.2f  avg %7.2f  max %7.2f us  (%d CTAs)\n", names[sl], mn, sum / n, mx, n);
        }
    }
    return e;
}

// A: fp16 [M][K], B: fp16 [N][K] (16-B aligned rows: K % 8 == 0); out[n * out_cs + m]
// split_ws: workspace of gemm_dense_ksplit(...) * N * M floats, or nullptr (no split)
cudaError_t launch_gemm_dense(const __half * A, const __half * B, float * out, long long out_cs, int M, int N, int K, int sms, float * split_ws, cudaStream_t st,
                              const float * res_src, float * res_dst, void * act_out) {
    const int ksplit = split_ws ? gemm_dense_ksplit(M, N, K, sms) : 1;
    if (res_dst && ksplit == 1) return cudaErrorInvalidValue;
    cudaError_t e;
    if (gemm_dense_pair()) {
        e = launch_pair(&A, &out, &out_cs, &M, 1, B, N, K, sms, ksplit, split_ws, st);
    } else {
        const int bn = gemm_dense_bn(M, N, sms);
        CUtensorMap tmA, tmB;
        if (!make_map_operand(&tmA, A, M, K, 2, GM_BM) || !make_map_operand(&tmB, B, N, K, 2, bn)) return cudaErrorInvalidValue;
        GemmDenseArgs p{};
        p.out = out; p.out_cs = out_cs; p.M = M; p.N = N; p.K = K;
        p.mtiles = (M + GM_BM - 1) / GM_BM; p.ntiles = (N + bn - 1) / bn;
        p.ksplit = ksplit;
        if (p.ksplit > 1) { p.out = split_ws; p.out_cs = M; p.split_stride = (long long) M * N; }
        const int work = p.mtiles * p.ntiles * p.ksplit;
        const int grid = work < sms ? work : sms;
        e = bn == 256 ? gemm_dense_launch_t<256>(tmA, tmB, p, grid, st) : gemm_dense_launch_t<128>(tmA, tmB, p, grid, st);
    }
    if (e != cudaSuccess || ksplit == 1) return e;
    return launch_reduce_splits(split_ws, ksplit, (long long) M * N, M, out, out_cs, 0, st, res_src, res_dst, act_out);
}

// Several weight matrices over the same activations in ONE launch of the pair kernel (no split-K): 2 <= nmat <= 3.
// Returns cudaErrorNotSupported when the 1-CTA kernel is selected (the caller then launches the matrices one by one).
cudaError_t launch_gemm_dense_group(const __half * const * A, float * const * out, const long long * out_cs, const int * M, int nmat,
                                    const __half * B, int N, int K, int sms, cudaStream_t st) {
    if (!gemm_dense_pair() || nmat < 1 || nmat > GD_MAXMAT) return cudaErrorNotSupported;
    return launch_pair(A, out, out_cs, M, nmat, B, N, K, sms, 1, nullptr, st);
}

}  // namespace b200

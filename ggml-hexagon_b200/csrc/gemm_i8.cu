// gemm_i8.cu — prefill MUL_MAT (many activation columns) on the 5th-generation tensor cores: int8 x int8 -> int32
// tcgen05.mma with TMEM accumulators, operands staged by TMA, exact per-32-block integer sums scaled in fp32.
//
// Takes the seat of the reference's blocked mul_mat loop for src1 with many columns
// (ggml_compute_forward_mul_mat_one_chunk, ggml-dsp.c:1091-1189 / ggml-cpu.c:6655-6743; vec_dot kernels
// ggml-cpu-quants.c:4004-4012 (q8_0), :2591-2604 (q4_0)).
//
// Math (identical to the CPU reference block by block):
//     dst[m, n] = sum_b  (d_w[m, b] * d_a[n, b]) * sumi_b[m, n],      sumi_b = sum_{k in block b} w_q[m, k] * a_q[n, k]   (int32, exact)
// Block scales change every 32 k, so an int32 accumulator cannot run over the whole K: ONE tcgen05.mma (M128 x N128 x K32)
// produces the 128 x 128 exact sums of one block into one of four TMEM accumulator buffers, and the epilogue warps
// fold it into fp32 registers while the tensor core already works on the next blocks.
//
// CTA (1 per SM, persistent over 128 x 128 output tiles), 18 warps:
//   warp 0      TMA producer: A = weights int8 [M][K] (the SoA quant array of q8_0 as it lies in HBM, or the unpacked
//               scratch of q4_0), B = activations int8 [N][K]; 128 x 128-byte boxes, 128-B swizzle, 6-stage ring
//   warp 1      TMEM allocation + MMA issue (one elected lane): 4 MMAs per stage, one per block, each into its own buffer
//   warps 2..17 epilogue: tcgen05.ld (32 lanes x 32 columns per warp), acc += float(sumi) * (d_w * d_a) in packed
//               fp32x2 math, final store
#include "tc_common.cuh"

#include <stdio.h>

namespace b200 {

// KIND selects the tensor-core data path of one exact block sum:
//   0  kind::i8  : int8 operands, s32 sums; one M128 x N128 x K32 MMA per 32-element block; the epilogue converts s32 -> f32
//   1  kind::f16 : the SAME integers stored as fp16 (|q| <= 127 is exact), f32 sums (every partial sum is an integer below 2^24,
//                  so the f32 accumulation is exact too); two K16 MMAs per 32-element block; no conversion in the epilogue
//   2  kind::f16 with 16-element blocks (q6_K: one scale per 16 weights): one K16 MMA per block, no zero padding
// Measured on B200 (4096 x 4096 x 512, q8_0): kind 0 is bound by the epilogue's issue slots (60.8 us; 32 us
// with the math removed = the TMA/L2 feed of 128 MB at 4 TB/s); kind 1 drops the 32 conversions per block and warp but doubles
// the operand bytes and lands on the TMA/L2 feed instead (60.9 us + 12 us more preparation).  So kind 0 is the default for the
// 32-element-block types and kind 2 for q6_K, where it saves the zero padding of kind 0 (119 vs 140 us).
__host__ __device__ constexpr int gm_bps(int kind) { return kind == 1 ? 2 : 4; }      // blocks per stage
__host__ __device__ constexpr int gm_kpb(int kind) { return kind == 2 ? 16 : 32; }    // K elements per block
__host__ __device__ constexpr int gm_mpb(int kind) { return kind == 1 ? 2 : 1; }      // MMA instructions per block
constexpr int GM_STAGES = 6;
constexpr int GM_NBUF = 4;                                  // TMEM accumulator buffers of GM_BN columns
constexpr int GM_EPI_WARPS = 16;                           // 4 TMEM lane quarters x 4 column groups of 32
constexpr int GM_THREADS = 64 + GM_EPI_WARPS * 32;
constexpr int GM_STAGE_A = GM_BM * GM_BK, GM_STAGE_B = GM_BN * GM_BK;
constexpr int GM_STAGE_S1 = 4 * GM_BN * 4;                  // activation block scales of a stage: 4 blocks x 128 columns f32
constexpr int GM_OFF_A = 0, GM_OFF_B = GM_STAGES * GM_STAGE_A, GM_OFF_S = GM_OFF_B + GM_STAGES * GM_STAGE_B;
constexpr int GM_STAGE_S = GM_STAGE_S1;
constexpr int GM_OFF_BAR = GM_OFF_S + GM_STAGES * GM_STAGE_S;
constexpr int GM_SMEM = GM_OFF_BAR + 256 + 1024;            // + barriers + slack to align the base to 1024 B
static_assert(GM_SMEM <= 232448, "shared memory");

struct GemmI8Args {
    const void *  dA; long long dA_stride;      // weight block scales [M][dA_stride]: fp16 (WS = 0) or f32 (WS = 1)
    int acc_init;                               // 1: dst already holds a partial result (q4_K block-minimum term) to add to
    float * out; long long out_cs;              // dst[n * out_cs + m]
    int M, N, K;
    int mtiles, ntiles;
    int ksplit; long long split_stride;        // split-K: work item = (tile, split); split sp accumulates its K range into out + sp * split_stride
};


// ------------------------------------------------------------------------------------------------ kernel
// WS: weight block scales are fp16 (0: q8_0 / q4_0, read where they lie in HBM) or f32 (1: products d * sc of the K-quants)
template <int WS, int KIND>
__global__ void __launch_bounds__(GM_THREADS, 1)
gemm_i8_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmS,
               const GemmI8Args p) {
    constexpr int BPS = gm_bps(KIND), KPB = gm_kpb(KIND), MPB = gm_mpb(KIND), KSTAGE = BPS * KPB;      // KSTAGE: K elements per stage
    constexpr uint32_t TX_S = BPS * GM_BN * 4u;                                   // scale bytes per stage
    extern __shared__ uint8_t gm_smem_raw[];
    const uint32_t base = (gm_smem_u32(gm_smem_raw) + 1023u) & ~1023u;           // 128-B swizzle atoms want 1024-B alignment
    uint8_t * smem = gm_smem_raw + (base - gm_smem_u32(gm_smem_raw));
    const uint32_t sA = base + GM_OFF_A, sB = base + GM_OFF_B, sS = base + GM_OFF_S, bars = base + GM_OFF_BAR;
    // full: TMA landed (A, B, scales); empty: the stage's MMAs are done with A/B; sempty: the epilogue is done with the scales
    const uint32_t full0 = bars, empty0 = bars + GM_STAGES * 8, sempty0 = bars + 2 * GM_STAGES * 8;
    const uint32_t tfull0 = bars + 3 * GM_STAGES * 8, tempty0 = tfull0 + GM_NBUF * 8;
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (smem + GM_OFF_BAR + 3 * GM_STAGES * 8 + 2 * GM_NBUF * 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nks = (p.K + KSTAGE - 1) / KSTAGE, KB = p.K / KPB, nwork = p.mtiles * p.ntiles * p.ksplit;

    if (threadIdx.x == 0) {
        for (int i = 0; i < GM_STAGES; i++) { gm_mbar_init(full0 + i * 8, 1); gm_mbar_init(empty0 + i * 8, 1); gm_mbar_init(sempty0 + i * 8, GM_EPI_WARPS); }
        for (int i = 0; i < GM_NBUF; i++) { gm_mbar_init(tfull0 + i * 8, 1); gm_mbar_init(tempty0 + i * 8, GM_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {            // all 512 TMEM columns: four 128-column int32 accumulator buffers
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(gm_smem_u32((const void *) tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    gm_fence_before();
    __syncthreads();
    gm_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmA) : "memory");
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmB) : "memory");
            asm volatile("prefetch.tensormap [%0];" :: "l"(&tmS) : "memory");
            uint32_t it = 0;
            for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
                const int tile = work / p.ksplit, sp = work - tile * p.ksplit;
                const int mt = tile / p.ntiles, nt = tile - mt * p.ntiles;
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % GM_STAGES, ph = (it / GM_STAGES) & 1;
                    gm_mbar_wait(empty0 + s * 8, ph ^ 1);
                    gm_mbar_wait(sempty0 + s * 8, ph ^ 1);
                    gm_mbar_expect_tx(full0 + s * 8, GM_STAGE_A + GM_STAGE_B + TX_S);
                    gm_tma_2d(sA + s * GM_STAGE_A, &tmA, full0 + s * 8, ks * KSTAGE, mt * GM_BM);
                    gm_tma_2d(sB + s * GM_STAGE_B, &tmB, full0 + s * 8, ks * KSTAGE, nt * GM_BN);
                    gm_tma_2d(sS + s * GM_STAGE_S, &tmS, full0 + s * 8, nt * GM_BN, ks * BPS);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one lane, one tcgen05.mma per 32-element block, each into its own accumulator buffer
        if (lane == 0) {
            // instruction descriptor: D format (bits 4-5: 1 = f32, 2 = s32), A / B format (bits 7-9 / 10-12: kind::i8 1 = s8,
            // kind::f16 0 = fp16), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
            const uint32_t idesc = (KIND == 0 ? (2u << 4) | (1u << 7) | (1u << 10) : (1u << 4)) |
                                   ((uint32_t) (GM_BN >> 3) << 17) | ((uint32_t) (GM_BM >> 4) << 24);
            uint32_t it = 0, bc = 0;
            for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
                const int sp = work % p.ksplit;
                const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
                for (int ks = ks0; ks < ks1; ks++, it++) {
                    const uint32_t s = it % GM_STAGES, ph = (it / GM_STAGES) & 1;
                    gm_mbar_wait(full0 + s * 8, ph);
                    gm_fence_after();
                    const int nb = min(BPS, KB - ks * BPS);
                    for (int j = 0; j < nb; j++, bc++) {
                        const uint32_t buf = bc % GM_NBUF, bph = (bc / GM_NBUF) & 1;
                        gm_mbar_wait(tempty0 + buf * 8, bph ^ 1);
                        gm_fence_after();
                        #pragma unroll
                        for (int mm = 0; mm < MPB; mm++) {        // every MMA consumes 32 bytes of K per row (32 int8 or 16 fp16)
                            const uint32_t off = (uint32_t) (j * MPB + mm) * 32u;
                            const uint64_t ad = gm_smem_desc(sA + s * GM_STAGE_A + off), bd = gm_smem_desc(sB + s * GM_STAGE_B + off);
                            if (KIND == 0) gm_mma_i8(tmem + buf * GM_BN, ad, bd, idesc, 0u);
                            else           gm_mma_f16(tmem + buf * GM_BN, ad, bd, idesc, mm == 0 ? 0u : 1u);
                        }
                        gm_commit(tfull0 + buf * 8);          // arrives when this block's sums are in TMEM
                    }
                    gm_commit(empty0 + s * 8);                // arrives when the stage's MMAs have read shared memory
                }
            }
        }
    } else {
        // ===== epilogue: warp w reads TMEM lanes 32 * (w % 4) .. +31 (hardware rule), columns 32 * ((w - 2) / 4) .. +31
        const int q = warp & 3, cg = (warp - 2) >> 2;
        const int m_local = q * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t) (q * 32) << 16) + cg * 32;
        uint32_t bc = 0, it = 0;
        for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
            const int tile = work / p.ksplit, sp = work - tile * p.ksplit;
            const int mt = tile / p.ntiles, nt = tile - mt * p.ntiles;
            const int ks0 = (int) ((long long) nks * sp / p.ksplit), ks1 = (int) ((long long) nks * (sp + 1) / p.ksplit);
            float * const outp = p.out + (long long) sp * p.split_stride;
            const int m = mt * GM_BM + m_local;
            const long long row_off = (long long) min(m, p.M - 1) * p.dA_stride;
            float2 acc[16];
            #pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = make_float2(0.f, 0.f);
            if (p.acc_init && m < p.M) {                                    // start from what dst holds (coalesced: 32 consecutive m)
                const int n0 = nt * GM_BN + cg * 32;
                #pragma unroll
                for (int i = 0; i < 16; i++) {
                    if (n0 + 2 * i     < p.N) acc[i].x = outp[(long long) (n0 + 2 * i)     * p.out_cs + m];
                    if (n0 + 2 * i + 1 < p.N) acc[i].y = outp[(long long) (n0 + 2 * i + 1) * p.out_cs + m];
                }
            }
            // weight scales of this row: the BPS blocks of one stage per load, fetched one stage ahead.  WS == 0 keeps them as
            // the packed fp16 they are in HBM until use (registers are tight).
            constexpr int DWB = BPS * (WS == 0 ? 2 : 4);                                  // bytes of weight scales per stage and row
            const bool dw_vec = (p.dA_stride % BPS) == 0 && ((uintptr_t) p.dA % DWB) == 0;     // else (ragged K): scalar loads
            auto load_dw = [&](int ks) -> float4 {
                float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
                if (WS == 0) {
                    const __half * dw_row = (const __half *) p.dA + row_off + ks * BPS;
                    uint32_t u[2] = {0u, 0u};
                    if (dw_vec) {
                        if (BPS == 4) { const uint2 t = *(const uint2 *) dw_row; u[0] = t.x; u[1] = t.y; }
                        else          u[0] = *(const uint32_t *) dw_row;
                    } else {
                        unsigned short h[4] = {0, 0, 0, 0};
                        for (int j = 0; j < BPS && ks * BPS + j < KB; j++) h[j] = *(const unsigned short *) (dw_row + j);
                        u[0] = (uint32_t) h[0] | ((uint32_t) h[1] << 16); u[1] = (uint32_t) h[2] | ((uint32_t) h[3] << 16);
                    }
                    r.x = __uint_as_float(u[0]); r.y = __uint_as_float(u[1]);
                } else {
                    const float * dw_row = (const float *) p.dA + row_off + ks * BPS;
                    if (dw_vec) {
                        if (BPS == 4) r = *(const float4 *) dw_row;
                        else { const float2 t = *(const float2 *) dw_row; r.x = t.x; r.y = t.y; }
                    } else {
                        float f[4] = {0.f, 0.f, 0.f, 0.f};
                        for (int j = 0; j < BPS && ks * BPS + j < KB; j++) f[j] = dw_row[j];
                        r = make_float4(f[0], f[1], f[2], f[3]);
                    }
                }
                return r;
            };
            auto cvt_dw = [&](float4 r) -> float4 {
                if (WS != 0) return r;
                const uint32_t a = __float_as_uint(r.x), b = __float_as_uint(r.y);
                const __half2 h01 = *(const __half2 *) &a, h23 = *(const __half2 *) &b;
                return make_float4(__low2float(h01), __high2float(h01), __low2float(h23), __high2float(h23));
            };
            float4 dwn = load_dw(ks0);
            for (int ks = ks0; ks < ks1; ks++, it++) {
                const uint32_t s = it % GM_STAGES, ph = (it / GM_STAGES) & 1;
                const float4 dwc = cvt_dw(dwn);
                if (ks + 1 < ks1) dwn = load_dw(ks + 1);
                const float dws[4] = {dwc.x, dwc.y, dwc.z, dwc.w};
                gm_mbar_wait(full0 + s * 8, ph);                            // scales of this stage are in shared memory
                const float4 * sc = (const float4 *) (smem + GM_OFF_S + s * GM_STAGE_S) + cg * 8;
                const int nb = min(BPS, KB - ks * BPS);
                // acc += float(sumi) * (d_w * d_a), as ggml_vec_dot_q8_0_q8_0 does per block (ggml-cpu-quants.c:4004-4012);
                // |sumi| <= 32 * 127 * 127 < 2^24: exact as f32 either way.
                #pragma unroll
                for (int j = 0; j < BPS; j++) {
                    if (j < nb) {
                        const uint32_t buf = bc % GM_NBUF, bph = (bc / GM_NBUF) & 1;
                        gm_mbar_wait(tfull0 + buf * 8, bph);
                        gm_fence_after();
                        int v[32];
                        gm_tmem_ld32(tbase + buf * GM_BN, v);
                        gm_tmem_ld_wait();
                        gm_fence_before();
                        __syncwarp();
                        if (lane == 0) gm_mbar_arrive(tempty0 + buf * 8);  // buffer free for the MMA issuer
                        bc++;
                        const float2 dw2 = make_float2(dws[j], dws[j]);
                        #pragma unroll
                        for (int i = 0; i < 8; i++) {
                            const float4 d = sc[j * (GM_BN / 4) + i];     // activation scales of block j, columns 4i..4i+3 of this group
                            // int -> float on the conversion unit: one issue slot per element.  (The magic-number route, as_float(
                            // 0x4B400000 + sumi) - 12582912.f, costs 1.5 slots and measured 15 % slower: the loop is issue-bound.)
                            // kind::f16 delivers the sums as f32 already
                            const float2 f0 = KIND == 0 ? make_float2(__int2float_rn(v[4 * i + 0]), __int2float_rn(v[4 * i + 1]))
                                                        : make_float2(__int_as_float(v[4 * i + 0]), __int_as_float(v[4 * i + 1]));
                            const float2 f1 = KIND == 0 ? make_float2(__int2float_rn(v[4 * i + 2]), __int2float_rn(v[4 * i + 3]))
                                                        : make_float2(__int_as_float(v[4 * i + 2]), __int_as_float(v[4 * i + 3]));
                            acc[2 * i + 0] = __ffma2_rn(f0, __fmul2_rn(dw2, make_float2(d.x, d.y)), acc[2 * i + 0]);
                            acc[2 * i + 1] = __ffma2_rn(f1, __fmul2_rn(dw2, make_float2(d.z, d.w)), acc[2 * i + 1]);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) gm_mbar_arrive(sempty0 + s * 8);            // scale tile of this stage may be overwritten
            }
            if (m < p.M) {
                const int n0 = nt * GM_BN + cg * 32;
                #pragma unroll
                for (int i = 0; i < 16; i++) {                              // 32 lanes = 32 consecutive m: coalesced
                    if (n0 + 2 * i     < p.N) outp[(long long) (n0 + 2 * i)     * p.out_cs + m] = acc[i].x;
                    if (n0 + 2 * i + 1 < p.N) outp[(long long) (n0 + 2 * i + 1) * p.out_cs + m] = acc[i].y;
                }
            }
        }
    }
    gm_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

// ------------------------------------------------------------------------------------------------ operand preparation
// Activations for the GEMM: q8_1 quantisation (= quantize_row_q8_1_ref, ggml-quants.c:220-253) into a plain row-major
// int8 matrix [N][K] (TMA operand B) and block-major scales dB[K/32][ldb] (f32 value of the fp16-rounded d).
// 8 lanes per 32-element block, one float4 each.
// MODE 0: as above.  MODE 1 (q4_K weights): also sB[b][row] = d * isum (isum = exact sum of the block's quants), stored
// KB * ldb floats after dB.  MODE 2 (q6_K weights, scales per 16 elements): every 16-element half block becomes its own
// 32-byte MMA block [16 quants, 16 zeros] in q [N][2K], dB [K/16][ldb] -- the weights are laid out alike (unpack_q6_K).
template <int MODE>
__global__ void __launch_bounds__(256)
quantize_rows_i8_kernel(const char * __restrict__ x, const long long row_stride, const int K, const int N,
                        int8_t * __restrict__ q, float * __restrict__ dB, const long long ldb) {
    const int NB = K / 32;
    const long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    const long long blk = gid >> 3;
    if (blk >= (long long) NB * N) return;
    const int row = (int) (blk / NB), b = (int) (blk - (long long) row * NB), sub = (int) (gid & 7);
    const float4 v = *(const float4 *) (x + row * row_stride + ((long long) b * 32 + sub * 4) * 4);
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int) roundf(__fmul_rn(v.x, id)), q1 = (int) roundf(__fmul_rn(v.y, id));
    const int q2 = (int) roundf(__fmul_rn(v.z, id)), q3 = (int) roundf(__fmul_rn(v.w, id));
    const uint32_t packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
    const float dh = __half2float(__float2half_rn(d));
    if (MODE == 2) {
        const int g = 2 * b + (sub >> 2);
        uint32_t * o = (uint32_t *) (q + (long long) row * 2 * K + (long long) g * 32 + (sub & 3) * 4);
        o[0] = packed; o[4] = 0u;
        if ((sub & 3) == 0) dB[(long long) g * ldb + row] = dh;
    } else {
        *(uint32_t *) (q + (long long) row * K + b * 32 + sub * 4) = packed;
        if (MODE == 1) {
            int isum = q0 + q1 + q2 + q3;
            isum += __shfl_xor_sync(0xffffffffu, isum, 1);
            isum += __shfl_xor_sync(0xffffffffu, isum, 2);
            isum += __shfl_xor_sync(0xffffffffu, isum, 4);
            if (sub == 0) dB[((long long) NB + b) * ldb + row] = __fmul_rn(dh, (float) isum);
        }
        if (sub == 0) dB[(long long) b * ldb + row] = dh;
    }
}

// q4_0 nibbles (SoA quant array, 16 B per block: low nibbles = elements 0..15, high = 16..31, ggml-common.h:167-172)
// -> signed int8 (q - 8) row-major [M][K], the TMA operand A.  One thread per block.
__global__ void __launch_bounds__(256)
unpack_q4_0_i8_kernel(const uint4 * __restrict__ qs, const long long nblocks, int8_t * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nblocks; i += (long long) gridDim.x * blockDim.x) {
        const uint4 v = qs[i];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t lo[4], hi[4];
        #pragma unroll
        for (int j = 0; j < 4; j++) {
            // per-byte (nibble - 8): nibbles are 0..15, so the subtraction never borrows across bytes after the +0x80 bias trick
            lo[j] = __vsub4(w[j] & 0x0F0F0F0Fu, 0x08080808u);
            hi[j] = __vsub4((w[j] >> 4) & 0x0F0F0F0Fu, 0x08080808u);
        }
        uint4 * o = (uint4 *) (out + i * 32);
        o[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        o[1] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    }
}

// q4_K (SoA: qs 128 B + header {f16 d, f16 dmin, 12 B packed 6-bit scales/mins} per 256-element super-block,
// ggml-common.h:285-296; nibble order and get_scale_min_k4 as dequantize_row_q4_K, ggml-quants.c:1352-1374)
// -> A int8 [M][K] = the raw 4-bit quants 0..15, wS f32 = d * sc_g, wM f32 = -(dmin * m_g) per 32-element sub-block g.
// One thread per 64 elements (one 32-byte run: low nibbles = sub-block 2j, high nibbles = sub-block 2j + 1).
__global__ void __launch_bounds__(256)
unpack_q4_K_kernel(const uint4 * __restrict__ qs, const uint4 * __restrict__ hdr, const long long nsuper,
                   int8_t * __restrict__ out, float * __restrict__ wS, float * __restrict__ wM) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 4; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 2; const int j = (int) (i & 3);
        const uint4 v0 = qs[sb * 8 + 2 * j], v1 = qs[sb * 8 + 2 * j + 1];
        uint4 * o = (uint4 *) (out + sb * 256 + j * 64);
        o[0] = make_uint4(v0.x & 0x0F0F0F0Fu, v0.y & 0x0F0F0F0Fu, v0.z & 0x0F0F0F0Fu, v0.w & 0x0F0F0F0Fu);
        o[1] = make_uint4(v1.x & 0x0F0F0F0Fu, v1.y & 0x0F0F0F0Fu, v1.z & 0x0F0F0F0Fu, v1.w & 0x0F0F0F0Fu);
        o[2] = make_uint4((v0.x >> 4) & 0x0F0F0F0Fu, (v0.y >> 4) & 0x0F0F0F0Fu, (v0.z >> 4) & 0x0F0F0F0Fu, (v0.w >> 4) & 0x0F0F0F0Fu);
        o[3] = make_uint4((v1.x >> 4) & 0x0F0F0F0Fu, (v1.y >> 4) & 0x0F0F0F0Fu, (v1.z >> 4) & 0x0F0F0F0Fu, (v1.w >> 4) & 0x0F0F0F0Fu);
        const uint4 h = hdr[sb];
        const __half2 dd = *(const __half2 *) &h.x;
        const float d = __low2float(dd), dmin = __high2float(dd);
        const uint32_t w[3] = {h.y, h.z, h.w};
        auto byte = [&](int k) -> uint32_t { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; };
        #pragma unroll
        for (int e = 0; e < 2; e++) {
            const int g = 2 * j + e;
            uint32_t sc, mn;
            if (g < 4) { sc = byte(g) & 63u; mn = byte(g + 4) & 63u; }
            else       { sc = (byte(g + 4) & 0xFu) | ((byte(g - 4) >> 6) << 4); mn = (byte(g + 4) >> 4) | ((byte(g) >> 6) << 4); }
            wS[sb * 8 + g] = __fmul_rn(d, (float) sc);
            wM[sb * 8 + g] = -__fmul_rn(dmin, (float) mn);
        }
    }
}

// q6_K (SoA: ql 128 B, qh 64 B, 16 int8 scales, f16 d per super-block, ggml-common.h:320-326; bit layout as
// dequantize_row_q6_K, ggml-quants.c:1681-1710) -> A int8 [M][2K]: each 16-element scale group g becomes one 32-byte MMA block
// [16 x (q - 32), 16 x 0]; wS f32 [M][K/16] = d * sc_g.  16 threads per super-block: thread t handles l = 4 (t & 7) .. +3
// of half (t >> 3) (4 quants for each of the 4 rows of the bit layout) and writes scale t.
__global__ void __launch_bounds__(256)
unpack_q6_K_kernel(const uint32_t * __restrict__ ql, const uint32_t * __restrict__ qh, const int8_t * __restrict__ sc,
                   const __half * __restrict__ dd, const long long nsuper, int8_t * __restrict__ out, float * __restrict__ wS) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 16; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 4; const int t = (int) (i & 15), half = t >> 3, l4 = t & 7;
        const uint32_t a = ql[sb * 32 + half * 16 + l4], b = ql[sb * 32 + half * 16 + 8 + l4], h = qh[sb * 16 + half * 8 + l4];
        const uint32_t v[4] = {
            (a & 0x0F0F0F0Fu)        | ((h << 4) & 0x30303030u),
            (b & 0x0F0F0F0Fu)        | ((h << 2) & 0x30303030u),
            ((a >> 4) & 0x0F0F0F0Fu) | ( h       & 0x30303030u),
            ((b >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u)};
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            const int e = half * 128 + r * 32 + l4 * 4;          // element index inside the super-block
            const int g = e >> 4;
            uint32_t * o = (uint32_t *) (out + (sb * 16 + g) * 32 + (e & 15));
            o[0] = __vsub4(v[r], 0x20202020u);
            o[4] = 0u;
        }
        wS[sb * 16 + t] = __fmul_rn(__half2float(dd[sb]), (float) sc[sb * 16 + t]);
    }
}

// ---- fp16 operand preparation (KIND 1 / 2): the same integers, stored as fp16
// 4 unsigned bytes -> 4 halfs (value - bias): 0x6400 | b is the half 1024 + b (ulp 1 in [1024, 2048)), then one packed subtract
__device__ __forceinline__ uint2 gm_h4(uint32_t w, float bias) {
    const uint32_t lo = __byte_perm(w, 0x64646464u, 0x4140), hi = __byte_perm(w, 0x64646464u, 0x4342);
    const __half2 off = __float2half2_rn(1024.f + bias);
    const __half2 a = __hsub2(*(const __half2 *) &lo, off), b = __hsub2(*(const __half2 *) &hi, off);
    return make_uint2(*(const uint32_t *) &a, *(const uint32_t *) &b);
}
__device__ __forceinline__ void gm_store_h16(__half * out, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, float bias) {
    const uint2 a = gm_h4(w0, bias), b = gm_h4(w1, bias), c = gm_h4(w2, bias), d = gm_h4(w3, bias);
    ((uint4 *) out)[0] = make_uint4(a.x, a.y, b.x, b.y);
    ((uint4 *) out)[1] = make_uint4(c.x, c.y, d.x, d.y);
}
// q8_0 quants (signed bytes; b ^ 0x80 = b + 128 as unsigned): one thread per 16 elements
__global__ void __launch_bounds__(256)
unpack_q8_0_f16_kernel(const uint4 * __restrict__ qs, const long long n16, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (long long) gridDim.x * blockDim.x) {
        const uint4 v = qs[i];
        gm_store_h16(out + i * 16, v.x ^ 0x80808080u, v.y ^ 0x80808080u, v.z ^ 0x80808080u, v.w ^ 0x80808080u, 128.f);
    }
}
__global__ void __launch_bounds__(256)
unpack_q4_0_f16_kernel(const uint4 * __restrict__ qs, const long long nblocks, __half * __restrict__ out) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nblocks; i += (long long) gridDim.x * blockDim.x) {
        const uint4 v = qs[i];
        const uint32_t m = 0x0F0F0F0Fu;
        gm_store_h16(out + i * 32,      v.x & m, v.y & m, v.z & m, v.w & m, 8.f);
        gm_store_h16(out + i * 32 + 16, (v.x >> 4) & m, (v.y >> 4) & m, (v.z >> 4) & m, (v.w >> 4) & m, 8.f);
    }
}
__global__ void __launch_bounds__(256)
unpack_q4_K_f16_kernel(const uint4 * __restrict__ qs, const uint4 * __restrict__ hdr, const long long nsuper,
                       __half * __restrict__ out, float * __restrict__ wS, float * __restrict__ wM) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 4; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 2; const int j = (int) (i & 3);
        const uint4 v0 = qs[sb * 8 + 2 * j], v1 = qs[sb * 8 + 2 * j + 1];
        const uint32_t m = 0x0F0F0F0Fu;
        __half * o = out + sb * 256 + j * 64;
        gm_store_h16(o,      v0.x & m, v0.y & m, v0.z & m, v0.w & m, 0.f);
        gm_store_h16(o + 16, v1.x & m, v1.y & m, v1.z & m, v1.w & m, 0.f);
        gm_store_h16(o + 32, (v0.x >> 4) & m, (v0.y >> 4) & m, (v0.z >> 4) & m, (v0.w >> 4) & m, 0.f);
        gm_store_h16(o + 48, (v1.x >> 4) & m, (v1.y >> 4) & m, (v1.z >> 4) & m, (v1.w >> 4) & m, 0.f);
        const uint4 h = hdr[sb];
        const __half2 dd = *(const __half2 *) &h.x;
        const float d = __low2float(dd), dmin = __high2float(dd);
        const uint32_t w[3] = {h.y, h.z, h.w};
        auto byte = [&](int k) -> uint32_t { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; };
        #pragma unroll
        for (int e = 0; e < 2; e++) {
            const int g = 2 * j + e;
            uint32_t sc, mn;
            if (g < 4) { sc = byte(g) & 63u; mn = byte(g + 4) & 63u; }
            else       { sc = (byte(g + 4) & 0xFu) | ((byte(g - 4) >> 6) << 4); mn = (byte(g + 4) >> 4) | ((byte(g) >> 6) << 4); }
            wS[sb * 8 + g] = __fmul_rn(d, (float) sc);
            wM[sb * 8 + g] = -__fmul_rn(dmin, (float) mn);
        }
    }
}
// q6_K without padding: the 16-element scale groups are the MMA blocks of KIND 2
__global__ void __launch_bounds__(256)
unpack_q6_K_f16_kernel(const uint32_t * __restrict__ ql, const uint32_t * __restrict__ qh, const int8_t * __restrict__ sc,
                       const __half * __restrict__ dd, const long long nsuper, __half * __restrict__ out, float * __restrict__ wS) {
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nsuper * 16; i += (long long) gridDim.x * blockDim.x) {
        const long long sb = i >> 4; const int t = (int) (i & 15), half = t >> 3, l4 = t & 7;
        const uint32_t a = ql[sb * 32 + half * 16 + l4], b = ql[sb * 32 + half * 16 + 8 + l4], h = qh[sb * 16 + half * 8 + l4];
        const uint32_t v[4] = {
            (a & 0x0F0F0F0Fu)        | ((h << 4) & 0x30303030u),
            (b & 0x0F0F0F0Fu)        | ((h << 2) & 0x30303030u),
            ((a >> 4) & 0x0F0F0F0Fu) | ( h       & 0x30303030u),
            ((b >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u)};
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            const int e = half * 128 + r * 32 + l4 * 4;
            *(uint2 *) (out + sb * 256 + e) = gm_h4(v[r], 32.f);
        }
        wS[sb * 16 + t] = __fmul_rn(__half2float(dd[sb]), (float) sc[sb * 16 + t]);
    }
}
// activations: q8_1 quants as fp16 [N][K]; MODE 1: + d * isum behind dB; MODE 2: dB per 16-element group ([K/16][ldb])
template <int MODE>
__global__ void __launch_bounds__(256)
quantize_rows_f16_kernel(const char * __restrict__ x, const long long row_stride, const int K, const int N,
                         __half * __restrict__ q, float * __restrict__ dB, const long long ldb) {
    const int NB = K / 32;
    const long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    const long long blk = gid >> 3;
    if (blk >= (long long) NB * N) return;
    const int row = (int) (blk / NB), b = (int) (blk - (long long) row * NB), sub = (int) (gid & 7);
    const float4 v = *(const float4 *) (x + row * row_stride + ((long long) b * 32 + sub * 4) * 4);
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const float r0 = roundf(__fmul_rn(v.x, id)), r1 = roundf(__fmul_rn(v.y, id)), r2 = roundf(__fmul_rn(v.z, id)), r3 = roundf(__fmul_rn(v.w, id));
    const __half2 h01 = __floats2half2_rn(r0, r1), h23 = __floats2half2_rn(r2, r3);
    *(uint2 *) (q + (long long) row * K + b * 32 + sub * 4) = make_uint2(*(const uint32_t *) &h01, *(const uint32_t *) &h23);
    const float dh = __half2float(__float2half_rn(d));
    if (MODE == 2) {
        if ((sub & 3) == 0) dB[(long long) (2 * b + (sub >> 2)) * ldb + row] = dh;
    } else {
        if (MODE == 1) {
            int isum = (int) r0 + (int) r1 + (int) r2 + (int) r3;
            isum += __shfl_xor_sync(0xffffffffu, isum, 1);
            isum += __shfl_xor_sync(0xffffffffu, isum, 2);
            isum += __shfl_xor_sync(0xffffffffu, isum, 4);
            if (sub == 0) dB[((long long) NB + b) * ldb + row] = __fmul_rn(dh, (float) isum);
        }
        if (sub == 0) dB[(long long) b * ldb + row] = dh;
    }
}

cudaError_t launch_quantize_rows_f16(const float * x, long long row_stride_bytes, int K, int N, __half * q, float * dB, long long ldb, int mode, cudaStream_t st) {
    const long long threads = (long long) (K / 4) * N;
    const unsigned grid = (unsigned) ((threads + 255) / 256);
    if (mode == 0)      quantize_rows_f16_kernel<0><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    else if (mode == 1) quantize_rows_f16_kernel<1><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    else                quantize_rows_f16_kernel<2><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    return cudaGetLastError();
}
// weights of `type` (SoA field pointers w, nblocks native blocks) -> fp16 integers [M][K]; K-quants also get wS (and q4_K wM)
cudaError_t launch_unpack_f16(int type, const WPtrs & w, long long nblocks, __half * out, float * wS, float * wM, cudaStream_t st) {
    auto grid_for = [](long long threads) { long long b = (threads + 255) / 256; return (unsigned) (b > 148 * 32 ? 148 * 32 : (b < 1 ? 1 : b)); };
    switch (type) {
        case T_Q8_0: unpack_q8_0_f16_kernel<<<grid_for(nblocks * 2), 256, 0, st>>>((const uint4 *) w.p[0], nblocks * 2, out); break;
        case T_Q4_0: unpack_q4_0_f16_kernel<<<grid_for(nblocks), 256, 0, st>>>((const uint4 *) w.p[0], nblocks, out); break;
        case T_Q4_K: unpack_q4_K_f16_kernel<<<grid_for(nblocks * 4), 256, 0, st>>>((const uint4 *) w.p[0], (const uint4 *) w.p[1], nblocks, out, wS, wM); break;
        case T_Q6_K: unpack_q6_K_f16_kernel<<<grid_for(nblocks * 16), 256, 0, st>>>((const uint32_t *) w.p[0], (const uint32_t *) w.p[1], (const int8_t *) w.p[2],
                                                                                   (const __half *) w.p[3], nblocks, out, wS); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_quantize_rows_i8(const float * x, long long row_stride_bytes, int K, int N, int8_t * q, float * dB, long long ldb, int mode, cudaStream_t st) {
    const long long threads = (long long) (K / 4) * N;
    const unsigned grid = (unsigned) ((threads + 255) / 256);
    if (mode == 0)      quantize_rows_i8_kernel<0><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    else if (mode == 1) quantize_rows_i8_kernel<1><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    else                quantize_rows_i8_kernel<2><<<grid, 256, 0, st>>>((const char *) x, row_stride_bytes, K, N, q, dB, ldb);
    return cudaGetLastError();
}
cudaError_t launch_unpack_q4_0_i8(const void * qs, long long nblocks, int8_t * out, cudaStream_t st) {
    long long blocks = (nblocks + 255) / 256; if (blocks > 148 * 32) blocks = 148 * 32;
    unpack_q4_0_i8_kernel<<<(unsigned) blocks, 256, 0, st>>>((const uint4 *) qs, nblocks, out);
    return cudaGetLastError();
}
cudaError_t launch_unpack_q4_K_i8(const WPtrs & w, long long nsuper, int8_t * out, float * wS, float * wM, cudaStream_t st) {
    long long blocks = (nsuper * 4 + 255) / 256; if (blocks > 148 * 32) blocks = 148 * 32;
    unpack_q4_K_kernel<<<(unsigned) blocks, 256, 0, st>>>((const uint4 *) w.p[0], (const uint4 *) w.p[1], nsuper, out, wS, wM);
    return cudaGetLastError();
}
cudaError_t launch_unpack_q6_K_i8(const WPtrs & w, long long nsuper, int8_t * out, float * wS, cudaStream_t st) {
    long long blocks = (nsuper * 16 + 255) / 256; if (blocks > 148 * 32) blocks = 148 * 32;
    unpack_q6_K_kernel<<<(unsigned) blocks, 256, 0, st>>>((const uint32_t *) w.p[0], (const uint32_t *) w.p[1], (const int8_t *) w.p[2],
                                                           (const __half *) w.p[3], nsuper, out, wS);
    return cudaGetLastError();
}


// activation block scales f32 [KB][ldb]: box = 128 columns x (blocks per stage), no swizzle
static bool make_map_scales(CUtensorMap * m, const void * base, long long KB, long long ldb, int bps) {
    encode_tiled_t enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t) ldb, (cuuint64_t) KB};
    const cuuint64_t strides[1] = {(cuuint64_t) ldb * 4};
    const cuuint32_t box[2] = {(cuuint32_t) GM_BN, (cuuint32_t) bps};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int gemm_i8_ldb(int N) { return (N + GM_BN - 1) / GM_BN * GM_BN; }      // leading dimension the caller must give dB

// ------------------------------------------------------------------------------------------------ q4_K block-minimum term
// dst[n * cs + m] = sum_g mA[m][g] * sB[g][n]    (mA = -(dmin * m_g), sB = d_a * isum_a; ggml-cpu-quants.c:7535-7590 folds the
// same products into its per-super-block sum).  A plain f32 product with a short contraction (K / 32 terms, 1/32 of the main
// GEMM's multiply-adds) on the CUDA cores: 64 x 64 tile per CTA, 4 x 4 per thread; the tensor-core kernel then starts its
// accumulators from dst.  (Folding the term into the tensor-core epilogue, one more packed FFMA and a second scale tile per
// block, measured 121 us against 68 us for the 4096 x 4096 x 512 kernel: that loop has no slack.)
__global__ void __launch_bounds__(256)
min_term_kernel(const float * __restrict__ mA, const float * __restrict__ sB, const long long ldb, float * __restrict__ out,
                const long long out_cs, const int M, const int N, const int KB) {
    __shared__ float As[32][64 + 4];       // [g][m]
    __shared__ float Bs[32][64];           // [g][n]
    const int m0 = blockIdx.x * 64, n0 = blockIdx.y * 64, tid = threadIdx.x;
    const int tm = (tid & 15) * 4, tn = (tid >> 4) * 4;
    float acc[4][4];                        // [n][m]
    #pragma unroll
    for (int i = 0; i < 4; i++)
        #pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    for (int g0 = 0; g0 < KB; g0 += 32) {
        // mA tile: 64 rows x 32 g, read along g (coalesced), stored transposed
        #pragma unroll
        for (int r = 0; r < 8; r++) {
            const int row = (tid >> 5) + r * 8, g = tid & 31;
            const int m = m0 + row;
            As[g][row] = (m < M && g0 + g < KB) ? mA[(long long) m * KB + g0 + g] : 0.f;
        }
        #pragma unroll
        for (int r = 0; r < 8; r++) {
            const int g = (tid >> 6) + r * 4, n = tid & 63;
            Bs[g][n] = (g0 + g < KB && n0 + n < N) ? sB[(long long) (g0 + g) * ldb + n0 + n] : 0.f;
        }
        __syncthreads();
        #pragma unroll 8
        for (int g = 0; g < 32; g++) {
            const float4 a = *(const float4 *) &As[g][tm], b = *(const float4 *) &Bs[g][tn];
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
            #pragma unroll
            for (int i = 0; i < 4; i++)
                #pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(av[j], bv[i], acc[i][j]);
        }
        __syncthreads();
    }
    #pragma unroll
    for (int i = 0; i < 4; i++) {
        const int n = n0 + tn + i;
        if (n >= N) continue;
        #pragma unroll
        for (int j = 0; j < 4; j++) {
            const int m = m0 + tm + j;
            if (m < M) out[(long long) n * out_cs + m] = acc[i][j];
        }
    }
}

// A: [M][K] int8 (kind 0) or fp16 (kind 1, 2), 16-B aligned rows; dA_h: fp16 [M][KB] or dA_f: f32 [M][KB] (exactly one), KB = K / 32
// (kind 2: K / 16); mA: f32 [M][KB] or nullptr (q4_K: the block-minimum term, with d * isum in the [KB][ldb] floats behind dB);
// B: [N][K] like A; dB: f32 [KB][ldb]; out[n * out_cs + m].
// K is the MMA-side contraction length (kind 0 with q6_K: 2 x the tensor's, zero padded).
template <int WS, int KIND>
static cudaError_t gemm_i8_launch_t(const CUtensorMap & tmA, const CUtensorMap & tmB, const CUtensorMap & tmS,
                                    const GemmI8Args & p, int grid, cudaStream_t st) {
    static bool attr_done[64] = {false};          // per device
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    if (!attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_i8_kernel<WS, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_SMEM);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    gemm_i8_kernel<WS, KIND><<<(unsigned) grid, GM_THREADS, GM_SMEM, st>>>(tmA, tmB, tmS, p);
    return cudaGetLastError();
}

// split-K for shapes with few output tiles (e.g. the 1024-row k / v projections of grouped-query models: 32 tiles for 148 SMs):
// every (tile, split) is a work item, the partial results go to a workspace [S][N][M] and a second kernel adds them up in split
// order (deterministic).
int gemm_i8_ksplit(int M, int N, int K, int kind, int sms) {
    const int tiles = ((M + GM_BM - 1) / GM_BM) * ((N + GM_BN - 1) / GM_BN);
    const int nks = (K + gm_bps(kind) * gm_kpb(kind) - 1) / (gm_bps(kind) * gm_kpb(kind));
    int s = sms / (tiles > 0 ? tiles : 1);
    if (s > nks / 8) s = nks / 8;
    if (s > 8) s = 8;
    return s < 2 ? 1 : s;
}
__global__ void __launch_bounds__(256)
reduce_splits_kernel(const float * __restrict__ ws, const int S, const long long MN, const int M, float * __restrict__ out, const long long out_cs,
                     const int add_to_out) {
    const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= MN) return;
    const long long n = i / M; const int m = (int) (i - n * M);
    float * o = out + n * out_cs + m;
    float v = add_to_out ? *o : 0.f;
    for (int sp = 0; sp < S; sp++) v += ws[(long long) sp * MN + i];
    *o = v;
}

// four consecutive m per thread (M, out_cs multiples of 4, 16-byte aligned buffers): 128-bit loads and stores, no division per element
__global__ void __launch_bounds__(256)
reduce_splits4_kernel(const float4 * __restrict__ ws, const int S, const long long MN4, const int M4, float4 * __restrict__ out, const long long out_cs4,
                      const int add_to_out, const float4 * __restrict__ res_src, float4 * __restrict__ res_dst /* fused residual ADD: [N][M] contiguous, or null */,
                      uint2 * __restrict__ act_out /* with res_dst, M % 32 == 0: fp16(q8_1(res_dst)) for the next dense GEMM, or null */) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; ; i += (long long) gridDim.x * blockDim.x) {
        const bool active = i < MN4;
        const unsigned mask = __ballot_sync(0xffffffffu, active);       // (whole groups of 8 lanes = one 32-element block: MN4 % 8 == 0 with act_out)
        if (mask == 0) break;
        if (!active) continue;
        const long long n = i / M4; const int m = (int) (i - n * M4);
        float4 * o = out + n * out_cs4 + m;
        float4 v = add_to_out ? *o : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int sp = 0; sp < S; sp++) {               // split order: deterministic
            const float4 w = ws[(long long) sp * MN4 + i];
            v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
        }
        *o = v;
        if (res_dst) {                                  // the GGML_OP_ADD node that consumes this MUL_MAT (same IEEE add as add_f32_*)
            const float4 r = res_src[i];
            const float4 z = make_float4(r.x + v.x, r.y + v.y, r.z + v.z, r.w + v.w);
            res_dst[i] = z;
            if (act_out) {
                // the sum is the next MUL_MAT's src1 (its K runs along m): convert it here as dequant_act_f16_kernel would
                // (quantize_row_q8_1_ref ggml-quants.c:220-253, then d * q rounded to fp16), 8 lanes per 32-element block
                float amax = fmaxf(fmaxf(fabsf(z.x), fabsf(z.y)), fmaxf(fabsf(z.z), fabsf(z.w)));
                amax = fmaxf(amax, __shfl_xor_sync(mask, amax, 1));
                amax = fmaxf(amax, __shfl_xor_sync(mask, amax, 2));
                amax = fmaxf(amax, __shfl_xor_sync(mask, amax, 4));
                const float d  = __fdiv_rn(amax, 127.0f);
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
                const float dh = __half2float(__float2half_rn(d));
                const float r0 = dh * roundf(__fmul_rn(z.x, id)), r1 = dh * roundf(__fmul_rn(z.y, id));
                const float r2 = dh * roundf(__fmul_rn(z.z, id)), r3 = dh * roundf(__fmul_rn(z.w, id));
                uint2 h;
                asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.x) : "f"(r1), "f"(r0));
                asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h.y) : "f"(r3), "f"(r2));
                act_out[i] = h;
            }
        }
    }
}

bool reduce_splits_can_fuse_residual(long long M, long long out_cs, const void * out, const void * res_src, const void * res_dst) {
    return M % 4 == 0 && out_cs % 4 == 0 && (((uintptr_t) out | (uintptr_t) res_src | (uintptr_t) res_dst) & 15) == 0;
}
cudaError_t launch_reduce_splits(const float * ws, int S, long long MN, int M, float * out, long long out_cs, int add_to_out, cudaStream_t st,
                                 const float * res_src, float * res_dst, void * act_out) {
    if (res_dst && !reduce_splits_can_fuse_residual(M, out_cs, out, res_src, res_dst)) return cudaErrorInvalidValue;
    if (act_out && (!res_dst || M % 32 != 0 || ((uintptr_t) act_out & 7) != 0)) return cudaErrorInvalidValue;
    if (M % 4 == 0 && out_cs % 4 == 0 && (((uintptr_t) ws | (uintptr_t) out) & 15) == 0) {
        const long long MN4 = MN / 4;
        long long blocks = (MN4 + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned) blocks); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
        return cudaLaunchKernelEx(&cfg, reduce_splits4_kernel, (const float4 *) ws, S, MN4, M / 4, (float4 *) out, out_cs / 4, add_to_out,
                                  (const float4 *) res_src, (float4 *) res_dst, (uint2 *) act_out);
    }
    reduce_splits_kernel<<<(unsigned) ((MN + 255) / 256), 256, 0, st>>>(ws, S, MN, M, out, out_cs, add_to_out);
    return cudaGetLastError();
}

// split_ws: workspace of gemm_i8_ksplit(...) * N * M floats, or nullptr (no split)
cudaError_t launch_gemm_i8(int kind, const void * A, const __half * dA_h, const float * dA_f, const float * mA, const void * B, const float * dB,
                           long long ldb, float * out, long long out_cs, int M, int N, int K, int sms, float * split_ws, cudaStream_t st) {
    CUtensorMap tmA, tmB, tmS;
    const long long KB = K / gm_kpb(kind);
    const int esize = kind == 0 ? 1 : 2, bps = gm_bps(kind);
    if (!make_map_operand(&tmA, A, M, K, esize) || !make_map_operand(&tmB, B, N, K, esize) || !make_map_scales(&tmS, dB, KB, ldb, bps)) return cudaErrorInvalidValue;
    GemmI8Args p{};
    p.dA = dA_h ? (const void *) dA_h : (const void *) dA_f; p.dA_stride = KB; p.out = out; p.out_cs = out_cs;
    p.M = M; p.N = N; p.K = K;
    p.mtiles = (M + GM_BM - 1) / GM_BM; p.ntiles = (N + GM_BN - 1) / GM_BN;
    p.ksplit = split_ws ? gemm_i8_ksplit(M, N, K, kind, sms) : 1;
    if (mA) {
        min_term_kernel<<<dim3((unsigned) ((M + 63) / 64), (unsigned) ((N + 63) / 64)), 256, 0, st>>>(mA, dB + KB * ldb, ldb, out, out_cs, M, N, (int) KB);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        p.acc_init = 1;
    }
    const int add_to_out = p.acc_init;
    if (p.ksplit > 1) { p.out = split_ws; p.out_cs = M; p.split_stride = (long long) M * N; p.acc_init = 0; }
    const int work = p.mtiles * p.ntiles * p.ksplit;
    const int grid = work < sms ? work : sms;
    cudaError_t e;
    if (kind == 0)      e = dA_h ? gemm_i8_launch_t<0, 0>(tmA, tmB, tmS, p, grid, st) : gemm_i8_launch_t<1, 0>(tmA, tmB, tmS, p, grid, st);
    else if (kind == 1) e = dA_h ? gemm_i8_launch_t<0, 1>(tmA, tmB, tmS, p, grid, st) : gemm_i8_launch_t<1, 1>(tmA, tmB, tmS, p, grid, st);
    else if (dA_f && !mA) e = gemm_i8_launch_t<1, 2>(tmA, tmB, tmS, p, grid, st);
    else e = cudaErrorInvalidValue;
    if (e != cudaSuccess || p.ksplit == 1) return e;
    return launch_reduce_splits(split_ws, p.ksplit, (long long) M * N, M, out, out_cs, add_to_out, st);
}

}  // namespace b200

// tc_common.cuh — PTX wrappers shared by the tensor-core kernels (gemm_i8.cu, gemm_dense.cu): mbarriers, TMA tile loads,
// tcgen05 shared-memory descriptors / MMA / commit / TMEM loads, and the host-side tensor-map encoder.
#pragma once
#include "b200_internal.cuh"

#include <cuda.h>

namespace b200 {

constexpr int GM_BM = 128, GM_BN = 128, GM_BK = 128;        // tile; BK = bytes of K per stage and row = one 128-B swizzle row
constexpr long long GM_TIMEOUT_CYCLES = 8000000000ll;

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t gm_smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void gm_mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count)); }
__device__ __forceinline__ void gm_mbar_arrive(uint32_t bar) { asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(bar) : "memory"); }
__device__ __forceinline__ void gm_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool gm_mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void gm_mbar_wait(uint32_t bar, uint32_t parity) {
    if (gm_mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    while (!gm_mbar_try(bar, parity)) { if (clock64() - t0 > GM_TIMEOUT_CYCLES) __trap(); }
}
__device__ __forceinline__ void gm_tma_2d(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// K-major operand tile, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO), LBO unused, descriptor version 1
__device__ __forceinline__ uint64_t gm_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t) ((saddr & 0x3FFFF) >> 4);             // start address, bits [0,14)
    d |= (uint64_t) (1024 >> 4) << 32;                    // stride byte offset, bits [32,46)
    d |= (uint64_t) 1 << 46;                              // version = 1 (sm_100)
    d |= (uint64_t) 2 << 61;                              // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void gm_mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void gm_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void gm_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
// Whole-warp forms for a convergent MMA-issue loop: every lane executes the statement, elect.sync picks the (always same) lane
// that issues.  With `if (lane == 0) { ... tcgen05.mma ... }` the operands are per-thread values for the compiler, and every MMA
// costs an ELECT / R2UR.BROADCAST / BRA.U.ANY "waterfall" of ~27 dependent instructions in a one-thread loop: ~190 clocks per
// 64-clock MMA (ncu source page, profiles/r01_l: the issuer never waits for data, the TMA producer waits for the issuer).
// Convergent code keeps descriptors in uniform registers and the loop at a few instructions per MMA.
__device__ __forceinline__ void gm_mma_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void gm_mbar_expect_tx_elect(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gm_tma_2d_elect(uint32_t dst, const CUtensorMap * map, uint32_t bar, int c0, int c1) {
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void gm_commit_elect(uint32_t bar) {
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void gm_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gm_tmem_ld32(uint32_t taddr, int (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void gm_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ host: tensor maps + launch
typedef CUresult (*encode_tiled_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline encode_tiled_t encode_tiled() {
    static encode_tiled_t fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_t) p;
    }();
    return fn;
}
// operand matrix [rows][K], K contiguous, 1-byte (int8) or 2-byte (fp16) elements; box = 128 bytes of K x 128 rows, 128-B swizzle,
// out-of-range rows / k read as zero
static inline bool make_map_operand(CUtensorMap * m, const void * base, long long rows, long long K, int esize, int box_rows = GM_BM) {
    encode_tiled_t enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t) K, (cuuint64_t) rows};
    const cuuint64_t strides[1] = {(cuuint64_t) K * esize};
    const cuuint32_t box[2] = {(cuuint32_t) (GM_BK / esize), (cuuint32_t) box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, esize == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace b200

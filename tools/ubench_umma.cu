// ubench_umma.cu — dense tcgen05.mma peak of this box, no loads in the loop: every CTA pair (cluster of 2, as gemm_dense2_kernel) issues
// back-to-back tcgen05.mma.cta_group::2 M = 256, N = 256 on operands that sit in shared memory (K-major, 128-byte swizzle), accumulating
// in TMEM; kind::f16 (K = 16 per instruction) and kind::i8 (K = 32).  The figure is the tensor-pipe ceiling a kernel with shared-memory
// operands can reach under this box's power management (clocks are not touched) -- the denominators bench.py quotes the prefill GEMM against.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ggml-hexagon_b200/csrc -o /tmp/ubench_umma tools/ubench_umma.cu
#include "tc_common.cuh"
#include <stdio.h>
#include <stdlib.h>
using namespace b200;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

template <int KIND>   // 0: i8, 1: f16
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) umma_kernel(const int n_mma, const uint32_t seed) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (gm_smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t * smem = smem_raw + (base - gm_smem_u32(smem_raw));
    const uint32_t sA = base, sB = base + 16384, bar = base + 32768;
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (smem + 32768 + 16);
    uint32_t rank; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    // operands: pseudo-random small values (fp16: bytes 0x30..0x3b -> |v| in [0.13, 1); i8: any byte)
    uint32_t x = seed + blockIdx.x * 977u + threadIdx.x * 31u;
    for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) {
        x = x * 1664525u + 1013904223u;
        uint32_t v = x;
        if (KIND == 1) v = (x & 0x0BFF0BFFu) | 0x30003000u;
        ((uint32_t *) smem)[i] = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) { gm_mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(gm_smem_u32((const void *) tmem_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    gm_fence_before();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    gm_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (rank == 0 && threadIdx.x < 32) {
        const uint32_t idesc = (KIND == 0 ? (2u << 4) | (1u << 7) | (1u << 10) : (1u << 4)) | ((uint32_t) (256 >> 3) << 17) | ((uint32_t) (256 >> 4) << 24);
        const uint64_t ad = gm_smem_desc(sA), bd = gm_smem_desc(sB);
        for (int i = 0; i < n_mma; i++) {
            const uint32_t j = (uint32_t) (i & 3), buf = (uint32_t) ((i >> 2) & 1);        // 4 K steps of a 128-byte row, two accumulators
            if (KIND == 0)
                asm volatile("{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|e, 0xffffffff;\n\t@e tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                             :: "r"(tmem + buf * 256), "l"(ad + 2u * j), "l"(bd + 2u * j), "r"(idesc), "r"(i > 7 ? 1u : 0u) : "memory");
            else
                asm volatile("{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|e, 0xffffffff;\n\t@e tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             :: "r"(tmem + buf * 256), "l"(ad + 2u * j), "l"(bd + 2u * j), "r"(idesc), "r"(i > 7 ? 1u : 0u) : "memory");
        }
        asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                     "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" :: "r"(bar), "h"((uint16_t) 3) : "memory");
    }
    gm_mbar_wait(bar, 0);
    gm_fence_after();
    __syncthreads();
    gm_fence_before();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(512) : "memory");
}

template <int KIND>
static int run(const char * name, int kper, int sms) {
    const int smem = 32768 + 64 + 1024, n_mma = 40000, grid = sms / 2 * 2;
    CK(cudaFuncSetAttribute(umma_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    umma_kernel<KIND><<<grid, 128, smem>>>(1000, 1u);
    CK(cudaDeviceSynchronize());
    double best = 0, sum = 0; const int reps = 20;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0));
        umma_kernel<KIND><<<grid, 128, smem>>>(n_mma, (uint32_t) r);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double tops = 2.0 * 256 * 256 * kper * (double) n_mma * (grid / 2) / (ms * 1e-3) / 1e12;
        if (tops > best) best = tops;
        sum += tops;
    }
    printf("{\"kind\": \"%s\", \"shape\": \"cta_group::2 M=256 N=256 K=%d, operands in shared memory\", \"pairs\": %d, \"mma_per_pair\": %d, \"TOPS_best\": %.1f, \"TOPS_mean_of_%d_back_to_back\": %.1f}\n",
           name, kper, grid / 2, n_mma, best, reps, sum / reps);
    return 0;
}
int main() {
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    if (run<1>("f16", 16, sms)) return 1;
    if (run<0>("i8", 32, sms)) return 1;
    return 0;
}

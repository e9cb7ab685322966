// ubench_ring.cu — steady-state HBM read bandwidth of per-warp cp.async.bulk rings in a persistent kernel (one CTA per SM), as a
// function of warps, ring slots and bytes per copy; and what a dependent L2-hit load costs WHILE that stream runs (the activation
// vector of a decode GEMV sits behind exactly that).  Informs the ring geometry of gemv6.cu.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/ubench_ring tools/ubench_ring.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

// mode 0: wait + re-arm only; 1: every lane also reads its share of the slot (LDS.128) before re-arming
// probe: warp 0 of every CTA does not stream; its lane 0 chases dependent ld.global.cg loads through a small (L2-resident) buffer
__global__ void ring_kernel(const uint8_t * __restrict__ src, size_t bytes_per_cta, int S, int U, int mode, int probe,
                            const uint32_t * __restrict__ chase, unsigned long long * lat_sum, unsigned long long * lat_n, uint32_t * sink) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int sw = probe ? warp - 1 : warp, nsw = probe ? nwarps - 1 : nwarps;        // streaming warp index
    if (probe && warp == 0) {
        if (lane == 0) {
            uint32_t idx = blockIdx.x * 37u; unsigned long long sum = 0, n = 0;
            // run until the streaming warps are done (flag in shared memory)
            volatile int * done = (volatile int *) sm;
            while (*done < nsw) {
                const long long t0 = clock64();
                uint32_t v; asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(chase + (idx & 4095)));
                idx = v + idx * 0 + (idx + 97u);       // dependent on the loaded value (v == 0)
                const long long t1 = clock64();
                sum += (unsigned long long) (t1 - t0); n++;
            }
            atomicAdd(lat_sum, sum); atomicAdd(lat_n, n);
        }
        return;
    }
    if (threadIdx.x == (probe ? 32 : 0)) *(volatile int *) sm = 0;
    uint8_t * ring = sm + 128 + (size_t) sw * S * U;
    const uint32_t bar0 = smem_u32(sm + 128 + (size_t) nsw * S * U) + sw * S * 8;
    const uint8_t * base = src + (size_t) blockIdx.x * bytes_per_cta;
    const int nunits = (int) (bytes_per_cta / U);
    const int nmine = nunits > sw ? (nunits - sw + nsw - 1) / nsw : 0;
    if (lane == 0) {
        for (int s = 0; s < S; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar0 + s * 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int i = 0; i < S && i < nmine; i++) {
            asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar0 + i * 8), "r"(U) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(smem_u32(ring + (size_t) i * U)), "l"(base + (size_t) (sw + i * nsw) * U), "r"(U), "r"(bar0 + i * 8) : "memory");
        }
    }
    __syncwarp();
    uint32_t acc = 0; int slot = 0; uint32_t phase = 0;
    for (int i = 0; i < nmine; i++) {
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar0 + slot * 8), "r"(phase) : "memory");
        if (mode == 1) {
            const uint4 * p = (const uint4 *) (ring + (size_t) slot * U);
            for (int k = lane; k < U / 16; k += 32) { const uint4 v = p[k]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
        }
        __syncwarp();
        if (lane == 0 && i + S < nmine) {
            asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar0 + slot * 8), "r"(U) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(smem_u32(ring + (size_t) slot * U)), "l"(base + (size_t) (sw + (i + S) * nsw) * U), "r"(U), "r"(bar0 + slot * 8) : "memory");
        }
        if (++slot == S) { slot = 0; phase ^= 1u; }
    }
    if (acc == 0x12345678u) sink[blockIdx.x] = acc;
    if (lane == 0) atomicAdd((int *) sm, 1);
}

// plain 16-byte loads, persistent, same chunking: the LSU path next to the bulk-copy path
__global__ void ldg_kernel(const uint4 * __restrict__ src, size_t n16_per_cta, int unroll_dummy, uint32_t * sink) {
    const uint4 * p = src + (size_t) blockIdx.x * n16_per_cta;
    uint32_t acc = 0;
    size_t i = threadIdx.x;
    for (; i + 7 * blockDim.x < n16_per_cta; i += 8 * blockDim.x) {
        uint4 v[8];
        #pragma unroll
        for (int u = 0; u < 8; u++) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p + i + u * blockDim.x));
        #pragma unroll
        for (int u = 0; u < 8; u++) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) sink[blockIdx.x] = acc;
}

int main(int argc, char ** argv) {
    const size_t pool = (size_t) 8 << 30;
    uint8_t * buf; uint32_t * sink; uint32_t * chase; unsigned long long * lat;
    CK(cudaMalloc(&buf, pool)); CK(cudaMalloc(&sink, 1 << 20)); CK(cudaMalloc(&chase, 4096 * 4)); CK(cudaMalloc(&lat, 16));
    CK(cudaMemset(buf, 1, pool)); CK(cudaMemset(chase, 0, 4096 * 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    printf("cps warps  S      U  KB/SM mode probe |    GB/s   L2-hit load under stream (cycles)\n");
    const int Us[] = {2304, 4608, 9216, 18432, 36864};
    for (int cps : {1, 2})
    for (int nw : {4, 8, 16})
    for (int S : {2, 3, 4, 8})
    for (int U : Us)
    for (int mode : {0, 1})
    for (int probe : {0, 1}) {
        if (probe && !(mode == 1 && cps == 1)) continue;
        const int nsw = nw, warps = nw + (probe ? 1 : 0);
        const size_t smem = 128 + (size_t) nsw * S * U + nsw * S * 8 + 64;
        if (smem * cps > 226 * 1024 || warps * 32 * cps > 2048) continue;
        const int inflight_kb = (int) ((size_t) nsw * S * U * cps / 1024);
        if (inflight_kb < 30) continue;
        if (mode == 0 && !(inflight_kb >= 60)) continue;
        const int grid = sms * cps;
        size_t per_cta = ((size_t) 4 << 30) / grid; per_cta = per_cta / ((size_t) U * nsw) * ((size_t) U * nsw);
        CK(cudaMemset(lat, 0, 16));
        for (int rep = 0; rep < 2; rep++) {
            if (rep == 1) CK(cudaEventRecord(e0));
            ring_kernel<<<grid, warps * 32, smem>>>(buf + (rep ? ((size_t) 4 << 30) : 0), per_cta, S, U, mode, probe, chase, lat, lat + 1, sink);
            if (rep == 1) CK(cudaEventRecord(e1));
        }
        CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        unsigned long long h[2]; CK(cudaMemcpy(h, lat, 16, cudaMemcpyDeviceToHost));
        printf("%3d %5d %2d %6d %6d %4d %5d | %7.1f   %s", cps, nw, S, U, inflight_kb, mode, probe, (double) per_cta * grid / ms / 1e6, "");
        if (probe && h[1]) printf("%.0f", (double) h[0] / (double) h[1] / 2.0);     // (two warm + timed launches accumulate: mean over both)
        printf("\n");
    }
    for (int threads : {256, 512, 1024}) for (int cps : {1, 2}) {
        if (threads * cps > 2048) continue;
        const int grid = sms * cps;
        size_t n16 = ((size_t) 4 << 30) / 16 / grid; n16 = n16 / (8 * threads) * (8 * threads);
        for (int rep = 0; rep < 2; rep++) {
            if (rep == 1) CK(cudaEventRecord(e0));
            ldg_kernel<<<grid, threads>>>((const uint4 *) (buf + (rep ? ((size_t) 4 << 30) : 0)), n16, 0, sink);
            if (rep == 1) CK(cudaEventRecord(e1));
        }
        CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("ldg persistent: %d x %d threads, 8 x 16 B per thread in flight (%d KB/SM): %7.1f GB/s\n", grid, threads, threads * cps * 128 / 1024, (double) n16 * 16 * grid / ms / 1e6);
    }
    return 0;
}

// ubench_tc.cu — two B200 micro-benchmarks behind the prefill GEMM design:
//   (1) tcgen05.ld throughput per SM (TMEM -> registers), shapes 32x32b.x32 / .x64 / 16x256b.x8 and .pack::16b
//   (2) legacy mma.sync.m16n8k32 s8 x s8 -> s32 throughput per SM (accumulators in registers, no TMEM)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tc tools/ubench_tc.cu ; run on one GPU.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

template <int SHAPE>
__device__ __forceinline__ int tm_ld(uint32_t taddr) {
    int r[32];
    if (SHAPE == 0) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr) : "memory");
    } else if (SHAPE == 1) {     // 64 columns packed to 32 registers (low 16 bits of two adjacent columns per register)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr) : "memory");
    } else {                     // 16 lanes x 256 bit, x8: 16 rows x 64 columns -> 32 registers
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr) : "memory");
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    int s = 0;
    #pragma unroll
    for (int i = 0; i < 32; i++) s ^= r[i];
    return s;
}

template <int SHAPE>
__global__ void __launch_bounds__(512, 1) tmem_ld_kernel(int iters, long long * cycles, int * sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"((uint32_t) __cvta_generic_to_shared(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    const uint32_t taddr = tmem + ((uint32_t) ((warp & 3) * 32) << 16) + (uint32_t) (((warp >> 2) * 64) & 511);
    int s = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) s ^= tm_ld<SHAPE>(taddr + (i & 1) * 32);
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    if (s == 0x12345678) sink[0] = s;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

// ---- legacy tensor path: each warp keeps NACC independent m16n8k32 accumulators
template <int NACC>
__global__ void __launch_bounds__(512, 1) mma_sync_kernel(int iters, long long * cycles, int * sink) {
    int acc[NACC][4];
    #pragma unroll
    for (int j = 0; j < NACC; j++) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0; }
    uint32_t a[4] = {threadIdx.x * 0x01010101u, 0x01020304u, 0x7f7f7f7fu, threadIdx.x}, b[2] = {0x01010101u, threadIdx.x * 3u};
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        #pragma unroll
        for (int j = 0; j < NACC; j++) {
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                         : "+r"(acc[j][0]), "+r"(acc[j][1]), "+r"(acc[j][2]), "+r"(acc[j][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    int s = 0;
    #pragma unroll
    for (int j = 0; j < NACC; j++) s ^= acc[j][0] ^ acc[j][1] ^ acc[j][2] ^ acc[j][3];
    if (s == 0x12345678) sink[0] = s;
}

int main() {
    long long * cyc; int * sink;
    CK(cudaMalloc(&cyc, 1024 * 8)); CK(cudaMalloc(&sink, 4));
    long long h[148];
    const int iters = 4000;
    const char * names[3] = {"32x32b.x32 (4 B cells)", "32x32b.x32.pack::16b", "16x256b.x8"};
    for (int threads = 128; threads <= 512; threads *= 2) {
        for (int shape = 0; shape < 3; shape++) {
            if (shape == 0) tmem_ld_kernel<0><<<148, threads>>>(iters, cyc, sink);
            if (shape == 1) tmem_ld_kernel<1><<<148, threads>>>(iters, cyc, sink);
            if (shape == 2) tmem_ld_kernel<2><<<148, threads>>>(iters, cyc, sink);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
            const double regs_bytes = (double) iters * (threads / 32) * 32 * 32 * 4;          // bytes landed in registers per SM
            const double cells = shape == 1 ? 2.0 : 1.0;                                       // TMEM cells read per register
            printf("tcgen05.ld %-26s warps %2d : %8lld cycles, %6.1f register B/clk/SM, %6.1f TMEM cell B/clk/SM\n", names[shape], threads / 32, h[0],
                   regs_bytes / h[0], regs_bytes * cells / h[0]);
        }
    }
    for (int threads = 128; threads <= 512; threads *= 2) {
        mma_sync_kernel<8><<<148, threads>>>(iters, cyc, sink);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
        const double macs = (double) iters * 8 * (threads / 32) * 16 * 8 * 32;
        printf("mma.sync.m16n8k32.s8 warps %2d x 8 accumulators: %8lld cycles, %7.1f MAC/clk/SM (= %.0f dense TOPS on 148 SMs at 1.965 GHz)\n",
               threads / 32, h[0], macs / h[0], macs / h[0] * 2 * 148 * 1.965e9 / 1e12);
    }
    return 0;
}

// ubench_stream.cu — what a launch of the decode GEMVs' size can reach at all: a kernel that only READS n bytes (16-byte loads,
// XOR-reduced, one store per CTA), launched back to back over distinct buffers (no L2 reuse), for the byte counts of the
// Llama-2-7B q4_0 launches (9.4 / 25.4 / 28.3 / 50.7 MB) and two large ones.  Gives the per-launch "speed of light" next to
// which DESIGN.md section 4 puts the GEMV times.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/ubench_stream tools/ubench_stream.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

// each thread: UNROLL independent 16-byte loads per iteration, grid-stride
// PDL: like the GEMVs -- let the next launch start at once, issue the first loads, then wait for the previous grid
template <int UNROLL, bool PDL>
__global__ void __launch_bounds__(256) read_kernel(const uint4 * __restrict__ p, size_t n16, uint32_t * out) {
    uint32_t acc = 0;
    if (PDL) cudaTriggerProgrammaticLaunchCompletion();
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < n16; i += UNROLL * stride) {
        uint4 v[UNROLL];
        #pragma unroll
        for (int u = 0; u < UNROLL; u++)
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p + i + u * stride));
        if (PDL) cudaGridDependencySynchronize();        // (returns at once after the first time)
        #pragma unroll
        for (int u = 0; u < UNROLL; u++) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    for (; i < n16; i += stride) { const uint4 v = p[i]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    if (acc == 0x12345678u) out[blockIdx.x] = acc;
}

int main() {
    const size_t pool = (size_t) 6 << 30;                 // 6 GB of distinct data
    uint8_t * buf; uint32_t * out;
    CK(cudaMalloc(&buf, pool)); CK(cudaMalloc(&out, 1 << 20));
    CK(cudaMemset(buf, 1, pool));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double sizes_mb[] = {9.44, 25.4, 28.3, 50.7, 73.7, 256.0, 1024.0};
    for (double mb : sizes_mb) {
        const size_t bytes = ((size_t) (mb * 1e6) + 255) & ~(size_t) 255;
        const int nbuf = (int) (pool / bytes) < 64 ? (int) (pool / bytes) : 64;
        for (int grid_mult : {4, 8}) {
            const int grid = 148 * grid_mult;
            const int reps = 3;
            for (int pdl = 0; pdl < 2; pdl++) {
                cudaLaunchConfig_t cfg{};
                cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = st;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at; cfg.numAttrs = pdl;
                auto launch = [&](int b) {
                    const uint4 * p = (const uint4 *) (buf + (size_t) b * bytes); const size_t n16 = bytes / 16;
                    return pdl ? cudaLaunchKernelEx(&cfg, read_kernel<8, true>, p, n16, out) : cudaLaunchKernelEx(&cfg, read_kernel<8, false>, p, n16, out);
                };
                for (int b = 0; b < nbuf; b++) CK(launch(b));
                CK(cudaStreamSynchronize(st));
                CK(cudaEventRecord(e0, st));
                for (int r = 0; r < reps; r++)
                    for (int b = 0; b < nbuf; b++) CK(launch(b));
                CK(cudaEventRecord(e1, st));
                CK(cudaStreamSynchronize(st));
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
                const double us = ms * 1e3 / (reps * nbuf);
                printf("read %7.2f MB per launch, grid %4d x 256, %2d buffers, %s: %7.2f us per launch, %7.1f GB/s\n", mb, grid, nbuf,
                       pdl ? "PDL   " : "no PDL", us, bytes / us / 1e3);
            }
        }
    }
    return 0;
}

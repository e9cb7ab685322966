// norm_ops.cu — the three ops ggml-hexagon's cDSP table lists beside ADD / MUL_MAT and leaves as empty stubs
// (caps table ggml-hexagon.cpp:619, 639, 649; stubs kernels/ggml-dsp.c:1353-1372): RMS_NORM, SOFT_MAX, POOL_2D, as sm_100a kernels.
// Specifications: ggml_compute_forward_rms_norm_f32 (ggml-cpu/ggml-cpu.c:6254-6300), ggml_compute_forward_soft_max_f32 (:8261-8357),
// ggml_compute_forward_pool_2d (:9758-9830).  All three are HBM / latency bound row-wise work: one CTA per row (one thread per output
// for the pooling), 128-bit loads where the row is 16-byte aligned, warp-shuffle reductions.
#include "b200_internal.cuh"

#include <math_constants.h>

namespace b200 {

template <typename T>
__device__ __forceinline__ T block_reduce(T v, T * sh, const bool is_max) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const T w = __shfl_xor_sync(0xffffffffu, v, o); v = is_max ? (w > v ? w : v) : v + w; }
    __syncthreads();                       // sh may still be read from the previous reduction
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    T r = sh[0];
    for (int i = 1; i < nw; i++) r = is_max ? (sh[i] > r ? sh[i] : r) : r + sh[i];      // same order in every thread: one value for the CTA
    return r;
}

// ---- RMS_NORM: y = x / sqrt(mean(x^2) + eps) per row; the sum of the f32 squares in double, as the reference
__global__ void __launch_bounds__(256)
rms_norm_f32_kernel(const char * __restrict__ x, char * __restrict__ y, const int64_t ne00, const int64_t ne01, const int64_t ne02,
                    const int64_t nb01, const int64_t nb02, const int64_t nb03, const int64_t nbd1, const int64_t nbd2, const int64_t nbd3, const float eps) {
    __shared__ double sh[8];
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
#endif
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % ne01, i2 = (row / ne01) % ne02, i3 = row / (ne01 * ne02);
    const float * xr = (const float *) (x + i1 * nb01 + i2 * nb02 + i3 * nb03);
    float * yr = (float *) (y + i1 * nbd1 + i2 * nbd2 + i3 * nbd3);
    const bool vec = (((uintptr_t) xr | (uintptr_t) yr) & 15) == 0 && ne00 % 4 == 0;
    double sum = 0.0;
    if (vec) {
        for (int64_t i = threadIdx.x; i < ne00 / 4; i += blockDim.x) {
            const float4 v = ((const float4 *) xr)[i];
            sum += (double) __fmul_rn(v.x, v.x) + (double) __fmul_rn(v.y, v.y) + (double) __fmul_rn(v.z, v.z) + (double) __fmul_rn(v.w, v.w);
        }
    } else {
        for (int64_t i = threadIdx.x; i < ne00; i += blockDim.x) sum += (double) __fmul_rn(xr[i], xr[i]);
    }
    sum = block_reduce<double>(sum, sh, false);
    const float mean = (float) (sum / (double) ne00);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    if (vec) {
        for (int64_t i = threadIdx.x; i < ne00 / 4; i += blockDim.x) {
            const float4 v = ((const float4 *) xr)[i];
            ((float4 *) yr)[i] = make_float4(__fmul_rn(v.x, scale), __fmul_rn(v.y, scale), __fmul_rn(v.z, scale), __fmul_rn(v.w, scale));
        }
    } else {
        for (int64_t i = threadIdx.x; i < ne00; i += blockDim.x) yr[i] = __fmul_rn(xr[i], scale);
    }
}
cudaError_t launch_rms_norm_f32(const void * x, void * y, const int64_t * ne, const int64_t * nbs, const int64_t * nbd, float eps, cudaStream_t st) {
    const int64_t rows = ne[1] * ne[2] * ne[3];
    if (rows == 0 || ne[0] == 0) return cudaSuccess;
    if (rows > 0x7fffffff) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) rows); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, rms_norm_f32_kernel, (const char *) x, (char *) y, ne[0], ne[1], ne[2], nbs[1], nbs[2], nbs[3], nbd[1], nbd[2], nbd[3], eps);
}

// ---- SOFT_MAX(x * scale + slope(head) * mask) per row of contiguous src0 / dst; mask f32 or f16 [ne00, >= ne01], broadcast over dims 2, 3
template <bool MASK_F16>
__global__ void __launch_bounds__(256)
soft_max_f32_kernel(const float * __restrict__ x, const void * __restrict__ mask, float * __restrict__ y, const int64_t ne00, const int64_t ne01,
                    const int64_t ne02, const float scale, const float max_bias, const float m0, const float m1, const uint32_t n_head_log2) {
    __shared__ double shd[8];
    __shared__ float shf[8];
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
#endif
    const int64_t row = blockIdx.x;
    const uint32_t h = (uint32_t) ((row / ne01) % ne02);
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
    const float * sp = x + row * ne00;
    float * dp = y + row * ne00;
    const float * mf = mask && !MASK_F16 ? (const float *) mask + (row % ne01) * ne00 : nullptr;
    const __half * mh = mask && MASK_F16 ? (const __half *) mask + (row % ne01) * ne00 : nullptr;
    auto wp = [&](int64_t i) {
        float v = __fmul_rn(sp[i], scale);
        if (mf) v = __fadd_rn(v, __fmul_rn(slope, mf[i]));
        if (mh) v = __fadd_rn(v, __fmul_rn(slope, __half2float(mh[i])));
        return v;
    };
    float mx = -CUDART_INF_F;
    for (int64_t i = threadIdx.x; i < ne00; i += blockDim.x) mx = fmaxf(mx, wp(i));
    mx = block_reduce<float>(mx, shf, true);
    double sum = 0.0;
    for (int64_t i = threadIdx.x; i < ne00; i += blockDim.x) { const float v = expf(__fsub_rn(wp(i), mx)); dp[i] = v; sum += (double) v; }
    sum = block_reduce<double>(sum, shd, false);
    const float inv = (float) (1.0 / sum);
    for (int64_t i = threadIdx.x; i < ne00; i += blockDim.x) dp[i] = __fmul_rn(dp[i], inv);      // (each thread rescales what it wrote itself)
}
// rows of up to 1024 elements (attention scores at prefill: thousands of short rows): one warp per row, the row held in registers between
// the three passes; same operations per element as the block kernel, the double sum reduced by shuffles
template <bool MASK_F16, int NT>
__global__ void __launch_bounds__(256)
soft_max_f32_warp_kernel(const float * __restrict__ x, const void * __restrict__ mask, float * __restrict__ y, const int64_t ne00, const int64_t ne01,
                         const int64_t ne02, const int64_t rows, const float scale, const float max_bias, const float m0, const float m1, const uint32_t n_head_log2) {
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
#endif
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const uint32_t h = (uint32_t) ((row / ne01) % ne02);
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
    const float * sp = x + row * ne00;
    float * dp = y + row * ne00;
    const float * mf = mask && !MASK_F16 ? (const float *) mask + (row % ne01) * ne00 : nullptr;
    const __half * mh = mask && MASK_F16 ? (const __half *) mask + (row % ne01) * ne00 : nullptr;
    float v[NT];                       // NT x 32 >= ne00
    float mx = -CUDART_INF_F;
    #pragma unroll
    for (int t = 0; t < NT; t++) {
        const int64_t i = lane + 32 * t;
        v[t] = -CUDART_INF_F;
        if (i < ne00) {
            float w = __fmul_rn(sp[i], scale);
            if (mf) w = __fadd_rn(w, __fmul_rn(slope, mf[i]));
            if (mh) w = __fadd_rn(w, __fmul_rn(slope, __half2float(mh[i])));
            v[t] = w; mx = fmaxf(mx, w);
        }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    double sum = 0.0;
    #pragma unroll
    for (int t = 0; t < NT; t++) if (lane + 32 * t < ne00) { v[t] = expf(__fsub_rn(v[t], mx)); sum += (double) v[t]; }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = (float) (1.0 / sum);
    #pragma unroll
    for (int t = 0; t < NT; t++) if (lane + 32 * t < ne00) dp[lane + 32 * t] = __fmul_rn(v[t], inv);
}
cudaError_t launch_soft_max_f32(const float * x, const void * mask, bool mask_f16, float * y, const int64_t * ne, float scale, float max_bias, cudaStream_t st) {
    const int64_t rows = ne[1] * ne[2] * ne[3];
    if (rows == 0 || ne[0] == 0) return cudaSuccess;
    if (rows > 0x7fffffff) return cudaErrorInvalidValue;
    const uint32_t n_head = (uint32_t) ne[2];
    const uint32_t n_head_log2 = 1u << (uint32_t) floor(log2((double) n_head));
    const float m0 = powf(2.0f, -(max_bias) / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) rows); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    if (ne[0] <= 1024 && rows >= 512) {        // (fewer rows: the block kernel, whose order the one-column attention kernel of graph_ops.cu reproduces)
        cfg.gridDim = dim3((unsigned) ((rows + 7) / 8));
#define B200_SM_WARP(NT) (mask_f16 ? cudaLaunchKernelEx(&cfg, soft_max_f32_warp_kernel<true, NT>, x, mask, y, ne[0], ne[1], ne[2], rows, scale, max_bias, m0, m1, n_head_log2) \
                                   : cudaLaunchKernelEx(&cfg, soft_max_f32_warp_kernel<false, NT>, x, mask, y, ne[0], ne[1], ne[2], rows, scale, max_bias, m0, m1, n_head_log2))
        if (ne[0] <= 256) return B200_SM_WARP(8);
        if (ne[0] <= 512) return B200_SM_WARP(16);
        return B200_SM_WARP(32);
#undef B200_SM_WARP
    }
    if (mask_f16) return cudaLaunchKernelEx(&cfg, soft_max_f32_kernel<true>, x, mask, y, ne[0], ne[1], ne[2], scale, max_bias, m0, m1, n_head_log2);
    return cudaLaunchKernelEx(&cfg, soft_max_f32_kernel<false>, x, mask, y, ne[0], ne[1], ne[2], scale, max_bias, m0, m1, n_head_log2);
}

// ---- POOL_2D: one thread per output element; taps summed ky-outer / kx-inner as the reference (bit-exact averages)
template <bool SRC_F16>
__global__ void __launch_bounds__(256)
pool_2d_kernel(const char * __restrict__ x, float * __restrict__ y, const int64_t ne0, const int64_t ne1, const int64_t planes, const int64_t nb1, const int64_t nb2,
               const int64_t px, const int64_t py, const int op, const int k0, const int k1, const int s0, const int s1, const int p0, const int p1) {
    const int64_t total = planes * px * py;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t ox = i % px, oy = (i / px) % py, pl = i / (px * py);
        const char * src = x + pl * nb2;
        float out = op == 1 ? 0.0f : -3.402823466e+38f;
        const int64_t ix = -p0 + ox * s0, iy = -p1 + oy * s1;
        for (int ky = 0; ky < k1; ky++) {
            if (iy + ky < 0 || iy + ky >= ne1) continue;
            const char * srow = src + nb1 * (iy + ky);
            for (int kx = 0; kx < k0; kx++) {
                const int64_t j = ix + kx;
                if (j < 0 || j >= ne0) continue;
                const float v = SRC_F16 ? __half2float(((const __half *) srow)[j]) : ((const float *) srow)[j];
                if (op == 1) out = __fadd_rn(out, v); else if (v > out) out = v;
            }
        }
        if (op == 1) out = __fdiv_rn(out, (float) (k0 * k1));
        y[i] = out;
    }
}
cudaError_t launch_pool_2d(const void * x, bool src_f16, float * y, const int64_t * ne, const int64_t * nb, int64_t px, int64_t py, const int32_t * opts, cudaStream_t st) {
    const int64_t planes = ne[2] * ne[3], total = planes * px * py;
    if (total == 0) return cudaSuccess;
    int64_t blocks = (total + 255) / 256; if (blocks > 148 * 16) blocks = 148 * 16;
    if (src_f16) pool_2d_kernel<true><<<(unsigned) blocks, 256, 0, st>>>((const char *) x, y, ne[0], ne[1], planes, nb[1], nb[2], px, py, opts[0], opts[1], opts[2], opts[3], opts[4], opts[5], opts[6]);
    else         pool_2d_kernel<false><<<(unsigned) blocks, 256, 0, st>>>((const char *) x, y, ne[0], ne[1], planes, nb[1], nb[2], px, py, opts[0], opts[1], opts[2], opts[3], opts[4], opts[5], opts[6]);
    return cudaGetLastError();
}

}  // namespace b200

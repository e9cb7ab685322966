// graph_ops.cu — the ops that sit BETWEEN the MUL_MATs of a Llama layer in the graphs llama.cpp builds (src/llama-model.cpp:4191-4358,
// src/llama-graph.cpp:1126-1216, :1306-1420): MUL (norm weight, gate*up), UNARY SILU, ROPE, CPY / CONT (KV-cache store, head merge),
// GET_ROWS f32 (output rows), and the f16 x f32 MUL_MAT of attention (K and V live in the cache as f16).  The cDSP backend leaves all of
// them to ggml-cpu; here that would cost a scheduler split with a host round trip after every MUL_MAT (8(f) rank 2-3 of SURVEY.md), so
// they run on the device as plain HBM-bound kernels and the whole layer stack stays one split.  Also here: the kernels ops_api.cu's graph-level
// fusions launch instead of pairs / triples of the above (RMS_NORM * weight, SILU * up, CONT, each with the q8_1 write for the vector MUL_MAT that
// follows; ROPE with the f16 K-cache store; K^T q -> soft_max -> V p of one query column), every one of them bit-equal to what it replaces.
// Specifications (ggml-cpu): binary-ops.cpp:50-117 (MUL), unary-ops.cpp / ggml_silu_f32, ggml_compute_forward_rope_f32
// (ggml-cpu.c:8708-8889; cache init :8634-8648, YaRN :8610-8632), ggml_compute_forward_dup (:3164-4010: element i of src in logical order
// -> element i of dst in logical order), ggml_compute_forward_get_rows_f32 (:7853-7892), ggml_compute_forward_mul_mat with an f16 src0
// (src1 rounded to f16 first: vec_dot_type, ggml-cpu.c:256-270 and :6807-6842; f32 accumulation).
#include "b200_internal.cuh"

#include <mma.h>
#include <math_constants.h>

namespace b200 {

// Every kernel here starts with the programmatic-dependent-launch pair (its grid may be scheduled while the previous kernel drains; nothing is
// read or written before cudaGridDependencySynchronize returns, which waits for the whole previous grid), and is launched through pdl_launch.
#define PDL_PROLOGUE() do { cudaTriggerProgrammaticLaunchCompletion(); cudaGridDependencySynchronize(); } while (0)
template <typename... KArgs, typename... Args>
static cudaError_t pdl_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// q8_1 quantisation of one 32-element block held by 8 consecutive lanes (one float4 each), written into the two-plane SoA scratch the GEMV
// kernels read: the body of quantize_q8_1_kernel (kernels.cu; = quantize_row_q8_1_ref, ggml-quants.c:220-253), for producers that quantise
// what they have just computed.  All 32 lanes of the warp must call it.
__device__ __forceinline__ void q8_1_soa_emit(const float4 v, const int sub, const int64_t row, const int64_t b, const int64_t NB,
                                              int8_t * __restrict__ qs, ActDS * __restrict__ ds) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int) roundf(__fmul_rn(v.x, id)), q1 = (int) roundf(__fmul_rn(v.y, id));
    const int q2 = (int) roundf(__fmul_rn(v.z, id)), q3 = (int) roundf(__fmul_rn(v.w, id));
    int sum = q0 + q1 + q2 + q3;
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    const int other = __shfl_xor_sync(0xffffffffu, sum, 4);
    const uint32_t packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
    int8_t * col = qs + row * NB * 32;
    *(uint32_t *) (col + ((int64_t) (sub >> 2) * NB + b) * 16 + (sub & 3) * 4) = packed;
    if (sub == 0) { ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other; ds[row * NB + b] = o; }
}

template <typename TS, typename TD> __device__ __forceinline__ TD cvt_el(TS v);
template <> __device__ __forceinline__ float  cvt_el<float, float>(float v)    { return v; }
template <> __device__ __forceinline__ __half cvt_el<float, __half>(float v)   { return __float2half_rn(v); }
template <> __device__ __forceinline__ float  cvt_el<__half, float>(__half v)  { return __half2float(v); }
template <> __device__ __forceinline__ __half cvt_el<__half, __half>(__half v) { return v; }

// ------------------------------------------------------------------------------------------------ UNARY (SILU)
__global__ void __launch_bounds__(256) silu_f32_kernel(const float * __restrict__ x, float * __restrict__ y, const int64_t n) {
    PDL_PROLOGUE();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float v = x[i];
        y[i] = __fdiv_rn(v, __fadd_rn(1.0f, expf(-v)));                 // ggml_silu_f32: x / (1 + expf(-x))
    }
}
// SILU(gate) * up in one pass: the pair every SwiGLU FFN issues around the up projection (llama-graph.cpp build_ffn, LLM_FFN_SILU +
// LLM_FFN_PAR); same two roundings as the separate kernels.  QUANT: rows of K (K % 128 == 0) are also written as q8_1 for the MUL_MAT that
// consumes the product next (ffn_down).
template <bool QUANT>
__global__ void __launch_bounds__(256) silu_mul_f32_kernel(const float4 * __restrict__ g, const float4 * __restrict__ u, float4 * __restrict__ y, const int64_t n4,
                                                           const int64_t K, int8_t * __restrict__ qs, ActDS * __restrict__ ds) {
    PDL_PROLOGUE();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 a = g[i], b = u[i];
        float4 r;
        r.x = __fmul_rn(__fdiv_rn(a.x, __fadd_rn(1.0f, expf(-a.x))), b.x); r.y = __fmul_rn(__fdiv_rn(a.y, __fadd_rn(1.0f, expf(-a.y))), b.y);
        r.z = __fmul_rn(__fdiv_rn(a.z, __fadd_rn(1.0f, expf(-a.z))), b.z); r.w = __fmul_rn(__fdiv_rn(a.w, __fadd_rn(1.0f, expf(-a.w))), b.w);
        y[i] = r;
        if (QUANT) {                                            // n4 % 32 == 0: whole warps leave the loop together
            const int64_t e = i * 4, row = e / K, b32 = (e - row * K) >> 5;
            q8_1_soa_emit(r, (int) (i & 7), row, b32, K / 32, qs, ds);
        }
    }
}
static unsigned ew_blocks(int64_t n) { int64_t b = (n + 255) / 256; if (b > 148 * 16) b = 148 * 16; return (unsigned) (b < 1 ? 1 : b); }
cudaError_t launch_silu_f32(const float * x, float * y, int64_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    return pdl_launch(silu_f32_kernel, dim3(ew_blocks(n)), dim3(256), st, x, y, n);
}
// n % 4 == 0, 16-byte aligned pointers; qs != NULL: also quantise rows of K elements (K % 128 == 0)
cudaError_t launch_silu_mul_f32(const float * g, const float * u, float * y, int64_t n, int64_t K, int8_t * qs, ActDS * ds, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const int64_t n4 = n / 4;
    if (qs) return pdl_launch(silu_mul_f32_kernel<true>, dim3(ew_blocks(n4)), dim3(256), st, (const float4 *) g, (const float4 *) u, (float4 *) y, n4, K, qs, ds);
    return pdl_launch(silu_mul_f32_kernel<false>, dim3(ew_blocks(n4)), dim3(256), st, (const float4 *) g, (const float4 *) u, (float4 *) y, n4, K, qs, ds);
}

// ------------------------------------------------------------------------------------------------ RMS_NORM * weight (+ q8_1)
// RMS_NORM followed by the in-place MUL with the norm weight (build_norm, llama-graph.cpp): one CTA per row; y = fl(fl(x * scale) * w), the
// two roundings of the separate kernels.  QUANT (ne00 % 128 == 0): the row is also written as q8_1 for the MUL_MATs that read it next.
template <bool QUANT>
__global__ void __launch_bounds__(256)
rms_norm_mul_kernel(const float * __restrict__ x, const float * __restrict__ w, float * __restrict__ y, const int64_t ne00, const int64_t xs, const int64_t ys,
                    const float eps, int8_t * __restrict__ qs, ActDS * __restrict__ ds) {
    __shared__ double sh[8];
    PDL_PROLOGUE();
    const int64_t row = blockIdx.x;
    const float4 * xr = (const float4 *) (x + row * xs);
    float4 * yr = (float4 *) (y + row * ys);
    const int64_t n4 = ne00 / 4;
    double sum = 0.0;
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 v = xr[i];
        sum += (double) __fmul_rn(v.x, v.x) + (double) __fmul_rn(v.y, v.y) + (double) __fmul_rn(v.z, v.z) + (double) __fmul_rn(v.w, v.w);
    }
    {   // block sum, same order in every thread (as block_reduce of norm_ops.cu)
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (lane == 0) sh[warp] = sum;
        __syncthreads();
        sum = sh[0];
        for (int i = 1; i < 8; i++) sum += sh[i];
    }
    const float mean = (float) (sum / (double) ne00);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 v = xr[i], g = ((const float4 *) w)[i];
        const float4 r = make_float4(__fmul_rn(__fmul_rn(v.x, scale), g.x), __fmul_rn(__fmul_rn(v.y, scale), g.y),
                                     __fmul_rn(__fmul_rn(v.z, scale), g.z), __fmul_rn(__fmul_rn(v.w, scale), g.w));
        yr[i] = r;
        if (QUANT) q8_1_soa_emit(r, (int) (i & 7), row, i >> 3, ne00 / 32, qs, ds);
    }
}
cudaError_t launch_rms_norm_mul(const float * x, const float * w, float * y, int64_t ne00, int64_t rows, int64_t xs, int64_t ys, float eps,
                                int8_t * qs, ActDS * ds, cudaStream_t st) {
    if (rows == 0 || ne00 == 0) return cudaSuccess;
    if (rows > 0x7fffffff) return cudaErrorInvalidValue;
    if (qs) return pdl_launch(rms_norm_mul_kernel<true>, dim3((unsigned) rows), dim3(256), st, x, w, y, ne00, xs, ys, eps, qs, ds);
    return pdl_launch(rms_norm_mul_kernel<false>, dim3((unsigned) rows), dim3(256), st, x, w, y, ne00, xs, ys, eps, qs, ds);
}

// ------------------------------------------------------------------------------------------------ ROPE (f32; normal and NeoX pairing)
// one thread per rotated pair (or per passed-through pair beyond n_dims).  theta of pair k = pos * theta_scale^k, built by k successive
// f32 multiplications exactly as the reference's cache loop does (ggml-cpu.c:8638-8647), so the angle is the reference's bit for bit.
// The angle depends on (position, pair) only: one thread computes it once and rotates that pair in ROPE_HEADS heads (blockIdx.y walks the
// head chunks), so the sin / cos and the multiplication chain are not repeated per head.
#define ROPE_HEADS 8
// T = float, or __half (the in-place rotation of f16 K-cache rows when llama.cpp shifts its context: convert, rotate in f32, round back, as
// ggml_compute_forward_rope_f16 does)
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(const RopeArgs p) {
    PDL_PROLOGUE();
    const int64_t half = p.ne[0] / 2;
    const int64_t total = half * p.ne[2] * p.ne[3];
    const int64_t h0 = (int64_t) blockIdx.y * ROPE_HEADS, h1 = h0 + ROPE_HEADS < p.ne[1] ? h0 + ROPE_HEADS : p.ne[1];
    __half * const st = (__half *) (p.st_dyn ? *p.st_dyn : p.st_y);
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i;
        const int64_t k = t % half; t /= half;
        const int64_t i2 = t % p.ne[2]; const int64_t i3 = t / p.ne[2];
        const char * s = p.x + i2 * p.nbs[2] + i3 * p.nbs[3];
        char * d = p.y + i2 * p.nbd[2] + i3 * p.nbd[3];
        const int64_t i0 = 2 * k;
        if (i0 >= p.n_dims) {                                            // channels past n_dims are copied through
            for (int64_t i1 = h0; i1 < h1; i1++) {
                const T * sr = (const T *) (s + i1 * p.nbs[1]); T * dr = (T *) (d + i1 * p.nbd[1]);
                dr[i0] = sr[i0]; dr[i0 + 1] = sr[i0 + 1];
                if (st) {
                    const int64_t row = ((i3 * p.ne[2] + i2) * p.ne[1] + i1) * p.ne[0];
                    st[row + i0] = __float2half_rn(cvt_el<T, float>(sr[i0])); st[row + i0 + 1] = __float2half_rn(cvt_el<T, float>(sr[i0 + 1]));
                }
            }
            continue;
        }
        float theta = (float) p.pos[i2];
        for (int64_t j = 0; j < k; j++) theta = __fmul_rn(theta, p.theta_scale);
        const float theta_extrap = p.ff ? __fdiv_rn(theta, p.ff[k]) : theta;     // (theta / freq_factors[i0 / 2], ggml-cpu.c:8640-8642)
        const float theta_interp = __fmul_rn(p.freq_scale, theta_extrap);
        float th = theta_interp, mscale = p.attn_factor;
        if (p.ext_factor != 0.0f) {
            const float y = ((float) (i0 / 2) - p.corr0) / fmaxf(0.001f, p.corr1 - p.corr0);
            const float ramp = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * p.ext_factor;
            th = theta_interp * (1.0f - ramp) + theta_extrap * ramp;
            mscale *= 1.0f + 0.1f * logf(1.0f / p.freq_scale);
        }
        const float c = __fmul_rn(cosf(th), mscale), sn = __fmul_rn(sinf(th), mscale);
        const int64_t a = p.neox ? k : i0, b = p.neox ? k + p.n_dims / 2 : i0 + 1;
        for (int64_t i1 = h0; i1 < h1; i1++) {
            const T * sr = (const T *) (s + i1 * p.nbs[1]); T * dr = (T *) (d + i1 * p.nbd[1]);
            const float x0 = cvt_el<T, float>(sr[a]), x1 = cvt_el<T, float>(sr[b]);
            const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, sn)), r1 = __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c));
            dr[a] = cvt_el<float, T>(r0);
            dr[b] = cvt_el<float, T>(r1);
            if (st) {                                                    // the f16 cache row, element (a | b, i1, i2, i3) at its logical index
                const int64_t row = ((i3 * p.ne[2] + i2) * p.ne[1] + i1) * p.ne[0];
                st[row + a] = __float2half_rn(r0); st[row + b] = __float2half_rn(r1);
            }
        }
    }
}
cudaError_t launch_rope_f32(const RopeArgs & p, bool f16, cudaStream_t st) {
    const int64_t total = p.ne[0] / 2 * p.ne[2] * p.ne[3];
    if (total == 0 || p.ne[1] == 0) return cudaSuccess;
    const int64_t chunks = (p.ne[1] + ROPE_HEADS - 1) / ROPE_HEADS;
    if (chunks > 65535) return cudaErrorInvalidValue;
    const int block = total < 256 ? 64 : 256;
    int64_t gx = (total + block - 1) / block; if (gx > 148 * 16) gx = 148 * 16;
    if (f16) return pdl_launch(rope_kernel<__half>, dim3((unsigned) gx, (unsigned) chunks), dim3(block), st, p);
    return pdl_launch(rope_kernel<float>, dim3((unsigned) gx, (unsigned) chunks), dim3(block), st, p);
}

// ------------------------------------------------------------------------------------------------ CPY / CONT / DUP
// element i (logical order of src) -> element i (logical order of dst); any byte strides on both sides; f32 / f16 either side
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) cpy_kernel(const CpyArgs p, const int64_t n) {
    PDL_PROLOGUE();
    char * const y = p.dyn ? *p.dyn : p.y;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i;
        const int64_t s0 = t % p.nes[0]; t /= p.nes[0];
        const int64_t s1 = t % p.nes[1]; t /= p.nes[1];
        const int64_t s2 = t % p.nes[2]; const int64_t s3 = t / p.nes[2];
        t = i;
        const int64_t d0 = t % p.ned[0]; t /= p.ned[0];
        const int64_t d1 = t % p.ned[1]; t /= p.ned[1];
        const int64_t d2 = t % p.ned[2]; const int64_t d3 = t / p.ned[2];
        const TS v = *(const TS *) (p.x + s0 * p.nbs[0] + s1 * p.nbs[1] + s2 * p.nbs[2] + s3 * p.nbs[3]);
        *(TD *) (y + d0 * p.nbd[0] + d1 * p.nbd[1] + d2 * p.nbd[2] + d3 * p.nbd[3]) = cvt_el<TS, TD>(v);
    }
}
// fast paths of the shapes a Llama graph stores with (everything else takes the generic kernel above):
//  flat      both sides contiguous in the same logical order (K rows into the cache): 4 elements per thread
//  transpose 2-D, src contiguous along dim 1, dst along dim 0 (V^T into the cache at prefill): 32 x 32 tiles through shared memory
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) cpy_flat4_kernel(const CpyArgs p, const int64_t n4) {
    PDL_PROLOGUE();
    char * const y = p.dyn ? *p.dyn : p.y;
    const bool al = ((uintptr_t) y & (sizeof(TD) * 4 - 1)) == 0;        // (a dynamic destination's alignment is only known now)
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        TS v[4]; TD o[4];
        if (sizeof(TS) == 4) *(float4 *) v = ((const float4 *) p.x)[i]; else *(uint2 *) v = ((const uint2 *) p.x)[i];
        #pragma unroll
        for (int k = 0; k < 4; k++) o[k] = cvt_el<TS, TD>(v[k]);
        if (!al) { for (int k = 0; k < 4; k++) ((TD *) y)[i * 4 + k] = o[k]; }
        else if (sizeof(TD) == 4) ((float4 *) y)[i] = *(float4 *) o; else ((uint2 *) y)[i] = *(uint2 *) o;
    }
}
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) cpy_transpose_kernel(const CpyArgs p) {
    __shared__ float tile[32][33];
    PDL_PROLOGUE();
    char * const y = p.dyn ? *p.dyn : p.y;
    const int64_t i0 = (int64_t) blockIdx.x * 32, i1 = (int64_t) blockIdx.y * 32;      // dims 0 and 1 of the (shared) logical shape
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {           // read: tx walks dim 1 (contiguous in src)
        const int64_t a0 = i0 + r, a1 = i1 + tx;
        if (a0 < p.nes[0] && a1 < p.nes[1]) tile[r][tx] = cvt_el<TS, float>(*(const TS *) (p.x + a0 * p.nbs[0] + a1 * p.nbs[1]));
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {           // write: tx walks dim 0 (contiguous in dst)
        const int64_t a0 = i0 + tx, a1 = i1 + r;
        if (a0 < p.nes[0] && a1 < p.nes[1]) *(TD *) (y + a0 * p.nbd[0] + a1 * p.nbd[1]) = cvt_el<float, TD>(tile[tx][r]);
    }
}
cudaError_t launch_cont_quant(const CpyArgs & p, int64_t K, int8_t * qs, ActDS * ds, cudaStream_t st);
template <typename T> static bool logical_contig(const int64_t * ne, const int64_t * nb) {
    int64_t s = sizeof(T);
    for (int i = 0; i < 4; i++) { if (ne[i] != 1 && nb[i] != s) return false; s *= ne[i]; }
    return true;
}
template <typename TS, typename TD>
static cudaError_t launch_cpy_t(const CpyArgs & p, int64_t n, cudaStream_t st) {
    if (n % 4 == 0 && logical_contig<TS>(p.nes, p.nbs) && logical_contig<TD>(p.ned, p.nbd) && ((uintptr_t) p.x & 15) == 0)
        return pdl_launch(cpy_flat4_kernel<TS, TD>, dim3(ew_blocks(n / 4)), dim3(256), st, p, n / 4);
    if (p.nes[2] == 1 && p.nes[3] == 1 && p.ned[2] == 1 && p.ned[3] == 1 && p.nes[0] == p.ned[0] && p.nes[1] == p.ned[1] &&
        p.nbs[1] == (int64_t) sizeof(TS) && p.nbd[0] == (int64_t) sizeof(TD) && p.nes[0] >= 32 && p.nes[1] >= 32 && (p.nes[1] + 31) / 32 <= 65535)
        return pdl_launch(cpy_transpose_kernel<TS, TD>, dim3((unsigned) ((p.nes[0] + 31) / 32), (unsigned) ((p.nes[1] + 31) / 32)), dim3(256), st, p);
    if (sizeof(TS) == 4 && sizeof(TD) == 4 && !p.dyn && logical_contig<TD>(p.ned, p.nbd) && p.nbs[0] == 4 && p.nes[0] % 4 == 0 &&
        (((uintptr_t) p.x | (uintptr_t) p.y | (uintptr_t) p.nbs[1] | (uintptr_t) p.nbs[2] | (uintptr_t) p.nbs[3]) & 15) == 0)
        return launch_cont_quant(p, 0, nullptr, nullptr, st);        // contiguous destination, rows of src gathered 16 bytes at a time (head merge)
    return pdl_launch(cpy_kernel<TS, TD>, dim3(ew_blocks(n)), dim3(256), st, p, n);
}
cudaError_t launch_cpy(const CpyArgs & p, bool src_f16, bool dst_f16, cudaStream_t st) {
    const int64_t n = p.nes[0] * p.nes[1] * p.nes[2] * p.nes[3];
    if (n == 0) return cudaSuccess;
    if (!src_f16 && !dst_f16) return launch_cpy_t<float, float>(p, n, st);
    if (!src_f16 && dst_f16)  return launch_cpy_t<float, __half>(p, n, st);
    if (src_f16 && !dst_f16)  return launch_cpy_t<__half, float>(p, n, st);
    return launch_cpy_t<__half, __half>(p, n, st);
}

// CONT of f32 into contiguous rows of K elements that a MUL_MAT reads next (the head merge in front of the attention output projection):
// float4 along dim 0 of both sides (src ne[0] % 4 == 0), the row also written as q8_1 (K % 128 == 0)
template <bool QUANT>
__global__ void __launch_bounds__(256) cont_quant_kernel(const CpyArgs p, const int64_t n4, const int64_t K, int8_t * __restrict__ qs, ActDS * __restrict__ ds) {
    PDL_PROLOGUE();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i * 4;
        const int64_t s0 = t % p.nes[0]; t /= p.nes[0];
        const int64_t s1 = t % p.nes[1]; t /= p.nes[1];
        const int64_t s2 = t % p.nes[2]; const int64_t s3 = t / p.nes[2];
        const float4 v = *(const float4 *) (p.x + s0 * 4 + s1 * p.nbs[1] + s2 * p.nbs[2] + s3 * p.nbs[3]);
        ((float4 *) p.y)[i] = v;
        if (QUANT) {
            const int64_t e = i * 4, row = e / K, b32 = (e - row * K) >> 5;
            q8_1_soa_emit(v, (int) (i & 7), row, b32, K / 32, qs, ds);
        }
    }
}
cudaError_t launch_cont_quant(const CpyArgs & p, int64_t K, int8_t * qs, ActDS * ds, cudaStream_t st) {
    const int64_t n = p.nes[0] * p.nes[1] * p.nes[2] * p.nes[3];
    if (n == 0) return cudaSuccess;
    if (qs) return pdl_launch(cont_quant_kernel<true>, dim3(ew_blocks(n / 4)), dim3(256), st, p, n / 4, K, qs, ds);
    return pdl_launch(cont_quant_kernel<false>, dim3(ew_blocks(n / 4)), dim3(256), st, p, n / 4, K, qs, ds);
}

// ------------------------------------------------------------------------------------------------ GET_ROWS (f32 rows, i32 ids)
__global__ void __launch_bounds__(256) get_rows_f32_kernel(const GetRowsArgs p, const int64_t n) {
    PDL_PROLOGUE();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i;
        const int64_t i0 = t % p.ne00; t /= p.ne00;
        const int64_t i10 = t % p.ne10; t /= p.ne10;
        const int64_t i11 = t % p.ne11; const int64_t i12 = t / p.ne11;
        const int64_t r = *(const int32_t *) (p.ids + i10 * p.nbi[0] + i11 * p.nbi[1] + i12 * p.nbi[2]);
        if (r < 0 || r >= p.ne01) continue;                              // (the reference asserts; never write from outside src0)
        *(float *) (p.y + i0 * 4 + i10 * p.nbd[1] + i11 * p.nbd[2] + i12 * p.nbd[3]) =
            *(const float *) (p.x + i0 * 4 + r * p.nbs[1] + i11 * p.nbs[2] + i12 * p.nbs[3]);
    }
}
cudaError_t launch_get_rows_f32(const GetRowsArgs & p, cudaStream_t st) {
    const int64_t n = p.ne00 * p.ne10 * p.ne11 * p.ne12;
    if (n == 0) return cudaSuccess;
    return pdl_launch(get_rows_f32_kernel, dim3(ew_blocks(n)), dim3(256), st, p, n);
}

// ------------------------------------------------------------------------------------------------ MUL_MAT, f16 src0 x f32 src1 -> f32
// dst[m, n, i12, i13] = sum_k src0[k, m, i12 / r2, i13 / r3] * f16(src1[k, n, i12, i13]); K contiguous on both sides, every other stride free
// (K / V cache views: rows n_embd_k_gqa apart, heads interleaved; V transposed with rows n_ctx apart).
__device__ __forceinline__ float rnd_h(float v) { return __half2float(__float2half_rn(v)); }

// few columns (decode): one warp per output row, the row's f16 values read once and used for up to 8 columns
template <int NC>
__global__ void __launch_bounds__(128) mm_f16_vec_kernel(const MmF16Args p) {
    PDL_PROLOGUE();
    const int lane = threadIdx.x & 31;
    const int64_t m = (int64_t) blockIdx.x * 4 + (threadIdx.x >> 5);
    if (m >= p.M) return;
    const int64_t i12 = blockIdx.y % p.ne12, i13 = blockIdx.y / p.ne12;
    const __half * w = (const __half *) (p.a + m * p.nba[1] + (i12 / p.r2) * p.nba[2] + (i13 / p.r3) * p.nba[3]);
    const char * xb = p.b + i12 * p.nbb[2] + i13 * p.nbb[3];
    char * db = p.d + i12 * p.nbd[2] + i13 * p.nbd[3];
    for (int64_t n0 = 0; n0 < p.N; n0 += NC) {
        float acc[NC];
        #pragma unroll
        for (int c = 0; c < NC; c++) acc[c] = 0.f;
        const bool vec = ((uintptr_t) w & 3) == 0 && ((((uintptr_t) xb) | (uintptr_t) p.nbb[1]) & 7) == 0 && p.K % 2 == 0;
        if (vec) {
            for (int64_t k2 = lane; k2 < p.K / 2; k2 += 32) {
                const float2 wf = __half22float2(((const __half2 *) w)[k2]);
                #pragma unroll
                for (int c = 0; c < NC; c++) {
                    if (n0 + c < p.N) {
                        const float2 xv = *(const float2 *) (xb + (n0 + c) * p.nbb[1] + k2 * 8);
                        acc[c] = fmaf(wf.x, rnd_h(xv.x), acc[c]); acc[c] = fmaf(wf.y, rnd_h(xv.y), acc[c]);
                    }
                }
            }
        } else {
            for (int64_t k = lane; k < p.K; k += 32) {
                const float wf = __half2float(w[k]);
                #pragma unroll
                for (int c = 0; c < NC; c++)
                    if (n0 + c < p.N) acc[c] = fmaf(wf, rnd_h(*(const float *) (xb + (n0 + c) * p.nbb[1] + k * 4)), acc[c]);
            }
        }
        #pragma unroll
        for (int c = 0; c < NC; c++) {
            float s = acc[c];
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && n0 + c < p.N) *(float *) (db + (n0 + c) * p.nbd[1] + m * 4) = s;
        }
    }
}

// many columns (prefill): 128 x 128 output tile per CTA of 8 warps (4 x 2, each 32 x 64 = 2 x 4 HMMA fragments with f32 accumulators), K in steps
// of 32 staged through shared memory as f16 -- src1 is rounded to f16 on the way in, which is what the reference does before its dot products
// -- with the next step's 16-byte global loads in flight while the current one is multiplied.  This is the attention GEMM beside the hot path
// (2 x 2.1 GFLOP per layer at pp512 against 26 GFLOP of weight GEMMs), not the weight GEMM.
#define MMT 128
#define MMK 32
#define MMLD (MMK + 8)
struct MmStage { uint4 a[2]; uint4 b[2]; };
__device__ __forceinline__ uint4 mm_load_a(const char * row, const int64_t k, const int64_t K, const bool ok, const bool vec) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (!ok || k >= K) return r;
    if (vec && k + 8 <= K) return *(const uint4 *) (row + k * 2);
    __half h[8];
    #pragma unroll
    for (int i = 0; i < 8; i++) h[i] = k + i < K ? *(const __half *) (row + (k + i) * 2) : __float2half_rn(0.f);
    return *(uint4 *) h;
}
__device__ __forceinline__ uint4 mm_load_b(const char * row, const int64_t k, const int64_t K, const bool ok, const bool vec) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (!ok || k >= K) return r;
    float f[8];
    if (vec && k + 8 <= K) { *(float4 *) f = *(const float4 *) (row + k * 4); *(float4 *) (f + 4) = *(const float4 *) (row + k * 4 + 16); }
    else {
        #pragma unroll
        for (int i = 0; i < 8; i++) f[i] = k + i < K ? *(const float *) (row + (k + i) * 4) : 0.f;
    }
    __half2 h[4];
    #pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    return *(uint4 *) h;
}
// TN = columns per tile: 128, or 64 where 128-wide tiles would leave SMs without a CTA (V p at prefill: M = head size)
template <int TN>
__global__ void __launch_bounds__(256) mm_f16_tile_kernel(const MmF16Args p) {
    using namespace nvcuda;
    __shared__ __align__(32) __half As[MMT * MMLD];
    __shared__ __align__(32) __half Bs[TN * MMLD];
    constexpr int NB = TN / 64;            // B chunks per thread per K step (TN rows x 4 chunks / 256 threads)
    constexpr int NF = TN / 32;            // 16-wide fragments per warp along N
    PDL_PROLOGUE();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t m0 = (int64_t) blockIdx.x * MMT, n0 = (int64_t) blockIdx.y * TN;
    const int64_t i12 = blockIdx.z % p.ne12, i13 = blockIdx.z / p.ne12;
    const char * ab = p.a + (i12 / p.r2) * p.nba[2] + (i13 / p.r3) * p.nba[3];
    const char * bb = p.b + i12 * p.nbb[2] + i13 * p.nbb[3];
    char * db = p.d + i12 * p.nbd[2] + i13 * p.nbd[3];
    const bool veca = (((uintptr_t) ab | (uintptr_t) p.nba[1]) & 15) == 0, vecb = (((uintptr_t) bb | (uintptr_t) p.nbb[1]) & 15) == 0;
    // each thread moves two 8-element chunks of each operand per K step: chunk c -> row c / 4, k offset (c % 4) * 8
    int rr[2], kc[2]; const char * arow[2]; const char * brow[2]; bool aok[2], bok[2];
    #pragma unroll
    for (int j = 0; j < 2; j++) {
        const int c = threadIdx.x + j * 256;
        rr[j] = c >> 2; kc[j] = (c & 3) * 8;
        aok[j] = m0 + rr[j] < p.M; bok[j] = j < NB && n0 + rr[j] < p.N;
        arow[j] = ab + (m0 + rr[j]) * p.nba[1]; brow[j] = bb + (n0 + rr[j]) * p.nbb[1];
    }
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2][NF];
    #pragma unroll
    for (int i = 0; i < 2; i++)
        #pragma unroll
        for (int j = 0; j < NF; j++) wmma::fill_fragment(acc[i][j], 0.0f);
    const int wm = (warp >> 1) * 32, wn = (warp & 1) * (TN / 2);
    MmStage st;
    #pragma unroll
    for (int j = 0; j < 2; j++) { st.a[j] = mm_load_a(arow[j], kc[j], p.K, aok[j], veca); if (j < NB) st.b[j] = mm_load_b(brow[j], kc[j], p.K, bok[j], vecb); }
    for (int64_t k0 = 0; k0 < p.K; k0 += MMK) {
        #pragma unroll
        for (int j = 0; j < 2; j++) { *(uint4 *) (As + rr[j] * MMLD + kc[j]) = st.a[j]; if (j < NB) *(uint4 *) (Bs + rr[j] * MMLD + kc[j]) = st.b[j]; }
        __syncthreads();
        if (k0 + MMK < p.K) {
            #pragma unroll
            for (int j = 0; j < 2; j++) { st.a[j] = mm_load_a(arow[j], k0 + MMK + kc[j], p.K, aok[j], veca); if (j < NB) st.b[j] = mm_load_b(brow[j], k0 + MMK + kc[j], p.K, bok[j], vecb); }
        }
        #pragma unroll
        for (int kk = 0; kk < MMK; kk += 16) {
            wmma::fragment<wmma::matrix_a, 16, 16, 16, __half, wmma::row_major> fa[2];
            wmma::fragment<wmma::matrix_b, 16, 16, 16, __half, wmma::col_major> fb[NF];
            #pragma unroll
            for (int i = 0; i < 2; i++) wmma::load_matrix_sync(fa[i], As + (wm + i * 16) * MMLD + kk, MMLD);
            #pragma unroll
            for (int j = 0; j < NF; j++) wmma::load_matrix_sync(fb[j], Bs + (wn + j * 16) * MMLD + kk, MMLD);
            #pragma unroll
            for (int i = 0; i < 2; i++)
                #pragma unroll
                for (int j = 0; j < NF; j++) wmma::mma_sync(acc[i][j], fa[i], fb[j], acc[i][j]);
        }
        __syncthreads();
    }
    // C fragment (m, n) -> dst[n][m]: staged per warp (in the A tile's memory, free now) so that the store walks m, contiguous in dst, and respects the edges
    float * Cs = (float *) As + warp * 256;
    #pragma unroll
    for (int i = 0; i < 2; i++)
        #pragma unroll
        for (int j = 0; j < NF; j++) {
            wmma::store_matrix_sync(Cs, acc[i][j], 16, wmma::mem_col_major);      // Cs[n * 16 + m]
            __syncwarp();
            for (int e = lane; e < 256; e += 32) {
                const int nn = e >> 4, mm = e & 15;
                const int64_t m = m0 + wm + i * 16 + mm, n = n0 + wn + j * 16 + nn;
                if (m < p.M && n < p.N) *(float *) (db + n * p.nbd[1] + m * 4) = Cs[e];
            }
            __syncwarp();
        }
}
// ------------------------------------------------------------------------------------------------ one-column attention (decode)
// K^T q -> soft_max(scale * . + mask) -> V p for ONE query column, one CTA of 8 warps per head: the three launches a decode step spends per
// layer on attention become one.  Every loop is the loop of the kernel it replaces (mm_f16_vec_kernel<1> for both products, the block
// soft_max of norm_ops.cu), same lane / thread assignment and reduction order, so the results are the unfused ones bit for bit; the
// probabilities are written out too (soft_max ran in place on the scores, so they are what the graph can observe).
__global__ void __launch_bounds__(1024) attn_decode_kernel(const AttnDecodeArgs p) {
    extern __shared__ float prob[];                  // n_kv scores -> probabilities
    __shared__ double shd[8];
    __shared__ float shf[8];
    PDL_PROLOGUE();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int64_t h = blockIdx.x, hk = h / p.r2;
    const char * kb = p.k + hk * p.nbk2;
    const char * vb = p.v + hk * p.nbv2;
    const char * qb = p.q + h * p.nbq2;
    float * pr = (float *) (p.prob + h * p.nbp2);
    // scores: row j of K (hd f16, contiguous) . q -- 4 rows in flight per warp, each row's sum in the order of mm_f16_vec_kernel<1>
    for (int64_t j0 = (int64_t) warp * 4; j0 < p.n_kv; j0 += (int64_t) nw * 4) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int64_t k2 = lane; k2 < p.hd / 2; k2 += 32) {
            const float2 xv = *(const float2 *) (qb + k2 * 8);
            const float x0 = rnd_h(xv.x), x1 = rnd_h(xv.y);
            #pragma unroll
            for (int r = 0; r < 4; r++) {
                if (j0 + r < p.n_kv) {
                    const float2 wf = __half22float2(((const __half2 *) (kb + (j0 + r) * p.nbk1))[k2]);
                    acc[r] = fmaf(wf.x, x0, acc[r]); acc[r] = fmaf(wf.y, x1, acc[r]);
                }
            }
        }
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            float a = acc[r];
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0 && j0 + r < p.n_kv) prob[j0 + r] = a;
        }
    }
    __syncthreads();
    // soft_max over the n_kv scores (max_bias = 0: slope 1): the first 256 threads in the roles of the block kernel's 256
    const float * mf = p.mask;
    const bool act = threadIdx.x < 256;
    auto wp = [&](int64_t i) { float v = __fmul_rn(prob[i], p.scale); if (mf) v = __fadd_rn(v, __fmul_rn(1.0f, mf[i])); return v; };
    float mx = -CUDART_INF_F;
    if (act) for (int64_t i = threadIdx.x; i < p.n_kv; i += 256) mx = fmaxf(mx, wp(i));
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const float w = __shfl_xor_sync(0xffffffffu, mx, o); mx = w > mx ? w : mx; }
    if (lane == 0 && warp < 8) shf[warp] = mx;
    __syncthreads();
    mx = shf[0];
    for (int i = 1; i < 8; i++) mx = shf[i] > mx ? shf[i] : mx;
    double sum = 0.0;
    float e[AD_MAX_PER_THREAD];
    int ne = 0;
    if (act) for (int64_t i = threadIdx.x; i < p.n_kv; i += 256) { const float v = expf(__fsub_rn(wp(i), mx)); e[ne++] = v; sum += (double) v; }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0 && warp < 8) shd[warp] = sum;
    __syncthreads();                                  // (also: every score has been read before the probabilities overwrite them)
    sum = shd[0];
    for (int i = 1; i < 8; i++) sum += shd[i];
    const float inv = (float) (1.0 / sum);
    ne = 0;
    if (act) for (int64_t i = threadIdx.x; i < p.n_kv; i += 256) { const float v = __fmul_rn(e[ne++], inv); prob[i] = v; pr[i] = v; }
    __syncthreads();
    // V p: row d of V^T (n_kv f16, contiguous) . probabilities, 4 rows in flight per warp
    float * out = (float *) (p.out + h * p.nbo2);
    for (int64_t d0 = (int64_t) warp * 4; d0 < p.hd; d0 += (int64_t) nw * 4) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int64_t k2 = lane; k2 < p.n_kv / 2; k2 += 32) {
            const float x0 = rnd_h(prob[2 * k2]), x1 = rnd_h(prob[2 * k2 + 1]);
            #pragma unroll
            for (int r = 0; r < 4; r++) {
                if (d0 + r < p.hd) {
                    const float2 wf = __half22float2(((const __half2 *) (vb + (d0 + r) * p.nbv1))[k2]);
                    acc[r] = fmaf(wf.x, x0, acc[r]); acc[r] = fmaf(wf.y, x1, acc[r]);
                }
            }
        }
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            float a = acc[r];
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0 && d0 + r < p.hd) out[d0 + r] = a;
        }
    }
}
cudaError_t launch_attn_decode(const AttnDecodeArgs & p, int n_head, cudaStream_t st) {
    if (n_head == 0 || p.n_kv == 0) return cudaSuccess;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) n_head); cfg.blockDim = dim3(1024); cfg.stream = st; cfg.dynamicSmemBytes = (size_t) p.n_kv * sizeof(float);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, attn_decode_kernel, p);
}

cudaError_t launch_mm_f16(const MmF16Args & p, cudaStream_t st) {
    const int64_t batch = p.ne12 * p.ne13;
    if (p.M == 0 || p.N == 0 || batch == 0) return cudaSuccess;
    if (batch > 65535) return cudaErrorInvalidValue;
    if (p.N <= 16) {
        dim3 grid((unsigned) ((p.M + 3) / 4), (unsigned) batch);
        if (p.N == 1) return pdl_launch(mm_f16_vec_kernel<1>, grid, dim3(128), st, p);
        if (p.N <= 4) return pdl_launch(mm_f16_vec_kernel<4>, grid, dim3(128), st, p);
        return pdl_launch(mm_f16_vec_kernel<8>, grid, dim3(128), st, p);
    } else {
        const int64_t gx = (p.M + MMT - 1) / MMT;
        const bool narrow = gx * ((p.N + MMT - 1) / MMT) * batch < 2 * 148;        // fewer than two CTAs per SM with 128-wide tiles
        const int64_t gy = (p.N + (narrow ? 64 : MMT) - 1) / (narrow ? 64 : MMT);
        if (gy > 65535) return cudaErrorInvalidValue;
        dim3 grid((unsigned) gx, (unsigned) gy, (unsigned) batch);
        if (narrow) return pdl_launch(mm_f16_tile_kernel<64>, grid, dim3(256), st, p);
        return pdl_launch(mm_f16_tile_kernel<128>, grid, dim3(256), st, p);
    }
    return cudaGetLastError();
}

}  // namespace b200

// kernels.cu — hand-written sm_100a kernels of the MUL_MAT / ADD hot path (decode side).
//
//   quantize_q8_1   : f32 activations -> int8 + scale   (= quantize_row_q8_1_ref, ggml-quants.c:220-253)
//   gemv_q          : dequant-fused GEMV, 1..8 activation columns, q4_0 / q8_0 / q4_K / q6_K weights
//                     (takes the seat of ggml_compute_forward_mul_mat_one_chunk + vec_dot, ggml-dsp.c:1091-1189,
//                      ggml-cpu-quants.c:2591-2604 / :4004-4012 / :7535-7590 / :9423-9465)
//   add_f32         : dst = a + b with ggml broadcast       (ggml-dsp.c:951-1066, binary-ops.cpp:50-117)
//   repack          : native ggml blocks <-> SoA HBM layout (b200_internal.cuh)
//
// All integer work (per-group dot products) is exact; float combines follow oracle_vec_dot_q8_1.
#include "b200_internal.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------ helpers
__device__ __forceinline__ uint4 ldg_stream(const void * p) {     // weights: read once, do not pollute L1
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ldg_stream8(const void * p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldg_stream_h(const void * p) {
    unsigned short v;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(v) : "l"(p));
    return __half2float(__ushort_as_half(v));
}
__device__ __forceinline__ uint4 ldg_cached(const void * p) {     // activations: re-read by every warp, keep in L1
    return __ldg((const uint4 *) p);
}
__device__ __forceinline__ int dp4a_ss(uint32_t a, uint32_t b, int c) { return __dp4a((int) a, (int) b, c); }   // s8 x s8
// weights unpacked to 0..63 are also valid non-negative s8, so the signed form serves both cases
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) { return __dp4a((int) a, (int) b, c); }

__device__ __forceinline__ int dot16_us(const uint4 & w, const uint4 & a) {   // 16 x (u8 * s8)
    int s = dp4a_us(w.x, a.x, 0); s = dp4a_us(w.y, a.y, s); s = dp4a_us(w.z, a.z, s); return dp4a_us(w.w, a.w, s);
}
__device__ __forceinline__ int dot16_ss(const uint4 & w, const uint4 & a) {   // 16 x (s8 * s8)
    int s = dp4a_ss(w.x, a.x, 0); s = dp4a_ss(w.y, a.y, s); s = dp4a_ss(w.z, a.z, s); return dp4a_ss(w.w, a.w, s);
}
__device__ __forceinline__ uint4 lo_nib(const uint4 & q) {
    return make_uint4(q.x & 0x0F0F0F0Fu, q.y & 0x0F0F0F0Fu, q.z & 0x0F0F0F0Fu, q.w & 0x0F0F0F0Fu);
}
__device__ __forceinline__ uint4 hi_nib(const uint4 & q) {
    return make_uint4((q.x >> 4) & 0x0F0F0F0Fu, (q.y >> 4) & 0x0F0F0F0Fu, (q.z >> 4) & 0x0F0F0F0Fu, (q.w >> 4) & 0x0F0F0F0Fu);
}

// activation accessors on the two-plane scratch (b200_internal.cuh)
struct ActCol {
    const uint4 * p0; const uint4 * p1; const int2 * ds;
    __device__ __forceinline__ ActCol(const ActView & v, int c) {
        const int8_t * q = v.qs + (int64_t) c * v.NB * 32;
        p0 = (const uint4 *) q; p1 = p0 + v.NB; ds = (const int2 *) (v.ds + (int64_t) c * v.NB);
    }
};
struct ADesc { float d; int isum0, isum1; };
__device__ __forceinline__ ADesc load_adesc(const int2 * p) {
    int2 v = __ldg(p);
    ADesc a; a.d = __int_as_float(v.x); a.isum0 = (short) (v.y & 0xFFFF); a.isum1 = v.y >> 16;
    return a;
}

// ------------------------------------------------------------------------------------------------ weight types
// Every type defines: STEP (elements one lane consumes per step), Frag (the weight registers of a step),
// load(w, step index within the tensor), idots(frag, act..., out ints) and dot(frag, act col, step in row) -> float.

template <int TYPE> struct WT;

// ---- q4_0: one step = one 32-element block.  sumi = sum (nib-8)*a = sum nib*a - 8*isum  (ggml-cpu-quants.c:2596-2603)
template <> struct WT<T_Q4_0> {
    static constexpr int STEP = 32;
    static constexpr int NINTS = 1;
    struct Frag { uint4 q; float d; };
    static __device__ __forceinline__ Frag load(const WPtrs & w, int64_t blk) {
        Frag f; f.q = ldg_stream(w.p[0] + blk * 16); f.d = ldg_stream_h(w.p[1] + blk * 2); return f;
    }
    static __device__ __forceinline__ void idots(const Frag & f, const ActCol & a, int s, int * out) {
        const uint4 a0 = ldg_cached(a.p0 + s), a1 = ldg_cached(a.p1 + s);
        const ADesc  ad = load_adesc(a.ds + s);
        out[0] = dot16_us(lo_nib(f.q), a0) + dot16_us(hi_nib(f.q), a1) - 8 * (ad.isum0 + ad.isum1);
    }
    static __device__ __forceinline__ float dot(const Frag & f, const ActCol & a, int s) {
        const uint4 a0 = ldg_cached(a.p0 + s), a1 = ldg_cached(a.p1 + s);
        const ADesc  ad = load_adesc(a.ds + s);
        const int sumi = dot16_us(lo_nib(f.q), a0) + dot16_us(hi_nib(f.q), a1) - 8 * (ad.isum0 + ad.isum1);
        return (f.d * ad.d) * (float) sumi;
    }
};

// ---- q8_0: one step = one 32-element block (ggml-cpu-quants.c:4004-4012)
template <> struct WT<T_Q8_0> {
    static constexpr int STEP = 32;
    static constexpr int NINTS = 1;
    struct Frag { uint4 q0, q1; float d; };
    static __device__ __forceinline__ Frag load(const WPtrs & w, int64_t blk) {
        Frag f; f.q0 = ldg_stream(w.p[0] + blk * 32); f.q1 = ldg_stream(w.p[0] + blk * 32 + 16);
        f.d = ldg_stream_h(w.p[1] + blk * 2); return f;
    }
    static __device__ __forceinline__ void idots(const Frag & f, const ActCol & a, int s, int * out) {
        out[0] = dot16_ss(f.q0, ldg_cached(a.p0 + s)) + dot16_ss(f.q1, ldg_cached(a.p1 + s));
    }
    static __device__ __forceinline__ float dot(const Frag & f, const ActCol & a, int s) {
        const int sumi = dot16_ss(f.q0, ldg_cached(a.p0 + s)) + dot16_ss(f.q1, ldg_cached(a.p1 + s));
        const ADesc ad = load_adesc(a.ds + s);
        return (f.d * ad.d) * (float) sumi;
    }
};

// ---- q4_K: one step = 64 elements = 32 bytes of nibbles = groups 2c (low nibbles) and 2c+1 (high nibbles)
//      of super-block sb = step/4, c = step%4   (layout ggml-quants.c:1292-1299; dot ggml-cpu-quants.c:7553-7588;
//      q8_1 algebra ggml-cuda/vecdotq.cuh:357-379)
__device__ __forceinline__ uint32_t sel3(uint32_t w0, uint32_t w1, uint32_t w2, int i) {   // byte i of 12
    const uint32_t w = i < 4 ? w0 : (i < 8 ? w1 : w2);
    return (w >> ((i & 3) * 8)) & 0xFFu;
}
__device__ __forceinline__ void k4_scale_min(uint32_t w0, uint32_t w1, uint32_t w2, int j, int & sc, int & mn) {
    // get_scale_min_k4, ggml-quants.c:631-638
    if (j < 4) { sc = sel3(w0, w1, w2, j) & 63; mn = sel3(w0, w1, w2, j + 4) & 63; }
    else {
        const uint32_t hi = sel3(w0, w1, w2, j + 4);
        sc = (hi & 0xF) | ((sel3(w0, w1, w2, j - 4) >> 6) << 4);
        mn = (hi >> 4)  | ((sel3(w0, w1, w2, j) >> 6) << 4);
    }
}
template <> struct WT<T_Q4_K> {
    static constexpr int STEP = 64;
    static constexpr int NINTS = 2;
    struct Frag { uint4 q0, q1; uint4 hdr; };
    static __device__ __forceinline__ Frag load(const WPtrs & w, int64_t step) {
        Frag f;
        f.q0 = ldg_stream(w.p[0] + step * 32); f.q1 = ldg_stream(w.p[0] + step * 32 + 16);
        f.hdr = __ldg((const uint4 *) (w.p[1] + (step >> 2) * 16));     // shared by the 4 lanes of a super-block
        return f;
    }
    static __device__ __forceinline__ void idots(const Frag & f, const ActCol & a, int s, int * out) {
        const int b0 = 2 * s, b1 = 2 * s + 1;
        out[0] = dot16_us(lo_nib(f.q0), ldg_cached(a.p0 + b0)) + dot16_us(lo_nib(f.q1), ldg_cached(a.p1 + b0));
        out[1] = dot16_us(hi_nib(f.q0), ldg_cached(a.p0 + b1)) + dot16_us(hi_nib(f.q1), ldg_cached(a.p1 + b1));
    }
    static __device__ __forceinline__ float dot(const Frag & f, const ActCol & a, int s) {
        int d2[2]; idots(f, a, s, d2);
        const int c = s & 3;
        int sc0, m0, sc1, m1;
        k4_scale_min(f.hdr.y, f.hdr.z, f.hdr.w, 2 * c,     sc0, m0);
        k4_scale_min(f.hdr.y, f.hdr.z, f.hdr.w, 2 * c + 1, sc1, m1);
        const ADesc a0 = load_adesc(a.ds + 2 * s), a1 = load_adesc(a.ds + 2 * s + 1);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&f.hdr.x));
        const float sd = a0.d * (float) (d2[0] * sc0) + a1.d * (float) (d2[1] * sc1);
        const float sm = a0.d * (float) ((a0.isum0 + a0.isum1) * m0) + a1.d * (float) ((a1.isum0 + a1.isum1) * m1);
        return dm.x * sd - dm.y * sm;
    }
};

// ---- q6_K: one step = 64 elements: sb = step/4, half h = (step%4)/2, j = step%2; four 16-element runs
//      r = 0..3 at element 128h + 32r + 16j, each exactly one scale group (index 8h + 2r + j)
//      (layout ggml-quants.c:1701-1717; dot ggml-cpu-quants.c:9438-9460; q8_1 algebra vecdotq.cuh:476-496)
template <> struct WT<T_Q6_K> {
    static constexpr int STEP = 64;
    static constexpr int NINTS = 4;
    struct Frag { uint4 l0, l1, hq; uint2 sc; float d; };
    static __device__ __forceinline__ Frag load(const WPtrs & w, int64_t step) {
        const int64_t sb = step >> 2; const int h = (int) (step & 3) >> 1, j = (int) step & 1;
        Frag f;
        f.l0 = ldg_stream(w.p[0] + sb * 128 + 64 * h + 16 * j);
        f.l1 = ldg_stream(w.p[0] + sb * 128 + 64 * h + 32 + 16 * j);
        f.hq = ldg_stream(w.p[1] + sb * 64 + 32 * h + 16 * j);
        f.sc = ldg_stream8(w.p[2] + sb * 16 + 8 * h);
        f.d  = ldg_stream_h(w.p[3] + sb * 2);
        return f;
    }
    static __device__ __forceinline__ void idots(const Frag & f, const ActCol & a, int s, int * out) {
        const int sbh = (s >> 1) * 4;            // first activation block of this half: sb*8 + 4h
        const int j = s & 1;
        const uint4 * pl = j ? a.p1 : a.p0;
        #pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint4 & l = (r & 1) ? f.l1 : f.l0;
            uint4 v;
            if (r < 2) v = lo_nib(l); else v = hi_nib(l);
            // two high bits: qh >> 2r, moved to bit 4..5 of each byte
            if (r == 0) { v.x |= (f.hq.x << 4) & 0x30303030u; v.y |= (f.hq.y << 4) & 0x30303030u; v.z |= (f.hq.z << 4) & 0x30303030u; v.w |= (f.hq.w << 4) & 0x30303030u; }
            if (r == 1) { v.x |= (f.hq.x << 2) & 0x30303030u; v.y |= (f.hq.y << 2) & 0x30303030u; v.z |= (f.hq.z << 2) & 0x30303030u; v.w |= (f.hq.w << 2) & 0x30303030u; }
            if (r == 2) { v.x |= (f.hq.x     ) & 0x30303030u; v.y |= (f.hq.y     ) & 0x30303030u; v.z |= (f.hq.z     ) & 0x30303030u; v.w |= (f.hq.w     ) & 0x30303030u; }
            if (r == 3) { v.x |= (f.hq.x >> 2) & 0x30303030u; v.y |= (f.hq.y >> 2) & 0x30303030u; v.z |= (f.hq.z >> 2) & 0x30303030u; v.w |= (f.hq.w >> 2) & 0x30303030u; }
            const ADesc ad = load_adesc(a.ds + sbh + r);
            out[r] = dot16_us(v, ldg_cached(pl + sbh + r)) - 32 * (j ? ad.isum1 : ad.isum0);
        }
    }
    static __device__ __forceinline__ float dot(const Frag & f, const ActCol & a, int s) {
        int d4[4]; idots(f, a, s, d4);
        const int sbh = (s >> 1) * 4, j = s & 1;
        // scale bytes j, j+2, j+4, j+6 of the 8 loaded
        const uint32_t lo = f.sc.x >> (8 * j), hi = f.sc.y >> (8 * j);
        const int sc0 = (int8_t) (lo & 0xFF), sc1 = (int8_t) ((lo >> 16) & 0xFF);
        const int sc2 = (int8_t) (hi & 0xFF), sc3 = (int8_t) ((hi >> 16) & 0xFF);
        float sd;
        sd  = __int_as_float(__ldg(a.ds + sbh + 0).x) * (float) (d4[0] * sc0);
        sd += __int_as_float(__ldg(a.ds + sbh + 1).x) * (float) (d4[1] * sc1);
        sd += __int_as_float(__ldg(a.ds + sbh + 2).x) * (float) (d4[2] * sc2);
        sd += __int_as_float(__ldg(a.ds + sbh + 3).x) * (float) (d4[3] * sc3);
        return f.d * sd;
    }
};

// ------------------------------------------------------------------------------------------------ GEMV
// One warp owns RPW consecutive weight rows; its lanes stride over the steps of the row, so consecutive lanes
// read consecutive 16 B of the quant array.  Weights of the next iteration are requested before the current one
// is consumed (register double buffer).  Grid = ceil(M / (WARPS*RPW)).
constexpr int GEMV_WARPS = 4;

template <int TYPE, int NCOLS, int RPW>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
gemv_q_kernel(const WPtrs w, const int64_t M, const int steps, const ActView act,
              float * __restrict__ dst, const int64_t dst_col_stride) {
    using W = WT<TYPE>;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row0 = ((int64_t) blockIdx.x * GEMV_WARPS + warp) * RPW;
    if (row0 >= M) return;

    int64_t rbase[RPW];
    #pragma unroll
    for (int r = 0; r < RPW; r++) rbase[r] = (row0 + r < M ? row0 + r : M - 1) * (int64_t) steps;

    float acc[RPW][NCOLS];
    #pragma unroll
    for (int r = 0; r < RPW; r++)
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) acc[r][c] = 0.f;

    typename W::Frag cur[RPW], nxt[RPW];
    int s = lane;
    if (s < steps) {
        #pragma unroll
        for (int r = 0; r < RPW; r++) cur[r] = W::load(w, rbase[r] + s);
    }
#if __CUDA_ARCH__ >= 900
    // weights do not depend on the previous kernel; activations do (programmatic dependent launch).
    // Trigger right away: the next kernel may start fetching ITS weights while we compute.
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    for (; s < steps; s += 32) {
        const bool more = s + 32 < steps;
        if (more) {
            #pragma unroll
            for (int r = 0; r < RPW; r++) nxt[r] = W::load(w, rbase[r] + s + 32);
        }
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const ActCol a(act, c);
            #pragma unroll
            for (int r = 0; r < RPW; r++) acc[r][c] += W::dot(cur[r], a, s);
        }
        if (more) {
            #pragma unroll
            for (int r = 0; r < RPW; r++) cur[r] = nxt[r];
        }
    }
    #pragma unroll
    for (int r = 0; r < RPW; r++)
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            float v = acc[r][c];
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && row0 + r < M) dst[(int64_t) c * dst_col_stride + row0 + r] = v;
        }
}

template <int TYPE, int NCOLS, int RPW>
static cudaError_t launch_gemv_t(WPtrs w, int64_t M, int64_t K, ActView act, float * dst, int64_t dcs,
                                 cudaStream_t st, bool pdl) {
    const int steps = (int) (K / WT<TYPE>::STEP);
    const int64_t rows_per_cta = GEMV_WARPS * RPW;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) ((M + rows_per_cta - 1) / rows_per_cta));
    cfg.blockDim = dim3(GEMV_WARPS * 32);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, gemv_q_kernel<TYPE, NCOLS, RPW>, w, M, steps, act, dst, dcs);
}

template <int TYPE>
static cudaError_t launch_gemv_type(WPtrs w, int64_t M, int64_t K, ActView act, int ncols, float * dst, int64_t dcs,
                                    cudaStream_t st, bool pdl) {
    switch (ncols) {
        case 1: return launch_gemv_t<TYPE, 1, 2>(w, M, K, act, dst, dcs, st, pdl);
        case 2: return launch_gemv_t<TYPE, 2, 2>(w, M, K, act, dst, dcs, st, pdl);
        case 3: return launch_gemv_t<TYPE, 3, 2>(w, M, K, act, dst, dcs, st, pdl);
        case 4: return launch_gemv_t<TYPE, 4, 2>(w, M, K, act, dst, dcs, st, pdl);
        case 5: return launch_gemv_t<TYPE, 5, 1>(w, M, K, act, dst, dcs, st, pdl);
        case 6: return launch_gemv_t<TYPE, 6, 1>(w, M, K, act, dst, dcs, st, pdl);
        case 7: return launch_gemv_t<TYPE, 7, 1>(w, M, K, act, dst, dcs, st, pdl);
        case 8: return launch_gemv_t<TYPE, 8, 1>(w, M, K, act, dst, dcs, st, pdl);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_gemv_q(int type, WPtrs w, int64_t M, int64_t K, ActView act, int ncols,
                          float * dst, int64_t dcs, cudaStream_t st, bool pdl) {
    switch (type) {
        case T_Q4_0: return launch_gemv_type<T_Q4_0>(w, M, K, act, ncols, dst, dcs, st, pdl);
        case T_Q8_0: return launch_gemv_type<T_Q8_0>(w, M, K, act, ncols, dst, dcs, st, pdl);
        case T_Q4_K: return launch_gemv_type<T_Q4_K>(w, M, K, act, ncols, dst, dcs, st, pdl);
        case T_Q6_K: return launch_gemv_type<T_Q6_K>(w, M, K, act, ncols, dst, dcs, st, pdl);
    }
    return cudaErrorInvalidValue;
}

// ---- f32 weights (the only MUL_MAT the cDSP path runs by default: ggml_vec_dot_f32, ggml-dsp.c:360-372).
// One warp per (row, column); float4 loads when 16-B aligned; fp32 accumulation.
__global__ void __launch_bounds__(128)
gemv_f32_kernel(const char * __restrict__ w, int64_t wrs, int64_t M, int64_t K,
                const char * __restrict__ x, int64_t xcs, int ncols, float * __restrict__ dst, int64_t dcs) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t) blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= M) return;
    const float * wr = (const float *) (w + row * wrs);
    for (int c = 0; c < ncols; c++) {
        const float * xc = (const float *) (x + c * xcs);
        float s = 0.f;
        const bool vec = ((((uintptr_t) wr) | ((uintptr_t) xc)) & 15) == 0;
        int64_t i0 = 0;
        if (vec) {
            const int64_t k4 = K / 4;
            for (int64_t i = lane; i < k4; i += 32) {
                const float4 a = __ldg((const float4 *) wr + i), b = __ldg((const float4 *) xc + i);
                s += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
            }
            i0 = k4 * 4;
        }
        for (int64_t i = i0 + lane; i < K; i += 32) s += wr[i] * xc[i];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) dst[(int64_t) c * dcs + row] = s;
    }
}
cudaError_t launch_gemv_f32(const float * w, int64_t wrs, int64_t M, int64_t K, const float * x, int64_t xcs, int ncols,
                            float * dst, int64_t dcs, cudaStream_t st) {
    gemv_f32_kernel<<<(unsigned) ((M + 3) / 4), 128, 0, st>>>((const char *) w, wrs, M, K, (const char *) x, xcs, ncols, dst, dcs);
    return cudaGetLastError();
}

// ---- exact integer group dots of ONE row (parity instrument; same idots() code as the GEMV)
template <int TYPE>
__global__ void block_dots_kernel(const WPtrs w, const int steps, const int8_t * __restrict__ a_qs, int32_t * __restrict__ out,
                                  int8_t * scratch_qs, ActDS * scratch_ds) {
    using W = WT<TYPE>;
    // build the two-plane view of the plain int8[k] activations in scratch (d = 1, exact isums)
    const int NB = steps * W::STEP / 32;
    for (int b = threadIdx.x; b < NB; b += blockDim.x) {
        int s0 = 0, s1 = 0;
        for (int i = 0; i < 16; i++) {
            const int8_t v0 = a_qs[b * 32 + i], v1 = a_qs[b * 32 + 16 + i];
            scratch_qs[b * 16 + i] = v0; scratch_qs[(NB + b) * 16 + i] = v1; s0 += v0; s1 += v1;
        }
        scratch_ds[b].d = 1.f; scratch_ds[b].isum0 = (short) s0; scratch_ds[b].isum1 = (short) s1;
    }
    __syncthreads();
    ActView av{scratch_qs, scratch_ds, NB};
    const ActCol a(av, 0);
    for (int s = threadIdx.x; s < steps; s += blockDim.x) {
        typename W::Frag f = W::load(w, s);
        int o[W::NINTS];
        W::idots(f, a, s, o);
        if (TYPE == T_Q6_K) {
            // run r of step s (sb, h, j) is scale group 16*sb + 8h + 2r + j
            const int sb = s >> 2, h = (s & 3) >> 1, j = s & 1;
            for (int r = 0; r < 4; r++) out[16 * sb + 8 * h + 2 * r + j] = o[r];
        } else {
            for (int i = 0; i < W::NINTS; i++) out[s * W::NINTS + i] = o[i];
        }
    }
}
// a_qs must be followed by scratch provided by the caller: we allocate it inside ops_api (see b200_block_dots)
cudaError_t launch_block_dots_impl(int type, WPtrs w, int64_t K, const int8_t * a_qs, int32_t * out,
                                   int8_t * sq, ActDS * sd, cudaStream_t st) {
    switch (type) {
        case T_Q4_0: block_dots_kernel<T_Q4_0><<<1, 256, 0, st>>>(w, (int) (K / 32), a_qs, out, sq, sd); break;
        case T_Q8_0: block_dots_kernel<T_Q8_0><<<1, 256, 0, st>>>(w, (int) (K / 32), a_qs, out, sq, sd); break;
        case T_Q4_K: block_dots_kernel<T_Q4_K><<<1, 256, 0, st>>>(w, (int) (K / 64), a_qs, out, sq, sd); break;
        case T_Q6_K: block_dots_kernel<T_Q6_K><<<1, 256, 0, st>>>(w, (int) (K / 64), a_qs, out, sq, sd); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ quantize_q8_1
// 8 lanes per 32-element block, one float4 each.  Exactly quantize_row_q8_1_ref (ggml-quants.c:220-253):
//   d = amax / 127 (IEEE div), id = d ? 1/d : 0, q = roundf(x * id), s = fp16(sum(q) * d)
// IEEE-rounded intrinsics keep the compiler from contracting or approximating anything.
template <bool AOS>
__global__ void __launch_bounds__(256)
quantize_q8_1_kernel(const char * __restrict__ x, const int64_t row_stride, const int64_t K, const int64_t nrows,
                     int8_t * __restrict__ qs, ActDS * __restrict__ ds, uint8_t * __restrict__ aos) {
    const int64_t NB = K / 32;
    const int64_t gid = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;      // one per 4 elements
    const int64_t blk = gid >> 3;                                              // global block index
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    if (blk >= NB * nrows) return;             // whole 8-lane groups exit together
    const int64_t row = blk / NB, b = blk - row * NB;
    const int sub = (int) (gid & 7);
    const float4 v = *(const float4 *) (x + row * row_stride + (b * 32 + sub * 4) * 4);
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
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);          // lanes 0-3: plane 0 sum, lanes 4-7: plane 1 sum
    const int other = __shfl_xor_sync(0xffffffffu, sum, 4);
    const uint32_t packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
    if (AOS) {
        uint8_t * blkp = aos + blk * 36;
        *(uint32_t *) (blkp + 4 + sub * 4) = packed;
        if (sub == 0) {
            const __half hd = __float2half_rn(d);
            const __half hs = __float2half_rn(__fmul_rn((float) (sum + other), d));
            *(__half *) blkp = hd; *(__half *) (blkp + 2) = hs;
        }
    } else {
        // plane = sub / 4; 16 B per (plane, block)
        int8_t * col = qs + row * NB * 32;
        *(uint32_t *) (col + ((int64_t) (sub >> 2) * NB + b) * 16 + (sub & 3) * 4) = packed;
        if (sub == 0) {
            ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other;
            ds[row * NB + b] = o;
        }
    }
}

static cudaError_t launch_quant(bool aos, const float * x, int64_t rs, int64_t K, int64_t nrows, int8_t * qs, ActDS * ds,
                                void * y, cudaStream_t st, bool pdl) {
    const int64_t threads = K / 4 * nrows;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) ((threads + 255) / 256));
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (aos) return cudaLaunchKernelEx(&cfg, quantize_q8_1_kernel<true>,  (const char *) x, rs, K, nrows, qs, ds, (uint8_t *) y);
    return      cudaLaunchKernelEx(&cfg, quantize_q8_1_kernel<false>, (const char *) x, rs, K, nrows, qs, ds, (uint8_t *) y);
}
cudaError_t launch_quantize_q8_1_soa(const float * x, int64_t rs, int64_t K, int64_t nrows, int8_t * qs, ActDS * ds, cudaStream_t st) {
    return launch_quant(false, x, rs, K, nrows, qs, ds, nullptr, st, true);
}
cudaError_t launch_quantize_q8_1_aos(const float * x, int64_t K, int64_t nrows, void * y, cudaStream_t st) {
    return launch_quant(true, x, K * 4, K, nrows, nullptr, nullptr, y, st, false);
}

// ------------------------------------------------------------------------------------------------ ADD
// Fast path: everything contiguous, src1 either same shape or one row repeated -> float4 grid-stride.
__global__ void __launch_bounds__(256)
add_f32_vec_kernel(const float4 * __restrict__ a, const float4 * __restrict__ b, float4 * __restrict__ d,
                   const int64_t n4, const int64_t nb4 /* period of b in float4, or n4 */) {
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
#endif
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 x = a[i], y = b[nb4 == n4 ? i : i % nb4];
        d[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    }
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
#endif
}
// General path: any byte strides, src1 repeated in every dim (i % ne1x), one thread per element.
__global__ void __launch_bounds__(256) add_f32_gen_kernel(const AddArgs p, const int64_t n) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i;
        const int64_t i0 = t % p.ne[0]; t /= p.ne[0];
        const int64_t i1 = t % p.ne[1]; t /= p.ne[1];
        const int64_t i2 = t % p.ne[2]; const int64_t i3 = t / p.ne[2];
        const float x = *(const float *) (p.a + i0 * p.nba[0] + i1 * p.nba[1] + i2 * p.nba[2] + i3 * p.nba[3]);
        const float y = *(const float *) (p.b + (i0 % p.neb[0]) * p.nbb[0] + (i1 % p.neb[1]) * p.nbb[1] +
                                                 (i2 % p.neb[2]) * p.nbb[2] + (i3 % p.neb[3]) * p.nbb[3]);
        *(float *) (p.d + i0 * p.nbd[0] + i1 * p.nbd[1] + i2 * p.nbd[2] + i3 * p.nbd[3]) = x + y;
    }
}

static bool contiguous_f32(const int64_t * ne, const int64_t * nb) {
    int64_t s = 4;
    for (int i = 0; i < 4; i++) { if (ne[i] != 1 && nb[i] != s) return false; s *= ne[i]; }
    return true;
}

cudaError_t launch_add_f32(const AddArgs & p, cudaStream_t st) {
    const int64_t n = p.ne[0] * p.ne[1] * p.ne[2] * p.ne[3];
    if (n == 0) return cudaSuccess;
    const int64_t nb_el = p.neb[0] * p.neb[1] * p.neb[2] * p.neb[3];
    bool same = true;
    for (int i = 0; i < 4; i++) same = same && p.ne[i] == p.neb[i];
    // b periodic over the flattened index: b's shape is a prefix of a's shape (trailing dims 1)
    bool prefix = true; { bool ended = false; for (int i = 0; i < 4; i++) { if (p.neb[i] != p.ne[i]) ended = true; if (ended && p.neb[i] != 1) prefix = false; } }
    const bool fast = contiguous_f32(p.ne, p.nba) && contiguous_f32(p.ne, p.nbd) && contiguous_f32(p.neb, p.nbb) &&
                      (same || prefix) && n % 4 == 0 && nb_el % 4 == 0 &&
                      ((((uintptr_t) p.a) | ((uintptr_t) p.b) | ((uintptr_t) p.d)) & 15) == 0;
    if (fast) {
        const int64_t n4 = n / 4;
        int64_t blocks = (n4 + 255) / 256; if (blocks > 148 * 16) blocks = 148 * 16;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned) blocks); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, add_f32_vec_kernel, (const float4 *) p.a, (const float4 *) p.b, (float4 *) p.d, n4, nb_el / 4);
    }
    int64_t blocks = (n + 255) / 256; if (blocks > 148 * 16) blocks = 148 * 16;
    add_f32_gen_kernel<<<(unsigned) blocks, 256, 0, st>>>(p, n);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ repack
// One thread per byte of the native block range; runs once per upload/download, never in the hot path.
__global__ void __launch_bounds__(256)
repack_kernel(const WFields f, const bool to_soa, uint8_t * __restrict__ soa, const int64_t T, const int64_t first,
              const int64_t n, uint8_t * __restrict__ native) {
    const int64_t total = n * f.block_bytes;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t b = i / f.block_bytes; const int o = (int) (i - b * f.block_bytes);
        int64_t base = 0; int64_t dst = -1;
        #pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < f.n) {
                if (o >= f.native_off[k] && o < f.native_off[k] + f.size[k]) dst = base + (first + b) * f.size[k] + (o - f.native_off[k]);
                base += T * f.size[k];
            }
        }
        if (to_soa) soa[dst] = native[i]; else native[i] = soa[dst];
    }
}
cudaError_t launch_repack(int type, bool to_soa, uint8_t * soa, int64_t T, int64_t first, int64_t n, uint8_t * native, cudaStream_t st) {
    const WFields f = wfields(type);
    if (f.n == 0) return cudaErrorInvalidValue;
    if (n == 0) return cudaSuccess;
    int64_t blocks = (n * f.block_bytes + 255) / 256; if (blocks > 148 * 64) blocks = 148 * 64;
    repack_kernel<<<(unsigned) blocks, 256, 0, st>>>(f, to_soa, soa, T, first, n, native);
    return cudaGetLastError();
}

__global__ void fill_kernel(uint4 * p, size_t n16, uint32_t v) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t) gridDim.x * blockDim.x)
        p[i] = make_uint4(v, v, v, v);
}
cudaError_t launch_fill(void * p, size_t n_bytes, uint32_t v, cudaStream_t st) {
    fill_kernel<<<148 * 8, 256, 0, st>>>((uint4 *) p, n_bytes / 16, v);
    return cudaGetLastError();
}

}  // namespace b200

// gemv.cu — dequant-fused GEMV for decode (1..8 activation columns) over q4_0 / q8_0 / q4_K / q6_K weights.
//
// Takes the seat of the reference's hot loop: ggml_compute_forward_mul_mat_one_chunk + vec_dot
// (ggml-hexagon/kernels/ggml-dsp.c:1091-1189; ggml-cpu/ggml-cpu.c:6655-6743) including the activation
// quantisation that precedes it (ggml-dsp.c:1257-1285, quantize_row_q8_1_ref ggml-quants.c:220-253).
//
// Design (HBM-bound):
//  * one warp owns RPW consecutive weight rows; lanes stride over the row so consecutive lanes read consecutive
//    16 B of the SoA quant array (coalesced 128-bit streaming loads, L1 no-allocate);
//  * every lane keeps DEPTH x RPW weight fragments in flight in registers (rolling refill), so a warp has
//    DEPTH*RPW*512 B outstanding: enough bytes in flight per SM to cover HBM latency at ~6.5 TB/s;
//  * the first DEPTH fragments are requested BEFORE cudaGridDependencySynchronize(): weights do not depend on the
//    previous kernel, only the activations do, so under programmatic dependent launch the next GEMV streams its
//    weights while the previous kernel drains;
//  * FUSEQ: the CTA quantises the f32 activations to q8_1 into shared memory itself (no separate kernel, no HBM
//    round trip); otherwise it reads the pre-quantised two-plane scratch;
//  * up to 3 matrices that share the activation vector (q/k/v, gate/up) run as ONE launch (CTA ranges);
//  * optional fused residual: add_dst[row] = add_src[row] + dst[row]  (GGML_OP_ADD that follows the MUL_MAT).
#include "b200_internal.cuh"
#include "wtypes.cuh"

#include <stdio.h>
#include <stdlib.h>

namespace b200 {

// 8 consecutive lanes quantise one 32-float block, 4 floats each (same arithmetic as quantize_q8_1_kernel).
__device__ __forceinline__ void quantize8(const float4 v, uint32_t & packed, float & d_h, int & sum_mine, int & sum_other) {
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
    sum_mine = sum;
    sum_other = __shfl_xor_sync(0xffffffffu, sum, 4);
    packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
    d_h = __half2float(__float2half_rn(d));
}

template <int TYPE, int NCOLS, int RPW, int DEPTH, int WARPS, bool FUSEQ>
__global__ void __launch_bounds__(WARPS * 32)
gemv2_kernel(const GemvArgs a) {
    using W = WT<TYPE>;
    extern __shared__ uint4 smem_u4[];
    constexpr int T = WARPS * 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int steps = a.steps;

    // which matrix of the group does this CTA serve?
    int m = 0, cta = blockIdx.x;
    if (a.nmat > 1 && cta >= a.cta_end[0]) { m = 1; if (a.nmat > 2 && cta >= a.cta_end[1]) m = 2; }
    if (m > 0) cta -= a.cta_end[m - 1];
    const GemvMat & mt = a.mat[m];
    const int64_t M = mt.M;
    const WPtrs w = mt.w;
    const int64_t row0 = ((int64_t) cta * WARPS + warp) * RPW;
    const bool active = row0 < M;

    int64_t rbase[RPW];
    #pragma unroll
    for (int r = 0; r < RPW; r++) rbase[r] = (row0 + r < M ? row0 + r : M - 1) * (int64_t) steps;

    // ---- weight prefetch: DEPTH fragments per row, before the dependency wait
    typename W::Frag f[DEPTH][RPW];
    if (active) {
        #pragma unroll
        for (int d = 0; d < DEPTH; d++) {
            const int s = lane + 32 * d;
            if (s < steps) {
                #pragma unroll
                for (int r = 0; r < RPW; r++) f[d][r] = W::load(w, rbase[r] + s);
            }
        }
    }
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
#endif

    // ---- activations
    const int64_t NB = a.K / 32;
    const uint4 * ap0; const int2 * ads; int64_t col_u4, col_ds;   // per-column strides in uint4 / int2
    if (FUSEQ) {
        // smem per column: [plane0 NB x 16 B][plane1 NB x 16 B][ds NB x 8 B]
        col_u4 = 2 * NB + (NB + 1) / 2; col_ds = col_u4 * 2;
        const int64_t n4 = a.K / 4;
        #pragma unroll 1
        for (int c = 0; c < NCOLS; c++) {
            const float4 * xc = (const float4 *) ((const char *) a.x + (int64_t) c * a.x_col_stride);
            uint32_t * qs = (uint32_t *) (smem_u4 + c * col_u4);
            int2 * ds = (int2 *) (smem_u4 + c * col_u4 + 2 * NB);
            for (int64_t i0 = 0; i0 < n4; i0 += 4 * T) {
                float4 v[4];
                #pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int64_t i = i0 + u * T + tid;
                    v[u] = i < n4 ? xc[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                #pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int64_t i = i0 + u * T + tid;
                    uint32_t packed; float dh; int s_mine, s_other;
                    quantize8(v[u], packed, dh, s_mine, s_other);
                    if (i < n4) {
                        const int64_t b = i >> 3; const int sub = (int) (i & 7);
                        qs[((int64_t) (sub >> 2) * NB + b) * 4 + (sub & 3)] = packed;
                        if (sub == 0) ds[b] = make_int2(__float_as_int(dh), (s_mine & 0xFFFF) | (s_other << 16));
                    }
                }
            }
        }
        __syncthreads();
        ap0 = smem_u4; ads = (const int2 *) (smem_u4 + 2 * NB);
    } else {
        col_u4 = 2 * NB; col_ds = NB;
        ap0 = (const uint4 *) a.act.qs; ads = (const int2 *) a.act.ds;
    }
    if (!active) return;

    float acc[RPW][NCOLS];
    #pragma unroll
    for (int r = 0; r < RPW; r++)
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) acc[r][c] = 0.f;

    // ---- main loop: consume fragment d, refill it with the step DEPTH iterations ahead
    for (int base = 0; base < steps; base += 32 * DEPTH) {
        #pragma unroll
        for (int d = 0; d < DEPTH; d++) {
            const int s = base + lane + 32 * d;
            if (s < steps) {
                #pragma unroll
                for (int c = 0; c < NCOLS; c++) {
                    const ActCol ac(ap0 + c * col_u4, ap0 + c * col_u4 + NB, ads + c * col_ds);
                    #pragma unroll
                    for (int r = 0; r < RPW; r++) acc[r][c] += W::dot(f[d][r], ac, s);
                }
                const int sn = s + 32 * DEPTH;
                if (sn < steps) {
                    #pragma unroll
                    for (int r = 0; r < RPW; r++) f[d][r] = W::load(w, rbase[r] + sn);
                }
            }
        }
    }

    #pragma unroll
    for (int r = 0; r < RPW; r++)
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            float v = acc[r][c];
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && row0 + r < M) {
                const int64_t o = (int64_t) c * a.dcs + row0 + r;
                mt.dst[o] = v;
                if (mt.add_dst) mt.add_dst[o] = mt.add_src[o] + v;
            }
        }
}

// ------------------------------------------------------------------------------------------------ launch
template <int TYPE, int NCOLS, int RPW, int DEPTH, int WARPS, bool FUSEQ>
static cudaError_t launch_cfg(GemvArgs & a, cudaStream_t st, bool pdl) {
    auto kern = gemv2_kernel<TYPE, NCOLS, RPW, DEPTH, WARPS, FUSEQ>;
    const int64_t NB = a.K / 32;
    const size_t smem = FUSEQ ? (size_t) NCOLS * (2 * NB + (NB + 1) / 2) * 16 : 0;
    static bool attr_done = false;          // one flag per instantiation
    if (smem > 48 * 1024 && !attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    const int64_t rows_per_cta = WARPS * RPW;
    int total = 0;
    for (int i = 0; i < a.nmat; i++) { total += (int) ((a.mat[i].M + rows_per_cta - 1) / rows_per_cta); a.cta_end[i] = total; }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) total);
    cfg.blockDim = dim3(WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, a);
}

// tuning knob (development only): B200_GEMV_CFG="rpw,depth,warps" selects among the compiled q4_0 N=1 variants
struct Tune { int rpw = 2, depth = 4, warps = 4; bool pdl = true; };
static const Tune & tune() {
    static Tune t = [] {
        Tune x;
        if (const char * e = getenv("B200_GEMV_CFG")) sscanf(e, "%d,%d,%d", &x.rpw, &x.depth, &x.warps);
        if (const char * e = getenv("B200_PDL")) x.pdl = atoi(e) != 0;
        return x;
    }();
    return t;
}
bool gemv_pdl_enabled() { return tune().pdl; }

template <int TYPE, int NCOLS, bool FUSEQ>
static cudaError_t launch_ncols(GemvArgs & a, cudaStream_t st, bool pdl) {
    if constexpr (TYPE == T_Q4_0 && NCOLS == 1 && FUSEQ) {
        const Tune & t = tune();
#define V(R, D, Wp) if (t.rpw == R && t.depth == D && t.warps == Wp) return launch_cfg<TYPE, NCOLS, R, D, Wp, FUSEQ>(a, st, pdl)
        V(1, 4, 4); V(1, 8, 4); V(2, 2, 4); V(2, 4, 4); V(2, 8, 4); V(4, 2, 4); V(4, 4, 4);
        V(1, 4, 8); V(1, 8, 8); V(2, 2, 8); V(2, 4, 8); V(2, 8, 8); V(4, 2, 8); V(4, 4, 8);
#undef V
    }
    constexpr int RPW = NCOLS <= 2 ? 2 : 1;
    constexpr int DEPTH = (TYPE == T_Q6_K || NCOLS > 4) ? 2 : 4;
    return launch_cfg<TYPE, NCOLS, RPW, DEPTH, 4, FUSEQ>(a, st, pdl);
}

template <int TYPE>
static cudaError_t launch_type(GemvArgs & a, int ncols, bool fuseq, cudaStream_t st, bool pdl) {
    if (fuseq) {
        switch (ncols) {
            case 1: return launch_ncols<TYPE, 1, true>(a, st, pdl);
            case 2: return launch_ncols<TYPE, 2, true>(a, st, pdl);
        }
        return cudaErrorInvalidValue;
    }
    switch (ncols) {
        case 1: return launch_ncols<TYPE, 1, false>(a, st, pdl);
        case 2: return launch_ncols<TYPE, 2, false>(a, st, pdl);
        case 3: return launch_ncols<TYPE, 3, false>(a, st, pdl);
        case 4: return launch_ncols<TYPE, 4, false>(a, st, pdl);
        case 5: return launch_ncols<TYPE, 5, false>(a, st, pdl);
        case 6: return launch_ncols<TYPE, 6, false>(a, st, pdl);
        case 7: return launch_ncols<TYPE, 7, false>(a, st, pdl);
        case 8: return launch_ncols<TYPE, 8, false>(a, st, pdl);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_gemv2(int type, GemvArgs & a, int ncols, bool fuseq, cudaStream_t st) {
    a.steps = (int) (a.K / (type == T_Q4_K || type == T_Q6_K ? 64 : 32));
    const bool pdl = tune().pdl;
    switch (type) {
        case T_Q4_0: return launch_type<T_Q4_0>(a, ncols, fuseq, st, pdl);
        case T_Q8_0: return launch_type<T_Q8_0>(a, ncols, fuseq, st, pdl);
        case T_Q4_K: return launch_type<T_Q4_K>(a, ncols, fuseq, st, pdl);
        case T_Q6_K: return launch_type<T_Q6_K>(a, ncols, fuseq, st, pdl);
    }
    return cudaErrorInvalidValue;
}

}  // namespace b200

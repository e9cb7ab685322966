// gemv.cu — dequant-fused GEMV for decode (1..8 activation columns) over q4_0 / q8_0 / q4_K / q6_K weights.
//
// Takes the seat of the reference's hot loop: ggml_compute_forward_mul_mat_one_chunk + vec_dot
// (ggml-hexagon/kernels/ggml-dsp.c:1091-1189; ggml-cpu/ggml-cpu.c:6655-6743).
//
// Design (HBM-bound; DESIGN.md has the numbers):
//  * a CTA owns R consecutive weight rows and the whole K; warp w owns the K-slice of 32 "steps" [32w, 32w+32)
//    (a step = 32 elements for q4_0/q8_0, 64 for the K-quants), lane l owns step 32w+l of EVERY row of the tile;
//  * so a lane loads the q8_1 activations of its step ONCE into registers and afterwards only streams weights:
//    per 16 B of q4_0 weights the inner loop is 1 LDG.128 + 1 LDG.U16 + 8 LOP3 + 8 dp4a + a handful of float ops;
//  * consecutive lanes read consecutive 16 B of the SoA quant array (coalesced 128-bit, L1 no-allocate), and all
//    R rows of the tile are requested back to back BEFORE cudaGridDependencySynchronize(): weights do not depend on
//    the previous kernel, so under programmatic dependent launch a GEMV streams its weights while the kernels
//    before it (the previous GEMV, the activation quantiser) are still running;
//  * per-row sums: butterfly transpose-reduce across the 32 lanes (R-1+log2(32/R) shuffles for R rows), then the
//    per-warp partials meet in shared memory;
//  * up to 3 matrices that share the activation vector (q/k/v, gate/up) run as ONE launch (CTA ranges);
//  * optional fused residual: add_dst[row] = add_src[row] + dst[row]  (the GGML_OP_ADD that follows the MUL_MAT).
#include "b200_internal.cuh"
#include "wtypes.cuh"

#include <stdio.h>
#include <stdlib.h>

namespace b200 {

// Quantise on exit (GemvArgs::qx_*), called by the whole warp 0 once lane r < rows holds result v of row row0 + r (rows divides 32).
// The CTAs that share a 32-element block of the result vector hand their values to the block's LAST CTA through {value, 1} slots
// (one 8-byte store each, no fence, no atomic: a CTA that is not the last one leaves at once); the last CTA -- dispatched after the
// others, which are therefore running or done -- polls the slots, clears them for the next launch, and writes the block as
// quantize_q8_1_kernel would (quantize_row_q8_1_ref, ggml-quants.c:220-253; SoA planes of wtypes.cuh).
__device__ __forceinline__ void gemv_quant_on_exit(const GemvArgs & a, const int m, const int64_t row0, const int rows, const float v, const int64_t M) {
    if (!a.qx_qs || m != a.qx_mat) return;
    const int lane = threadIdx.x & 31, off = (int) (row0 & 31);
    const int64_t blk = row0 >> 5, NB = M >> 5;
    uint2 * slots = a.qx_slots + blk * 32;
    if (off != 32 - rows) {
        if (lane < rows) asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(slots + off + lane), "r"(__float_as_uint(v)), "r"(1u) : "memory");
        return;
    }
    float x = __shfl_sync(0xffffffffu, v, lane >= off ? lane - off : 0);      // this CTA's rows sit at the end of the block
    if (lane < off) {
        uint32_t val, flag;
        const long long t0 = clock64();
        do {
            asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(val), "=r"(flag) : "l"(slots + lane) : "memory");
            if (!flag && clock64() - t0 > 8000000000ll) __trap();
        } while (!flag);
        asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(slots + lane), "r"(0u), "r"(0u) : "memory");
        x = __uint_as_float(val);
    }
    float amax = fabsf(x);
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q = (int) roundf(__fmul_rn(x, id));
    int sum = q;
    #pragma unroll
    for (int o = 8; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);      // lanes 0-15: plane 0, lanes 16-31: plane 1
    const int other = __shfl_xor_sync(0xffffffffu, sum, 16);
    a.qx_qs[((int64_t) (lane >> 4) * NB + blk) * 16 + (lane & 15)] = (int8_t) q;
    if (lane == 0) { ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other; a.qx_ds[blk] = o; }
}

// ITERS > 1: the row has more than 32*nwarps steps; warp w then owns slices w, w+nwarps, ... and the fragments of the
// next slice are requested before the current one is consumed (DB = register double buffer).
template <int TYPE, int NCOLS, int R, bool DB>
__global__ void __launch_bounds__(512)
b200_gemv4_kernel(const GemvArgs a) {
    using W = WT<TYPE>;
    using A = WR<TYPE>;
    __shared__ float part[16][R * NCOLS + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int steps = a.steps;
#if __CUDA_ARCH__ >= 900
    // let the next kernel of the stream launch now: it only prefetches ITS weights until its own
    // cudaGridDependencySynchronize() returns, which waits for this whole grid to finish
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    int m = 0, cta = blockIdx.x;
    if (a.nmat > 1 && cta >= a.cta_end[0]) { m = 1; if (a.nmat > 2 && cta >= a.cta_end[1]) m = 2; }
    if (m > 0) cta -= a.cta_end[m - 1];
    const GemvMat & mt = a.mat[m];
    const int64_t M = mt.M;
    const WPtrs w = mt.w;
    const int64_t row0 = (int64_t) cta * R;
    int64_t rbase[R];
    #pragma unroll
    for (int r = 0; r < R; r++) rbase[r] = (row0 + r < M ? row0 + r : M - 1) * (int64_t) steps;

    int s = warp * 32 + lane;                    // this lane's step within every row (first slice)
    const int stride = nwarps * 32;
    // ---- all R weight fragments of the first slice, requested before the dependency wait
    typename W::Frag f[R], g[DB ? R : 1];
    if (s < steps) {
        #pragma unroll
        for (int r = 0; r < R; r++) f[r] = W::load(w, rbase[r] + s);
    }
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();             // activations (and add_src) come from the kernels before us
#endif
    float acc[NCOLS][R];
    #pragma unroll
    for (int c = 0; c < NCOLS; c++)
        #pragma unroll
        for (int r = 0; r < R; r++) acc[c][r] = 0.f;
    for (; s < steps; s += stride) {
        const int sn = s + stride;
        if (DB && sn < steps) {
            #pragma unroll
            for (int r = 0; r < R; r++) g[r] = W::load(w, rbase[r] + sn);
        }
        #pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const ActCol ac(a.act, c);
            const typename A::Act av = A::load_act(ac, s);
            #pragma unroll
            for (int r = 0; r < R; r++) acc[c][r] += wr_dot<TYPE>(f[r], av, s);
        }
        if (sn < steps) {
            #pragma unroll
            for (int r = 0; r < R; r++) { if (DB) f[r] = g[r]; else f[r] = W::load(w, rbase[r] + sn); }
        }
    }
    // ---- reduce: lanes (butterfly), then warps (shared memory)
    constexpr int SH = 5 - Log2<R>::v;
    #pragma unroll
    for (int c = 0; c < NCOLS; c++) {
        const float v = reduce_rows<R>(acc[c], lane);
        if ((lane & ((1 << SH) - 1)) == 0) part[warp][c * R + (lane >> SH)] = v;
    }
    __syncthreads();
    if (warp == 0) {
      float qv = 0.f;
      if (tid < R * NCOLS) {
        float v = 0.f;
        for (int q = 0; q < nwarps; q++) v += part[q][tid];
        const int c = tid / R, r = tid - c * R;
        if (row0 + r < M) {
            const int64_t o = (int64_t) c * a.dcs + row0 + r;
            mt.dst[o] = v;
            float s = 0.f;
            if (mt.add_dst) { s = mt.add_src[o] + v; mt.add_dst[o] = s; }
            qv = mt.add_dst ? s : v;
            if (a.ll_slots) {                                 // row split, LL protocol: {value, seq} in one 8-byte NVLink store per peer
                const uint32_t seq = *a.ll_count + a.ll_op;
                const uint2 * slot = a.ll_slots + (size_t) (seq & 1u) * LL_HALF_ELEMS + a.ll_off[m] + (int64_t) a.ll_rank * M + row0 + r;
                for (int pi = 0; pi < a.npeers; pi++)
                    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"((const char *) slot + a.peer_delta[pi]), "r"(__float_as_uint(v)), "r"(seq) : "memory");
            } else
            for (int pi = 0; pi < a.npeers; pi++) {           // row split: the same values into every peer's dst (NVLink stores)
                *(float *) ((char *) (mt.dst + o) + a.peer_delta[pi]) = v;
                if (mt.add_dst) *(float *) ((char *) (mt.add_dst + o) + a.peer_delta[pi]) = s;
            }
        }
      }
      if (NCOLS == 1) gemv_quant_on_exit(a, m, row0, R, qv, M);
    }
}

// ------------------------------------------------------------------------------------------------ gemv5: TMA-staged tiles
// Same lane / warp roles as gemv4, but the weights of the CTA's RT rows do not travel through registers: thread 0 issues one
// cp.async.bulk per SoA field (the rows of a tile are contiguous in every field array) into shared memory, BEFORE the
// dependency wait, and the lanes read their fragments from shared memory once the mbarrier flips.  What this buys is bytes
// in flight: a resident CTA holds its whole tile (18-50 KB) in flight at the cost of ~48 registers per thread, so an SM has
// ~200 KB outstanding instead of ~80 KB -- the difference between 3.5-4.2 TB/s and what a plain read kernel of the same
// launch size reaches under programmatic dependent launch (tools/ubench_stream.cu: 25 MB in 3.9 us, 51 MB in 8.9 us).
struct Gemv5Tile { int nfields; int rowbytes[4]; int foff[4]; int tile_bytes; };

__device__ __forceinline__ uint32_t g5_smem(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

template <int TYPE, int RT, int ITERS>
__global__ void __launch_bounds__(512)
b200_gemv5_kernel(const GemvArgs a, const Gemv5Tile t) {
    using W = WT<TYPE>;
    using A = WR<TYPE>;
    extern __shared__ __align__(128) uint8_t g5_sm[];
    float (*part)[RT + 1] = (float (*)[RT + 1]) (g5_sm + t.tile_bytes);
    const uint32_t mbar = g5_smem(g5_sm + t.tile_bytes + 16 * (RT + 1) * 4);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int steps = a.steps;
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    int m = 0, cta = blockIdx.x;
    if (a.nmat > 1 && cta >= a.cta_end[0]) { m = 1; if (a.nmat > 2 && cta >= a.cta_end[1]) m = 2; }
    if (m > 0) cta -= a.cta_end[m - 1];
    const GemvMat & mt = a.mat[m];
    const int64_t M = mt.M;
    const int64_t row0 = (int64_t) cta * RT;
    const int nrows = (int) (M - row0 < RT ? M - row0 : RT);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mbar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // (static indices into the kernel parameters: a run-time index makes the compiler copy them to local memory)
        uint32_t total = 0;
        #pragma unroll
        for (int i = 0; i < 4; i++) total += (uint32_t) (nrows * t.rowbytes[i]);           // unused fields have rowbytes 0
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(mbar), "r"(total) : "memory");
        #pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < t.nfields) {
                const uint8_t * src = mt.w.p[i] + row0 * t.rowbytes[i];
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(g5_smem(g5_sm + t.foff[i])), "l"(src), "r"((uint32_t) (nrows * t.rowbytes[i])), "r"(mbar) : "memory");
            }
        }
    }
    __syncthreads();                             // the barrier is initialised for everyone
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();             // activations (and add_src) come from the kernels before us
#endif
    // this lane's activations: step s of slice warp (+ nwarps) of the row
    const ActCol ac(a.act, 0);
    typename A::Act av[ITERS];
    int sidx[ITERS];
    #pragma unroll
    for (int it = 0; it < ITERS; it++) {
        sidx[it] = (warp + it * nwarps) * 32 + lane;
        if (sidx[it] < steps) av[it] = A::load_act(ac, sidx[it]);
    }
    WPtrs sw;
    #pragma unroll
    for (int i = 0; i < 4; i++) sw.p[i] = g5_sm + t.foff[i];
    {   // wait for the tile
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(mbar) : "memory");
    }
    float acc[RT];
    #pragma unroll
    for (int r = 0; r < RT; r++) acc[r] = 0.f;
    #pragma unroll
    for (int it = 0; it < ITERS; it++) {
        const int s = sidx[it];
        if (s < steps) {
            #pragma unroll
            for (int r = 0; r < RT; r++)
                if (r < nrows) acc[r] += wr_dot<TYPE>(W::loads(sw, r * steps + s), av[it], s);
        }
    }
    constexpr int SH = 5 - Log2<RT>::v;
    {
        const float v = reduce_rows<RT>(acc, lane);
        if ((lane & ((1 << SH) - 1)) == 0) part[warp][lane >> SH] = v;
    }
    __syncthreads();
    if (warp == 0) {
      float qv = 0.f;
      if (tid < RT) {
        float v = 0.f;
        for (int q = 0; q < nwarps; q++) v += part[q][tid];
        const int r = tid;
        if (row0 + r < M) {
            const int64_t o = row0 + r;
            mt.dst[o] = v;
            float s = 0.f;
            if (mt.add_dst) { s = mt.add_src[o] + v; mt.add_dst[o] = s; }
            qv = mt.add_dst ? s : v;
            if (a.ll_slots) {
                const uint32_t seq = *a.ll_count + a.ll_op;
                const uint2 * slot = a.ll_slots + (size_t) (seq & 1u) * LL_HALF_ELEMS + a.ll_off[m] + (int64_t) a.ll_rank * M + row0 + r;
                for (int pi = 0; pi < a.npeers; pi++)
                    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"((const char *) slot + a.peer_delta[pi]), "r"(__float_as_uint(v)), "r"(seq) : "memory");
            } else
            for (int pi = 0; pi < a.npeers; pi++) {
                *(float *) ((char *) (mt.dst + o) + a.peer_delta[pi]) = v;
                if (mt.add_dst) *(float *) ((char *) (mt.add_dst + o) + a.peer_delta[pi]) = s;
            }
        }
      }
      gemv_quant_on_exit(a, m, row0, RT, qv, M);
    }
}

// ------------------------------------------------------------------------------------------------ gemv5n: several columns per tile
// 2 .. 31 activation columns (small batches, speculative decoding; test-backend-ops n = 2..8): the tile is fetched ONCE and
// stays in shared memory while the CTA walks over all columns, NC at a time (a weight fragment is read from shared memory once
// per NC columns; the lane keeps NC activation blocks in registers).  The register-path gemv4 re-reads the weights from
// HBM / L2 for every chunk of 8 columns and the activations of every column for every group of R rows.
template <int TYPE, int RT, int ITERS, int NC>
__global__ void __launch_bounds__(512)
b200_gemv5n_kernel(const GemvArgs a, const Gemv5Tile t, const int ncols) {
    using W = WT<TYPE>;
    using A = WR<TYPE>;
    extern __shared__ __align__(128) uint8_t g5_sm[];
    float (*part)[NC * RT + 1] = (float (*)[NC * RT + 1]) (g5_sm + t.tile_bytes);
    const uint32_t mbar = g5_smem(g5_sm + t.tile_bytes + 16 * (NC * RT + 1) * 4);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int steps = a.steps;
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    const GemvMat & mt = a.mat[0];
    const int64_t M = mt.M;
    const int64_t row0 = (int64_t) blockIdx.x * RT;
    const int nrows = (int) (M - row0 < RT ? M - row0 : RT);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mbar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        uint32_t total = 0;
        #pragma unroll
        for (int i = 0; i < 4; i++) total += (uint32_t) (nrows * t.rowbytes[i]);
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(mbar), "r"(total) : "memory");
        #pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < t.nfields) {
                const uint8_t * src = mt.w.p[i] + row0 * t.rowbytes[i];
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(g5_smem(g5_sm + t.foff[i])), "l"(src), "r"((uint32_t) (nrows * t.rowbytes[i])), "r"(mbar) : "memory");
            }
        }
    }
    __syncthreads();
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
#endif
    WPtrs sw;
    #pragma unroll
    for (int i = 0; i < 4; i++) sw.p[i] = g5_sm + t.foff[i];
    int sidx[ITERS];
    #pragma unroll
    for (int it = 0; it < ITERS; it++) sidx[it] = (warp + it * nwarps) * 32 + lane;
    {
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(mbar) : "memory");
    }
    constexpr int SH = 5 - Log2<RT>::v;
    for (int c0 = 0; c0 < ncols; c0 += NC) {
        float acc[NC][RT];
        #pragma unroll
        for (int c = 0; c < NC; c++)
            #pragma unroll
            for (int r = 0; r < RT; r++) acc[c][r] = 0.f;
        #pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const int s = sidx[it];
            if (s < steps) {
                typename A::Act av[NC];
                #pragma unroll
                for (int c = 0; c < NC; c++) {
                    const ActCol ac(a.act, c0 + c < ncols ? c0 + c : c0);       // (past the end: a valid column, result unused)
                    av[c] = A::load_act(ac, s);
                }
                #pragma unroll
                for (int r = 0; r < RT; r++) {
                    if (r < nrows) {
                        const typename W::Frag f = W::loads(sw, r * steps + s);
                        #pragma unroll
                        for (int c = 0; c < NC; c++) acc[c][r] += wr_dot<TYPE>(f, av[c], s);
                    }
                }
            }
        }
        #pragma unroll
        for (int c = 0; c < NC; c++) {
            const float v = reduce_rows<RT>(acc[c], lane);
            if ((lane & ((1 << SH) - 1)) == 0) part[warp][c * RT + (lane >> SH)] = v;
        }
        __syncthreads();
        if (tid < NC * RT) {
            float v = 0.f;
            for (int q = 0; q < nwarps; q++) v += part[q][tid];
            const int c = tid / RT, r = tid - c * RT;
            if (row0 + r < M && c0 + c < ncols) mt.dst[(int64_t) (c0 + c) * a.dcs + row0 + r] = v;
        }
        __syncthreads();                         // part is reused by the next column group
    }
}

// ------------------------------------------------------------------------------------------------ launch
struct Tune { int rows = 0; bool pdl = true; };    // rows 0 = per-shape heuristic
static const Tune & tune() {
    static Tune t = [] {
        Tune x;
        if (const char * e = getenv("B200_GEMV_ROWS")) x.rows = atoi(e);     // development knob
        if (const char * e = getenv("B200_PDL")) x.pdl = atoi(e) != 0;
        return x;
    }();
    return t;
}
bool gemv_pdl_enabled() { return tune().pdl; }

template <int TYPE, int NCOLS, int R>
static cudaError_t launch_cfg(GemvArgs & a, cudaStream_t st, bool pdl) {
    const int nslices = (a.steps + 31) / 32;
    const int iters = (nslices + 15) / 16;                  // at most 16 warps (512 threads) per CTA
    const int nwarps = (nslices + iters - 1) / iters;
    int total = 0;
    for (int i = 0; i < a.nmat; i++) { total += (int) ((a.mat[i].M + R - 1) / R); a.cta_end[i] = total; }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) total);
    cfg.blockDim = dim3(nwarps * 32);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    // the register double buffer only where it fits without spills (ptxas -v): q4_0 / q8_0, one column
    if constexpr ((TYPE == T_Q4_0 || TYPE == T_Q8_0) && NCOLS == 1) {
        if (iters > 1) return cudaLaunchKernelEx(&cfg, b200_gemv4_kernel<TYPE, NCOLS, R, true>, a);
    }
    return cudaLaunchKernelEx(&cfg, b200_gemv4_kernel<TYPE, NCOLS, R, false>, a);
}

// gemv5 eligibility: one column, pre-quantised activations, every field's row is a multiple of 16 bytes and 16-byte aligned
// (bulk copies), at most 2 slices per warp, no in-kernel completion signal
template <int TYPE, int RT, int ITERS>
static cudaError_t launch_g5(GemvArgs & a, const Gemv5Tile & t, int nwarps, cudaStream_t st, bool pdl) {
    static int smem_set[64] = {0};                // per device: the attribute belongs to the device's copy of the function
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    const int smem = t.tile_bytes + 16 * (RT + 1) * 4 + 16;
    if (smem > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(b200_gemv5_kernel<TYPE, RT, ITERS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        smem_set[dev] = smem;
    }
    int total = 0;
    for (int i = 0; i < a.nmat; i++) { total += (int) ((a.mat[i].M + RT - 1) / RT); a.cta_end[i] = total; }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) total); cfg.blockDim = dim3(nwarps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, b200_gemv5_kernel<TYPE, RT, ITERS>, a, t);
}
template <int TYPE>
static bool try_gemv5(GemvArgs & a, cudaStream_t st, bool pdl, cudaError_t * err) {
    static const int mode = getenv("B200_GEMV5") ? atoi(getenv("B200_GEMV5")) : 1;       // 0: off (gemv4 only)
    if (!mode) return false;
    const WFields f = wfields(TYPE);
    const int64_t bpr = a.K / (TYPE == T_Q4_K || TYPE == T_Q6_K ? 256 : 32);            // native blocks per row
    Gemv5Tile t{};
    t.nfields = f.n;
    int rowsum = 0;
    for (int i = 0; i < f.n; i++) {
        const int64_t rb = bpr * f.size[i];
        if (rb % 16 != 0 || rb > (1 << 20)) return false;
        t.rowbytes[i] = (int) rb; rowsum += (int) rb;
        for (int mi = 0; mi < a.nmat; mi++) if (((uintptr_t) a.mat[mi].w.p[i] & 15) != 0) return false;
    }
    const int nslices = (a.steps + 31) / 32;
    const int iters = (nslices + 15) / 16;
    if (iters > 2) return false;
    const int nwarps = (nslices + iters - 1) / iters;
    // small launches (a 4096 x 4096 q4_0 matrix is 64 KB per SM) are pure latency: the register path of gemv4 is quicker there
    const double min_mb = getenv("B200_GEMV5_MIN_MB") ? atof(getenv("B200_GEMV5_MIN_MB")) : 12.0;     // (read per launch: tests switch it)
    int64_t mtot = 0; for (int i = 0; i < a.nmat; i++) mtot += a.mat[i].M;
    if ((double) mtot * rowsum < min_mb * 1e6) return false;
    const int rt_env = getenv("B200_GEMV5_ROWS") ? atoi(getenv("B200_GEMV5_ROWS")) : 0;
    int rt = rt_env == 4 || rt_env == 8 || rt_env == 16 ? rt_env : 8;                                                         // B200 sweep: 8 rows (18-65 KB tiles)
    if (rt * rowsum > 72 * 1024) rt = 4;
    if (rt * rowsum > 72 * 1024) return false;
    int off = 0;
    for (int i = 0; i < f.n; i++) { t.foff[i] = off; off += rt * t.rowbytes[i]; }
    t.tile_bytes = (off + 15) & ~15;
    if (rt == 16)     *err = iters == 1 ? launch_g5<TYPE, 16, 1>(a, t, nwarps, st, pdl) : launch_g5<TYPE, 16, 2>(a, t, nwarps, st, pdl);
    else if (rt == 8) *err = iters == 1 ? launch_g5<TYPE, 8, 1>(a, t, nwarps, st, pdl)  : launch_g5<TYPE, 8, 2>(a, t, nwarps, st, pdl);
    else              *err = iters == 1 ? launch_g5<TYPE, 4, 1>(a, t, nwarps, st, pdl)  : launch_g5<TYPE, 4, 2>(a, t, nwarps, st, pdl);
    return true;
}

// several columns in one launch (gemv5n); false = not eligible, the caller falls back to 8-column gemv4 chunks
template <int TYPE, int RT, int ITERS, int NC>
static cudaError_t launch_g5n(GemvArgs & a, const Gemv5Tile & t, int nwarps, int ncols, cudaStream_t st, bool pdl) {
    static int smem_set[64] = {0};
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    const int smem = t.tile_bytes + 16 * (NC * RT + 1) * 4 + 16;
    if (smem > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(b200_gemv5n_kernel<TYPE, RT, ITERS, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        smem_set[dev] = smem;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) ((a.mat[0].M + RT - 1) / RT)); cfg.blockDim = dim3(nwarps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, b200_gemv5n_kernel<TYPE, RT, ITERS, NC>, a, t, ncols);
}
template <int TYPE>
static bool try_gemv5n(GemvArgs & a, int ncols, cudaStream_t st, bool pdl, cudaError_t * err) {
    const int mode = getenv("B200_GEMV5N") ? atoi(getenv("B200_GEMV5N")) : 1;            // 0: off (gemv4 chunks)
    // (2 and 3 columns: the register path is as fast or faster -- 13.1 / 14.7 us against 14.8 / 14.9 us for q4_0 4096 x 14336)
    if (!mode || a.nmat != 1 || a.npeers > 0 || a.mat[0].add_dst || ncols < (mode == 2 ? 2 : 4)) return false;
    const WFields f = wfields(TYPE);
    const int64_t bpr = a.K / (TYPE == T_Q4_K || TYPE == T_Q6_K ? 256 : 32);
    Gemv5Tile t{};
    t.nfields = f.n;
    int rowsum = 0;
    for (int i = 0; i < f.n; i++) {
        const int64_t rb = bpr * f.size[i];
        if (rb % 16 != 0 || rb > (1 << 20) || ((uintptr_t) a.mat[0].w.p[i] & 15) != 0) return false;
        t.rowbytes[i] = (int) rb; rowsum += (int) rb;
    }
    const int nslices = (a.steps + 31) / 32;
    const int iters = (nslices + 15) / 16;
    if (iters > 2) return false;
    const int nwarps = (nslices + iters - 1) / iters;
    int rt = 8;
    if (rt * rowsum > 72 * 1024) rt = 4;
    if (rt * rowsum > 72 * 1024) return false;
    int off = 0;
    for (int i = 0; i < f.n; i++) { t.foff[i] = off; off += rt * t.rowbytes[i]; }
    t.tile_bytes = (off + 15) & ~15;
    constexpr bool KQ = TYPE == T_Q4_K || TYPE == T_Q6_K;
    constexpr int NC = KQ ? 2 : 4;                                                        // columns per pass (registers)
    if (rt == 8) *err = iters == 1 ? launch_g5n<TYPE, 8, 1, NC>(a, t, nwarps, ncols, st, pdl) : launch_g5n<TYPE, 8, 2, NC>(a, t, nwarps, ncols, st, pdl);
    else         *err = iters == 1 ? launch_g5n<TYPE, 4, 1, NC>(a, t, nwarps, ncols, st, pdl) : launch_g5n<TYPE, 4, 2, NC>(a, t, nwarps, ncols, st, pdl);
    return true;
}
bool launch_gemv_cols(int type, GemvArgs & a, int ncols, cudaStream_t st, cudaError_t * err) {
    a.steps = (int) (a.K / (type == T_Q4_K || type == T_Q6_K ? 64 : 32));
    const bool pdl = tune().pdl;
    switch (type) {
        case T_Q4_0: return try_gemv5n<T_Q4_0>(a, ncols, st, pdl, err);
        case T_Q8_0: return try_gemv5n<T_Q8_0>(a, ncols, st, pdl, err);
        case T_Q4_K: return try_gemv5n<T_Q4_K>(a, ncols, st, pdl, err);
        case T_Q6_K: return try_gemv5n<T_Q6_K>(a, ncols, st, pdl, err);
    }
    return false;
}

template <int TYPE, int NCOLS>
static cudaError_t launch_ncols(GemvArgs & a, cudaStream_t st, bool pdl) {
    constexpr bool KQ = TYPE == T_Q4_K || TYPE == T_Q6_K;
    if constexpr (NCOLS == 1) {
        cudaError_t e5 = cudaSuccess;
        if (try_gemv5<TYPE>(a, st, pdl, &e5)) return e5;
    }
    if constexpr (NCOLS == 1) {
        // rows per CTA, from the B200 sweep (profiles/r01_b_gemv_rows_sweep.txt): tall matrices want 8, long rows 2, else 4
        int64_t mtot = 0; for (int i = 0; i < a.nmat; i++) mtot += a.mat[i].M;
        const int r = tune().rows > 0 ? tune().rows : (mtot >= 8192 ? 8 : (a.K >= 8192 ? 2 : 4));
        if constexpr (TYPE == T_Q4_0) {
            if (r == 2) return launch_cfg<TYPE, 1, 2>(a, st, pdl);
            if (r == 4) return launch_cfg<TYPE, 1, 4>(a, st, pdl);
            return launch_cfg<TYPE, 1, 8>(a, st, pdl);
        } else {
            if (r == 2) return launch_cfg<TYPE, 1, 2>(a, st, pdl);
            return launch_cfg<TYPE, 1, 4>(a, st, pdl);
        }
    } else if constexpr (NCOLS == 2) {
        return launch_cfg<TYPE, NCOLS, (KQ ? 2 : 4)>(a, st, pdl);
    } else if constexpr (NCOLS <= 4) {
        return launch_cfg<TYPE, NCOLS, 2>(a, st, pdl);
    } else {
        return launch_cfg<TYPE, NCOLS, (KQ ? 1 : 2)>(a, st, pdl);
    }
}

template <int TYPE>
static cudaError_t launch_type(GemvArgs & a, int ncols, cudaStream_t st, bool pdl) {
    switch (ncols) {
        case 1: return launch_ncols<TYPE, 1>(a, st, pdl);
        case 2: return launch_ncols<TYPE, 2>(a, st, pdl);
        case 3: return launch_ncols<TYPE, 3>(a, st, pdl);
        case 4: return launch_ncols<TYPE, 4>(a, st, pdl);
        case 5: return launch_ncols<TYPE, 5>(a, st, pdl);
        case 6: return launch_ncols<TYPE, 6>(a, st, pdl);
        case 7: return launch_ncols<TYPE, 7>(a, st, pdl);
        case 8: return launch_ncols<TYPE, 8>(a, st, pdl);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_gemv2(int type, GemvArgs & a, int ncols, cudaStream_t st) {
    a.steps = (int) (a.K / (type == T_Q4_K || type == T_Q6_K ? 64 : 32));
    const bool pdl = tune().pdl;
    switch (type) {
        case T_Q4_0: return launch_type<T_Q4_0>(a, ncols, st, pdl);
        case T_Q8_0: return launch_type<T_Q8_0>(a, ncols, st, pdl);
        case T_Q4_K: return launch_type<T_Q4_K>(a, ncols, st, pdl);
        case T_Q6_K: return launch_type<T_Q6_K>(a, ncols, st, pdl);
    }
    return cudaErrorInvalidValue;
}

}  // namespace b200

// gemv6.cu — decode GEMV as a persistent ring kernel with the q8_1 activation quantiser inside it.
//
// Takes the seat of ggml_compute_forward_mul_mat_one_chunk + vec_dot for ONE activation column
// (ggml-hexagon/kernels/ggml-dsp.c:1091-1189, 1262-1285: from_float of src1, then the dot loop), all four weight types.
//
// Why (profiles/r01_o_*): traffic already equals the algorithmic bytes; what was left on the table was
// latency: (1) every decode MUL_MAT paid TWO kernel boundaries on its dependency chain (quantize_q8_1 launch, then the
// GEMV), (2) 18 KB CTAs that fetch one tile, wait, dot and exit keep HBM busy only through CTA turnover.
//
// Design:
//  * one CTA per SM (grid = SM count), each owns a contiguous range of row units; every WARP runs its own pipeline:
//    lane 0 issues cp.async.bulk copies of whole row units (RU rows x this warp's K segment, every SoA field) into the
//    warp's private ring of S shared-memory slots, each with its own mbarrier; the warp dots a slot as soon as it lands
//    and re-arms it at once.  No CTA-wide synchronisation in the loop;
//  * lane l owns steps l, l+32, ... of the warp's K segment in EVERY row, so the q8_1 activations of those steps are
//    loaded into registers once per kernel; a row's dot product is finished by one butterfly reduction (KS = 1), or by
//    the KS warps of a team through shared memory and a named barrier (long rows: K split over KS warps);
//  * the ring is filled BEFORE cudaGridDependencySynchronize(): weights depend on nothing, so under programmatic
//    dependent launch the next MUL_MAT's first ~100 KB per SM stream in while the previous kernel still runs;
//  * after the dependency wait the CTA quantises the f32 activation vector itself (same code as quantize_q8_1_kernel,
//    bit-exact with quantize_row_q8_1_ref, ggml-quants.c:220-253) into shared memory: 148 x K x 4 bytes of L2 reads
//    instead of a kernel launch + boundary on the critical path;
//  * q / k / v and gate / up run as ONE launch (units run through up to 3 matrices); optional fused residual ADD;
//    row-split peers (LL slots or plain peer stores) as in gemv4 / gemv5.
#include "b200_internal.cuh"
#include "wtypes.cuh"

#include <stdio.h>
#include <stdlib.h>

namespace b200 {

struct G6Cfg {
    int nfields;
    int b4[4];            // bytes per 4 steps of each SoA field (q4_0: 64, 8; q8_0: 128, 8; q4_K: 128, 16; q6_K: 128, 64, 16, 2)
    int foff[4];          // offset of each field inside a ring slot
    int slot_bytes;
    int S;                // ring slots per warp
    int KS;               // warps per team (K segments per row)
    int seg_steps;        // steps per K segment (multiple of 32); the last segment may be shorter
    int steps;            // steps per row
    int unit_end[3];      // running unit count over the matrices of the launch
    int total_units;
    int act_off, part_off, bar_off;
    long long * trace;    // development: 4 x %globaltimer per CTA
};

__device__ __forceinline__ uint32_t g6_smem(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ long long g6_now() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

// row split over the GPUs of one box (rare path, kept out of line): the result also goes into every peer's memory over NVLink, as
// {value, sequence number} in one 8-byte store (LL protocol) or as plain stores into the peer's dst / add_dst
__device__ __noinline__ void g6_peer_stores(const GemvArgs & a, int m, int64_t o, float v, float s, float * dst, float * add_dst) {
    if (a.ll_slots) {
        const int64_t M = m == 0 ? a.mat[0].M : m == 1 ? a.mat[1].M : a.mat[2].M;
        const int lloff = m == 0 ? a.ll_off[0] : m == 1 ? a.ll_off[1] : a.ll_off[2];
        const uint32_t seq = *a.ll_count + a.ll_op;
        const uint2 * sl = a.ll_slots + (size_t) (seq & 1u) * LL_HALF_ELEMS + lloff + (int64_t) a.ll_rank * M + o;
        #pragma unroll 1
        for (int pi = 0; pi < a.npeers; pi++)
            asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"((const char *) sl + a.peer_delta[pi]), "r"(__float_as_uint(v)), "r"(seq) : "memory");
    } else {
        #pragma unroll 1
        for (int pi = 0; pi < a.npeers; pi++) {
            *(float *) ((char *) (dst + o) + a.peer_delta[pi]) = v;
            if (add_dst) *(float *) ((char *) (add_dst + o) + a.peer_delta[pi]) = s;
        }
    }
}

// FULL: every K segment has exactly 32 x NJ steps (K = 4096 q4_0 with NJ = 4, ...): segment length, row pitch and field offsets
// inside a slot are compile-time constants, every shared-memory load of the dot loop has an immediate offset and no lane is masked
#define G6_REGS(TEAM, FULL, NJ, KQ) ((TEAM) ? ((NJ) >= 6 ? 112 : 96) : ((FULL) && !((KQ) && (NJ) > 1) ? 64 : 80))
template <int TYPE, int NJ, int RU, bool TEAM, bool FULL>
__global__ void __maxnreg__(G6_REGS(TEAM, FULL, NJ, TYPE == T_Q4_K || TYPE == T_Q6_K))
b200_gemv6_kernel(const GemvArgs a, const G6Cfg g) {
    using W = WT<TYPE>;
    using A = WR<TYPE>;
    constexpr bool KQ = TYPE == T_Q4_K || TYPE == T_Q6_K;
    constexpr int SEG = 32 * NJ;
    extern __shared__ __align__(128) uint8_t g6_sm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();     // the next kernel may launch and fill ITS ring while we run
#endif
    if (g.trace && tid == 0) g.trace[blockIdx.x * 4 + 0] = g6_now();
    const int KS = TEAM ? g.KS : 1;
    const int team = TEAM ? warp / KS : warp, kseg = TEAM ? warp - team * KS : 0, nteams = TEAM ? nwarps / KS : nwarps;
    const int steps = g.steps;
    const int seg0 = kseg * g.seg_steps;
    const int seglen = FULL ? SEG : TEAM ? (steps - seg0 < g.seg_steps ? steps - seg0 : g.seg_steps) : steps;
    // bytes per 4 steps of every SoA field, offsets of the fields inside a ring slot (compile-time when FULL)
    constexpr int NF = TYPE == T_Q6_K ? 4 : 2;
    int b4[4], foff[4];
    #pragma unroll
    for (int f = 0; f < 4; f++) { b4[f] = f < NF ? (KQ ? wfields(TYPE).size[f] : wfields(TYPE).size[f] * 4) : 0; }
    #pragma unroll
    for (int f = 0, off = 0; f < 4; f++) { foff[f] = FULL ? off : g.foff[f]; off += RU * (SEG / 4) * b4[f]; }
    // this CTA's units, this team's share of them: u0 + team, u0 + team + nteams, ...
    const int u0 = (int) ((long long) g.total_units * blockIdx.x / gridDim.x);
    const int u1 = (int) ((long long) g.total_units * (blockIdx.x + 1) / gridDim.x);
    const int nmine = u1 - u0 > team ? (u1 - u0 - team + nteams - 1) / nteams : 0;
    uint8_t * ring = g6_sm + (size_t) warp * g.S * g.slot_bytes;
    const uint32_t bar0 = g6_smem(g6_sm + g.bar_off) + warp * g.S * 8;

    // unit u -> matrix, first row, rows (static indices into the kernel parameters only)
    auto locate = [&](int u, int & m, int64_t & row0, int & nrows) {
        m = 0; int ub = 0;
        if (a.nmat > 1 && u >= g.unit_end[0]) { m = 1; ub = g.unit_end[0]; if (a.nmat > 2 && u >= g.unit_end[1]) { m = 2; ub = g.unit_end[1]; } }
        const int64_t M = m == 0 ? a.mat[0].M : m == 1 ? a.mat[1].M : a.mat[2].M;
        row0 = (int64_t) (u - ub) * RU;
        nrows = (int) (M - row0 < RU ? M - row0 : RU);
    };
    // lane 0: fetch unit i of this warp into slot i % S
    auto issue = [&](int i, int slot) {
        int m, nrows; int64_t row0;
        locate(u0 + team + i * nteams, m, row0, nrows);
        const uint32_t dst = g6_smem(ring + (size_t) slot * g.slot_bytes), bar = bar0 + slot * 8;
        uint32_t total = 0;
        #pragma unroll
        for (int f = 0; f < 4; f++) if (f < NF) total += (uint32_t) (nrows * (seglen >> 2) * b4[f]);
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar), "r"(total) : "memory");
        #pragma unroll
        for (int f = 0; f < 4; f++) {
            if (f < NF) {
                const uint8_t * base = m == 0 ? a.mat[0].w.p[f] : m == 1 ? a.mat[1].w.p[f] : a.mat[2].w.p[f];
                const uint32_t segb = (uint32_t) ((seglen >> 2) * b4[f]);
                if (!TEAM) {         // whole rows: the unit is one contiguous range of the field array
                    const uint8_t * src = base + ((row0 * steps) >> 2) * b4[f];
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst + foff[f]), "l"(src), "r"((uint32_t) nrows * segb), "r"(bar) : "memory");
                } else {
                    for (int r = 0; r < nrows; r++) {
                        const uint8_t * src = base + (((row0 + r) * steps + seg0) >> 2) * b4[f];
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     :: "r"(dst + foff[f] + r * segb), "l"(src), "r"(segb), "r"(bar) : "memory");
                    }
                }
            }
        }
    };

    if (lane == 0) {
        for (int s = 0; s < g.S; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar0 + s * 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int i = 0; i < g.S && i < nmine; i++) issue(i, i);
    }
    __syncwarp();
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();               // the activations (and add_src) come from the kernels before us
#endif
    if (g.trace && tid == 0) g.trace[blockIdx.x * 4 + 1] = g6_now();

    // ---- quantize_row_q8_1_ref of the whole activation vector, once per CTA: 8 lanes per 32-element block
    const int NB = (int) (a.K >> 5);
    {
        int8_t * aq = (int8_t *) (g6_sm + g.act_off);
        ActDS * ads = (ActDS *) (g6_sm + g.act_off + (size_t) NB * 32);
        const float4 * x4 = (const float4 *) a.x;
        const int K4 = (int) (a.K >> 2);
        // 4 lanes per 32-element block, 8 elements (one 32-byte sector) per lane.  All loads of a round are in flight together:
        // the vector sits on the critical path of the whole decode step and an L2 round trip under load is ~1 us (measured)
        constexpr int QB = 3;
        const int nthr = blockDim.x, K8 = K4 >> 1;             // K % 256 == 0: whole warps enter and leave together
        for (int base = 0; base < K8; base += nthr * QB) {
            float4 va[QB], vb[QB];
            #pragma unroll
            for (int q = 0; q < QB; q++) { const int i = base + q * nthr + tid; if (i < K8) { va[q] = __ldcg(x4 + 2 * i); vb[q] = __ldcg(x4 + 2 * i + 1); } }
            #pragma unroll
            for (int q = 0; q < QB; q++) {
                const int i = base + q * nthr + tid;
                if (i < K8) {
                    const float4 v = va[q], w = vb[q];
                    const int blk = i >> 2, sub = i & 3;           // sub 0, 1: elements 0..15 (plane 0); 2, 3: plane 1
                    float amax = fmaxf(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))),
                                       fmaxf(fmaxf(fabsf(w.x), fabsf(w.y)), fmaxf(fabsf(w.z), fabsf(w.w))));
                    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
                    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
                    const float d  = __fdiv_rn(amax, 127.0f);
                    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
                    const int q0 = (int) roundf(__fmul_rn(v.x, id)), q1 = (int) roundf(__fmul_rn(v.y, id));
                    const int q2 = (int) roundf(__fmul_rn(v.z, id)), q3 = (int) roundf(__fmul_rn(v.w, id));
                    const int q4 = (int) roundf(__fmul_rn(w.x, id)), q5 = (int) roundf(__fmul_rn(w.y, id));
                    const int q6 = (int) roundf(__fmul_rn(w.z, id)), q7 = (int) roundf(__fmul_rn(w.w, id));
                    int sum = (q0 + q1) + (q2 + q3) + (q4 + q5) + (q6 + q7);
                    sum += __shfl_xor_sync(0xffffffffu, sum, 1);           // lanes 0, 1: plane 0 sum; lanes 2, 3: plane 1 sum
                    const int other = __shfl_xor_sync(0xffffffffu, sum, 2);
                    uint2 pk;
                    pk.x = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
                    pk.y = (uint32_t) (q4 & 0xFF) | ((uint32_t) (q5 & 0xFF) << 8) | ((uint32_t) (q6 & 0xFF) << 16) | ((uint32_t) (q7 & 0xFF) << 24);
                    *(uint2 *) (aq + ((size_t) (sub >> 1) * NB + blk) * 16 + (sub & 1) * 8) = pk;
                    if (sub == 0) { ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other; ads[blk] = o; }
                }
            }
        }
    }
    __syncthreads();
    if (g.trace && tid == 0) g.trace[blockIdx.x * 4 + 2] = g6_now();
    // this lane's activations: steps seg0 + lane + 32 j of the row
    // (steps past the end of the segment: clamped to a valid step, their products are dropped -- no branch in the dot loop)
    typename A::Act av[NJ];
    int sidx[NJ];
    {
        const uint4 * p0 = (const uint4 *) (g6_sm + g.act_off);
        const ActCol ac(p0, p0 + NB, (const int2 *) (g6_sm + g.act_off + (size_t) NB * 32));
        #pragma unroll
        for (int j = 0; j < NJ; j++) {
            const int s = lane + 32 * j;
            sidx[j] = FULL || s < seglen ? s : seglen - 1;
            av[j] = A::load_act(ac, seg0 + sidx[j]);
        }
    }
    float * part = (float *) (g6_sm + g.part_off);
    constexpr int SH = 5 - Log2<RU>::v;
    int slot = 0; uint32_t phase = 0;
    for (int i = 0; i < nmine; i++) {
        int m, nrows; int64_t row0;
        locate(u0 + team + i * nteams, m, row0, nrows);
        const bool holder = (lane & ((1 << SH) - 1)) == 0;
        const int r = lane >> SH;
        float * add_dst = m == 0 ? a.mat[0].add_dst : m == 1 ? a.mat[1].add_dst : a.mat[2].add_dst;
        float addv = 0.f;                               // the residual of the fused ADD: requested now, used after the dots
        if (kseg == 0 && holder && r < nrows && add_dst) {
            const float * add_src = m == 0 ? a.mat[0].add_src : m == 1 ? a.mat[1].add_src : a.mat[2].add_src;
            addv = __ldcg(add_src + row0 + r);
        }
        {   // the slot has landed
            uint32_t ok = 0; const uint32_t bar = bar0 + slot * 8;
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(phase) : "memory");
            if (!ok) {                                 // bounded: a copy that never lands traps instead of hanging the GPU
                const long long t0 = clock64();
                while (!ok) {
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(phase) : "memory");
                    if (!ok && clock64() - t0 > 4000000000ll) __trap();
                }
            }
        }
        WPtrs sw;
        #pragma unroll
        for (int f = 0; f < 4; f++) sw.p[f] = ring + (size_t) slot * g.slot_bytes + foff[f];
        float acc[RU];
        #pragma unroll
        for (int r = 0; r < RU; r++) acc[r] = 0.f;
        // rows past the end of the matrix (last unit only) are computed from whatever the slot holds and never stored
        #pragma unroll
        for (int j = 0; j < NJ; j++) {
            if (FULL || 32 * j < seglen) {                       // (warp-uniform: a short last segment skips whole steps)
                #pragma unroll
                for (int r = 0; r < RU; r++) {
                    const float d = wr_dot<TYPE>(W::loads(sw, r * seglen + sidx[j]), av[j], sidx[j]);
                    if (FULL) acc[r] += d; else acc[r] += lane + 32 * j < seglen ? d : 0.f;
                }
            }
        }
        float v = reduce_rows<RU>(acc, lane);       // (uses every loaded value: the slot is free once this returns)
        __syncwarp();
        if (lane == 0 && i + g.S < nmine) issue(i + g.S, slot);
        if (TEAM) {
            float * mine = part + ((size_t) (i & 1) * nwarps + warp) * RU;
            if (kseg != 0 && holder) mine[r] = v;
            asm volatile("bar.sync %0, %1;" :: "r"(1 + team), "r"(32 * KS) : "memory");
            if (kseg == 0 && holder)
                for (int q = 1; q < KS; q++) v += mine[q * RU + r];
        }
        if (kseg == 0 && holder && r < nrows) {
            const int64_t o = row0 + r;
            float * dst = m == 0 ? a.mat[0].dst : m == 1 ? a.mat[1].dst : a.mat[2].dst;
            dst[o] = v;
            float s = 0.f;
            if (add_dst) { s = addv + v; add_dst[o] = s; }
            if (a.npeers > 0) g6_peer_stores(a, m, o, v, s, dst, add_dst);
        }
        if (++slot == g.S) { slot = 0; phase ^= 1u; }
    }
    if (g.trace) { __syncthreads(); if (tid == 0) g.trace[blockIdx.x * 4 + 3] = g6_now(); }
}

// ------------------------------------------------------------------------------------------------ launch
struct G6Tune { int warps = 0, ru = 0, smem_kb = 100, cps = 1, maxnj = 0, slots = 0; long long * trace = nullptr; int trace_cap = 0, trace_n = 0; };
static G6Tune & g6_tune() {
    static G6Tune t = [] {
        G6Tune x;
        if (const char * e = getenv("B200_G6_WARPS")) x.warps = atoi(e);
        if (const char * e = getenv("B200_G6_RU")) x.ru = atoi(e);
        if (const char * e = getenv("B200_G6_SMEM_KB")) x.smem_kb = atoi(e);
        if (const char * e = getenv("B200_G6_CPS")) x.cps = atoi(e);
        if (const char * e = getenv("B200_G6_MAXNJ")) x.maxnj = atoi(e);
        if (const char * e = getenv("B200_G6_SLOTS")) x.slots = atoi(e);
        return x;
    }();
    return t;
}
// development: record 4 x %globaltimer per CTA for the next launches (tools/g6_trace.py); buf holds cap launches x 148 x 4 values
void gemv6_trace(long long * buf, int cap_launches) { G6Tune & t = g6_tune(); t.trace = buf; t.trace_cap = cap_launches; t.trace_n = 0; }
int gemv6_trace_count() { return g6_tune().trace_n; }

template <int TYPE, int NJ, int RU, bool TEAM, bool FULL>
static cudaError_t g6_go(const GemvArgs & a, const G6Cfg & g, int grid, int nwarps, int smem, cudaStream_t st, bool pdl) {
    static int smem_set[64] = {0};
    int dev = 0; cudaGetDevice(&dev); dev &= 63;
    if (smem > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(b200_gemv6_kernel<TYPE, NJ, RU, TEAM, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        smem_set[dev] = smem;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3(nwarps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, b200_gemv6_kernel<TYPE, NJ, RU, TEAM, FULL>, a, g);
}
template <int TYPE, int NJ>
static cudaError_t g6_pick(const GemvArgs & a, const G6Cfg & g, int ru, bool team, bool full, int grid, int nwarps, int smem, cudaStream_t st, bool pdl) {
    if (full) {
        if constexpr (NJ <= 4) {
            if (team) return ru == 2 ? g6_go<TYPE, NJ, 2, true, true>(a, g, grid, nwarps, smem, st, pdl)  : g6_go<TYPE, NJ, 1, true, true>(a, g, grid, nwarps, smem, st, pdl);
            return           ru == 2 ? g6_go<TYPE, NJ, 2, false, true>(a, g, grid, nwarps, smem, st, pdl) : g6_go<TYPE, NJ, 1, false, true>(a, g, grid, nwarps, smem, st, pdl);
        }
    }
    if (team) return ru == 2 ? g6_go<TYPE, NJ, 2, true, false>(a, g, grid, nwarps, smem, st, pdl)  : g6_go<TYPE, NJ, 1, true, false>(a, g, grid, nwarps, smem, st, pdl);
    return           ru == 2 ? g6_go<TYPE, NJ, 2, false, false>(a, g, grid, nwarps, smem, st, pdl) : g6_go<TYPE, NJ, 1, false, false>(a, g, grid, nwarps, smem, st, pdl);
}

template <int TYPE>
static bool try_gemv6(GemvArgs & a, int sms, cudaStream_t st, bool pdl, cudaError_t * err) {
    G6Tune & tn = g6_tune();
    constexpr bool KQ = TYPE == T_Q4_K || TYPE == T_Q6_K;
    // opt-in (B200_GEMV6=1) while it is slower than the quantize + gemv4 / gemv5 launches on the whole decode step (measured on B200:
    // 1.38 against 0.955 ms per Llama-2-7B token); read per launch: tests switch it
    const int on = getenv("B200_GEMV6") ? atoi(getenv("B200_GEMV6")) : 0;
    if (!on || !a.x || ((uintptr_t) a.x & 15) || a.K % 256) return false;
    const WFields wf = wfields(TYPE);
    const int steps = (int) (a.K / (KQ ? 64 : 32));
    G6Cfg g{};
    g.nfields = wf.n; g.steps = steps;
    for (int f = 0; f < wf.n; f++) {
        // bytes per 4 steps: q4_0 / q8_0: 4 blocks; K-quants: one 256-element super-block
        g.b4[f] = KQ ? wf.size[f] : wf.size[f] * 4;
        if (((int64_t) steps / 4 * g.b4[f]) % 16) return false;                          // every row of every field starts 16-B aligned
        for (int mi = 0; mi < a.nmat; mi++) if ((uintptr_t) a.mat[mi].w.p[f] & 15) return false;
    }
    // K segment per warp: NJ x 32 steps kept in registers (q4_0 / q8_0: 10 registers per step, K-quants 20 - 24)
    const int maxnj = tn.maxnj > 0 ? tn.maxnj : (KQ ? 2 : 6);
    const int nj_full = (steps + 31) / 32;
    int ks = (nj_full + maxnj - 1) / maxnj;
    int nj = (nj_full + ks - 1) / ks;
    if (!KQ && (nj & 1)) nj++;                                                            // instantiated: 2, 4, 6 (K-quants 1, 2)
    if (nj > (KQ ? 2 : 6)) return false;
    g.seg_steps = ks > 1 ? 32 * nj : steps;
    if (ks > 1) ks = (steps + g.seg_steps - 1) / g.seg_steps;
    if (ks > 8) return false;
    g.KS = ks;
    if (ks > 1) {
        const int last = steps - (ks - 1) * g.seg_steps;
        if (last <= 0) return false;
        for (int f = 0; f < wf.n; f++) if ((last / 4 * g.b4[f]) % 16 || last % 4 || (g.seg_steps / 4 * g.b4[f]) % 16) return false;
    }
    const int ru = tn.ru == 1 || tn.ru == 2 ? tn.ru : 2;
    int off = 0;
    const int segmax = ks > 1 ? g.seg_steps : steps;
    for (int f = 0; f < wf.n; f++) { g.foff[f] = off; off += ru * (segmax / 4) * g.b4[f]; }
    g.slot_bytes = (off + 15) & ~15;
    // warps and ring depth from the shared-memory budget (two kernels of this kind are meant to fit one SM together)
    const int act_bytes = (int) (a.K / 32 * 40);
    // registers: 64 per thread when FULL, else 80 (teams 96, 112 with 6 steps per lane): 16 / 12 (10, 8) warps keep the kernel under half of the SM's register file
    const bool full = steps == ks * 32 * nj && nj <= 4;              // every segment exactly 32 x NJ steps
    const bool team_ = ks > 1;
    int nw = tn.warps > 0 ? tn.warps : (ks > 1 ? (nj >= 6 ? 8 : 10) : full && !(KQ && nj > 1) ? 16 : 12);
    nw = nw / ks * ks; if (nw < ks) nw = ks;
    { const int cap = !team_ && full ? 24 : 16; if (nw > cap) nw = cap / ks * ks; }
    if (ks > 1 && nw / ks > 15) nw = 15 * ks;                                             // named barriers 1 .. 15
    const int budget = tn.smem_kb * 1024 - act_bytes - 1024;
    int S = tn.slots > 0 ? tn.slots : budget / (nw * g.slot_bytes);
    while (S < 2 && nw > ks) { nw -= ks; S = budget / (nw * g.slot_bytes); }
    if (S < 2) S = 2;
    if (S > 8) S = 8;
    g.S = S;
    int64_t units = 0;
    for (int i = 0; i < a.nmat; i++) { units += (a.mat[i].M + ru - 1) / ru; if (units > 0x7fffffff) return false; g.unit_end[i] = (int) units; }
    g.total_units = (int) units;
    int grid = sms * (tn.cps > 0 ? tn.cps : 1);
    if (grid > g.total_units) grid = g.total_units;
    if (grid < 1) return false;
    const int ring_bytes = nw * S * g.slot_bytes;
    g.act_off = (ring_bytes + 15) & ~15;
    g.part_off = g.act_off + ((act_bytes + 15) & ~15);
    g.bar_off = g.part_off + 2 * nw * ru * 4;
    g.bar_off = (g.bar_off + 15) & ~15;
    const int smem = g.bar_off + nw * S * 8;
    if (smem > 225 * 1024) return false;
    if (tn.trace && tn.trace_n < tn.trace_cap) { g.trace = tn.trace + (size_t) tn.trace_n * 4 * 1024; tn.trace_n++; }
    const bool team = ks > 1;
    if constexpr (KQ) {
        *err = nj <= 1 ? g6_pick<TYPE, 1>(a, g, ru, team, full, grid, nw, smem, st, pdl) : g6_pick<TYPE, 2>(a, g, ru, team, full, grid, nw, smem, st, pdl);
    } else {
        *err = nj <= 2 ? g6_pick<TYPE, 2>(a, g, ru, team, full, grid, nw, smem, st, pdl) :
               nj <= 4 ? g6_pick<TYPE, 4>(a, g, ru, team, full, grid, nw, smem, st, pdl) : g6_pick<TYPE, 6>(a, g, ru, team, full, grid, nw, smem, st, pdl);
    }
    return true;
}

// one f32 activation column (a.x), 1 .. 3 matrices; false: not eligible (*err untouched), the caller quantises and runs gemv4 / gemv5
bool launch_gemv6(int type, GemvArgs & a, int sms, cudaStream_t st, cudaError_t * err) {
    const bool pdl = gemv_pdl_enabled();
    switch (type) {
        case T_Q4_0: return try_gemv6<T_Q4_0>(a, sms, st, pdl, err);
        case T_Q8_0: return try_gemv6<T_Q8_0>(a, sms, st, pdl, err);
        case T_Q4_K: return try_gemv6<T_Q4_K>(a, sms, st, pdl, err);
        case T_Q6_K: return try_gemv6<T_Q6_K>(a, sms, st, pdl, err);
    }
    return false;
}

}  // namespace b200

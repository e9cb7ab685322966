// decode_stream.cu — a whole decode node list (every MUL_MAT + ADD of one token) as ONE persistent kernel.
//
// Takes the seat of the reference's per-node dispatch loop + per-op RPC (ggmlhexagon_backend_graph_compute_general ->
// ggmlhexagon_compute -> ggmlop_dsp_mulmat / _add, ggml-hexagon.cpp:5555-5574, :4975-5060; ggml-dsp.c:1069-1351).
//
// Why: a batch-1 MUL_MAT is 1-4 us of HBM traffic; launched one by one the stream is idle between kernels (launch
// ramp, tail wave, dependency) as long as it is busy.  Weights never depend on anything, so here HBM never stops:
//   * grid = one CTA per SM (cooperative launch), 16 consumer warps + 1 producer warp per CTA;
//   * the PRODUCER lane walks the whole program and streams this CTA's share of every matrix (balanced contiguous
//     row ranges) through a 10-slot x 18 KB shared-memory ring with 1-D bulk async copies (cp.async.bulk ->
//     mbarrier complete_tx, L2 evict-first), running ahead across stage boundaries: up to 185 KB per SM in flight;
//   * the CONSUMER warps wait for the stage's inputs (grid-wide arrive counter in L2), quantise the activation
//     vector to q8_1 straight into shared memory (= quantize_row_q8_1_ref, ggml-quants.c:220-253), keep their
//     K-slice of it in registers and then only read weights from the ring: per 16 B of q4_0 weights
//     1 LDS.128 + 1 LDS.U16 + 8 LOP3 + 8 dp4a, exact integer block dots (ggml-cpu-quants.c:2591-2604 etc., wtypes.cuh);
//   * per-row sums: butterfly over lanes, partials of the K-slice warps meet in shared memory, fused residual ADD.
// Spin waits are bounded: a wait that exceeds DS_TIMEOUT_CYCLES traps instead of hanging the GPU.
#include "b200_internal.cuh"
#include "wtypes.cuh"

#include <stdio.h>

namespace b200 {

constexpr int DS_CWARPS   = 16;                    // consumer warps
constexpr int DS_CTHREADS = DS_CWARPS * 32;
constexpr int DS_THREADS  = DS_CTHREADS + 32;      // + producer warp
constexpr long long DS_TIMEOUT_CYCLES = 8000000000ll;      // ~4 s at 2 GHz (clock64 is cheap; %globaltimer is not)

// shared memory map (dynamic, 128-B aligned)
constexpr int DS_OFF_RING = 0;
constexpr int DS_OFF_ACT  = DS_OFF_RING + DS_NSLOTS * DS_SLOT_BYTES;
constexpr int DS_OFF_PART = DS_OFF_ACT + DS_ACT_BYTES;             // float [2][32 rows][DS_CWARPS]
constexpr int DS_OFF_BAR  = DS_OFF_PART + 2 * 32 * DS_CWARPS * 4;  // uint64 full[NSLOTS], empty[NSLOTS]
constexpr int DS_DESC_WORDS = (int) ((sizeof(DsStage) + 15) / 16 * 4);
constexpr int DS_OFF_DESC = DS_OFF_BAR + 2 * DS_NSLOTS * 8;        // two DsStage copies
constexpr int DS_SMEM     = DS_OFF_DESC + 2 * DS_DESC_WORDS * 4;
static_assert(DS_SLOT_BYTES % 128 == 0 && DS_ACT_BYTES % 128 == 0, "alignment");
static_assert(DS_SMEM <= 232448, "shared memory");

int ds_smem_bytes() { return DS_SMEM; }
int ds_block_threads() { return DS_THREADS; }

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        if (clock64() - t0 > DS_TIMEOUT_CYCLES) __trap();
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}
__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, %0;" :: "n"(DS_CTHREADS) : "memory"); }   // consumer warps only
__device__ __forceinline__ uint32_t ld_acquire(const uint32_t * p) {
    uint32_t v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t * p) {
    uint32_t v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void red_release(uint32_t * p) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(p) : "memory");
}

// this CTA's contiguous share [lo, hi) of M rows
__device__ __forceinline__ void row_share(int M, int & lo, int & hi) {
    lo = (int) ((long long) M * blockIdx.x / gridDim.x);
    hi = (int) ((long long) M * (blockIdx.x + 1) / gridDim.x);
}

// ------------------------------------------------------------------------------------------------ producer
__device__ __noinline__ void ds_producer(const DsStage * __restrict__ prog, const int nstages, uint8_t * smem) {
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    const uint32_t ring = smem_u32(smem + DS_OFF_RING), full0 = smem_u32(smem + DS_OFF_BAR), empty0 = full0 + DS_NSLOTS * 8;
    uint32_t cnt = 0;
    for (int s = 0; s < nstages; s++) {
        const DsStage * st = prog + s;
        if (st->kind != DS_GEMV) continue;
        const int R = st->R, nf = st->nfields;
        int rb[4], fo[4], fl[4];
        #pragma unroll
        for (int f = 0; f < 4; f++) { rb[f] = st->rowbytes[f]; fo[f] = st->foff[f]; fl[f] = st->flead[f]; }
        for (int o = 0; o < st->nops; o++) {
            int lo, hi; row_share(st->op[o].M, lo, hi);
            const uint8_t * fp[4];
            #pragma unroll
            for (int f = 0; f < 4; f++) fp[f] = st->op[o].f[f];
            for (int r0 = lo; r0 < hi; r0 += R, cnt++) {
                const uint32_t slot = cnt % DS_NSLOTS, use = cnt / DS_NSLOTS;
                if (use > 0) mbar_wait(empty0 + slot * 8, (use - 1) & 1);
                const int n = min(R, hi - r0);
                uint32_t total = 0;
                const uint8_t * src[4]; uint32_t sz[4];
                #pragma unroll
                for (int f = 0; f < 4; f++) {
                    if (f < nf) {
                        const uint8_t * a = fp[f] + (long long) r0 * rb[f];
                        const uint32_t lead = fl[f] ? (uint32_t) ((uintptr_t) a & 15) : 0u;
                        src[f] = a - lead; sz[f] = (lead + (uint32_t) (n * rb[f]) + 15u) & ~15u;
                        total += sz[f];
                    }
                }
                const uint32_t fb = full0 + slot * 8;
                mbar_expect_tx(fb, total);
                #pragma unroll
                for (int f = 0; f < 4; f++)
                    if (f < nf) bulk_g2s(ring + slot * DS_SLOT_BYTES + fo[f], src[f], sz[f], fb, policy);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ consumer pieces
// q8_1 quantisation of the stage's activation vector into shared memory (two planes + {d, isum0, isum1}); 8 lanes per
// block, two blocks per thread in flight.  Activations were written by other SMs during this launch: L2 loads (.cg).
__device__ __forceinline__ void ds_quant_block(const float4 v, const int b, const int NB, const int sub, int8_t * qs, ActDS * ds) {
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
    if (b < NB) {
        const uint32_t packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
        *(uint32_t *) (qs + ((sub >> 2) * NB + b) * 16 + (sub & 3) * 4) = packed;
        if (sub == 0) { ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other; ds[b] = o; }
    }
}
__device__ __forceinline__ void ds_quantize(const float * __restrict__ x, const int NB, int8_t * qs, ActDS * ds, const int tid) {
    const int sub = tid & 7;
    constexpr int BP = DS_CTHREADS / 8;            // blocks per pass
    for (int b = tid >> 3; b < NB; b += 2 * BP) {  // whole warps iterate together (b differs by < 4 inside a warp ... NB % 4 == 0)
        const int b1 = b + BP;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 v0 = __ldcg((const float4 *) (x + b * 32 + sub * 4));
        const float4 v1 = b1 < NB ? __ldcg((const float4 *) (x + b1 * 32 + sub * 4)) : z;
        ds_quant_block(v0, b, NB, sub, qs, ds);
        ds_quant_block(v1, b1, NB, sub, qs, ds);
    }
}

struct DsState { uint32_t cnt; int pbuf; };

// registers: how many weight fragments a lane keeps in flight (LDS issued back to back before the first dot)
template <int TYPE> struct DsChunk { static constexpr int v = 1; };
template <> struct DsChunk<T_Q4_0> { static constexpr int v = 4; };
template <> struct DsChunk<T_Q8_0> { static constexpr int v = 2; };
template <> struct DsChunk<T_Q4_K> { static constexpr int v = 2; };

template <int TYPE> __device__ __forceinline__ void ds_zero_act(typename WR<TYPE>::Act & a) {
    uint32_t * p = (uint32_t *) &a;
    #pragma unroll
    for (int i = 0; i < (int) (sizeof(a) / 4); i++) p[i] = 0u;
}

// The (slot, row) items of one group for one warp: NI items, item it -> slot it / RW, row rbase + it % RW.
// FULL: all NI items exist (no predicates in the unrolled body).
template <int TYPE, bool TWO, int RW, bool FULL>
__device__ __forceinline__ void ds_group_dots(float (&acc)[8], const int ni, const uint8_t * ring, const uint32_t full0, const uint32_t cnt,
                                              const int (&foff)[4], const int (&lead)[4], const int lead_step, const bool has_lead,
                                              const int idx0, const int idx1, const int steps,
                                              const typename WR<TYPE>::Act & a0, const typename WR<TYPE>::Act & a1, const int s0, const int s1) {
    using W = WT<TYPE>;
    constexpr int C = TWO ? (DsChunk<TYPE>::v + 1) / 2 : DsChunk<TYPE>::v;
    #pragma unroll
    for (int c0 = 0; c0 < 8; c0 += C) {
        typename W::Frag fa[C], fb[TWO ? C : 1];
        #pragma unroll
        for (int i = 0; i < C; i++) {
            const int it = c0 + i;
            if (FULL || it < ni) {
                const int j = it / RW, ir = it % RW;
                const uint32_t cn = cnt + j, slot = cn % DS_NSLOTS;
                if (ir == 0) mbar_wait(full0 + slot * 8, (cn / DS_NSLOTS) & 1);
                const uint8_t * sb = ring + slot * DS_SLOT_BYTES;
                WPtrs w;
                #pragma unroll
                for (int f = 0; f < 4; f++) w.p[f] = sb + foff[f] + (has_lead ? ((lead[f] + j * lead_step) & 15) : 0);
                fa[i] = W::loads(w, idx0 + ir * steps);
                if (TWO) fb[i] = W::loads(w, idx1 + ir * steps);
            }
        }
        #pragma unroll
        for (int i = 0; i < C; i++) {
            const int it = c0 + i;
            if (FULL || it < ni) {
                acc[it] = wr_dot<TYPE>(fa[i], a0, s0);
                if (TWO) acc[it] += wr_dot<TYPE>(fb[i], a1, s1);
            }
        }
    }
}

// One GEMV stage.  Ring slots are consumed in GROUPS of gs slots (up to 8 (slot, row) items per warp): all fragments of
// a chunk are requested first, then the dots run, ONE butterfly reduces the 8 row sums over the lanes, one consumer
// barrier per group hands the per-slice partials to the finisher threads.
// TWO: rows longer than 16 slices (K > 16384 for the 32-element types): every warp owns two K-slices.
// RW : rows of a slot one warp handles (1 or 2).
template <int TYPE, bool TWO, int RW>
__device__ __forceinline__ void ds_gemv_stage(const DsStage * st, uint8_t * smem, DsState & S, const int tid) {
    using A = WR<TYPE>;
    const int lane = tid & 31, warp = tid >> 5;
    const int K = st->K, steps = st->steps, R = st->R, NB = K / 32;
    int8_t * act_qs = (int8_t *) (smem + DS_OFF_ACT);
    ActDS *  act_ds = (ActDS *) (smem + DS_OFF_ACT + NB * 32);
    float *  part   = (float *) (smem + DS_OFF_PART);
    const uint8_t * ring = smem + DS_OFF_RING;
    const uint32_t full0 = smem_u32(smem + DS_OFF_BAR), empty0 = full0 + DS_NSLOTS * 8;

    ds_quantize(st->x, NB, act_qs, act_ds, tid);
    cbar();

    // ---- which (K-slice, rows of a slot) this warp owns.  nsl = 32-lane slices per row.
    const int nsl = (steps + 31) >> 5;
    int s0, s1 = 0, rbase, wslot, ncol;
    bool wactive = true;
    if (TWO) {
        s0 = warp * 32 + lane; s1 = (warp + DS_CWARPS) * 32 + lane;
        rbase = 0; wslot = warp; ncol = DS_CWARPS;
    } else {
        const int rg_n = R / RW;                                        // row groups (host: RW = R / min(R, pow2floor(16 / nsl)))
        wactive = warp < nsl * rg_n;
        const int sl = warp % nsl, rg = warp / nsl;
        s0 = sl * 32 + lane;
        rbase = rg * RW; wslot = sl; ncol = nsl;
    }
    // lanes past the end of the row (and idle warps) compute on block 0 with zero activations: no divergence in the loop
    const bool v0 = wactive && s0 < steps, v1 = TWO && s1 < steps;
    if (!v0) s0 = 0;
    if (!v1) s1 = 0;
    const ActCol ac((const uint4 *) act_qs, (const uint4 *) act_qs + NB, (const int2 *) act_ds);
    typename A::Act a0 = A::load_act(ac, s0), a1;
    if (!v0) ds_zero_act<TYPE>(a0);
    if (TWO) { a1 = A::load_act(ac, s1); if (!v1) ds_zero_act<TYPE>(a1); }
    const int idx0 = rbase * steps + s0, idx1 = rbase * steps + s1;
    // slots per group: 8 items per warp at most, 32 rows per group at most, and leave the producer room to run ahead
    const int gs_max = min(min(8 / RW, 32 / R), DS_NSLOTS - 2);
    int foff[4];
    bool has_lead = false;
    #pragma unroll
    for (int f = 0; f < 4; f++) { foff[f] = st->foff[f]; has_lead |= st->flead[f] != 0; }

    for (int o = 0; o < st->nops; o++) {
        const DsOp & op = st->op[o];
        int lo, hi; row_share(op.M, lo, hi);
        float * dst = op.dst; float * add_dst = op.add_dst; const float * add_src = op.add_src;
        for (int g0 = lo; g0 < hi; ) {
            const int gs = min(gs_max, (hi - g0 + R - 1) / R);          // slots in this group
            const int ni = gs * RW;                                     // (slot, row) items of this warp
            int lead[4] = {0, 0, 0, 0}; int lead_step = 0;
            if (has_lead) {                                             // 2-byte scale arrays whose rows are not 16-B multiples
                #pragma unroll
                for (int f = 0; f < 4; f++) if (st->flead[f]) {
                    lead[f] = (int) (((uintptr_t) op.f[f] + (long long) g0 * st->rowbytes[f]) & 15); lead_step = R * st->rowbytes[f];
                }
            }
            float acc[8];
            #pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = 0.f;
            if (ni == 8) ds_group_dots<TYPE, TWO, RW, true >(acc, ni, ring, full0, S.cnt, foff, lead, lead_step, has_lead, idx0, idx1, steps, a0, a1, s0, s1);
            else         ds_group_dots<TYPE, TWO, RW, false>(acc, ni, ring, full0, S.cnt, foff, lead, lead_step, has_lead, idx0, idx1, steps, a0, a1, s0, s1);
            const float tot = reduce_rows<8>(acc, lane);                // lanes 4i..4i+3 hold item i
            float * pp = part + S.pbuf * (32 * DS_CWARPS);
            if ((lane & 3) == 0 && wactive) {
                const int it = lane >> 2;
                if (it < ni) pp[((it / RW) * R + rbase + it % RW) * DS_CWARPS + wslot] = tot;
            }
            __syncwarp();
            if (lane < gs) mbar_arrive(empty0 + ((S.cnt + lane) % DS_NSLOTS) * 8);     // this warp is done with the group's slots
            cbar();
            const int nrows = min(gs * R, hi - g0);
            if (tid < nrows) {
                float v = 0.f;
                for (int q = 0; q < ncol; q++) v += pp[tid * DS_CWARPS + q];
                const int row = g0 + tid;
                if (add_dst) { const float a = __ldcg(add_src + row); dst[row] = v; add_dst[row] = a + v; }
                else dst[row] = v;
            }
            S.pbuf ^= 1; S.cnt += gs; g0 += gs * R;
        }
    }
}

__device__ __forceinline__ void ds_add_stage(const DsStage * st, const int tid) {
    const int n4 = st->op[0].M / 4;
    const int lo = (int) ((long long) n4 * blockIdx.x / gridDim.x), hi = (int) ((long long) n4 * (blockIdx.x + 1) / gridDim.x);
    const float4 * a = (const float4 *) st->x; const float4 * b = (const float4 *) st->op[0].add_src;
    float4 * d = (float4 *) st->op[0].dst;
    for (int i = lo + tid; i < hi; i += DS_CTHREADS) {
        const float4 x = __ldcg(a + i), y = __ldcg(b + i);
        d[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    }
}

// ------------------------------------------------------------------------------------------------ kernel
// MASK: what the program contains (bit 0 q4_0, 1 q8_0, 2 q4_K, 3 q6_K, 4 rows longer than 16 K-slices); a variant carries
// only those code paths, so a pure q4_0 model does not pay the register budget of the q6_K unpacking
template <int MASK>
__global__ void __launch_bounds__(DS_THREADS, 1)
decode_stream_kernel(const DsStage * __restrict__ prog, const int nstages, uint32_t * __restrict__ counters, const uint32_t epoch,
                     long long * __restrict__ trace) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x;
    constexpr int DW = (int) (sizeof(DsStage) / 4);
    uint32_t * desc = (uint32_t *) (smem + DS_OFF_DESC);        // two stage descriptors (this stage, next stage)
    if (tid == 0) {
        const uint32_t full0 = smem_u32(smem + DS_OFF_BAR), empty0 = full0 + DS_NSLOTS * 8;
        for (int i = 0; i < DS_NSLOTS; i++) { mbar_init(full0 + i * 8, 1); mbar_init(empty0 + i * 8, DS_CWARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < DW) desc[tid] = __ldg((const uint32_t *) prog + tid);
    __syncthreads();
    if (tid >= DS_CTHREADS) {                       // producer warp: one lane streams, the rest leave
        if (tid == DS_CTHREADS) ds_producer(prog, nstages, smem);
        return;
    }
    DsState S{0u, 0};
    const uint32_t target = epoch * gridDim.x;
    for (int s = 0; s < nstages; s++) {
        const DsStage * st = (const DsStage *) (desc + (s & 1) * DS_DESC_WORDS);
        // development trace (B200_STREAM_TRACE): per stage {start, inputs ready, done, published} SM clocks of thread 0
        long long * tr = trace ? trace + ((long long) blockIdx.x * nstages + s) * 4 : nullptr;
        if (tr && tid == 0) tr[0] = clock64();
        if (s + 1 < nstages && tid < DW) desc[((s + 1) & 1) * DS_DESC_WORDS + tid] = __ldg((const uint32_t *) (prog + s + 1) + tid);
        if (s > 0 && st->wait_prev) {
            // grid-wide dependency: every CTA has published stage s-1 (bar.sync + red.release below), i.e. every earlier stage
            if (tid == 0) {
                const long long t0 = clock64();
                while ((int) (ld_relaxed(counters + (s - 1)) - target) < 0) {
                    if (clock64() - t0 > DS_TIMEOUT_CYCLES) __trap();
                }
                (void) ld_acquire(counters + (s - 1));
            }
            cbar();
        }
        if (tr && tid == 0) tr[1] = clock64();
        if (st->kind == DS_ADD) ds_add_stage(st, tid);
        else {
            const bool two = (MASK & 16) && st->steps > 32 * DS_CWARPS;
            const bool rw2 = !two && st->rw == 2;                       // only the 18-byte type packs R * nsl = 32
            switch (st->type) {
                case T_Q4_0: if (MASK & 1) {
                    if (two) ds_gemv_stage<T_Q4_0, true, 1>(st, smem, S, tid);
                    else if (rw2) ds_gemv_stage<T_Q4_0, false, 2>(st, smem, S, tid);
                    else ds_gemv_stage<T_Q4_0, false, 1>(st, smem, S, tid);
                } break;
                case T_Q8_0: if (MASK & 2) { if (two) ds_gemv_stage<T_Q8_0, true, 1>(st, smem, S, tid); else ds_gemv_stage<T_Q8_0, false, 1>(st, smem, S, tid); } break;
                case T_Q4_K: if (MASK & 4) ds_gemv_stage<T_Q4_K, false, 1>(st, smem, S, tid); break;      // K <= DS_KMAX: at most 14 slices
                case T_Q6_K: if (MASK & 8) ds_gemv_stage<T_Q6_K, false, 1>(st, smem, S, tid); break;
            }
        }
        // all consumer threads are done with stage s: its global results are written (bar.sync orders them before the
        // release below), the activation tile and the old descriptor may be overwritten, the next descriptor is in place
        if (tr && tid == 0) tr[2] = clock64();
        cbar();
        if (tid == 0) red_release(counters + s);
        if (tr && tid == 0) tr[3] = clock64();
    }
}

typedef void (*ds_kernel_t)(const DsStage *, int, uint32_t *, uint32_t, long long *);
static ds_kernel_t ds_variant(int mask) {
    if (mask & 16) return (mask & ~17) == 0 ? decode_stream_kernel<17> : decode_stream_kernel<31>;
    if ((mask & ~1) == 0)  return decode_stream_kernel<1>;
    if ((mask & ~2) == 0)  return decode_stream_kernel<2>;
    if ((mask & ~3) == 0)  return decode_stream_kernel<3>;
    if ((mask & ~4) == 0)  return decode_stream_kernel<4>;
    if ((mask & ~8) == 0)  return decode_stream_kernel<8>;
    if ((mask & ~12) == 0) return decode_stream_kernel<12>;
    return decode_stream_kernel<15>;
}
int ds_type_bit(int type) { return type == T_Q4_0 ? 1 : type == T_Q8_0 ? 2 : type == T_Q4_K ? 4 : type == T_Q6_K ? 8 : 0; }

// CTAs of the cooperative grid: one per SM (0 if the kernel cannot be resident)
int ds_grid(int device) {
    int sms = 0, occ = 0, coop = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device) != cudaSuccess || !coop) return 0;
    const int masks[] = {1, 2, 3, 4, 8, 12, 15, 17, 31};
    for (int m : masks) {
        if (cudaFuncSetAttribute(ds_variant(m), cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM) != cudaSuccess) { cudaGetLastError(); return 0; }
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ds_variant(m), DS_THREADS, DS_SMEM) != cudaSuccess || occ < 1) { cudaGetLastError(); return 0; }
    }
    return sms;
}

// ds_grid(device) must have succeeded on this device before (it raises the kernels' shared-memory limit)
cudaError_t launch_decode_stream(const DsStage * prog, int nstages, uint32_t * counters, uint32_t epoch, int grid, int type_mask,
                                 long long * trace, cudaStream_t st) {
    void * args[] = { (void *) &prog, (void *) &nstages, (void *) &counters, (void *) &epoch, (void *) &trace };
    return cudaLaunchCooperativeKernel((const void *) ds_variant(type_mask), dim3((unsigned) grid), dim3(DS_THREADS), args, (size_t) DS_SMEM, st);
}

}  // namespace b200

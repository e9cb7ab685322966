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
#include "wtypes.cuh"

namespace b200 {

// ---- f32 weights (the only MUL_MAT the cDSP path runs by default: ggml_vec_dot_f32, ggml-dsp.c:360-372).
// One warp per (row, column); float4 loads when 16-B aligned; fp32 accumulation.
__global__ void __launch_bounds__(128)
b200_gemv_f32_kernel(const char * __restrict__ w, int64_t wrs, int64_t M, int64_t K,
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
    b200_gemv_f32_kernel<<<(unsigned) ((M + 3) / 4), 128, 0, st>>>((const char *) w, wrs, M, K, (const char *) x, xcs, ncols, dst, dcs);
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
    cudaTriggerProgrammaticLaunchCompletion();   // dependents only prefetch weights until their own sync returns
    cudaGridDependencySynchronize();
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
template <bool MUL>
__global__ void __launch_bounds__(256)
add_f32_vec_kernel(const float4 * __restrict__ a, const float4 * __restrict__ b, float4 * __restrict__ d,
                   const int64_t n4, const int64_t nb4 /* period of b in float4, or n4 */) {
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
#endif
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 x = a[i], y = b[nb4 == n4 ? i : i % nb4];
        d[i] = MUL ? make_float4(__fmul_rn(x.x, y.x), __fmul_rn(x.y, y.y), __fmul_rn(x.z, y.z), __fmul_rn(x.w, y.w))
                   : make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    }
}
// General path: any byte strides, src1 repeated in every dim (i % ne1x), one thread per element.
template <bool MUL>
__global__ void __launch_bounds__(256) add_f32_gen_kernel(const AddArgs p, const int64_t n) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t t = i;
        const int64_t i0 = t % p.ne[0]; t /= p.ne[0];
        const int64_t i1 = t % p.ne[1]; t /= p.ne[1];
        const int64_t i2 = t % p.ne[2]; const int64_t i3 = t / p.ne[2];
        const float x = *(const float *) (p.a + i0 * p.nba[0] + i1 * p.nba[1] + i2 * p.nba[2] + i3 * p.nba[3]);
        const float y = *(const float *) (p.b + (i0 % p.neb[0]) * p.nbb[0] + (i1 % p.neb[1]) * p.nbb[1] +
                                                 (i2 % p.neb[2]) * p.nbb[2] + (i3 % p.neb[3]) * p.nbb[3]);
        *(float *) (p.d + i0 * p.nbd[0] + i1 * p.nbd[1] + i2 * p.nbd[2] + i3 * p.nbd[3]) = MUL ? __fmul_rn(x, y) : x + y;
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
        return p.mul ? cudaLaunchKernelEx(&cfg, add_f32_vec_kernel<true>, (const float4 *) p.a, (const float4 *) p.b, (float4 *) p.d, n4, nb_el / 4)
                     : cudaLaunchKernelEx(&cfg, add_f32_vec_kernel<false>, (const float4 *) p.a, (const float4 *) p.b, (float4 *) p.d, n4, nb_el / 4);
    }
    int64_t blocks = (n + 255) / 256; if (blocks > 148 * 16) blocks = 148 * 16;
    if (p.mul) add_f32_gen_kernel<true><<<(unsigned) blocks, 256, 0, st>>>(p, n);
    else       add_f32_gen_kernel<false><<<(unsigned) blocks, 256, 0, st>>>(p, n);
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

// ------------------------------------------------------------------------------------------------ row-split exchange
// Runs after a row-split GEMV (stream order: its NVLink stores are complete).  Lane p: publish "op number n of this
// rank is in your memory" into peer p's flag slot, then wait until peer p has published the same op number here.
// Counters only grow, so CUDA-graph replays need no reset.  The waits are bounded (trap instead of hanging the GPU).
struct ExchangeArgs { long long peer_flag_addr[7]; int peer_rank[7]; int npeers; };
__global__ void rowsplit_exchange_kernel(uint32_t * local_flags, uint32_t * counter, const ExchangeArgs a) {
    const int p = threadIdx.x;
    const uint32_t n = *counter + 1;
    __syncwarp();
    if (p < a.npeers) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(a.peer_flag_addr[p]), "r"(n) : "memory");
        const long long t0 = clock64();
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(local_flags + a.peer_rank[p]) : "memory");
            if (clock64() - t0 > 8000000000ll) __trap();
        } while ((int) (v - n) < 0);
    }
    __syncwarp();
    if (p == 0) *counter = n;
}
// Row-split prefill: the dense GEMM wrote this rank's rows x N slice of dst (column stride dcs) into LOCAL memory; copy it into every
// peer's dst at the same heap offset over NVLink (plain 16-byte stores; the flag exchange that follows publishes them)
struct PushPeers { long long delta[7]; };
template <bool VEC>
__global__ void __launch_bounds__(256)
rowsplit_push_kernel(const float * __restrict__ src, const long long rows, const long long N, const long long dcs, const int npeers, const PushPeers pp) {
    const long long per = VEC ? rows / 4 : rows, total = per * N;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long) gridDim.x * blockDim.x) {
        const long long n = i / per, m = i - n * per;
        if (VEC) {
            const float4 * p = (const float4 *) (src + n * dcs) + m;
            const float4 v = *p;
            for (int q = 0; q < npeers; q++) *(float4 *) ((char *) p + pp.delta[q]) = v;
        } else {
            const float * p = src + n * dcs + m;
            const float v = *p;
            for (int q = 0; q < npeers; q++) *(float *) ((char *) p + pp.delta[q]) = v;
        }
    }
}
cudaError_t launch_rowsplit_push(const float * slice, long long rows, long long N, long long dcs, int npeers, const long long * peer_delta, cudaStream_t st) {
    if (npeers <= 0 || rows <= 0 || N <= 0) return cudaSuccess;
    PushPeers pp{};
    for (int i = 0; i < npeers && i < 7; i++) pp.delta[i] = peer_delta[i];
    const bool vec = rows % 4 == 0 && dcs % 4 == 0 && ((uintptr_t) slice & 15) == 0;
    const long long total = (vec ? rows / 4 : rows) * N;
    long long blocks = (total + 255) / 256; if (blocks > 148 * 8) blocks = 148 * 8;
    if (vec) rowsplit_push_kernel<true><<<(unsigned) blocks, 256, 0, st>>>(slice, rows, N, dcs, npeers, pp);
    else     rowsplit_push_kernel<false><<<(unsigned) blocks, 256, 0, st>>>(slice, rows, N, dcs, npeers, pp);
    return cudaGetLastError();
}

cudaError_t launch_rowsplit_exchange(uint32_t * local_flags, uint32_t * counter, int npeers, const long long * peer_flag_addr,
                                     const int * peer_rank, cudaStream_t st) {
    ExchangeArgs a{};
    a.npeers = npeers;
    for (int i = 0; i < npeers; i++) { a.peer_flag_addr[i] = peer_flag_addr[i]; a.peer_rank[i] = peer_rank[i]; }
    rowsplit_exchange_kernel<<<1, 32, 0, st>>>(local_flags, counter, a);
    return cudaGetLastError();
}

// ---- LL protocol receiver: one thread per (matrix, peer row); spins on the slot until the peer's store of THIS exchange (sequence
// number) has landed, copies the value into the local dst vector.  Launched with programmatic dependent launch right behind the
// GEMV, so it polls while the local GEMV still runs; its completion implies the GEMV's (grid dependency at the end).
__global__ void __launch_bounds__(1024)
ll_gather_kernel(const LlGatherArgs a) {
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();           // the next quantise / GEMV may launch and fetch its weights meanwhile
#endif
    const uint32_t seq = *a.ll_count + a.ll_op;
    const uint2 * half = a.ll_slots + (size_t) (seq & 1u) * LL_HALF_ELEMS;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    for (int m = 0; m < a.nmat; m++) {
        const int M = a.M[m], total = M * a.world;
        const uint2 * slots = half + a.ll_off[m];
        for (int i = gtid; i < total; i += gsize) {
            if (i / M == a.rank) continue;
            uint32_t v, f;
            const long long t0 = clock64();
            do {
                asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(f) : "l"(slots + i) : "memory");
                if (f != seq && clock64() - t0 > 8000000000ll) __trap();
            } while (f != seq);
            a.dst[m][i] = __uint_as_float(v);
        }
    }
#if __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();                      // the local slice (plain stores of the GEMV before us) is complete too
#endif
}
cudaError_t launch_ll_gather(const LlGatherArgs & a, cudaStream_t st) {
    long long total = 0;
    for (int m = 0; m < a.nmat; m++) total += (long long) a.M[m] * a.world;
    int blocks = (int) ((total + 1023) / 1024); if (blocks > 64) blocks = 64; if (blocks < 1) blocks = 1;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) blocks); cfg.blockDim = dim3(1024); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, ll_gather_kernel, a);
}
// The same gather with the work of the next two nodes folded in (nmat vectors of M x world elements each, M % 32 == 0): for vector f.qmat,
// optionally z = gathered + add_src (the residual GGML_OP_ADD that consumes it, stored to add_dst), then the q8_1 quantisation of the
// result into the activation scratch of the MUL_MAT that reads it next (quantize_q8_1_kernel's arithmetic and layout, 8 lanes per block).
// Per decode layer of a row split this leaves 2 launches per exchange (GEMV, gather) instead of 3 - 4 (+ ADD, + quantise).
// The scratch may still be read by the local GEMV while we poll: it is written only after the grid dependency has resolved.
__global__ void __launch_bounds__(1024)
ll_gather_fused_kernel(const LlGatherArgs a, const LlFuseArgs f) {
#if __CUDA_ARCH__ >= 900
    cudaTriggerProgrammaticLaunchCompletion();
#endif
    const uint32_t seq = *a.ll_count + a.ll_op;
    const uint2 * half = a.ll_slots + (size_t) (seq & 1u) * LL_HALF_ELEMS;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    bool synced = false;
    for (int m = 0; m < a.nmat; m++) {
        const int M = a.M[m], total4 = M * a.world / 4, NB = M * a.world / 32;
        const uint2 * slots = half + a.ll_off[m];
        for (int base = 0; base < total4; base += gsize) {           // (uniform trip count: the shuffles below want whole warps)
            const int i4 = base + gtid, i = i4 * 4;
            const bool on = i4 < total4, own = on && i / M == a.rank;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (on && !own) {
                uint4 s0, s1;
                const long long t0 = clock64();
                for (;;) {
                    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(s0.x), "=r"(s0.y), "=r"(s0.z), "=r"(s0.w) : "l"(slots + i) : "memory");
                    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(s1.x), "=r"(s1.y), "=r"(s1.z), "=r"(s1.w) : "l"(slots + i + 2) : "memory");
                    if (s0.y == seq && s0.w == seq && s1.y == seq && s1.w == seq) break;
                    if (clock64() - t0 > 8000000000ll) __trap();
                }
                v = make_float4(__uint_as_float(s0.x), __uint_as_float(s0.z), __uint_as_float(s1.x), __uint_as_float(s1.z));
                *(float4 *) (a.dst[m] + i) = v;
            }
#if __CUDA_ARCH__ >= 900
            if (!synced) { cudaGridDependencySynchronize(); synced = true; }      // the local GEMV is complete: its slice, and the scratch, are ours
#endif
            if (own) v = __ldcg((const float4 *) (a.dst[m] + i));
            if (m != f.qmat) continue;
            if (f.add_dst && on) {
                const float4 o = *(const float4 *) (f.add_src + i);
                v = make_float4(v.x + o.x, v.y + o.y, v.z + o.z, v.w + o.w);
                *(float4 *) (f.add_dst + i) = v;
            }
            const int sub = gtid & 7, b = i4 >> 3;
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
            if (on) {
                const uint32_t packed = (uint32_t) (q0 & 0xFF) | ((uint32_t) (q1 & 0xFF) << 8) | ((uint32_t) (q2 & 0xFF) << 16) | ((uint32_t) (q3 & 0xFF) << 24);
                *(uint32_t *) (f.qs + ((int64_t) (sub >> 2) * NB + b) * 16 + (sub & 3) * 4) = packed;
                if (sub == 0) { ActDS o; o.d = __half2float(__float2half_rn(d)); o.isum0 = (short) sum; o.isum1 = (short) other; f.ds[b] = o; }
            }
        }
    }
#if __CUDA_ARCH__ >= 900
    if (!synced) cudaGridDependencySynchronize();
#endif
}
cudaError_t launch_ll_gather_fused(const LlGatherArgs & a, const LlFuseArgs & f, cudaStream_t st) {
    long long most = 0;
    for (int m = 0; m < a.nmat; m++) { const long long t = (long long) a.M[m] * a.world / 4; if (t > most) most = t; if (a.M[m] % 32) return cudaErrorInvalidValue; }
    int blocks = (int) ((most + 1023) / 1024); if (blocks > 64) blocks = 64; if (blocks < 1) blocks = 1;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) blocks); cfg.blockDim = dim3(1024); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = gemv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, ll_gather_fused_kernel, a, f);
}
__global__ void ll_bump_kernel(uint32_t * ll_count, const uint32_t nops) { *ll_count += nops; }
cudaError_t launch_ll_bump(uint32_t * ll_count, uint32_t nops, cudaStream_t st) {
    ll_bump_kernel<<<1, 1, 0, st>>>(ll_count, nops);
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

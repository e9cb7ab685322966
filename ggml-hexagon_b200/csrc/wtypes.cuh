// wtypes.cuh — device-side views of the four quantised weight layouts (SoA, b200_internal.cuh) and the exact
// integer dot products against q8_1 activations.  Shared by the GEMV kernels (gemv.cu) and the parity
// instrument block_dots (kernels.cu).
#pragma once
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
__device__ __forceinline__ uint4 ldg_cached(const void * p) {     // activations: shared memory or L1-cached global scratch
    return *(const uint4 *) p;
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
    __device__ __forceinline__ ActCol(const uint4 * plane0, const uint4 * plane1, const int2 * d) : p0(plane0), p1(plane1), ds(d) {}
};
struct ADesc { float d; int isum0, isum1; };
__device__ __forceinline__ ADesc load_adesc(const int2 * p) {
    int2 v = *p;
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
    // same fragment from a shared-memory copy of the SoA fields (gemv5 / gemv5c / gemv5n tiles); plain loads -> LDS
    static __device__ __forceinline__ Frag loads(const WPtrs & w, int blk) {
        Frag f; f.q = *(const uint4 *) (w.p[0] + blk * 16); f.d = __half2float(*(const __half *) (w.p[1] + blk * 2)); return f;
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
    static __device__ __forceinline__ Frag loads(const WPtrs & w, int blk) {
        Frag f; f.q0 = *(const uint4 *) (w.p[0] + blk * 32); f.q1 = *(const uint4 *) (w.p[0] + blk * 32 + 16);
        f.d = __half2float(*(const __half *) (w.p[1] + blk * 2)); return f;
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
    static __device__ __forceinline__ Frag loads(const WPtrs & w, int step) {
        Frag f;
        f.q0 = *(const uint4 *) (w.p[0] + step * 32); f.q1 = *(const uint4 *) (w.p[0] + step * 32 + 16);
        f.hdr = *(const uint4 *) (w.p[1] + (step >> 2) * 16);
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
    static __device__ __forceinline__ Frag loads(const WPtrs & w, int step) {
        const int sb = step >> 2; const int h = (step & 3) >> 1, j = step & 1;
        Frag f;
        f.l0 = *(const uint4 *) (w.p[0] + sb * 128 + 64 * h + 16 * j);
        f.l1 = *(const uint4 *) (w.p[0] + sb * 128 + 64 * h + 32 + 16 * j);
        f.hq = *(const uint4 *) (w.p[1] + sb * 64 + 32 * h + 16 * j);
        f.sc = *(const uint2 *) (w.p[2] + sb * 16 + 8 * h);
        f.d  = __half2float(*(const __half *) (w.p[3] + sb * 2));
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
        sd  = __int_as_float(a.ds[sbh + 0].x) * (float) (d4[0] * sc0);
        sd += __int_as_float(a.ds[sbh + 1].x) * (float) (d4[1] * sc1);
        sd += __int_as_float(a.ds[sbh + 2].x) * (float) (d4[2] * sc2);
        sd += __int_as_float(a.ds[sbh + 3].x) * (float) (d4[3] * sc3);
        return f.d * sd;
    }
};

// ------------------------------------------------------------------------------------------------ register-resident activations
// GEMV v4 (gemv.cu): a lane owns ONE step of the row for the whole CTA tile, so it loads that step's activations into
// registers once (WR<T>::Act) and then only streams weights.  dp4a.u32.s32 lets the high nibbles stay in place:
// dot(q & 0xF0F0F0F0, a) = 16 * dot(q >> 4, a), exact, so no shift per word is needed.
__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
    int d; asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ int dot16_lo(const uint4 & q, const uint4 & a) {     // low nibbles
    int s = dp4a_u8s8(q.x & 0x0F0F0F0Fu, a.x, 0); s = dp4a_u8s8(q.y & 0x0F0F0F0Fu, a.y, s);
    s = dp4a_u8s8(q.z & 0x0F0F0F0Fu, a.z, s); return dp4a_u8s8(q.w & 0x0F0F0F0Fu, a.w, s);
}
__device__ __forceinline__ int dot16_hi16(const uint4 & q, const uint4 & a) {   // 16 x (high nibbles . a)
    int s = dp4a_u8s8(q.x & 0xF0F0F0F0u, a.x, 0); s = dp4a_u8s8(q.y & 0xF0F0F0F0u, a.y, s);
    s = dp4a_u8s8(q.z & 0xF0F0F0F0u, a.z, s); return dp4a_u8s8(q.w & 0xF0F0F0F0u, a.w, s);
}

template <int TYPE> struct WR;

template <> struct WR<T_Q4_0> {
    struct Act { uint4 a0, a1; float d; int isum8; };
    static __device__ __forceinline__ Act load_act(const ActCol & a, int s) {
        Act r; r.a0 = a.p0[s]; r.a1 = a.p1[s];
        const ADesc ad = load_adesc(a.ds + s); r.d = ad.d; r.isum8 = 8 * (ad.isum0 + ad.isum1);
        return r;
    }
    static __device__ __forceinline__ float dot(const WT<T_Q4_0>::Frag & f, const Act & a) {
        const int sumi = dot16_lo(f.q, a.a0) + (dot16_hi16(f.q, a.a1) >> 4) - a.isum8;
        return (f.d * a.d) * (float) sumi;
    }
};

template <> struct WR<T_Q8_0> {
    struct Act { uint4 a0, a1; float d; };
    static __device__ __forceinline__ Act load_act(const ActCol & a, int s) {
        Act r; r.a0 = a.p0[s]; r.a1 = a.p1[s]; r.d = load_adesc(a.ds + s).d; return r;
    }
    static __device__ __forceinline__ float dot(const WT<T_Q8_0>::Frag & f, const Act & a) {
        const int sumi = dot16_ss(f.q0, a.a0) + dot16_ss(f.q1, a.a1);
        return (f.d * a.d) * (float) sumi;
    }
};

template <> struct WR<T_Q4_K> {
    struct Act { uint4 a00, a01, a10, a11; float d0, d1; int isum0, isum1; };
    static __device__ __forceinline__ Act load_act(const ActCol & a, int s) {
        Act r; const int b0 = 2 * s, b1 = 2 * s + 1;
        r.a00 = a.p0[b0]; r.a01 = a.p1[b0]; r.a10 = a.p0[b1]; r.a11 = a.p1[b1];
        const ADesc x = load_adesc(a.ds + b0), y = load_adesc(a.ds + b1);
        r.d0 = x.d; r.d1 = y.d; r.isum0 = x.isum0 + x.isum1; r.isum1 = y.isum0 + y.isum1;
        return r;
    }
    // s = step index within the row (only s & 3 matters: which 64-element chunk of the super-block)
    static __device__ __forceinline__ float dot(const WT<T_Q4_K>::Frag & f, const Act & a, int s) {
        const int dlo = dot16_lo(f.q0, a.a00) + dot16_lo(f.q1, a.a01);
        const int dhi = (dot16_hi16(f.q0, a.a10) + dot16_hi16(f.q1, a.a11)) >> 4;
        const int c = s & 3;
        int sc0, m0, sc1, m1;
        k4_scale_min(f.hdr.y, f.hdr.z, f.hdr.w, 2 * c,     sc0, m0);
        k4_scale_min(f.hdr.y, f.hdr.z, f.hdr.w, 2 * c + 1, sc1, m1);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&f.hdr.x));
        const float sd = a.d0 * (float) (dlo * sc0) + a.d1 * (float) (dhi * sc1);
        const float sm = a.d0 * (float) (a.isum0 * m0) + a.d1 * (float) (a.isum1 * m1);
        return dm.x * sd - dm.y * sm;
    }
};

template <> struct WR<T_Q6_K> {
    struct Act { uint4 a[4]; float d[4]; int isum32[4]; };
    static __device__ __forceinline__ Act load_act(const ActCol & a, int s) {
        Act r; const int sbh = (s >> 1) * 4, j = s & 1;
        const uint4 * pl = j ? a.p1 : a.p0;
        #pragma unroll
        for (int q = 0; q < 4; q++) {
            r.a[q] = pl[sbh + q];
            const ADesc ad = load_adesc(a.ds + sbh + q);
            r.d[q] = ad.d; r.isum32[q] = 32 * (j ? ad.isum1 : ad.isum0);
        }
        return r;
    }
    static __device__ __forceinline__ float dot(const WT<T_Q6_K>::Frag & f, const Act & a, int s) {
        const int j = s & 1;
        const uint32_t M3 = 0x30303030u, ML = 0x0F0F0F0Fu;
        uint4 v0, v1, v2, v3;
        v0 = make_uint4((f.l0.x & ML) | ((f.hq.x << 4) & M3), (f.l0.y & ML) | ((f.hq.y << 4) & M3), (f.l0.z & ML) | ((f.hq.z << 4) & M3), (f.l0.w & ML) | ((f.hq.w << 4) & M3));
        v1 = make_uint4((f.l1.x & ML) | ((f.hq.x << 2) & M3), (f.l1.y & ML) | ((f.hq.y << 2) & M3), (f.l1.z & ML) | ((f.hq.z << 2) & M3), (f.l1.w & ML) | ((f.hq.w << 2) & M3));
        v2 = make_uint4(((f.l0.x >> 4) & ML) | (f.hq.x & M3), ((f.l0.y >> 4) & ML) | (f.hq.y & M3), ((f.l0.z >> 4) & ML) | (f.hq.z & M3), ((f.l0.w >> 4) & ML) | (f.hq.w & M3));
        v3 = make_uint4(((f.l1.x >> 4) & ML) | ((f.hq.x >> 2) & M3), ((f.l1.y >> 4) & ML) | ((f.hq.y >> 2) & M3), ((f.l1.z >> 4) & ML) | ((f.hq.z >> 2) & M3), ((f.l1.w >> 4) & ML) | ((f.hq.w >> 2) & M3));
        const int d0 = dot16_ss(v0, a.a[0]) - a.isum32[0], d1 = dot16_ss(v1, a.a[1]) - a.isum32[1];
        const int d2 = dot16_ss(v2, a.a[2]) - a.isum32[2], d3 = dot16_ss(v3, a.a[3]) - a.isum32[3];
        const uint32_t lo = f.sc.x >> (8 * j), hi = f.sc.y >> (8 * j);
        const int sc0 = (int8_t) (lo & 0xFF), sc1 = (int8_t) ((lo >> 16) & 0xFF);
        const int sc2 = (int8_t) (hi & 0xFF), sc3 = (int8_t) ((hi >> 16) & 0xFF);
        const float sd = a.d[0] * (float) (d0 * sc0) + a.d[1] * (float) (d1 * sc1) + a.d[2] * (float) (d2 * sc2) + a.d[3] * (float) (d3 * sc3);
        return f.d * sd;
    }
};
// uniform call shape for the kernel
template <int TYPE> __device__ __forceinline__ float wr_dot(const typename WT<TYPE>::Frag & f, const typename WR<TYPE>::Act & a, int s) {
    if constexpr (TYPE == T_Q4_0 || TYPE == T_Q8_0) return WR<TYPE>::dot(f, a); else return WR<TYPE>::dot(f, a, s);
}

// ------------------------------------------------------------------------------------------------ lane reductions
// v[0..R) per lane -> total over the 32 lanes of v[row], returned in every lane with row == (lane >> (5-log2 R)) & (R-1)
template <int R>
__device__ __forceinline__ float reduce_rows(float (&v)[R], const int lane) {
    int m = 16;
    #pragma unroll
    for (int n = R; n > 1; n >>= 1, m >>= 1) {
        const bool up = lane & m;
        #pragma unroll
        for (int i = 0; i < n / 2; i++) {
            const float keep = up ? v[i + n / 2] : v[i];
            const float send = up ? v[i] : v[i + n / 2];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
        }
    }
    for (; m > 0; m >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], m);
    return v[0];
}
template <int R> struct Log2 { static constexpr int v = 1 + Log2<R / 2>::v; };
template <> struct Log2<1> { static constexpr int v = 0; };

}  // namespace b200

/*
 * ggml_oracle.c — CPU restatement of the reference's MUL_MAT / ADD hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (libb200ops.so, libggml-b200.so)
 * links, loads or calls this file.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it, and only as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below
 *   (a) against the reference itself (oracle/_ref/<variant>/libggml-{base,cpu}.so, built by
 *       oracle/Makefile.ref from the unmodified sources under /root/reference), and
 *   (b) against the fixtures in tests/golden/ that tests/golden/make_golden.py produced by
 *       calling those same reference libraries.
 * The reference holds no stored known-answer vectors for this path.
 *
 * Each function cites the reference lines it restates.  Blocks are addressed as raw bytes
 * (little-endian), the wire format of ggml/src/ggml-common.h:167-334:
 *   q4_0 : 18 B / 32 w  = f16 d | 16 B nibbles (lo nibble = elem j, hi nibble = elem j+16)
 *   q8_0 : 34 B / 32 w  = f16 d | 32 x i8
 *   q8_1 : 36 B / 32 w  = f16 d | f16 s (= d * sum q) | 32 x i8
 *   q4_K : 144 B / 256 w = f16 d | f16 dmin | 12 B packed 6-bit (scale,min) x 8 | 128 B nibbles
 *   q6_K : 210 B / 256 w = 128 B ql | 64 B qh | 16 x i8 scales | f16 d
 *   q8_K : 292 B / 256 w = f32 d | 256 x i8 | 16 x i16 bsums
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { OT_F32 = 0, OT_Q4_0 = 2, OT_Q8_0 = 8, OT_Q8_1 = 9, OT_Q4_K = 12, OT_Q6_K = 14, OT_Q8_K = 15 };
/* numeric values = enum ggml_type, ggml/include/ggml.h:349-390 */

/* ------------------------------------------------------------------ fp16 <-> fp32 */
/* IEEE binary16 conversions; bit-exact equivalent of ggml_compute_fp16_to_fp32 /
 * ggml_compute_fp32_to_fp16 (ggml/src/ggml-impl.h:357-413, copy at ggml-dsp.c:541-594):
 * round-to-nearest-even, subnormals kept, overflow -> inf, NaN -> quiet NaN. */
float oracle_fp16_to_fp32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp  = (h >> 10) & 0x1f;
    uint32_t man  = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: renormalise */
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    float f; memcpy(&f, &bits, 4); return f;
}

uint16_t oracle_fp32_to_fp16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    uint32_t ax = x & 0x7fffffffu;
    if (ax > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);            /* NaN */
    if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);           /* rounds to >= 65520 -> inf */
    if (ax < 0x33000001u) return sign;                                  /* <= 2^-25 -> 0 (ties-to-even) */
    int32_t e = (int32_t)(ax >> 23) - 127;
    uint32_t m = (ax & 0x7fffffu) | 0x800000u;                          /* 24-bit significand */
    int shift;                                                          /* bits to drop */
    uint32_t hexp;
    if (e < -14) { shift = 13 + (-14 - e); hexp = 0; }                  /* subnormal half */
    else         { shift = 13;             hexp = (uint32_t)(e + 15); }
    uint32_t kept = m >> shift;
    uint32_t rem  = m & ((1u << shift) - 1);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (kept & 1))) kept++;
    uint32_t out;
    if (hexp == 0) out = kept;                          /* may carry into exponent 1: correct */
    else           out = ((hexp - 1) << 10) + kept;     /* kept has the implicit bit at 0x400 */
    return (uint16_t)(sign | out);
}

static inline float rd_f16(const uint8_t *p) { uint16_t h; memcpy(&h, p, 2); return oracle_fp16_to_fp32(h); }
static inline void  wr_f16(uint8_t *p, float f) { uint16_t h = oracle_fp32_to_fp16(f); memcpy(p, &h, 2); }

int64_t oracle_row_size(int type, int64_t k) {
    switch (type) {
        case OT_F32:  return k * 4;
        case OT_Q4_0: return k / 32 * 18;
        case OT_Q8_0: return k / 32 * 34;
        case OT_Q8_1: return k / 32 * 36;
        case OT_Q4_K: return k / 256 * 144;
        case OT_Q6_K: return k / 256 * 210;
        case OT_Q8_K: return k / 256 * 292;
    }
    return -1;
}

/* ------------------------------------------------------------------ weight quantisers (simple types) */
/* quantize_row_q4_0_ref, ggml/src/ggml-quants.c:31-66 */
void oracle_quantize_row_q4_0(const float *x, uint8_t *y, int64_t k) {
    for (int64_t b = 0; b < k / 32; b++, x += 32, y += 18) {
        float amax = 0.0f, vmax = 0.0f;
        for (int j = 0; j < 32; j++) if (amax < fabsf(x[j])) { amax = fabsf(x[j]); vmax = x[j]; }
        const float d  = vmax / -8;
        const float id = d ? 1.0f / d : 0.0f;
        wr_f16(y, d);
        for (int j = 0; j < 16; j++) {
            int lo = (int8_t)(x[j] * id + 8.5f);       if (lo > 15) lo = 15;
            int hi = (int8_t)(x[j + 16] * id + 8.5f);  if (hi > 15) hi = 15;
            y[2 + j] = (uint8_t)(lo | (hi << 4));
        }
    }
}

/* quantize_row_q8_0_ref, ggml/src/ggml-quants.c:194-217 */
void oracle_quantize_row_q8_0(const float *x, uint8_t *y, int64_t k) {
    for (int64_t b = 0; b < k / 32; b++, x += 32, y += 34) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) amax = fmaxf(amax, fabsf(x[j]));
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        wr_f16(y, d);
        for (int j = 0; j < 32; j++) ((int8_t *)y)[2 + j] = (int8_t)roundf(x[j] * id);
    }
}

/* ------------------------------------------------------------------ activation quantisers */
/* quantize_row_q8_1_ref, ggml/src/ggml-quants.c:220-253.  s = fp16(sum * d) with the
 * UNROUNDED f32 d (the CUDA quantize.cu:27-36 variant stores sum(x) instead; we follow _ref). */
void oracle_quantize_row_q8_1(const float *x, uint8_t *y, int64_t k) {
    for (int64_t b = 0; b < k / 32; b++, x += 32, y += 36) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) amax = fmaxf(amax, fabsf(x[j]));
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        int sum = 0;
        for (int j = 0; j < 32; j++) {
            int8_t q = (int8_t)roundf(x[j] * id);
            ((int8_t *)y)[4 + j] = q;
            sum += q;
        }
        wr_f16(y, d);
        wr_f16(y + 2, sum * d);
    }
}

/* nearest_int, ggml/src/ggml-quants.c:372-377 (round-half-even through the 1.5*2^23 trick) */
static inline int nearest_int_rne(float v) {
    float t = v + 12582912.f;
    int i; memcpy(&i, &t, 4);
    return (i & 0x007fffff) - 0x00400000;
}

/* quantize_row_q8_K_ref, ggml/src/ggml-quants.c:2479-2516 */
void oracle_quantize_row_q8_K(const float *x, uint8_t *y, int64_t k) {
    for (int64_t b = 0; b < k / 256; b++, x += 256, y += 292) {
        float vmax = 0, amax = 0;
        for (int j = 0; j < 256; j++) { float ax = fabsf(x[j]); if (ax > amax) { amax = ax; vmax = x[j]; } }
        int8_t  *qs = (int8_t *)(y + 4);
        if (!amax) { float z = 0; memcpy(y, &z, 4); memset(qs, 0, 256); continue; } /* bsums untouched, as in ref */
        const float iscale = -127.f / vmax;
        for (int j = 0; j < 256; j++) { int v = nearest_int_rne(iscale * x[j]); qs[j] = (int8_t)(v < 127 ? v : 127); }
        for (int g = 0; g < 16; g++) {
            int s = 0; for (int i = 0; i < 16; i++) s += qs[g * 16 + i];
            int16_t s16 = (int16_t)s; memcpy(y + 260 + 2 * g, &s16, 2);
        }
        float d = 1 / iscale; memcpy(y, &d, 4);
    }
}

/* ------------------------------------------------------------------ K-quant scale unpack */
/* get_scale_min_k4, ggml/src/ggml-quants.c:631-638: 8 x (6-bit scale, 6-bit min) in 12 bytes */
static inline void k4_scale_min(int j, const uint8_t *q, int *sc, int *mn) {
    if (j < 4) { *sc = q[j] & 63;                               *mn = q[j + 4] & 63; }
    else       { *sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); *mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

/* unpack one block's weights to the integers the dot products use (before scaling):
 *   q4_0: nibble-8  in [-8,7]     q8_0: i8      q4_K: nibble in [0,15]     q6_K: 6-bit-32 in [-32,31] */
static void unpack_ints(int type, const uint8_t *blk, int8_t *out /* 32 or 256 */) {
    if (type == OT_Q4_0) {                       /* ggml-cpu-quants.c:2596-2600 */
        for (int j = 0; j < 16; j++) { out[j] = (int8_t)((blk[2 + j] & 0xF) - 8); out[j + 16] = (int8_t)((blk[2 + j] >> 4) - 8); }
    } else if (type == OT_Q8_0) {
        memcpy(out, blk + 2, 32);
    } else if (type == OT_Q4_K) {                /* ggml-cpu-quants.c:7553-7558; ggml-quants.c:1292-1299 */
        const uint8_t *q = blk + 16;
        for (int c = 0; c < 4; c++, q += 32, out += 64)
            for (int l = 0; l < 32; l++) { out[l] = (int8_t)(q[l] & 0xF); out[l + 32] = (int8_t)(q[l] >> 4); }
    } else if (type == OT_Q6_K) {                /* ggml-cpu-quants.c:9438-9448; ggml-quants.c:1701-1717 */
        const uint8_t *ql = blk, *qh = blk + 128;
        for (int h = 0; h < 2; h++, ql += 64, qh += 32, out += 128)
            for (int l = 0; l < 32; l++) {
                out[l +  0] = (int8_t)(((ql[l]      & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32);
                out[l + 32] = (int8_t)(((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32);
                out[l + 64] = (int8_t)(((ql[l]      >>  4) | (((qh[l] >> 4) & 3) << 4)) - 32);
                out[l + 96] = (int8_t)(((ql[l + 32] >>  4) | (((qh[l] >> 6) & 3) << 4)) - 32);
            }
    }
}

/* ------------------------------------------------------------------ dequantisers */
/* dequantize_row_q4_0 / q8_0 / q4_K / q6_K, ggml/src/ggml-quants.c:255-273, 330-345(ff), 1280-1302, 1690-1719 */
void oracle_dequantize_row(int type, const uint8_t *x, float *y, int64_t k) {
    int8_t w[256];
    if (type == OT_F32) { memcpy(y, x, (size_t)k * 4); return; }
    if (type == OT_Q4_0 || type == OT_Q8_0) {
        const int bs = type == OT_Q4_0 ? 18 : 34;
        for (int64_t b = 0; b < k / 32; b++, x += bs, y += 32) {
            const float d = rd_f16(x);
            unpack_ints(type, x, w);
            for (int j = 0; j < 32; j++) y[j] = w[j] * d;
        }
    } else if (type == OT_Q4_K) {
        for (int64_t b = 0; b < k / 256; b++, x += 144, y += 256) {
            const float d = rd_f16(x), dmin = rd_f16(x + 2);
            unpack_ints(type, x, w);
            for (int g = 0; g < 8; g++) {
                int sc, mn; k4_scale_min(g, x + 4, &sc, &mn);
                const float dg = d * sc, mg = dmin * mn;
                for (int l = 0; l < 32; l++) y[g * 32 + l] = dg * w[g * 32 + l] - mg;
            }
        }
    } else if (type == OT_Q6_K) {
        for (int64_t b = 0; b < k / 256; b++, x += 210, y += 256) {
            const float d = rd_f16(x + 208);
            const int8_t *sc = (const int8_t *)(x + 192);
            unpack_ints(type, x, w);
            for (int j = 0; j < 256; j++) y[j] = d * sc[j / 16] * w[j];
        }
    }
}

/* ------------------------------------------------------------------ vec_dot, ggml-cpu semantics */
/* ggml_vec_dot_q4_0_q8_0 scalar tail, ggml-cpu-quants.c:2591-2604
 * ggml_vec_dot_q8_0_q8_0 scalar tail, ggml-cpu-quants.c:4004-4012
 *   per block: sumi (exact int32), sumf += sumi * (d_w * d_a)                     */
float oracle_vec_dot_q8_0(int type, int64_t k, const uint8_t *w, const uint8_t *a /* q8_0 */) {
    const int bs = type == OT_Q4_0 ? 18 : 34;
    int8_t wi[32];
    float sumf = 0;
    for (int64_t b = 0; b < k / 32; b++, w += bs, a += 34) {
        unpack_ints(type, w, wi);
        const int8_t *q = (const int8_t *)(a + 2);
        int sumi = 0;
        for (int j = 0; j < 32; j++) sumi += wi[j] * q[j];
        if (type == OT_Q4_0) sumf += sumi * rd_f16(w) * rd_f16(a);      /* (sumi*d_w)*d_a, :2603 */
        else                 sumf += sumi * (rd_f16(w) * rd_f16(a));    /* sumi*(d_w*d_a), :4011 */
    }
    return sumf;
}

/* ggml_vec_dot_q4_K_q8_K generic, ggml-cpu-quants.c:7535-7590
 * ggml_vec_dot_q6_K_q8_K generic, ggml-cpu-quants.c:9423-9465
 *   8 float lanes `sums[l]`, lane l gathers elements with (index % 8) == l          */
float oracle_vec_dot_q8_K(int type, int64_t k, const uint8_t *w, const uint8_t *a /* q8_K */) {
    const int bs = type == OT_Q4_K ? 144 : 210;
    int8_t wi[256];
    float sums[8] = {0}, sumf = 0;
    for (int64_t b = 0; b < k / 256; b++, w += bs, a += 292) {
        int32_t aux32[8] = {0};
        float da; memcpy(&da, a, 4);
        const int8_t *q8 = (const int8_t *)(a + 4);
        unpack_ints(type, w, wi);
        if (type == OT_Q4_K) {
            int sumi = 0;
            for (int g = 0; g < 8; g++) {
                int sc, mn; k4_scale_min(g, w + 4, &sc, &mn);
                int16_t b0, b1; memcpy(&b0, a + 260 + 4 * g, 2); memcpy(&b1, a + 262 + 4 * g, 2);
                sumi += (b0 + b1) * mn;                                           /* :7566 */
                for (int j = 0; j < 32; j++) aux32[j % 8] += sc * (int16_t)(q8[g * 32 + j] * wi[g * 32 + j]);
            }
            const float d = rd_f16(w) * da;
            for (int l = 0; l < 8; l++) sums[l] += d * aux32[l];
            const float dmin = rd_f16(w + 2) * da;
            sumf -= dmin * sumi;
        } else {
            const int8_t *sc = (const int8_t *)(w + 192);
            for (int g = 0; g < 16; g++)
                for (int j = 0; j < 16; j++) aux32[j % 8] += sc[g] * (int16_t)(q8[g * 16 + j] * wi[g * 16 + j]);
            const float d = rd_f16(w + 208) * da;
            for (int l = 0; l < 8; l++) sums[l] += d * aux32[l];
        }
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* ------------------------------------------------------------------ per-32-group integer dots */
/* The exact integers every implementation must agree on, given the same int8 inputs:
 *   out[g] = sum_{j<32} W[32g+j] * A[32g+j], W = unpack_ints() above, A = int8 activation quants.
 * For q6_K (scale groups of 16) two values per 32-group: out[2g], out[2g+1].
 * This is the integer core of ggml-cpu-quants.c:2596-2601 / :4006-4009 / :7570-7582 / :9452-9460 and of the
 * reference's q8_1-based CUDA algebra (ggml-cuda/vecdotq.cuh:23-42,151-163,357-379,476-496).
 * `a_qs` is a plain int8[k] array. Returns number of ints written. */
int64_t oracle_block_dots(int type, int64_t k, const uint8_t *w, const int8_t *a_qs, int32_t *out) {
    int8_t wi[256];
    int64_t n = 0;
    if (type == OT_Q4_0 || type == OT_Q8_0) {
        const int bs = type == OT_Q4_0 ? 18 : 34;
        for (int64_t b = 0; b < k / 32; b++, w += bs, a_qs += 32) {
            unpack_ints(type, w, wi);
            int s = 0; for (int j = 0; j < 32; j++) s += wi[j] * a_qs[j];
            out[n++] = s;
        }
    } else {
        const int bs = type == OT_Q4_K ? 144 : 210;
        const int grp = type == OT_Q4_K ? 32 : 16;
        for (int64_t b = 0; b < k / 256; b++, w += bs, a_qs += 256) {
            unpack_ints(type, w, wi);
            for (int g = 0; g < 256 / grp; g++) {
                int s = 0; for (int j = 0; j < grp; j++) s += wi[g * grp + j] * a_qs[g * grp + j];
                out[n++] = s;
            }
        }
    }
    return n;
}

/* ------------------------------------------------------------------ vec_dot, q8_1 activation algebra */
/* The reference's own q8_1 formulation (ggml-cuda/vecdotq.cuh), restated with the min/offset
 * terms taken from the exact integer sum of the activation quants (isum = sum_j A[j]) instead of the
 * fp16 `s` field — this is what the B200 kernels compute, and for q4_0/q8_0 it equals the ggml-cpu
 * per-block expression sumi*d_w*d_a exactly (same integers, same two scales).
 * Accumulation order is block order, single float accumulator (order is NOT part of the contract;
 * tests compare with NMSE like tests/test-backend-ops.cpp:1982-1984). */
float oracle_vec_dot_q8_1(int type, int64_t k, const uint8_t *w, const uint8_t *a /* q8_1 AoS */) {
    int8_t wi[256];
    float acc = 0;
    if (type == OT_Q4_0 || type == OT_Q8_0) {
        const int bs = type == OT_Q4_0 ? 18 : 34;
        for (int64_t b = 0; b < k / 32; b++, w += bs, a += 36) {
            unpack_ints(type, w, wi);
            const int8_t *q = (const int8_t *)(a + 4);
            int s = 0; for (int j = 0; j < 32; j++) s += wi[j] * q[j];
            acc += (rd_f16(w) * rd_f16(a)) * (float)s;
        }
    } else if (type == OT_Q4_K) {
        for (int64_t b = 0; b < k / 256; b++, w += 144) {
            unpack_ints(type, w, wi);
            const float d = rd_f16(w), dmin = rd_f16(w + 2);
            float sd = 0, sm = 0;
            for (int g = 0; g < 8; g++, a += 36) {
                int sc, mn; k4_scale_min(g, w + 4, &sc, &mn);
                const int8_t *q = (const int8_t *)(a + 4);
                int dot = 0, isum = 0;
                for (int j = 0; j < 32; j++) { dot += wi[g * 32 + j] * q[j]; isum += q[j]; }
                const float d8 = rd_f16(a);
                sd += d8 * (float)(dot * sc);
                sm += d8 * (float)(isum * mn);
            }
            acc += d * sd - dmin * sm;
        }
    } else if (type == OT_Q6_K) {
        for (int64_t b = 0; b < k / 256; b++, w += 210) {
            unpack_ints(type, w, wi);
            const int8_t *sc = (const int8_t *)(w + 192);
            const float d = rd_f16(w + 208);
            float sd = 0;
            for (int g = 0; g < 8; g++, a += 36) {
                const int8_t *q = (const int8_t *)(a + 4);
                int d0 = 0, d1 = 0;
                for (int j = 0; j < 16; j++) { d0 += wi[g * 32 + j] * q[j]; d1 += wi[g * 32 + 16 + j] * q[16 + j]; }
                sd += rd_f16(a) * (float)(d0 * sc[2 * g] + d1 * sc[2 * g + 1]);
            }
            acc += d * sd;
        }
    }
    return acc;
}

/* ------------------------------------------------------------------ MUL_MAT */
/* ggml_compute_forward_mul_mat, ggml-cpu.c:6745-6937 (one_chunk :6655-6743; DSP twin ggml-dsp.c:1091-1351):
 *   dst[n*M + m] = vec_dot(K, W row m, quantised X row n)
 * W: M rows of type `type` (contiguous), X: N rows of K f32, dst: N columns of M f32 (not transposed).
 * mode 0 = ggml-cpu semantics: activations quantised to vec_dot_type (q8_0 for q4_0/q8_0, q8_K for K-quants;
 *          type_traits_cpu, ggml-cpu.c:268-344) — this is what test-backend-ops compares against.
 * mode 1 = q8_1 activation algebra (north_star's prescribed activation format).
 * f32 weights: plain float dot with float accumulation in index order (ggml_vec_dot_f32 scalar, vec.cpp). */
int oracle_mul_mat(int type, const uint8_t *w, const float *x, float *dst,
                   int64_t M, int64_t N, int64_t K, int mode) {
    const int64_t wrow = oracle_row_size(type, K);
    if (wrow < 0) return -1;
    if (type == OT_F32) {
        for (int64_t n = 0; n < N; n++)
            for (int64_t m = 0; m < M; m++) {
                const float *wr = (const float *)(w + m * wrow);
                double s = 0; /* ggml-dsp.c:360-372 accumulates in double (ggml_float) */
                for (int64_t i = 0; i < K; i++) s += (double)(wr[i] * x[n * K + i]);
                dst[n * M + m] = (float)s;
            }
        return 0;
    }
    const int kq = (type == OT_Q4_K || type == OT_Q6_K);
    const int atype = mode == 1 ? OT_Q8_1 : (kq ? OT_Q8_K : OT_Q8_0);
    const int64_t arow = oracle_row_size(atype, K);
    uint8_t *aq = (uint8_t *)calloc((size_t)arow, 1);
    if (!aq) return -2;
    for (int64_t n = 0; n < N; n++) {
        if (atype == OT_Q8_1)      oracle_quantize_row_q8_1(x + n * K, aq, K);
        else if (atype == OT_Q8_0) oracle_quantize_row_q8_0(x + n * K, aq, K);
        else                       oracle_quantize_row_q8_K(x + n * K, aq, K);
        for (int64_t m = 0; m < M; m++) {
            const uint8_t *wr = w + m * wrow;
            float v;
            if (atype == OT_Q8_1)      v = oracle_vec_dot_q8_1(type, K, wr, aq);
            else if (atype == OT_Q8_0) v = oracle_vec_dot_q8_0(type, K, wr, aq);
            else                       v = oracle_vec_dot_q8_K(type, K, wr, aq);
            dst[n * M + m] = v;
        }
    }
    free(aq);
    return 0;
}

/* ------------------------------------------------------------------ ADD */
/* ggml_compute_forward_add_f32 (ggml-dsp.c:991-1066) / binary-ops.cpp:50-117:
 * dst[i3,i2,i1,i0] = a[i3,i2,i1,i0] + b[i3%ne13, i2%ne12, i1%ne11, i0%ne10]; contiguous f32 tensors. */
void oracle_add_f32(const float *a, const float *b, float *dst, const int64_t ne0[4], const int64_t ne1[4]) {
    for (int64_t i3 = 0; i3 < ne0[3]; i3++)
    for (int64_t i2 = 0; i2 < ne0[2]; i2++)
    for (int64_t i1 = 0; i1 < ne0[1]; i1++) {
        const int64_t ra = ((i3 * ne0[2] + i2) * ne0[1] + i1) * ne0[0];
        const int64_t rb = (((i3 % ne1[3]) * ne1[2] + (i2 % ne1[2])) * ne1[1] + (i1 % ne1[1])) * ne1[0];
        for (int64_t i0 = 0; i0 < ne0[0]; i0++) dst[ra + i0] = a[ra + i0] + b[rb + i0 % ne1[0]];
    }
}

/* ------------------------------------------------------------------ RMS_NORM */
/* ggml_compute_forward_rms_norm_f32 (ggml-cpu/ggml-cpu.c:6254-6300); the op ggml-hexagon lists at ggml-hexagon.cpp:619 and leaves
 * as an empty stub in kernels/ggml-dsp.c:1353-1372.  Per row of ne00 floats: sum of squares of the f32 products in double,
 * mean = (float) (sum / ne00), scale = 1 / sqrtf(mean + eps), y = x * scale (ggml_vec_scale_f32). */
void oracle_rms_norm_f32(const float *x, float *y, int64_t ne00, int64_t nrows, float eps) {
    for (int64_t r = 0; r < nrows; r++) {
        const float *xr = x + r * ne00; float *yr = y + r * ne00;
        double sum = 0.0;
        for (int64_t i = 0; i < ne00; i++) sum += (double) (xr[i] * xr[i]);
        const float mean = (float) (sum / (double) ne00);
        const float scale = 1.0f / sqrtf(mean + eps);
        for (int64_t i = 0; i < ne00; i++) yr[i] = xr[i] * scale;
    }
}

/* ------------------------------------------------------------------ SOFT_MAX */
/* ggml_compute_forward_soft_max_f32 (ggml-cpu/ggml-cpu.c:8261-8357) with the scalar ggml_vec_soft_max_f32: per row of src0
 * [ne00, ne01, ne02, ne03]: wp = x * scale + slope * mask[row % ne01]  (mask: f32 [ne00, >= ne01], optional; ALiBi slope from max_bias
 * and the head index (row / ne01) % ne02), max, val = expf(wp - max), sum of the vals in double, y = val * (float) (1 / sum). */
void oracle_soft_max_f32(const float *x, const float *mask, float *y, int64_t ne00, int64_t ne01, int64_t ne02, int64_t ne03,
                         float scale, float max_bias) {
    const uint32_t n_head = (uint32_t) ne02;
    const uint32_t n_head_log2 = 1u << (uint32_t) floor(log2((double) n_head));
    const float m0 = powf(2.0f, -(max_bias) / n_head_log2);
    const float m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    const int64_t nr = ne01 * ne02 * ne03;
    float *wp = (float *) malloc(sizeof(float) * (size_t) (ne00 > 0 ? ne00 : 1));
    for (int64_t i1 = 0; i1 < nr; i1++) {
        const uint32_t h = (uint32_t) ((i1 / ne01) % ne02);
        const float slope = (max_bias > 0.0f) ? (h < n_head_log2 ? powf(m0, h + 1) : powf(m1, 2 * (h - n_head_log2) + 1)) : 1.0f;
        const float *sp = x + i1 * ne00; float *dp = y + i1 * ne00;
        const float *mp = mask ? mask + (i1 % ne01) * ne00 : NULL;
        for (int64_t i = 0; i < ne00; i++) wp[i] = sp[i] * scale;
        if (mp) for (int64_t i = 0; i < ne00; i++) wp[i] += slope * mp[i];
        float max = -INFINITY;
        for (int64_t i = 0; i < ne00; i++) if (wp[i] > max) max = wp[i];
        double sum = 0.0;
        for (int64_t i = 0; i < ne00; i++) { const float v = expf(wp[i] - max); sum += (double) v; dp[i] = v; }
        const float inv = (float) (1.0 / sum);
        for (int64_t i = 0; i < ne00; i++) dp[i] *= inv;
    }
    free(wp);
}

/* ------------------------------------------------------------------ POOL_2D */
/* ggml_compute_forward_pool_2d (ggml-cpu/ggml-cpu.c:9758-9830), f32 input: for every plane [ne00, ne01] of src (ne02 * ne03 of them) and
 * every output (ox, oy): op 0 = max (start -FLT_MAX), op 1 = average: sum in ky-outer / kx-inner order of the taps inside the input,
 * divided by k0 * k1 (padding taps count in the divisor).  Output plane [px, py]. */
void oracle_pool_2d_f32(const float *x, float *y, const int64_t ne[4], int64_t px, int64_t py, int op, int k0, int k1, int s0, int s1, int p0, int p1) {
    const int64_t planes = ne[2] * ne[3];
    for (int64_t pl = 0; pl < planes; pl++) {
        const float *src = x + pl * ne[0] * ne[1]; float *dst = y + pl * px * py;
        for (int64_t oy = 0; oy < py; oy++) for (int64_t ox = 0; ox < px; ox++) {
            float out = op == 1 ? 0.0f : -3.402823466e+38f;
            const int64_t ix = -p0 + ox * s0, iy = -p1 + oy * s1;
            for (int ky = 0; ky < k1; ky++) {
                if (iy + ky < 0 || iy + ky >= ne[1]) continue;
                for (int kx = 0; kx < k0; kx++) {
                    const int64_t j = ix + kx;
                    if (j < 0 || j >= ne[0]) continue;
                    const float v = src[(iy + ky) * ne[0] + j];
                    if (op == 1) out += v; else if (v > out) out = v;
                }
            }
            if (op == 1) out /= (float) (k0 * k1);
            dst[oy * px + ox] = out;
        }
    }
}

/* oracle/ggml_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, single-threaded restatement of the reference's arithmetic for the quantised
 * MUL_MAT / MUL_MAT_ID hot path. It is the CHECKER for the CUDA kernels:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it; nothing on
 * the product path may import, link or execute it.
 *
 * Parity status: PINNED. tests/test_oracle_vs_ref.py compares every function below with the
 * unmodified reference compiled into oracle/_ref (oracle/Makefile, oracle/ref_harness.cpp) and
 * tests/test_oracle_golden.py compares it with vectors the reference produced
 * (tests/golden/make_golden.py -> tests/golden/*.npz).
 *
 * Each function names the reference code (file:line under /root/reference) it restates.
 * Written from the format definitions; no reference source is copied.
 */
#include "ggml_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------ */
/* IEEE binary16 <-> binary32, round-to-nearest-even (ggml/src/ggml-impl.h:374-430;            */
/* on x86 the CPU backend uses F16C, which is the same IEEE conversion)                        */
/* ------------------------------------------------------------------------------------------ */
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float    u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1fu;
    uint32_t man = h & 0x3ffu;
    if (exp == 0x1f) return u2f(sign | 0x7f800000u | (man << 13));          /* inf / nan       */
    if (exp != 0)    return u2f(sign | ((exp + 112u) << 23) | (man << 13)); /* normal          */
    if (man == 0)    return u2f(sign);                                      /* +-0             */
    int e = -1;                                                             /* subnormal half  */
    do { man <<= 1; ++e; } while (!(man & 0x400u));
    return u2f(sign | ((uint32_t)(112 - e) << 23) | ((man & 0x3ffu) << 13));
}

uint16_t orc_fp32_to_fp16(float f) {
    const uint32_t u = f2u(f);
    const uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
    const uint32_t absu = u & 0x7fffffffu;
    if (absu > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);              /* nan             */
    if (absu >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);             /* rounds to inf   */
    if (absu < 0x33000001u) return sign;                                    /* rounds to zero  */
    int32_t exp = (int32_t)(absu >> 23) - 127 + 15;
    uint32_t man = (absu & 0x7fffffu) | 0x800000u;                          /* 24-bit mantissa */
    uint32_t shift = 13;
    if (exp <= 0) { shift += (uint32_t)(1 - exp); exp = 0; }                /* half subnormal  */
    const uint32_t halfbit = 1u << (shift - 1);
    const uint32_t rest = man & ((1u << shift) - 1u);
    uint32_t q = man >> shift;
    if (rest > halfbit || (rest == halfbit && (q & 1u))) ++q;               /* RNE             */
    /* q still holds the implicit bit for normals: adding it into the exponent field performs */
    /* the carry when the mantissa rounds up to the next binade.                               */
    const uint32_t out = (exp > 0) ? (((uint32_t)(exp - 1) << 10) + q) : q;
    return (uint16_t)(sign | out);
}

/* bf16, round-to-nearest-even (ggml-impl.h ggml_compute_fp32_to_bf16) -- fork's q8_2 scale  */
static uint16_t fp32_to_bf16_bits(float f) {
    uint32_t u = f2u(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 64u);  /* quiet the nan */
    return (uint16_t)((u + (0x7fffu + ((u >> 16) & 1u))) >> 16);
}
static float bf16_bits_to_fp32(uint16_t b) { return u2f((uint32_t) b << 16); }

/* ------------------------------------------------------------------------------------------ */
/* block formats (ggml/src/ggml-common.h:170-175, 219-237, 295-306, 330-344, 425-431)          */
/* ------------------------------------------------------------------------------------------ */
#define QK   32
#define QKK  256

#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } blk_q4_0;                              /*  18 B */
typedef struct { uint16_t d; int8_t qs[32]; } blk_q8_0;                               /*  34 B */
typedef struct { uint16_t d; uint16_t s; int8_t qs[32]; } blk_q8_1;                   /*  36 B */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qs[128]; } blk_q4_K; /* 144 B */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } blk_q5_K; /* 176 B, ggml-common.h:319-333 */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } blk_q6_K; /* 210 B */
typedef struct { float d; int8_t qs[256]; int16_t bsums[16]; } blk_q8_K;              /* 292 B */
typedef struct { uint16_t d; uint16_t scales_h; uint8_t scales_l[4]; uint8_t qs[128]; } blk_iq4_xs; /* 136 B */
#pragma pack(pop)

/* ggml-common.h:1088 */
static const int8_t k_iq4nl[16] = { -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 };

int orc_block_elems(int type) {
    switch (type) {
        case ORC_Q4_0: case ORC_Q8_0: case ORC_Q8_1: return QK;
        case ORC_Q4_K: case ORC_Q5_K: case ORC_Q6_K: case ORC_Q8_K: case ORC_IQ4_XS: return QKK;
        case ORC_F32: return 1;
    }
    return 0;
}
size_t orc_block_bytes(int type) {
    switch (type) {
        case ORC_Q4_0: return sizeof(blk_q4_0);   case ORC_Q8_0: return sizeof(blk_q8_0);
        case ORC_Q8_1: return sizeof(blk_q8_1);   case ORC_Q4_K: return sizeof(blk_q4_K);
        case ORC_Q6_K: return sizeof(blk_q6_K);   case ORC_Q8_K: return sizeof(blk_q8_K);
        case ORC_IQ4_XS: return sizeof(blk_iq4_xs); case ORC_F32: return 4;
        case ORC_Q5_K: return sizeof(blk_q5_K);
    }
    return 0;
}
size_t orc_row_size(int type, int64_t k) { return (size_t)(k / orc_block_elems(type)) * orc_block_bytes(type); }

/* type_traits_cpu[].vec_dot_type (ggml/src/ggml-cpu/ggml-cpu.c:217-370) */
int orc_vec_dot_type(int type) {
    switch (type) {
        case ORC_Q4_0: case ORC_Q8_0: return ORC_Q8_0;
        case ORC_Q4_K: case ORC_Q5_K: case ORC_Q6_K: case ORC_IQ4_XS: return ORC_Q8_K;
    }
    return -1;
}

/* 6-bit (scale, min) pair j of a Q4_K super-block (ggml-quants.c:703-711 get_scale_min_k4) */
static void q4k_scale_min(const uint8_t * p, int j, int * sc, int * mn) {
    if (j < 4) {
        *sc = p[j] & 63;
        *mn = p[j + 4] & 63;
    } else {
        *sc = (p[j + 4] & 0x0f) | ((p[j - 4] >> 6) << 4);
        *mn = (p[j + 4] >> 4)   | ((p[j]     >> 6) << 4);
    }
}

/* 6-bit scale of IQ4_XS sub-block ib, biased by 32 (ggml-quants.c:2541) */
static int iq4xs_ls(const blk_iq4_xs * b, int ib) {
    const int lo = (b->scales_l[ib >> 1] >> (4 * (ib & 1))) & 0x0f;
    const int hi = (b->scales_h >> (2 * ib)) & 3;
    return (lo | (hi << 4)) - 32;
}

/* signed 6-bit value e (0..255) of a Q6_K super-block (ggml-quants.c:1774-1783) */
/* 5-bit value l of sub-block sb (32 elements) of a Q5_K super-block: sub-blocks 2j / 2j+1 are the low / high nibbles of
   qs[32j .. 32j+31]; bit sb of qh[l] is the fifth bit (ggml-quants.c dequantize_row_q5_K, masks u1/u2 shifted by 2 per pair) */
static int q5k_value(const blk_q5_K * b, int sb, int l) {
    const uint8_t q = b->qs[32 * (sb >> 1) + l];
    return ((sb & 1) ? (q >> 4) : (q & 0x0f)) + (((b->qh[l] >> sb) & 1) << 4);
}

static int q6k_value(const blk_q6_K * b, int e) {
    const int half = e >> 7, r = e & 127, l = r & 31, quarter = r >> 5;
    const uint8_t * ql = b->ql + 64 * half;
    const uint8_t * qh = b->qh + 32 * half;
    const int lo = (quarter & 1) ? ql[l + 32] : ql[l];
    const int nib = (quarter & 2) ? (lo >> 4) : (lo & 0x0f);
    const int top = (qh[l] >> (2 * quarter)) & 3;
    return (nib | (top << 4)) - 32;
}

/* ------------------------------------------------------------------------------------------ */
/* dequantisers (ggml-quants.c:307-325 q4_0, 401-415 q8_0, 1352-1375 q4_K, 1762-1792 q6_K,      */
/* 2530-2551 iq4_xs). Float expression order follows the reference so results are bit-equal.   */
/* ------------------------------------------------------------------------------------------ */
void orc_dequantize_row(int type, const void * src, float * dst, int64_t k) {
    const int64_t nb = k / orc_block_elems(type);
    for (int64_t i = 0; i < nb; ++i) {
        switch (type) {
        case ORC_Q4_0: {
            const blk_q4_0 * b = (const blk_q4_0 *) src + i;
            const float d = orc_fp16_to_fp32(b->d);
            for (int j = 0; j < 16; ++j) {
                dst[i * QK + j]      = (float)((b->qs[j] & 0x0f) - 8) * d;
                dst[i * QK + j + 16] = (float)((b->qs[j] >> 4) - 8) * d;
            }
        } break;
        case ORC_Q8_0: {
            const blk_q8_0 * b = (const blk_q8_0 *) src + i;
            const float d = orc_fp16_to_fp32(b->d);
            for (int j = 0; j < QK; ++j) dst[i * QK + j] = (float) b->qs[j] * d;
        } break;
        case ORC_Q4_K: {
            const blk_q4_K * b = (const blk_q4_K *) src + i;
            const float d = orc_fp16_to_fp32(b->d), dmin = orc_fp16_to_fp32(b->dmin);
            for (int sb = 0; sb < 8; ++sb) {
                int sc, mn; q4k_scale_min(b->scales, sb, &sc, &mn);
                const float ds = d * (float) sc, ms = dmin * (float) mn;
                const uint8_t * q = b->qs + 32 * (sb >> 1);
                for (int l = 0; l < 32; ++l) {
                    const int nib = (sb & 1) ? (q[l] >> 4) : (q[l] & 0x0f);
                    dst[i * QKK + sb * 32 + l] = ds * (float) nib - ms;
                }
            }
        } break;
        case ORC_Q5_K: {   /* ggml-quants.c dequantize_row_q5_K: 4-bit nibble + bit sb of qh[l] as bit 4 */
            const blk_q5_K * b = (const blk_q5_K *) src + i;
            const float d = orc_fp16_to_fp32(b->d), dmin = orc_fp16_to_fp32(b->dmin);
            for (int sb = 0; sb < 8; ++sb) {
                int sc, mn; q4k_scale_min(b->scales, sb, &sc, &mn);
                const float ds = d * (float) sc, ms = dmin * (float) mn;
                for (int l = 0; l < 32; ++l)
                    dst[i * QKK + sb * 32 + l] = ds * (float) q5k_value(b, sb, l) - ms;
            }
        } break;
        case ORC_Q6_K: {
            const blk_q6_K * b = (const blk_q6_K *) src + i;
            const float d = orc_fp16_to_fp32(b->d);
            for (int e = 0; e < QKK; ++e)
                dst[i * QKK + e] = d * (float) b->scales[e >> 4] * (float) q6k_value(b, e);
        } break;
        case ORC_IQ4_XS: {
            const blk_iq4_xs * b = (const blk_iq4_xs *) src + i;
            const float d = orc_fp16_to_fp32(b->d);
            for (int ib = 0; ib < 8; ++ib) {
                const float dl = d * (float) iq4xs_ls(b, ib);
                for (int j = 0; j < 16; ++j) {
                    dst[i * QKK + ib * 32 + j]      = dl * (float) k_iq4nl[b->qs[16 * ib + j] & 0x0f];
                    dst[i * QKK + ib * 32 + j + 16] = dl * (float) k_iq4nl[b->qs[16 * ib + j] >> 4];
                }
            }
        } break;
        case ORC_Q8_K: {
            const blk_q8_K * b = (const blk_q8_K *) src + i;
            for (int j = 0; j < QKK; ++j) dst[i * QKK + j] = b->d * (float) b->qs[j];
        } break;
        default: break;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* activation quantisers                                                                       */
/* ------------------------------------------------------------------------------------------ */
static float block_amax(const float * x, int n) {
    float a = 0.0f;
    for (int j = 0; j < n; ++j) { const float v = fabsf(x[j]); if (v > a) a = v; }
    return a;
}

/* Q8_0 per 32: two definitions exist in the reference and they are NOT bit-identical:
 *   ORC_Q8_REF : quantize_row_q8_0_ref (ggml-quants.c:199-222)   d = amax/127, id = 1/d, roundf (half away)
 *   ORC_Q8_X86 : quantize_row_q8_0 AVX2 (arch/x86/quants.c:290-360) d = amax/127, id = 127/amax, RNE
 * test-backend-ops' CPU side runs the second one on x86. */
static int q8_round(float v, int flavour) {
    return flavour == ORC_Q8_REF ? (int) roundf(v) : (int) nearbyintf(v); /* default FP env: RNE */
}
static float q8_inverse(float amax, float d, int flavour) {
    if (flavour == ORC_Q8_REF) return d != 0.0f ? 1.0f / d : 0.0f;
    return amax != 0.0f ? 127.0f / amax : 0.0f;
}

void orc_quantize_row_q8_0(const float * x, void * vy, int64_t k, int flavour) {
    blk_q8_0 * y = (blk_q8_0 *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        const float amax = block_amax(x + i * QK, QK);
        const float d = amax / 127.0f;
        const float id = q8_inverse(amax, d, flavour);
        y[i].d = orc_fp32_to_fp16(d);
        for (int j = 0; j < QK; ++j) y[i].qs[j] = (int8_t) q8_round(x[i * QK + j] * id, flavour);
    }
}

/* Q8_1 = Q8_0 + s = d * sum(q) (ggml-quants.c:225-258 ref; arch/x86/quants.c:388-480 AVX2) */
void orc_quantize_row_q8_1(const float * x, void * vy, int64_t k, int flavour) {
    blk_q8_1 * y = (blk_q8_1 *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        const float amax = block_amax(x + i * QK, QK);
        const float d = amax / 127.0f;
        const float id = q8_inverse(amax, d, flavour);
        int sum = 0;
        for (int j = 0; j < QK; ++j) {
            const int q = q8_round(x[i * QK + j] * id, flavour);
            y[i].qs[j] = (int8_t) q;
            sum += (int8_t) q;
        }
        y[i].d = orc_fp32_to_fp16(d);
        /* ref: sum*d (int*float), x86: d * hsum -- the same single fp32 product */
        y[i].s = orc_fp32_to_fp16((float) sum * d);
    }
}

/* Q8_K per 256 (ggml-quants.c:2555-2591; the x86 CPU backend calls this same function,
 * arch/x86/quants.c:493-495). `max` is the FIRST element of largest magnitude, sign kept;
 * rounding is the magic-number nearest_int (ggml-quants.c:444-449) == RNE for |v| < 2^22.
 * The reference leaves bsums untouched for an all-zero block; we define them as 0. */
void orc_quantize_row_q8_K(const float * x, void * vy, int64_t k) {
    blk_q8_K * y = (blk_q8_K *) vy;
    for (int64_t i = 0; i < k / QKK; ++i, x += QKK) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QKK; ++j) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (amax == 0.0f) { memset(&y[i], 0, sizeof(y[i])); continue; }
        const float iscale = -127.0f / max;
        for (int j = 0; j < QKK; ++j) {
            const float v = iscale * x[j] + 12582912.0f;       /* 1.5 * 2^23 */
            const int q = (int)(f2u(v) & 0x007fffffu) - 0x00400000;
            y[i].qs[j] = (int8_t)(q > 127 ? 127 : q);
        }
        for (int g = 0; g < 16; ++g) {
            int s = 0;
            for (int j = 0; j < 16; ++j) s += y[i].qs[16 * g + j];
            y[i].bsums[g] = (int16_t) s;
        }
        y[i].d = 1.0f / iscale;
    }
}

void orc_quantize_row(int q8_type, const float * x, void * y, int64_t k, int flavour) {
    if (q8_type == ORC_Q8_0) orc_quantize_row_q8_0(x, y, k, flavour);
    else if (q8_type == ORC_Q8_1) orc_quantize_row_q8_1(x, y, k, flavour);
    else if (q8_type == ORC_Q8_K) orc_quantize_row_q8_K(x, y, k);
}

/* fork activation format for the ZYK Q4_0 kernels, x86 branch (repack.cpp:2386-2459):
 * d = amax/127 rounded to bf16 FIRST, id = 1/d(bf16), q = RNE(x*id); s = int16 sum of the
 * UNSATURATED quants, qs saturated to int8 by the pack instructions; groups of 4 blocks are
 * stored as {d[4], s[4], qs[4][32]}, the tail (nb % 4) as plain {d, s, qs[32]} blocks. */
void orc_quantize_row_q8_2_x4(const float * x, void * vy, int64_t k) {
    const int64_t nb = k / QK, nb4 = 4 * (nb / 4);
    uint8_t * base = (uint8_t *) vy;
    for (int64_t i = 0; i < nb; ++i) {
        const float amax = block_amax(x + i * QK, QK);
        const uint16_t dbits = fp32_to_bf16_bits(amax / 127.0f);
        const float d = bf16_bits_to_fp32(dbits);
        const float id = d > 0.0f ? 1.0f / d : 0.0f;
        int sum = 0; int8_t q8[QK];
        for (int j = 0; j < QK; ++j) {
            const int q = (int) nearbyintf(x[i * QK + j] * id);
            sum += q;
            q8[j] = (int8_t)(q > 127 ? 127 : (q < -128 ? -128 : q));
        }
        const int16_t s = (int16_t) sum;
        if (i < nb4) {
            uint8_t * g = base + (i / 4) * 144; const int r = (int)(i % 4);
            memcpy(g + 2 * r, &dbits, 2); memcpy(g + 8 + 2 * r, &s, 2); memcpy(g + 16 + 32 * r, q8, 32);
        } else {
            uint8_t * g = base + i * 36;
            memcpy(g, &dbits, 2); memcpy(g + 2, &s, 2); memcpy(g + 4, q8, 32);
        }
    }
}

/* CPU GEMM activation format of the repack path: four rows quantised to Q8_K at once and interleaved in
 * runs of eight bytes, block_q8_Kx4 = { float d[4]; int8 qs[1024]; int16 bsums[64] } (ggml-cpu/repack.h:55-61).
 * Follows ggml_quantize_mat_q8_K_4x8_generic (ggml-cpu/repack.cpp:180-229): per row and 256-block the signed
 * element of largest magnitude (first one wins), iscale = -127 / max, q = nearest_int(x * iscale) WITHOUT the
 * clamp to 127 that quantize_row_q8_K applies; qs[j] takes row (j % 32) / 8, element 8 * (j / 32) + j % 8;
 * bsums index ((j & 31) >> 3) * 4 + (j >> 8) * 16 + ((j >> 6) & 3), i.e. four sums of 16 per row, rows interleaved. */
void orc_quantize_mat_q8_K_4x8(const float * x, void * vy, int64_t k) {
    const int64_t nb = k / 256;
    uint8_t * out = (uint8_t *) vy;
    for (int64_t i = 0; i < nb; ++i) {
        uint8_t * blk = out + i * 1168;
        float * d = (float *) blk; int8_t * qs = (int8_t *) (blk + 16); int16_t * bsums = (int16_t *) (blk + 16 + 1024);
        float iscale[4];
        for (int r = 0; r < 4; ++r) {
            float amax = 0.0f, mx = 0.0f;
            for (int j = 0; j < 256; ++j) {
                const float v = x[r * k + i * 256 + j];
                if (amax < fabsf(v)) { amax = fabsf(v); mx = v; }
            }
            iscale[r] = amax ? -127.f / mx : 0.0f;
            d[r] = amax ? 1.0f / iscale[r] : 0.0f;
        }
        for (int j = 0; j < 64; ++j) bsums[j] = 0;
        for (int j = 0; j < 1024; ++j) {
            const int src_off = (j / 32) * 8 + (j % 8), src_id = (j % 32) / 8;
            const int index = (((j & 31) >> 3) << 2) + ((j >> 8) << 4) + ((j >> 6) & 3);
            const float v = x[src_id * k + i * 256 + src_off] * iscale[src_id] + 12582912.0f;       /* nearest_int, ggml-quants.c:444-449 */
            qs[j] = (int8_t) ((int) (f2u(v) & 0x007fffffu) - 0x00400000);
            bsums[index] = (int16_t) (bsums[index] + qs[j]);
        }
    }
}

/* The fork's ZYK Q4_0 path end to end (SURVEY 8 a9 + a16): activations to block_q8_2_x4 (above), then per weight row,
 * block after block in order, acc = fma(d_w * d_x, (float) (sum_j u4_j * q8_j - 8 * s), acc) with the UNSIGNED nibbles 0..15
 * and the block's int16 quant sum s as the bias -- arithmetic of ZykQ4_0_T::compute_block + mul_mat_q4_t_q8_0<.., -8, ..>
 * (ggml-cpu/arch/x86/repack.cpp:186-243, 460-508): one lane of the 512-bit accumulators is one weight row, so the fp32 order
 * of a row is exactly this loop. d_w is the block's fp16 scale, d_x the bf16 activation scale, their product one fp32 multiply.
 * w: canonical block_q4_0 rows (the reference re-lays them at load time, repack_q4_0_transpose_bl; values are unchanged). */
void orc_mul_mat_q4_0_zyk(int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y) {
    const int64_t nb = k / QK;
    uint8_t * xq = (uint8_t *) malloc((size_t) nb * 36 + 64);
    for (int64_t c = 0; c < n; ++c) {
        orc_quantize_row_q8_2_x4(x + c * k, xq, k);
        for (int64_t r = 0; r < m; ++r) {
            const blk_q4_0 * wr = (const blk_q4_0 *) ((const uint8_t *) w + (size_t) r * nb * sizeof(blk_q4_0));
            float acc = 0.0f;
            for (int64_t i = 0; i < nb; ++i) {
                uint16_t dbits; int16_t sx; const int8_t * q;
                if (i < 4 * (nb / 4)) { const uint8_t * g = xq + (i / 4) * 144; const int rr = (int) (i % 4); memcpy(&dbits, g + 2 * rr, 2); memcpy(&sx, g + 8 + 2 * rr, 2); q = (const int8_t *) (g + 16 + 32 * rr); }
                else { const uint8_t * g = xq + i * 36; memcpy(&dbits, g, 2); memcpy(&sx, g + 2, 2); q = (const int8_t *) (g + 4); }
                int isum = -8 * (int) sx;
                for (int j = 0; j < 16; ++j) isum += (wr[i].qs[j] & 0x0f) * q[j] + (wr[i].qs[j] >> 4) * q[j + 16];
                const float scale = orc_fp16_to_fp32(wr[i].d) * bf16_bits_to_fp32(dbits);
                acc = fmaf(scale, (float) isum, acc);
            }
            y[c * m + r] = acc;
        }
    }
    free(xq);
}

/* ------------------------------------------------------------------------------------------ */
/* dot products: integer-exact inside a block, fp32 across blocks in the order of the          */
/* reference's generic kernels (ggml-cpu/quants.c:115-150 q4_0, 305-335 q8_0, 550-625 q4_K,     */
/* 705-760 q6_K, 1137-1181 iq4_xs). The SIMD kernels the CPU backend really runs differ only   */
/* in fp32 summation order.                                                                    */
/* ------------------------------------------------------------------------------------------ */
float orc_vec_dot(int type, int64_t k, const void * w_row, const void * xq_row) {
    const int64_t nb = k / orc_block_elems(type);
    float sumf = 0.0f;
    switch (type) {
    case ORC_Q4_0: {
        const blk_q4_0 * w = (const blk_q4_0 *) w_row; const blk_q8_0 * x = (const blk_q8_0 *) xq_row;
        for (int64_t i = 0; i < nb; ++i) {
            int acc = 0;
            for (int j = 0; j < 16; ++j)
                acc += ((w[i].qs[j] & 0x0f) - 8) * x[i].qs[j] + ((w[i].qs[j] >> 4) - 8) * x[i].qs[j + 16];
            sumf += (float) acc * orc_fp16_to_fp32(w[i].d) * orc_fp16_to_fp32(x[i].d);
        }
    } break;
    case ORC_Q8_0: {
        const blk_q8_0 * w = (const blk_q8_0 *) w_row; const blk_q8_0 * x = (const blk_q8_0 *) xq_row;
        for (int64_t i = 0; i < nb; ++i) {
            int acc = 0;
            for (int j = 0; j < QK; ++j) acc += w[i].qs[j] * x[i].qs[j];
            sumf += (float) acc * (orc_fp16_to_fp32(w[i].d) * orc_fp16_to_fp32(x[i].d));
        }
    } break;
    case ORC_Q4_K: {
        const blk_q4_K * w = (const blk_q4_K *) w_row; const blk_q8_K * x = (const blk_q8_K *) xq_row;
        float lanes[8] = {0};
        for (int64_t i = 0; i < nb; ++i) {
            int32_t lane_acc[8] = {0}; int min_acc = 0;
            for (int sb = 0; sb < 8; ++sb) {
                int sc, mn; q4k_scale_min(w[i].scales, sb, &sc, &mn);
                min_acc += mn * (x[i].bsums[2 * sb] + x[i].bsums[2 * sb + 1]);
                const uint8_t * q = w[i].qs + 32 * (sb >> 1);
                for (int l = 0; l < 32; ++l) {
                    const int nib = (sb & 1) ? (q[l] >> 4) : (q[l] & 0x0f);
                    lane_acc[l & 7] += sc * (nib * x[i].qs[32 * sb + l]);
                }
            }
            const float d = orc_fp16_to_fp32(w[i].d) * x[i].d;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) lane_acc[l];
            sumf -= (orc_fp16_to_fp32(w[i].dmin) * x[i].d) * (float) min_acc;
        }
        for (int l = 0; l < 8; ++l) sumf += lanes[l];
    } break;
    case ORC_Q5_K: {   /* ggml-cpu/quants.c ggml_vec_dot_q5_K_q8_K_generic: the Q4_K scheme over 5-bit values */
        const blk_q5_K * w = (const blk_q5_K *) w_row; const blk_q8_K * x = (const blk_q8_K *) xq_row;
        float lanes[8] = {0};
        for (int64_t i = 0; i < nb; ++i) {
            int32_t lane_acc[8] = {0}; int min_acc = 0;
            for (int sb = 0; sb < 8; ++sb) {
                int sc, mn; q4k_scale_min(w[i].scales, sb, &sc, &mn);
                min_acc += mn * (x[i].bsums[2 * sb] + x[i].bsums[2 * sb + 1]);
                for (int l = 0; l < 32; ++l)
                    lane_acc[l & 7] += sc * (q5k_value(&w[i], sb, l) * x[i].qs[32 * sb + l]);
            }
            const float d = orc_fp16_to_fp32(w[i].d) * x[i].d;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) lane_acc[l];
            sumf -= (orc_fp16_to_fp32(w[i].dmin) * x[i].d) * (float) min_acc;
        }
        for (int l = 0; l < 8; ++l) sumf += lanes[l];
    } break;
    case ORC_Q6_K: {
        const blk_q6_K * w = (const blk_q6_K *) w_row; const blk_q8_K * x = (const blk_q8_K *) xq_row;
        float lanes[8] = {0};
        for (int64_t i = 0; i < nb; ++i) {
            int32_t lane_acc[8] = {0};
            for (int e = 0; e < QKK; ++e)
                lane_acc[e & 7] += (int) w[i].scales[e >> 4] * (q6k_value(&w[i], e) * x[i].qs[e]);
            const float d = orc_fp16_to_fp32(w[i].d) * x[i].d;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) lane_acc[l];
        }
        for (int l = 0; l < 8; ++l) sumf += lanes[l];
    } break;
    case ORC_IQ4_XS: {
        const blk_iq4_xs * w = (const blk_iq4_xs *) w_row; const blk_q8_K * x = (const blk_q8_K *) xq_row;
        for (int64_t i = 0; i < nb; ++i) {
            const float d = orc_fp16_to_fp32(w[i].d) * x[i].d;
            for (int ib = 0; ib < 8; ++ib) {
                int acc = 0;
                for (int j = 0; j < 16; ++j) {
                    acc += x[i].qs[32 * ib + j]      * k_iq4nl[w[i].qs[16 * ib + j] & 0x0f];
                    acc += x[i].qs[32 * ib + j + 16] * k_iq4nl[w[i].qs[16 * ib + j] >> 4];
                }
                sumf += (d * (float) iq4xs_ls(&w[i], ib)) * (float) acc;
            }
        }
    } break;
    default: break;
    }
    return sumf;
}

/* ------------------------------------------------------------------------------------------ */
/* MUL_MAT driver (ggml-cpu.c:1227-1427): quantise every activation row to vec_dot_type, then   */
/* one vec_dot per output element. Y is [n][m] (m contiguous), like ggml's dst.                 */
/* ------------------------------------------------------------------------------------------ */
void orc_mul_mat(int type, int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y, int flavour) {
    const int qt = orc_vec_dot_type(type);
    const size_t wrow = orc_row_size(type, k), xrow = orc_row_size(qt, k);
    uint8_t * xq = (uint8_t *) malloc(xrow ? xrow : 1);
    for (int64_t c = 0; c < n; ++c) {
        orc_quantize_row(qt, x + c * k, xq, k, flavour);
        for (int64_t r = 0; r < m; ++r)
            y[c * m + r] = orc_vec_dot(type, k, (const uint8_t *) w + r * wrow, xq);
    }
    free(xq);
}

/* MUL_MAT_ID (ggml.c:3217-3243 shapes; ggml-cpu.c:1500-1676 driver): for token t and slot s the
 * expert is ids[t][s]; its input row is b[t][s % nb1] (nb1 == 1 broadcasts one row to all slots). */
void orc_mul_mat_id(int type, int64_t m, int64_t k, int64_t n_as, int64_t n_used, int64_t nb1, int64_t n_tokens,
                    const void * as, const float * b, const int32_t * ids, float * y, int flavour) {
    const int qt = orc_vec_dot_type(type);
    const size_t wrow = orc_row_size(type, k), xrow = orc_row_size(qt, k);
    uint8_t * xq = (uint8_t *) malloc(xrow ? xrow : 1);
    for (int64_t t = 0; t < n_tokens; ++t) {
        for (int64_t s = 0; s < n_used; ++s) {
            const int32_t e = ids[t * n_used + s];
            float * out = y + (t * n_used + s) * m;
            if (e < 0 || e >= n_as) { for (int64_t r = 0; r < m; ++r) out[r] = 0.0f; continue; }
            orc_quantize_row(qt, b + (t * nb1 + (s % nb1)) * k, xq, k, flavour);
            const uint8_t * we = (const uint8_t *) as + (size_t) e * m * wrow;
            for (int64_t r = 0; r < m; ++r) out[r] = orc_vec_dot(type, k, we + r * wrow, xq);
        }
    }
    free(xq);
}

void orc_mul_mat_f64(int type, int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y) {
    const size_t wrow = orc_row_size(type, k);
    float * wf = (float *) malloc(sizeof(float) * (size_t) k);
    for (int64_t r = 0; r < m; ++r) {
        orc_dequantize_row(type, (const uint8_t *) w + r * wrow, wf, k);
        for (int64_t c = 0; c < n; ++c) {
            double acc = 0.0;
            for (int64_t i = 0; i < k; ++i) acc += (double) wf[i] * (double) x[c * k + i];
            y[c * m + r] = (float) acc;
        }
    }
    free(wf);
}

/* ------------------------------------------------------------------------------------------ */
/* B200 plane layout (our own; DESIGN.md "HBM layout"). A matrix of m rows x nb blocks keeps    */
/* its total byte size; fields of equal kind are gathered into planes, each plane row-major     */
/* [m][nb][field bytes], planes concatenated in the order below. With k % 256 == 0 (which        */
/* orc_b200_repack_supported() requires) every plane offset and every plane row is a multiple   */
/* of 16 B, so the kernels can use 128-bit loads throughout.                                     */
/*   Q4_0  : QS 16 | D 2                                                                        */
/*   Q8_0  : QS 32 | D 2                                                                        */
/*   Q4_K  : identity (144 B blocks are already 16 B aligned; header 16 B + 128 B nibbles)       */
/*   Q5_K  : identity (176 B = header 16 B + 32 B of fifth bits + 128 B nibbles)                 */
/*   Q6_K  : QL 128 | QH 64 | SC 16 | D 2                                                        */
/*   IQ4_XS: QS 128 | HDR 8 (d, scales_h, scales_l[4])                                           */
/* ------------------------------------------------------------------------------------------ */
typedef struct { int n; int off[4]; int len[4]; } plane_spec;

static int b200_planes(int type, plane_spec * ps) {
    switch (type) {
    case ORC_Q4_0:   *ps = (plane_spec){2, {2, 0}, {16, 2}}; return 1;
    case ORC_Q8_0:   *ps = (plane_spec){2, {2, 0}, {32, 2}}; return 1;
    case ORC_Q4_K:   *ps = (plane_spec){1, {0}, {144}}; return 1;
    case ORC_Q5_K:   *ps = (plane_spec){1, {0}, {176}}; return 1;
    case ORC_Q6_K:   *ps = (plane_spec){4, {0, 128, 192, 208}, {128, 64, 16, 2}}; return 1;
    case ORC_IQ4_XS: *ps = (plane_spec){2, {8, 0}, {128, 8}}; return 1;
    }
    return 0;
}

int orc_b200_repack_supported(int type, int64_t m, int64_t k) {
    plane_spec ps;
    if (!b200_planes(type, &ps)) return 0;
    return m > 0 && k > 0 && (k % 256) == 0;
}

static void b200_shuffle(int type, int64_t m, int64_t k, const uint8_t * src, uint8_t * dst, int inverse) {
    plane_spec ps; b200_planes(type, &ps);
    const int64_t nblk = m * (k / orc_block_elems(type));
    const size_t bsz = orc_block_bytes(type);
    size_t plane_base = 0;
    for (int p = 0; p < ps.n; ++p) {
        for (int64_t b = 0; b < nblk; ++b) {
            const size_t canon = (size_t) b * bsz + (size_t) ps.off[p];
            const size_t plane = plane_base + (size_t) b * (size_t) ps.len[p];
            if (!inverse) memcpy(dst + plane, src + canon, (size_t) ps.len[p]);
            else          memcpy(dst + canon, src + plane, (size_t) ps.len[p]);
        }
        plane_base += (size_t) nblk * (size_t) ps.len[p];
    }
}
void orc_b200_repack(int type, int64_t m, int64_t k, const void * canonical, void * planes) {
    b200_shuffle(type, m, k, (const uint8_t *) canonical, (uint8_t *) planes, 0);
}
void orc_b200_unrepack(int type, int64_t m, int64_t k, const void * planes, void * canonical) {
    b200_shuffle(type, m, k, (const uint8_t *) planes, (uint8_t *) canonical, 1);
}

/* ------------------------------------------------------------------------------------------ */
/* The reference CPU's own load-time repacks (rows a14 of SURVEY.md section 8a)                 */
/* ------------------------------------------------------------------------------------------ */

/* block_q4_Kx8 (repack.h:39-46): for each group of 8 rows and each super-block column:
 * d[8], dmin[8], 96 B of re-packed 6-bit scales/mins, 1024 B of nibbles taken 8 bytes at a time
 * round-robin over the 8 rows (repack.cpp:1501-1575 make_block_q4_Kx8 with interleave 8,
 * driver :1652-1681). The 96 scale bytes are 8 groups of 12: group g holds sub-block g of all
 * 8 rows, packed with the same 6-bit scheme as a Q4_K block but across rows instead of across
 * sub-blocks. */
int orc_cpu_repack_q4_K_8x8(int64_t m, int64_t k, const void * canonical, void * out) {
    if (m % 8 || k % QKK) return -1;
    const int64_t nb = k / QKK;
    const blk_q4_K * src = (const blk_q4_K *) canonical;
    uint8_t * dst = (uint8_t *) out;
    for (int64_t r0 = 0; r0 < m; r0 += 8) {
        for (int64_t c = 0; c < nb; ++c, dst += 1152) {
            const blk_q4_K * in[8];
            for (int r = 0; r < 8; ++r) in[r] = src + (r0 + r) * nb + c;
            for (int r = 0; r < 8; ++r) { memcpy(dst + 2 * r, &in[r]->d, 2); memcpy(dst + 16 + 2 * r, &in[r]->dmin, 2); }
            uint8_t * sc_out = dst + 32;
            for (int g = 0; g < 8; ++g) {
                int s[8], mn[8];
                for (int r = 0; r < 8; ++r) q4k_scale_min(in[r]->scales, g, &s[r], &mn[r]);
                uint8_t * o = sc_out + 12 * g;
                for (int r = 0; r < 4; ++r) {
                    o[r]     = (uint8_t)((s[r]  & 63) | ((s[r + 4]  & 0x30) << 2));
                    o[r + 4] = (uint8_t)((mn[r] & 63) | ((mn[r + 4] & 0x30) << 2));
                    o[r + 8] = (uint8_t)((s[r + 4] & 0x0f) | ((mn[r + 4] & 0x0f) << 4));
                }
            }
            uint8_t * q_out = dst + 128;
            for (int i = 0; i < 128; ++i)  /* 128 chunks of 8 bytes */
                memcpy(q_out + 8 * i, in[i % 8]->qs + 8 * (i / 8), 8);
        }
    }
    return 0;
}

/* fork ZYK layout for Q4_0 on x86 (repack.cpp:1746-1796 with interleave_block = 4 and
 * nrows_interleaved = N32B = 16 under AVX512-VNNI, repack.h:121-125): the tensor is rewritten as
 * a plane of all fp16 scales followed by a plane of nibbles. Rows are walked in chunks of
 * `stride` = m/4 (a further /4 when that exceeds 2048); inside a chunk, for each block column,
 * rows go in groups of 16: their 16 scales, then their nibble bytes 4 at a time round-robin. */
int orc_cpu_repack_q4_0_zyk(int64_t m, int64_t k, const void * canonical, void * out) {
    const int NR = 16, IL = 4;
    if (k % QK || m % 4) return -1;
    int64_t stride = m / 4; if (stride > 2048) stride /= 4;
    if (stride == 0 || m % stride || stride % NR) return -1;
    const int64_t nb = k / QK;
    const blk_q4_0 * src = (const blk_q4_0 *) canonical;
    uint8_t * dplane = (uint8_t *) out;
    uint8_t * qplane = dplane + 2 * nb * m;
    for (int64_t chunk = 0; chunk < m; chunk += stride) {
        for (int64_t c = 0; c < nb; ++c) {
            for (int64_t g = 0; g < stride; g += NR) {
                const blk_q4_0 * rows[16];
                for (int r = 0; r < NR; ++r) {
                    rows[r] = src + (chunk + g + r) * nb + c;
                    memcpy(dplane, &rows[r]->d, 2); dplane += 2;
                }
                for (int i = 0; i < 16 * NR / IL; ++i) {
                    memcpy(qplane, rows[i % NR]->qs + IL * (i / NR), IL); qplane += IL;
                }
            }
        }
    }
    return 0;
}

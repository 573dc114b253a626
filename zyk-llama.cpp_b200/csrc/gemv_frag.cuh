// gemv_frag.cuh -- per-format register fragments of the decode GEMV.
//
// One warp step covers 1024 consecutive k-elements of a weight row: lane l owns 32 of them as two
// runs of 16 ("run0", "run1"), loaded with 128-bit streaming loads from the plane layout
// (repack.cu). The integer arithmetic inside a block is the reference's, value for value
// (ggml-cpu/quants.c:115-150 q4_0, 305-335 q8_0, 550-625 q4_K, 705-760 q6_K, 1137-1181 iq4_xs):
// int8 x int8 products accumulated in int32 (dp4a), integer sub-block scales applied in int32,
// one fp32 multiply by d_w * d_x per (block, lane). Only the fp32 summation order across lanes and
// blocks differs from the CPU.
#pragma once
#include "common.cuh"

// the weight matrix as the kernels see it
struct GemvW {
    const uint8_t * p0;   // Q4_0/Q8_0/IQ4_XS: QS plane | Q6_K: QL plane | Q4_K, Q5_K: blocks
    const uint8_t * p1;   // Q4_0/Q8_0: D plane | Q6_K: QH plane | IQ4_XS: HDR plane
    const uint8_t * p2;   // Q6_K: SC plane
    const uint8_t * p3;   // Q6_K: D plane
    int64_t nb;           // blocks per row
};

__host__ __device__ inline GemvW gemv_w_make(int type, const void * w, int64_t m, int64_t k) {
    GemvW W;
    const uint8_t * b = (const uint8_t *) w;
    const int64_t be = (type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0) ? 32 : 256;
    W.nb = k / be;
    const int64_t nblk = m * W.nb;
    W.p0 = b; W.p1 = b; W.p2 = b; W.p3 = b;
    switch (type) {
        case B200_TYPE_Q4_0:   W.p1 = b + nblk * 16; break;
        case B200_TYPE_Q8_0:   W.p1 = b + nblk * 32; break;
        case B200_TYPE_Q6_K:   W.p1 = b + nblk * 128; W.p2 = b + nblk * 192; W.p3 = b + nblk * 208; break;
        case B200_TYPE_IQ4_XS: W.p1 = b + nblk * 128; break;
        default: break;
    }
    return W;
}

#ifdef __CUDACC__

// quantised activations of one column for one lane-step
struct XFrag {
    uint4 x0, x1;   // 16 int8 each: run0 / run1
    int bs0, bs1;   // sums of those 16 quants
    float dx;       // activation block scale
};

template <int TYPE> struct Fmt;

// element offsets (inside the lane's 256- or 32-element block) of run0 / run1
template <> struct Fmt<B200_TYPE_Q4_K>   { static constexpr int BLK = 256; __device__ static int off0(int g) { return 64 * (g >> 1) + 16 * (g & 1); } static constexpr int RUN1 = 32; };
template <> struct Fmt<B200_TYPE_Q5_K>   { static constexpr int BLK = 256; __device__ static int off0(int g) { return 64 * (g >> 1) + 16 * (g & 1); } static constexpr int RUN1 = 32; };
template <> struct Fmt<B200_TYPE_Q6_K>   { static constexpr int BLK = 256; __device__ static int off0(int g) { return 128 * (g >> 2) + 32 * ((g >> 1) & 1) + 16 * (g & 1); } static constexpr int RUN1 = 64; };
template <> struct Fmt<B200_TYPE_IQ4_XS> { static constexpr int BLK = 256; __device__ static int off0(int g) { return 32 * g; } static constexpr int RUN1 = 16; };
template <> struct Fmt<B200_TYPE_Q4_0>   { static constexpr int BLK = 32;  __device__ static int off0(int) { return 0; } static constexpr int RUN1 = 16; };
template <> struct Fmt<B200_TYPE_Q8_0>   { static constexpr int BLK = 32;  __device__ static int off0(int) { return 0; } static constexpr int RUN1 = 16; };

// blk = index of the lane's block inside the row (256-element blocks for K-quants, 32 otherwise),
// g = lane & 7 (which eighth of a 256-block)
template <int TYPE>
__device__ __forceinline__ void xfrag_load(XFrag & f, const int8_t * __restrict__ xqs, const float * __restrict__ xd,
                                           const int16_t * __restrict__ xbs, int64_t blk, int g, int act_block) {
    const int64_t e0 = blk * Fmt<TYPE>::BLK + Fmt<TYPE>::off0(g);
    const int64_t e1 = e0 + Fmt<TYPE>::RUN1;
    f.x0 = ld_cached16(xqs + e0);
    f.x1 = ld_cached16(xqs + e1);
    f.bs0 = __ldg(xbs + (e0 >> 4));
    f.bs1 = __ldg(xbs + (e1 >> 4));
    f.dx = __ldg(xd + (act_block == 256 ? (e0 >> 8) : (e0 >> 5)));
}

__device__ __forceinline__ int dot16(const uint4 & a, const uint4 & b) {
    int s = __dp4a((int) a.x, (int) b.x, 0);
    s = __dp4a((int) a.y, (int) b.y, s);
    s = __dp4a((int) a.z, (int) b.z, s);
    s = __dp4a((int) a.w, (int) b.w, s);
    return s;
}
// unsigned bytes of a  x  signed bytes of b, accumulated into c
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint4 and4(const uint4 & a, uint32_t m) { return make_uint4(a.x & m, a.y & m, a.z & m, a.w & m); }
__device__ __forceinline__ uint4 shr4(const uint4 & a, int s) { return make_uint4(a.x >> s, a.y >> s, a.z >> s, a.w >> s); }

template <int TYPE> struct WFrag;

// ---- Q4_K: canonical 144 B blocks {d, dmin, scales[12], qs[128]} (ggml-common.h:295-306) -------
template <> struct WFrag<B200_TYPE_Q4_K> {
    uint4 hdr, q;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int g) {
        const uint8_t * b = W.p0 + (row * W.nb + blk) * 144;
        hdr = ld_stream16(b);
        q = ld_stream16(b + 16 + 16 * g);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int g) const {
        // 6-bit scale/min pairs of sub-blocks j0 = g & 6 (low nibbles) and j0 + 1 (high nibbles): get_scale_min_k4
        // (ggml-quants.c:703-711) for two adjacent j at once, as 2 x 8-bit fields of a 16-bit value
        const int sh = (g & 2) << 3;                                   // j0 & 3 is 0 or 2 -> bit offset 0 or 16
        const uint32_t y16 = (hdr.y >> sh) & 0xffffu, z16 = (hdr.z >> sh) & 0xffffu, w16 = (hdr.w >> sh) & 0xffffu;
        const bool upper = g >= 4;                                     // j0 >= 4
        const uint32_t sc = upper ? ((w16 & 0x0f0fu) | ((y16 >> 2) & 0x3030u)) : (y16 & 0x3f3fu);
        const uint32_t mn = upper ? (((w16 >> 4) & 0x0f0fu) | ((z16 >> 2) & 0x3030u)) : (z16 & 0x3f3fu);
        const int sc0 = (int) (sc & 0xffu), sc1 = (int) (sc >> 8), m0 = (int) (mn & 0xffu), m1 = (int) (mn >> 8);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
        // high nibbles stay in place (value 16 * nibble as an unsigned byte): one AND instead of shift + AND, the
        // factor 16 leaves the int32 sum with an exact arithmetic shift
        const uint4 lo = and4(q, 0x0f0f0f0fu), hi = and4(q, 0xf0f0f0f0u);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            int dhi = dp4a_us(hi.x, x[c].x1.x, 0);
            dhi = dp4a_us(hi.y, x[c].x1.y, dhi);
            dhi = dp4a_us(hi.z, x[c].x1.z, dhi);
            dhi = dp4a_us(hi.w, x[c].x1.w, dhi);
            const int isum = sc0 * dot16(lo, x[c].x0) + sc1 * (dhi >> 4);
            const int msum = m0 * x[c].bs0 + m1 * x[c].bs1;
            acc[c] += x[c].dx * (dm.x * (float) isum - dm.y * (float) msum);
        }
    }
};

// ---- Q5_K: canonical 176 B blocks {d, dmin, scales[12], qh[32], qs[128]} (ggml-common.h:319-333) ----
// Same sub-block geometry as Q4_K; bit j of qh[l] is the fifth bit of element l of sub-block j
// (ggml-cpu/quants.c ggml_vec_dot_q5_K_q8_K_generic builds the same 0..31 values before the int8 products).
template <> struct WFrag<B200_TYPE_Q5_K> {
    uint4 hdr, h, q;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int g) {
        const uint8_t * b = W.p0 + (row * W.nb + blk) * 176;
        hdr = ld_stream16(b);
        h = ld_stream16(b + 16 + 16 * (g & 1));
        q = ld_stream16(b + 48 + 16 * g);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int g) const {
        const int sh = (g & 2) << 3;
        const uint32_t y16 = (hdr.y >> sh) & 0xffffu, z16 = (hdr.z >> sh) & 0xffffu, w16 = (hdr.w >> sh) & 0xffffu;
        const bool upper = g >= 4;
        const uint32_t sc = upper ? ((w16 & 0x0f0fu) | ((y16 >> 2) & 0x3030u)) : (y16 & 0x3f3fu);
        const uint32_t mn = upper ? (((w16 >> 4) & 0x0f0fu) | ((z16 >> 2) & 0x3030u)) : (z16 & 0x3f3fu);
        const int sc0 = (int) (sc & 0xffu), sc1 = (int) (sc >> 8), m0 = (int) (mn & 0xffu), m1 = (int) (mn >> 8);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
        const int j0 = g & 6;                                          // sub-block of the low nibbles; the high nibbles are j0 + 1
        const uint4 b0 = and4(shr4(h, j0), 0x01010101u), b1 = and4(shr4(h, j0 + 1), 0x01010101u);
        const uint4 lo = make_uint4((q.x & 0x0f0f0f0fu) | (b0.x << 4), (q.y & 0x0f0f0f0fu) | (b0.y << 4), (q.z & 0x0f0f0f0fu) | (b0.z << 4), (q.w & 0x0f0f0f0fu) | (b0.w << 4));
        const uint4 hi = make_uint4(((q.x >> 4) & 0x0f0f0f0fu) | (b1.x << 4), ((q.y >> 4) & 0x0f0f0f0fu) | (b1.y << 4), ((q.z >> 4) & 0x0f0f0f0fu) | (b1.z << 4), ((q.w >> 4) & 0x0f0f0f0fu) | (b1.w << 4));
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const int isum = sc0 * dot16(lo, x[c].x0) + sc1 * dot16(hi, x[c].x1);
            const int msum = m0 * x[c].bs0 + m1 * x[c].bs1;
            acc[c] += x[c].dx * (dm.x * (float) isum - dm.y * (float) msum);
        }
    }
};

// ---- Q6_K planes: QL 128 | QH 64 | SC 16 | D 2 (ggml-common.h:330-336) --------------------------
template <> struct WFrag<B200_TYPE_Q6_K> {
    uint4 ql, qh, sc; uint16_t d;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int g) {
        const int64_t bi = row * W.nb + blk;
        ql = ld_stream16(W.p0 + bi * 128 + 16 * g);
        qh = ld_stream16(W.p1 + bi * 64 + 32 * (g >> 2) + 16 * (g & 1));
        sc = ld_stream16(W.p2 + bi * 16);
        d = ld_stream2(W.p3 + bi * 2);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int g) const {
        const int sh = g & 2;                       // 2 * quarter of the low-nibble run
        const uint4 v0 = make_uint4((ql.x & 0x0f0f0f0fu) | (((qh.x >> sh) & 0x03030303u) << 4), (ql.y & 0x0f0f0f0fu) | (((qh.y >> sh) & 0x03030303u) << 4),
                                    (ql.z & 0x0f0f0f0fu) | (((qh.z >> sh) & 0x03030303u) << 4), (ql.w & 0x0f0f0f0fu) | (((qh.w >> sh) & 0x03030303u) << 4));
        const int sh1 = sh + 4;
        const uint4 v1 = make_uint4(((ql.x >> 4) & 0x0f0f0f0fu) | (((qh.x >> sh1) & 0x03030303u) << 4), ((ql.y >> 4) & 0x0f0f0f0fu) | (((qh.y >> sh1) & 0x03030303u) << 4),
                                    ((ql.z >> 4) & 0x0f0f0f0fu) | (((qh.z >> sh1) & 0x03030303u) << 4), ((ql.w >> 4) & 0x0f0f0f0fu) | (((qh.w >> sh1) & 0x03030303u) << 4));
        const int is = 8 * (g >> 2) + (g & 3);      // = (element offset of run0) / 16
        const int sc0 = (int) (int8_t) byte_of(sc, is), sc1 = (int) (int8_t) byte_of(sc, is + 4);
        const float dw = half_bits_to_float(d);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const int s0 = dot16(v0, x[c].x0) - 32 * x[c].bs0;
            const int s1 = dot16(v1, x[c].x1) - 32 * x[c].bs1;
            acc[c] += (x[c].dx * dw) * (float) (sc0 * s0 + sc1 * s1);
        }
    }
};

// ---- IQ4_XS planes: QS 128 | HDR 8 {d, scales_h, scales_l[4]} (ggml-common.h:425-431) -----------
// kvalues_iq4nl (ggml-common.h:1088) as four byte-permute registers
__device__ __forceinline__ uint32_t iq4nl_lut4(uint32_t nib) {     // four indices, one per byte (used by the GEMM dequantiser)
    const uint32_t sel = nib & 0x07070707u;
    const uint32_t s = sel | (sel >> 4);
    const uint32_t s16 = (s & 0xffu) | ((s >> 8) & 0xff00u);
    const uint32_t lo = __byte_perm(0xBFAD9881u, 0xF6EADDCFu, s16);
    const uint32_t hi = __byte_perm(0x26190D01u, 0x71594535u, s16);
    const uint32_t mask = ((nib >> 3) & 0x01010101u) * 0xffu;
    return (hi & mask) | (lo & ~mask);
}
// four PACKED indices (bits 0..15 of the selector words) -> four table bytes. The nibbles of a weight word already are the selector format
// of PRMT: bits 0..2 pick inside a table half (s), bit 3 picks the half with a third PRMT (sel = 0x3210 | 4 * bit 3 per nibble).
__device__ __forceinline__ uint32_t iq4nl_lut_sel(uint32_t s, uint32_t sel) {
    const uint32_t lo = __byte_perm(0xBFAD9881u, 0xF6EADDCFu, s);
    const uint32_t hi = __byte_perm(0x26190D01u, 0x71594535u, s);
    return __byte_perm(lo, hi, sel);
}
__device__ __forceinline__ uint32_t iq4nl_lut_packed(uint32_t h) { return iq4nl_lut_sel(h & 0x7777u, ((h >> 1) & 0x4444u) | 0x3210u); }
// the looked-up bytes of a weight word come out as (lo0, hi0, lo1, hi1) / (lo2, hi2, lo3, hi3): the ACTIVATIONS are interleaved to match
// (once per lane step, shared by all rows) instead of de-interleaving every weight word. Both selector words are built for the whole 32-bit
// word at once (PRMT reads only the low 16 bits): 5 logic / shift instructions + 6 PRMT + 2 dp4a per 8 weights.
__device__ __forceinline__ int iq4_word_dot(uint32_t q, uint32_t x0, uint32_t x1, int acc) {
    const uint32_t s = q & 0x77777777u, sel = ((q >> 1) & 0x44444444u) | 0x32103210u;
    acc = __dp4a((int) iq4nl_lut_sel(s, sel), (int) __byte_perm(x0, x1, 0x5140u), acc);
    return __dp4a((int) iq4nl_lut_sel(s >> 16, sel >> 16), (int) __byte_perm(x0, x1, 0x7362u), acc);
}
template <> struct WFrag<B200_TYPE_IQ4_XS> {
    uint4 q; uint2 hdr;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int g) {
        const int64_t bi = row * W.nb + blk;
        q = ld_stream16(W.p0 + bi * 128 + 16 * g);
        hdr = ld_stream8(W.p1 + bi * 8);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int g) const {
        const int ls = (int) (((hdr.y >> (4 * g)) & 0xfu) | ((((hdr.x >> 16) >> (2 * g)) & 3u) << 4)) - 32;
        const float dl = half_bits_to_float((uint16_t) (hdr.x & 0xffffu)) * (float) ls;
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            int s = iq4_word_dot(q.x, x[c].x0.x, x[c].x1.x, 0);
            s = iq4_word_dot(q.y, x[c].x0.y, x[c].x1.y, s);
            s = iq4_word_dot(q.z, x[c].x0.z, x[c].x1.z, s);
            s = iq4_word_dot(q.w, x[c].x0.w, x[c].x1.w, s);
            acc[c] += (x[c].dx * dl) * (float) s;
        }
    }
};

// ---- Q4_0 planes: QS 16 | D 2 (ggml-common.h:170-175) --------------------------------------------
template <> struct WFrag<B200_TYPE_Q4_0> {
    uint4 q; uint16_t d;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int) {
        const int64_t bi = row * W.nb + blk;
        q = ld_stream16(W.p0 + bi * 16);
        d = ld_stream2(W.p1 + bi * 2);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int) const {
        const uint4 lo = and4(q, 0x0f0f0f0fu), hi = and4(shr4(q, 4), 0x0f0f0f0fu);
        const float dw = half_bits_to_float(d);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const int isum = dot16(lo, x[c].x0) + dot16(hi, x[c].x1) - 8 * (x[c].bs0 + x[c].bs1);
            acc[c] += (x[c].dx * dw) * (float) isum;
        }
    }
};

// ---- Q8_0 planes: QS 32 | D 2 (ggml-common.h:219-224) --------------------------------------------
template <> struct WFrag<B200_TYPE_Q8_0> {
    uint4 q0, q1; uint16_t d;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int64_t blk, int) {
        const int64_t bi = row * W.nb + blk;
        q0 = ld_stream16(W.p0 + bi * 32);
        q1 = ld_stream16(W.p0 + bi * 32 + 16);
        d = ld_stream2(W.p1 + bi * 2);
    }
    template <int NCOLS>
    __device__ __forceinline__ void dot(const XFrag (&x)[NCOLS], float (&acc)[NCOLS], int) const {
        const float dw = half_bits_to_float(d);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c)
            acc[c] += (x[c].dx * dw) * (float) (dot16(q0, x[c].x0) + dot16(q1, x[c].x1));
    }
};

#endif // __CUDACC__

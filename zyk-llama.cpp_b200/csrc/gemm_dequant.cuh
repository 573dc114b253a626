// gemm_dequant.cuh -- A-operand producers of the prefill GEMM: one thread turns 64 consecutive
// k-elements of one quantised weight row into fp16 and writes them as one 128-byte row of a
// K-major SWIZZLE_128B UMMA tile (16-byte chunk c of row r lands at chunk position c ^ (r & 7)).
//
// Values follow the reference dequantisers (ggml/src/ggml-quants.c: dequantize_row_q4_0 :307-325,
// _q8_0 :401-415, _q4_K :1352-1375, _q6_K :1762-1792, _iq4_xs :2530-2551): w = scale * q (- min),
// with the integer q recovered exactly in fp16 through the 0x6400 | q trick (1024 + q) and the
// per-sub-block scale rounded once to fp16 (relative error 2^-11; the MUL_MAT tolerance of the
// reference is NMSE 5e-4, tests/test-backend-ops.cpp:3716-3718).
#pragma once
#include "gemv_frag.cuh"

#ifdef __CUDACC__

__device__ __forceinline__ __half2 u32_as_h2(uint32_t v) { return *reinterpret_cast<__half2 *>(&v); }
__device__ __forceinline__ uint32_t h2_as_u32(__half2 v) { return *reinterpret_cast<uint32_t *>(&v); }

// bytes b0..b3 of `v` (unsigned 0..255) -> fp16 pairs (b0 - bias, b1 - bias), (b2 - bias, b3 - bias); bias2 = 1024 + bias
__device__ __forceinline__ void bytes_to_h2(uint32_t v, __half2 bias2, __half2 & lo, __half2 & hi) {
    lo = __hsub2(u32_as_h2(__byte_perm(v, 0x64646464u, 0x4140)), bias2);
    hi = __hsub2(u32_as_h2(__byte_perm(v, 0x64646464u, 0x4342)), bias2);
}

// store chunk c (8 halves) of tile row r; row_addr = shared address of the row's 128 bytes
__device__ __forceinline__ void st_chunk(uint32_t row_addr, int r, int c, __half2 a, __half2 b, __half2 c2, __half2 d) {
    const uint32_t addr = row_addr + (uint32_t) (((c ^ r) & 7) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(h2_as_u32(a)), "r"(h2_as_u32(b)), "r"(h2_as_u32(c2)), "r"(h2_as_u32(d)) : "memory");
}

// two words (8 bytes, unsigned with bias) * s - c  -> one chunk
__device__ __forceinline__ void emit8(uint32_t row_addr, int r, int c, uint32_t w0, uint32_t w1, __half2 bias2, __half2 s, __half2 nc) {
    __half2 a, b, e, f;
    bytes_to_h2(w0, bias2, a, b);
    bytes_to_h2(w1, bias2, e, f);
    st_chunk(row_addr, r, c, __hfma2(a, s, nc), __hfma2(b, s, nc), __hfma2(e, s, nc), __hfma2(f, s, nc));
}


// ---- raw quantised tiles staged in shared memory by TMA (gemm.cu): 128 rows x one 256-element super-block ----
__device__ __forceinline__ uint4 raw_lds16(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ uint16_t raw_lds2(uint32_t a) { uint16_t r; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(a)); return r; }
// 16-byte chunk c of row r in a TMA SWIZZLE_128B region (128-byte rows) / SWIZZLE_64B region (64-byte rows)
__device__ __forceinline__ uint32_t sw128(uint32_t base, int r, int c) { return base + (uint32_t) (r * 128 + (((c ^ r) & 7) << 4)); }
__device__ __forceinline__ uint32_t sw64(uint32_t base, int r, int c) { return base + (uint32_t) (r * 64 + (((c ^ (r >> 1)) & 3) << 4)); }

// byte sizes of the raw stage regions per type (each region 1024-byte aligned): see RawLayout in gemm.cu
template <int TYPE> struct Deq;

// ---- Q4_K (canonical 144 B blocks) -------------------------------------------------------------------
template <> struct Deq<B200_TYPE_Q4_K> {
    uint4 hdr, q0, q1;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const uint8_t * b = W.p0 + (row * W.nb + (kb >> 2)) * 144;
        hdr = ld_cached16(b);
        q0 = ld_stream16(b + 16 + 32 * (kb & 3));
        q1 = ld_stream16(b + 32 + 32 * (kb & 3));
    }
    // raw: [128][144] no swizzle
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2) {
        const uint32_t b = raw + (uint32_t) (r * 144);
        hdr = raw_lds16(b); q0 = raw_lds16(b + 16 + 32 * j); q1 = raw_lds16(b + 32 + 32 * j);
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW &, int64_t, int) { return make_uint2(0u, 0u); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int kb) const {
        const int j0 = 2 * (kb & 3);
        int sc0, sc1, m0, m1;
        if (j0 < 4) {
            sc0 = byte_of(hdr, 4 + j0) & 63;      m0 = byte_of(hdr, 8 + j0) & 63;
            sc1 = byte_of(hdr, 5 + j0) & 63;      m1 = byte_of(hdr, 9 + j0) & 63;
        } else {
            const uint32_t a0 = byte_of(hdr, 8 + j0), a1 = byte_of(hdr, 9 + j0);
            sc0 = (a0 & 0xf) | ((byte_of(hdr, j0) >> 6) << 4);      m0 = (a0 >> 4) | ((byte_of(hdr, 4 + j0) >> 6) << 4);
            sc1 = (a1 & 0xf) | ((byte_of(hdr, 1 + j0) >> 6) << 4);  m1 = (a1 >> 4) | ((byte_of(hdr, 5 + j0) >> 6) << 4);
        }
        const float2 dm = __half22float2(u32_as_h2(hdr.x));
        const __half2 s0 = __float2half2_rn(dm.x * (float) sc0), s1 = __float2half2_rn(dm.x * (float) sc1);
        const __half2 c0 = __float2half2_rn(-dm.y * (float) m0), c1 = __float2half2_rn(-dm.y * (float) m1);
        const __half2 bias = __float2half2_rn(1024.0f);
        const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            emit8(row_addr, r, c,     w[2 * c] & 0x0f0f0f0fu,        w[2 * c + 1] & 0x0f0f0f0fu,        bias, s0, c0);
            emit8(row_addr, r, c + 4, (w[2 * c] >> 4) & 0x0f0f0f0fu, (w[2 * c + 1] >> 4) & 0x0f0f0f0fu, bias, s1, c1);
        }
    }
};

// ---- Q5_K (canonical 176 B blocks: hdr 16 | qh 32 | qs 128; dequantize_row_q5_K, ggml-quants.c) ----------
template <> struct Deq<B200_TYPE_Q5_K> {
    uint4 hdr, h0, h1, q0, q1;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const uint8_t * b = W.p0 + (row * W.nb + (kb >> 2)) * 176;
        hdr = ld_cached16(b); h0 = ld_cached16(b + 16); h1 = ld_cached16(b + 32);
        q0 = ld_stream16(b + 48 + 32 * (kb & 3));
        q1 = ld_stream16(b + 64 + 32 * (kb & 3));
    }
    // raw: [128][176] no swizzle (44-word rows: 8 consecutive rows x 16 B cover all 32 banks once)
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2) {
        const uint32_t b = raw + (uint32_t) (r * 176);
        hdr = raw_lds16(b); h0 = raw_lds16(b + 16); h1 = raw_lds16(b + 32);
        q0 = raw_lds16(b + 48 + 32 * j); q1 = raw_lds16(b + 64 + 32 * j);
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW &, int64_t, int) { return make_uint2(0u, 0u); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int kb) const {
        const int j0 = 2 * (kb & 3);
        int sc0, sc1, m0, m1;
        if (j0 < 4) {
            sc0 = byte_of(hdr, 4 + j0) & 63;      m0 = byte_of(hdr, 8 + j0) & 63;
            sc1 = byte_of(hdr, 5 + j0) & 63;      m1 = byte_of(hdr, 9 + j0) & 63;
        } else {
            const uint32_t a0 = byte_of(hdr, 8 + j0), a1 = byte_of(hdr, 9 + j0);
            sc0 = (a0 & 0xf) | ((byte_of(hdr, j0) >> 6) << 4);      m0 = (a0 >> 4) | ((byte_of(hdr, 4 + j0) >> 6) << 4);
            sc1 = (a1 & 0xf) | ((byte_of(hdr, 1 + j0) >> 6) << 4);  m1 = (a1 >> 4) | ((byte_of(hdr, 5 + j0) >> 6) << 4);
        }
        const float2 dm = __half22float2(u32_as_h2(hdr.x));
        const __half2 s0 = __float2half2_rn(dm.x * (float) sc0), s1 = __float2half2_rn(dm.x * (float) sc1);
        const __half2 c0 = __float2half2_rn(-dm.y * (float) m0), c1 = __float2half2_rn(-dm.y * (float) m1);
        const __half2 bias = __float2half2_rn(1024.0f);
        const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        const uint32_t hb[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t l0 = ((hb[2 * c] >> j0) & 0x01010101u) << 4, l1 = ((hb[2 * c + 1] >> j0) & 0x01010101u) << 4;
            const uint32_t u0 = ((hb[2 * c] >> (j0 + 1)) & 0x01010101u) << 4, u1 = ((hb[2 * c + 1] >> (j0 + 1)) & 0x01010101u) << 4;
            emit8(row_addr, r, c,     (w[2 * c] & 0x0f0f0f0fu) | l0,        (w[2 * c + 1] & 0x0f0f0f0fu) | l1,        bias, s0, c0);
            emit8(row_addr, r, c + 4, ((w[2 * c] >> 4) & 0x0f0f0f0fu) | u0, ((w[2 * c + 1] >> 4) & 0x0f0f0f0fu) | u1, bias, s1, c1);
        }
    }
};

// ---- Q6_K (planes: QL 128 | QH 64 | SC 16 | D 2) -------------------------------------------------------
template <> struct Deq<B200_TYPE_Q6_K> {
    uint4 ql[4], qh[2], sc; uint16_t d;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const int64_t bi = row * W.nb + (kb >> 2);
        const int h = (kb >> 1) & 1;                         // which 128-element half of the super-block
        const uint8_t * l = W.p0 + bi * 128 + 64 * h;
#pragma unroll
        for (int i = 0; i < 4; ++i) ql[i] = ld_stream16(l + 16 * i);
        qh[0] = ld_stream16(W.p1 + bi * 64 + 32 * h);
        qh[1] = ld_stream16(W.p1 + bi * 64 + 32 * h + 16);
        sc = ld_cached16(W.p2 + bi * 16);
        d = ld_stream2(W.p3 + bi * 2);
    }
    // raw: ql [128][128] SW128 @0 | qh [128][64] SW64 @16384 | sc [128][16] @24576 ; d comes through aux (global load)
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2 aux) {
        const int h = j >> 1;
#pragma unroll
        for (int i = 0; i < 4; ++i) ql[i] = raw_lds16(sw128(raw, r, 4 * h + i));
        qh[0] = raw_lds16(sw64(raw + 16384, r, 2 * h));
        qh[1] = raw_lds16(sw64(raw + 16384, r, 2 * h + 1));
        sc = raw_lds16(raw + 24576 + (uint32_t) (r * 16));
        d = (uint16_t) aux.x;
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW & W, int64_t row, int sb) { return make_uint2(ld_stream2(W.p3 + (row * W.nb + sb) * 2), 0u); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int kb) const {
        const int h = (kb >> 1) & 1, c2 = kb & 1;            // c2: 0 = low nibbles (elements 0..63 of the half), 1 = high
        const float dw = half_bits_to_float(d);
        const __half2 bias = __float2half2_rn(1024.0f + 32.0f);
        const __half2 zero = __float2half2_rn(0.0f);
        const uint32_t lw[16] = {ql[0].x, ql[0].y, ql[0].z, ql[0].w, ql[1].x, ql[1].y, ql[1].z, ql[1].w,
                                 ql[2].x, ql[2].y, ql[2].z, ql[2].w, ql[3].x, ql[3].y, ql[3].z, ql[3].w};
        const uint32_t hw[8] = {qh[0].x, qh[0].y, qh[0].z, qh[0].w, qh[1].x, qh[1].y, qh[1].z, qh[1].w};
#pragma unroll
        for (int c = 0; c < 8; ++c) {                         // chunk c = elements 8c..8c+7 of the k-block
            const int qd = c >> 2;                            // which 32-element quarter inside the k-block
            const int qi = 2 * c2 + qd;                       // quarter index inside the 128-half (0..3)
            const int is = 8 * h + 2 * qi + ((c >> 1) & 1);   // scale index, 16 elements per scale
            const __half2 s = __float2half2_rn(dw * (float) (int) (int8_t) byte_of(sc, is));
            uint32_t v[2];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const uint32_t l = lw[2 * c + t];             // ql bytes e = 8c + 4t .. +3 of this half
                const uint32_t hb = hw[2 * (c & 3) + t];      // qh bytes l = e % 32
                const uint32_t nib = c2 ? ((l >> 4) & 0x0f0f0f0fu) : (l & 0x0f0f0f0fu);
                v[t] = nib | (((hb >> (2 * qi)) & 0x03030303u) << 4);
            }
            emit8(row_addr, r, c, v[0], v[1], bias, s, zero);
        }
    }
};

// ---- Q8_0 (planes: QS 32 | D 2), two blocks per k-block ----------------------------------------------------
template <> struct Deq<B200_TYPE_Q8_0> {
    uint4 q[4]; uint16_t d[2];
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const int64_t bi = row * W.nb + 2 * kb;
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = ld_stream16(W.p0 + bi * 32 + 16 * i);
        d[0] = ld_stream2(W.p1 + bi * 2);
        d[1] = ld_stream2(W.p1 + bi * 2 + 2);
    }
    // raw: qs bytes 0..127 [128][128] SW128 @0 | qs bytes 128..255 [128][128] SW128 @16384 | d [128][16] @32768
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2) {
        const uint32_t reg = raw + (uint32_t) ((j >> 1) * 16384);
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = raw_lds16(sw128(reg, r, 4 * (j & 1) + i));
        d[0] = raw_lds2(raw + 32768 + (uint32_t) (r * 16 + (2 * j) * 2));
        d[1] = raw_lds2(raw + 32768 + (uint32_t) (r * 16 + (2 * j + 1) * 2));
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW &, int64_t, int) { return make_uint2(0u, 0u); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int) const {
        const __half2 bias = __float2half2_rn(1024.0f + 128.0f);
        const __half2 zero = __float2half2_rn(0.0f);
        const uint32_t w[16] = {q[0].x, q[0].y, q[0].z, q[0].w, q[1].x, q[1].y, q[1].z, q[1].w,
                                q[2].x, q[2].y, q[2].z, q[2].w, q[3].x, q[3].y, q[3].z, q[3].w};
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const __half hs = __ushort_as_half(d[c >> 2]);
            emit8(row_addr, r, c, w[2 * c] ^ 0x80808080u, w[2 * c + 1] ^ 0x80808080u, bias, __halves2half2(hs, hs), zero);
        }
    }
};

// ---- Q4_0 (planes: QS 16 | D 2), two blocks per k-block ----------------------------------------------------
template <> struct Deq<B200_TYPE_Q4_0> {
    uint4 q[2]; uint16_t d[2];
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const int64_t bi = row * W.nb + 2 * kb;
        q[0] = ld_stream16(W.p0 + bi * 16);
        q[1] = ld_stream16(W.p0 + bi * 16 + 16);
        d[0] = ld_stream2(W.p1 + bi * 2);
        d[1] = ld_stream2(W.p1 + bi * 2 + 2);
    }
    // raw: qs [128][128] SW128 @0 (8 blocks x 16 B) | d [128][16] @16384
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2) {
        q[0] = raw_lds16(sw128(raw, r, 2 * j));
        q[1] = raw_lds16(sw128(raw, r, 2 * j + 1));
        d[0] = raw_lds2(raw + 16384 + (uint32_t) (r * 16 + (2 * j) * 2));
        d[1] = raw_lds2(raw + 16384 + (uint32_t) (r * 16 + (2 * j + 1) * 2));
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW &, int64_t, int) { return make_uint2(0u, 0u); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int) const {
        const __half2 bias = __float2half2_rn(1024.0f + 8.0f);
        const __half2 zero = __float2half2_rn(0.0f);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const __half hs = __ushort_as_half(d[b]);
            const __half2 s = __halves2half2(hs, hs);
            const uint32_t w[4] = {q[b].x, q[b].y, q[b].z, q[b].w};
#pragma unroll
            for (int c = 0; c < 2; ++c) {                     // low nibbles: elements 0..15, high: 16..31 of the block
                emit8(row_addr, r, 4 * b + c,     w[2 * c] & 0x0f0f0f0fu,        w[2 * c + 1] & 0x0f0f0f0fu,        bias, s, zero);
                emit8(row_addr, r, 4 * b + c + 2, (w[2 * c] >> 4) & 0x0f0f0f0fu, (w[2 * c + 1] >> 4) & 0x0f0f0f0fu, bias, s, zero);
            }
        }
    }
};

// ---- IQ4_XS (planes: QS 128 | HDR 8), two 32-element sub-blocks per k-block ---------------------------------
template <> struct Deq<B200_TYPE_IQ4_XS> {
    uint4 q[2]; uint2 hdr;
    __device__ __forceinline__ void load(const GemvW & W, int64_t row, int kb) {
        const int64_t bi = row * W.nb + (kb >> 2);
        q[0] = ld_stream16(W.p0 + bi * 128 + 32 * (kb & 3));
        q[1] = ld_stream16(W.p0 + bi * 128 + 32 * (kb & 3) + 16);
        hdr = ld_stream8(W.p1 + bi * 8);
    }
    // raw: qs [128][128] SW128 @0 ; the 8-byte header comes through two aux words (global loads)
    __device__ __forceinline__ void load_smem(uint32_t raw, int r, int j, uint2 aux) {
        q[0] = raw_lds16(sw128(raw, r, 2 * j));
        q[1] = raw_lds16(sw128(raw, r, 2 * j + 1));
        hdr = aux;
    }
    static __device__ __forceinline__ uint2 aux_load(const GemvW & W, int64_t row, int sb) { return ld_stream8(W.p1 + (row * W.nb + sb) * 8); }
    __device__ __forceinline__ void store(uint32_t row_addr, int r, int kb) const {
        const __half2 bias = __float2half2_rn(1024.0f + 128.0f);
        const __half2 zero = __float2half2_rn(0.0f);
        const float dw = half_bits_to_float((uint16_t) (hdr.x & 0xffffu));
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int ib = 2 * (kb & 3) + b;
            const int ls = (int) (((hdr.y >> (4 * ib)) & 0xfu) | ((((hdr.x >> 16) >> (2 * ib)) & 3u) << 4)) - 32;
            const __half2 s = __float2half2_rn(dw * (float) ls);
            const uint32_t w[4] = {q[b].x, q[b].y, q[b].z, q[b].w};
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                emit8(row_addr, r, 4 * b + c,     iq4nl_lut4(w[2 * c] & 0x0f0f0f0fu) ^ 0x80808080u,        iq4nl_lut4(w[2 * c + 1] & 0x0f0f0f0fu) ^ 0x80808080u,        bias, s, zero);
                emit8(row_addr, r, 4 * b + c + 2, iq4nl_lut4((w[2 * c] >> 4) & 0x0f0f0f0fu) ^ 0x80808080u, iq4nl_lut4((w[2 * c + 1] >> 4) & 0x0f0f0f0fu) ^ 0x80808080u, bias, s, zero);
            }
        }
    }
};

#endif // __CUDACC__

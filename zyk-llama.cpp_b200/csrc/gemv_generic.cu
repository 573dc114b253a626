// gemv_generic.cu -- canonical-layout GEMV for weights that were not repacked into planes:
// k not a multiple of 256, strided rows (views), or tensors written piecemeal. Same integer
// arithmetic as gemv_frag.cuh, byte-granular loads straight from the ggml block structs
// (ggml-common.h:170-431). Correctness path, not the roofline path.
#include "gemv.h"

__device__ __forceinline__ int ld_u8(const uint8_t * p) { return (int) __ldg(p); }
__device__ __forceinline__ float ld_f16(const uint8_t * p) {
    return half_bits_to_float((uint16_t) (ld_u8(p) | (ld_u8(p + 1) << 8)));
}

// contribution of 32-element group `sb` of one weight row against one quantised activation row
template <int TYPE>
__device__ float group_dot(const uint8_t * __restrict__ wrow, int64_t sb, const int8_t * __restrict__ xq,
                           const float * __restrict__ xd, const int16_t * __restrict__ xbs, int act_block) {
    const int8_t * x = xq + sb * 32;
    const float dx = xd[act_block == 256 ? (sb >> 3) : sb];
    const int bsum = (int) xbs[2 * sb] + (int) xbs[2 * sb + 1];
    if (TYPE == B200_TYPE_Q4_0) {
        const uint8_t * b = wrow + sb * 18;
        int s = 0;
        for (int j = 0; j < 16; ++j) { const int q = ld_u8(b + 2 + j); s += (q & 15) * x[j] + (q >> 4) * x[j + 16]; }
        return (dx * ld_f16(b)) * (float) (s - 8 * bsum);
    } else if (TYPE == B200_TYPE_Q8_0) {
        const uint8_t * b = wrow + sb * 34;
        int s = 0;
        for (int j = 0; j < 32; ++j) s += (int) (int8_t) ld_u8(b + 2 + j) * x[j];
        return (dx * ld_f16(b)) * (float) s;
    } else if (TYPE == B200_TYPE_Q4_K) {
        const uint8_t * b = wrow + (sb >> 3) * 144;
        const int j = (int) (sb & 7);
        const uint8_t * sc = b + 4;
        int scale, mn;
        if (j < 4) { scale = ld_u8(sc + j) & 63; mn = ld_u8(sc + j + 4) & 63; }
        else { scale = (ld_u8(sc + j + 4) & 15) | ((ld_u8(sc + j - 4) >> 6) << 4); mn = (ld_u8(sc + j + 4) >> 4) | ((ld_u8(sc + j) >> 6) << 4); }
        const uint8_t * q = b + 16 + 32 * (j >> 1);
        int s = 0;
        for (int l = 0; l < 32; ++l) { const int v = ld_u8(q + l); s += ((j & 1) ? (v >> 4) : (v & 15)) * x[l]; }
        return dx * (ld_f16(b) * (float) (scale * s) - ld_f16(b + 2) * (float) (mn * bsum));
    } else if (TYPE == B200_TYPE_Q5_K) {
        const uint8_t * b = wrow + (sb >> 3) * 176;
        const int j = (int) (sb & 7);
        const uint8_t * sc = b + 4;
        int scale, mn;
        if (j < 4) { scale = ld_u8(sc + j) & 63; mn = ld_u8(sc + j + 4) & 63; }
        else { scale = (ld_u8(sc + j + 4) & 15) | ((ld_u8(sc + j - 4) >> 6) << 4); mn = (ld_u8(sc + j + 4) >> 4) | ((ld_u8(sc + j) >> 6) << 4); }
        const uint8_t * q = b + 48 + 32 * (j >> 1);
        int s = 0;
        for (int l = 0; l < 32; ++l) {
            const int v = ld_u8(q + l);
            s += (((j & 1) ? (v >> 4) : (v & 15)) | (((ld_u8(b + 16 + l) >> j) & 1) << 4)) * x[l];
        }
        return dx * (ld_f16(b) * (float) (scale * s) - ld_f16(b + 2) * (float) (mn * bsum));
    } else if (TYPE == B200_TYPE_Q6_K) {
        const uint8_t * b = wrow + (sb >> 3) * 210;
        const int q32 = (int) (sb & 7);            // which 32-element quarter of the 256 block
        const int half = q32 >> 2, quarter = q32 & 3;
        const uint8_t * ql = b + 64 * half + ((quarter & 1) ? 32 : 0);
        const uint8_t * qh = b + 128 + 32 * half;
        int s[2] = {0, 0};
        for (int l = 0; l < 32; ++l) {
            const int lo = ld_u8(ql + l);
            const int nib = (quarter & 2) ? (lo >> 4) : (lo & 15);
            const int v = (nib | (((ld_u8(qh + l) >> (2 * quarter)) & 3) << 4)) - 32;
            s[l >> 4] += v * x[l];
        }
        const int sc0 = (int) (int8_t) ld_u8(b + 192 + 2 * q32), sc1 = (int) (int8_t) ld_u8(b + 192 + 2 * q32 + 1);
        return (dx * ld_f16(b + 208)) * (float) (sc0 * s[0] + sc1 * s[1]);
    } else { // IQ4_XS
        const int8_t kv[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};
        const uint8_t * b = wrow + (sb >> 3) * 136;
        const int ib = (int) (sb & 7);
        const int sh = ld_u8(b + 2) | (ld_u8(b + 3) << 8);
        const int ls = (((ld_u8(b + 4 + (ib >> 1)) >> (4 * (ib & 1))) & 15) | (((sh >> (2 * ib)) & 3) << 4)) - 32;
        const uint8_t * q = b + 8 + 16 * ib;
        int s = 0;
        for (int j = 0; j < 16; ++j) { const int v = ld_u8(q + j); s += kv[v & 15] * x[j] + kv[v >> 4] * x[j + 16]; }
        return (dx * ld_f16(b)) * (float) (ls * s);
    }
}

// one warp per weight row; lanes stride over the row's 32-element groups
template <int TYPE>
__global__ void __launch_bounds__(128) gemv_generic_kernel(GemvGenericArgs a) {
    const int lane = threadIdx.x & 31;
    if (a.pdl) { pdl_launch_dependents(); pdl_wait(); }
    const int64_t row = (int64_t) blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= a.m) return;
    const uint8_t * w = a.w;
    const int8_t * xqs = a.xqs; const float * xd = a.xd; const int16_t * xbs = a.xbs;
    float * y = a.y;
    if (a.ids) {
        const int64_t t = blockIdx.y / a.n_used, s = blockIdx.y % a.n_used;
        const int e = a.ids[t * a.ids_row_stride + s];
        y += t * a.y_token_stride + s * a.ldy;
        if (e < 0 || e >= a.n_as) { if (lane == 0) y[row] = 0.0f; return; }
        w += (int64_t) e * a.expert_stride;
        const int64_t xr = t * a.nb1 + (s % a.nb1);
        xqs += xr * a.k; xd += xr * (a.k / a.act_block); xbs += xr * (a.k >> 4);
    }
    const uint8_t * wrow = w + row * a.w_row_stride;
    const int64_t ngroups = a.k >> 5;
    for (int c = 0; c < a.ncols; ++c) {
        float acc = 0.0f;
        for (int64_t sb = lane; sb < ngroups; sb += 32)
            acc += group_dot<TYPE>(wrow, sb, xqs + c * a.k, xd + c * (a.k / a.act_block), xbs + c * (a.k >> 4), a.act_block);
        acc = warp_sum(acc);
        if (lane == 0) y[c * a.ldy + row] = acc;
    }
}

int b200_gemv_generic_launch(int type, const GemvGenericArgs & a, int64_t pairs, cudaStream_t s) {
    dim3 grid((unsigned) ((a.m + 3) / 4), (unsigned) pairs);
#define B200_GEN(T) case T: B200_CHECK(b200_launch(gemv_generic_kernel<T>, grid, dim3(128), 0, s, a.pdl != 0, a)); return 0
    switch (type) {
        B200_GEN(B200_TYPE_Q4_0); B200_GEN(B200_TYPE_Q8_0); B200_GEN(B200_TYPE_Q4_K); B200_GEN(B200_TYPE_Q5_K); B200_GEN(B200_TYPE_Q6_K); B200_GEN(B200_TYPE_IQ4_XS);
    }
#undef B200_GEN
    b200_set_error("gemv_generic: unsupported weight type %d", type);
    return -2;
}

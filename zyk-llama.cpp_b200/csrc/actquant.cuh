// actquant.cuh -- activation quantiser shared by the C-ABI entry point (canonical ggml blocks,
// bit-exact with the reference) and the matmul paths (plane layout consumed by the GEMV kernels).
//
// Arithmetic follows, operation for operation in IEEE fp32 (explicit _rn intrinsics, so neither
// -use_fast_math nor FMA contraction can change a bit):
//   Q8_K : quantize_row_q8_K_ref            /root/reference/ggml/src/ggml-quants.c:2555-2591
//   Q8_0 : quantize_row_q8_0 (AVX2) / _ref   ggml-cpu/arch/x86/quants.c:290-360 / ggml-quants.c:199-222
//   Q8_1 : quantize_row_q8_1 (AVX2) / _ref   ggml-cpu/arch/x86/quants.c:388-480 / ggml-quants.c:225-258
#pragma once
#include "common.cuh"

// plane layout of a quantised activation matrix (n rows of k elements) inside a workspace:
//   qs [n][k]      int8
//   d  [n][k/B]    f32   (B = 32 for the Q8_0 family -- value of the fp16-rounded scale -- or 256 for Q8_K)
//   bs [n][k/16]   int16 (sums of 16 consecutive quants: Q8_K's bsums, also kept for Q8_0)
struct XqPlanes {
    int8_t *  qs;
    float *   d;
    int16_t * bs;
    int64_t   k;
    int       block; // 32 or 256
};

static inline size_t xq_align(size_t v) { return (v + 255) & ~(size_t) 255; }
static inline size_t xq_planes_bytes(int act_type, int64_t n, int64_t k) {
    const int B = act_type == B200_TYPE_Q8_K ? 256 : 32;
    return xq_align((size_t) n * k) + xq_align((size_t) n * (k / B) * 4) + xq_align((size_t) n * (k / 16) * 2);
}
static inline XqPlanes xq_planes_carve(void * ws, int act_type, int64_t n, int64_t k) {
    XqPlanes p;
    const int B = act_type == B200_TYPE_Q8_K ? 256 : 32;
    uint8_t * b = (uint8_t *) ws;
    p.qs = (int8_t *) b;            b += xq_align((size_t) n * k);
    p.d  = (float *) b;             b += xq_align((size_t) n * (k / B) * 4);
    p.bs = (int16_t *) b;
    p.k = k;
    p.block = B;
    return p;
}

#ifdef __CUDACC__

// Quantise the warp's 256-element chunk: lane l holds elements [8l, 8l+8) in v. Returns the lane's
// 8 quants, the scale of the lane's block as the format stores it (f32 for Q8_K, fp16-rounded for
// Q8_0/Q8_1) and, for Q8_1, s = d * sum(q). Warp-collective (shuffles): all 32 lanes must call it.
template <int QT, int FLAVOUR>
__device__ __forceinline__ void quant_chunk256(const float (&v)[8], int lane, int (&q)[8], float & dval, float & sval) {
    sval = 0.0f;
    if (QT == B200_TYPE_Q8_K) {
        // first element of largest magnitude, sign kept (strict '>' in the reference loop)
        float amax = 0.0f; int idx = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float ax = fabsf(v[j]); if (ax > amax) { amax = ax; idx = lane * 8 + j; } }
        if (amax == 0.0f) idx = 1 << 20;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
            const int   oi = __shfl_xor_sync(0xffffffffu, idx, o);
            if (oa > amax || (oa == amax && oi < idx)) { amax = oa; idx = oi; }
        }
        float mine = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) if ((idx & 7) == j) mine = v[j];
        const float maxv = __shfl_sync(0xffffffffu, mine, (idx >> 3) & 31);
        if (amax == 0.0f) {
#pragma unroll
            for (int j = 0; j < 8; ++j) q[j] = 0;
            dval = 0.0f;
        } else {
            const float iscale = __fdiv_rn(-127.0f, maxv);
#pragma unroll
            for (int j = 0; j < 8; ++j) { const int t = __float2int_rn(__fmul_rn(iscale, v[j])); q[j] = t > 127 ? 127 : t; }
            dval = __fdiv_rn(1.0f, iscale);
        }
    } else {
        float amax = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[j]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float d = __fdiv_rn(amax, 127.0f);
        float id;
        if (FLAVOUR == B200_Q8_X86) id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
        else                        id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float t = __fmul_rn(v[j], id);
            q[j] = FLAVOUR == B200_Q8_X86 ? __float2int_rn(t) : (int) roundf(t);
        }
        dval = __half2float(__float2half_rn(d));
        if (QT == B200_TYPE_Q8_1) {
            int s = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) s += (int)(int8_t) q[j];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            sval = __fmul_rn((float) s, d);
        }
    }
}

// One warp quantises one chunk of 256 consecutive elements of one row; lane l owns elements
// [8l, 8l+8). CANON selects canonical ggml block bytes (y) or the plane layout (pl).
// An activation source that is still to be gated: X = act(x) * x2 (GGML_OP_GLU with split operands; op 0 reglu, 1 geglu, 2 swiglu as in
// ggml_compute_forward_glu), so the GLU node in front of ffn_down needs no launch and no round trip of its own. op < 0: plain X.
struct ActGlu { const float * x2; int64_t ldx2; int op; int swapped; };
__device__ __forceinline__ float act_glu_fn(int op, float x) {
    if (op == 0) return x > 0.0f ? x : 0.0f;
    if (op == 1) return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x)));
    return x / (1.0f + expf(-x));
}
template <int QT, int FLAVOUR, bool CANON>
__global__ void __launch_bounds__(128) act_quant_kernel(const float * __restrict__ x, int64_t ldx, int64_t nrows, int64_t k,
                                                        uint8_t * __restrict__ y, XqPlanes pl, bool pdl, ActGlu glu) {
    const int lane = threadIdx.x & 31;
    const int64_t chunk = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nchunks = (k + 255) >> 8;
    if (pdl) { pdl_launch_dependents(); pdl_wait(); }
    if (chunk >= nchunks) return;
    const int64_t e0 = chunk * 256 + lane * 8;
    const bool live = e0 < k;               // k % 32 == 0, so a lane is entirely in or out

    for (int64_t row = blockIdx.y; row < nrows; row += gridDim.y) {
        const float * xr = x + row * ldx;
        float v[8];
        if (live) {
            if ((((uintptr_t)(xr + e0)) & 15) == 0) {
                const float4 a = *reinterpret_cast<const float4 *>(xr + e0);
                const float4 b = *reinterpret_cast<const float4 *>(xr + e0 + 4);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = xr[e0 + j];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.0f;
        }
        if (glu.op >= 0 && live) {
            const float * ur = glu.x2 + row * glu.ldx2 + e0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float u = ur[j]; v[j] = glu.swapped ? act_glu_fn(glu.op, u) * v[j] : act_glu_fn(glu.op, v[j]) * u; }
        }

        int q[8];
        float dval, sval;
        quant_chunk256<QT, FLAVOUR>(v, lane, q, dval, sval);

        int s8 = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) s8 += (int)(int8_t) q[j];
        const int s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);

        const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
        const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
        if (!live) continue;

        if (CANON) {
            if (QT == B200_TYPE_Q8_K) {
                uint8_t * blk = y + (row * (k >> 8) + chunk) * 292;
                if (lane == 0) *reinterpret_cast<float *>(blk) = dval;
                *reinterpret_cast<uint32_t *>(blk + 4 + 8 * lane) = w0;
                *reinterpret_cast<uint32_t *>(blk + 8 + 8 * lane) = w1;
                if ((lane & 1) == 0) *reinterpret_cast<int16_t *>(blk + 260 + lane) = (int16_t) s16;
            } else {
                const int hdr = QT == B200_TYPE_Q8_1 ? 4 : 2;
                const int bsz = 32 + hdr;
                uint8_t * blk = y + (row * (k >> 5) + chunk * 8 + (lane >> 2)) * bsz;
                if ((lane & 3) == 0) {
                    *reinterpret_cast<uint16_t *>(blk) = __half_as_ushort(__float2half_rn(dval));
                    if (QT == B200_TYPE_Q8_1) *reinterpret_cast<uint16_t *>(blk + 2) = __half_as_ushort(__float2half_rn(sval));
                }
                uint16_t * o = reinterpret_cast<uint16_t *>(blk + hdr + 8 * (lane & 3));
                o[0] = (uint16_t) w0; o[1] = (uint16_t)(w0 >> 16); o[2] = (uint16_t) w1; o[3] = (uint16_t)(w1 >> 16);
            }
        } else {
            *reinterpret_cast<uint2 *>(pl.qs + row * k + e0) = make_uint2(w0, w1);
            if ((lane & 1) == 0) pl.bs[row * (k >> 4) + (e0 >> 4)] = (int16_t) s16;
            if (QT == B200_TYPE_Q8_K) { if (lane == 0) pl.d[row * (k >> 8) + chunk] = dval; }
            else if ((lane & 3) == 0) pl.d[row * (k >> 5) + (e0 >> 5)] = dval;
        }
    }
}

#endif // __CUDACC__

// host launchers (quantize.cu)
int b200_act_quant_canonical(int q8_type, int flavour, const float * x, int64_t ldx, void * y, int64_t nrows, int64_t k, cudaStream_t s);
int b200_act_quant_planes(int act_type, int flavour, const float * x, int64_t ldx, const XqPlanes & pl, int64_t nrows, int64_t k,
                          cudaStream_t s, bool pdl, const ActGlu * glu = nullptr);

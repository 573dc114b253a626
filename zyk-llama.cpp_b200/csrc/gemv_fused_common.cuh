// gemv_fused_common.cuh -- scaffolding of the fused decode kernel (gemv_fused.cu; also used by the cp.async-staged variant kept
// under profiles/attic/gemv_staged.cu.txt): launch arguments, the shared-memory view of the quantised activations and the in-kernel
// activation quantiser.
#pragma once
#include "gemv_frag.cuh"
#include "actquant.cuh"
#include "gemv.h"
#include "sm100.cuh"

#include <stdlib.h>

namespace {
using namespace sm100;

constexpr int R = 2;                     // rows per warp step (share the activation fragments)

struct FusedMat {
    GemvW W;
    int type;
    int64_t m;
    float * y; int64_t ldy;
    int64_t pair_begin;                  // first row pair of this matrix in the concatenated pair space
};

struct FusedArgs {
    FusedMat mat[B200_MAX_GROUP];
    int nmat;
    int64_t total_pairs;
    int64_t k;
    const float * x; int64_t ldx;        // f32 activations, NCOLS rows
    int flavour;                         // enum b200_q8_flavour for the Q8_0 family
    int act_block;                       // 256 (Q8_K) or 32 (Q8_0)
    int pdl;
    XqPlanes xq_out;                     // CTA 0 also leaves the quantised X here (B200_MM_X_PREQUANTISED protocol)
    PeerGather peer;                     // row-split exchange fused into this launch (world <= 1: off)
};

__device__ __forceinline__ uint4 lds16(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ uint32_t lds4(uint32_t a) { uint32_t r; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a)); return r; }
__device__ __forceinline__ uint16_t lds2(uint32_t a) { uint16_t r; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(a)); return r; }
__device__ __forceinline__ void sts8(uint32_t a, uint32_t x, uint32_t y) { asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(a), "r"(x), "r"(y) : "memory"); }
__device__ __forceinline__ void sts4(uint32_t a, uint32_t x) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(x) : "memory"); }
__device__ __forceinline__ void sts2(uint32_t a, uint16_t x) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"(x) : "memory"); }

__device__ __forceinline__ int plane0_bytes(int type) {   // bytes per block in plane 0 of the plane layout (repack.cu)
    return type == B200_TYPE_Q4_K ? 144 : (type == B200_TYPE_Q6_K || type == B200_TYPE_IQ4_XS) ? 128 : type == B200_TYPE_Q8_0 ? 32 : 16;
}

// shared-memory view of the quantised activations
struct XS { uint32_t qs, d, bs; int k; int act_block; };

template <int TYPE>
__device__ __forceinline__ void xfrag_lds(XFrag & f, const XS & xs, int col, int blk, int g) {
    const int e0 = blk * Fmt<TYPE>::BLK + Fmt<TYPE>::off0(g);
    const int e1 = e0 + Fmt<TYPE>::RUN1;
    const uint32_t qc = xs.qs + (uint32_t) (col * xs.k);
    f.x0 = lds16(qc + e0);
    f.x1 = lds16(qc + e1);
    const uint32_t bc = xs.bs + (uint32_t) (col * (xs.k >> 4)) * 2;
    f.bs0 = (int) (int16_t) lds2(bc + (e0 >> 4) * 2);
    f.bs1 = (int) (int16_t) lds2(bc + (e1 >> 4) * 2);
    f.dx = __uint_as_float(lds4(xs.d + (uint32_t) (col * (xs.k / xs.act_block) + (xs.act_block == 256 ? (e0 >> 8) : (e0 >> 5))) * 4));
}


// fused activation quantiser: X f32 -> int8 planes in shared memory (whole CTA; ends with __syncthreads)
template <int NCOLS, int NWARPS>
__device__ __forceinline__ void quantise_x(const FusedArgs & a, const XS & xs, int warp, int lane) {
    const int k = xs.k;
    const int cpr = k >> 8;                          // 256-element chunks per activation row
    constexpr int QU = 2;                            // chunks whose loads are in flight together
    for (int c0 = warp; c0 < NCOLS * cpr; c0 += NWARPS * QU) {
        float v[QU][8];
#pragma unroll
        for (int u = 0; u < QU; ++u) {
            const int ci = c0 + u * NWARPS;
            if (ci < NCOLS * cpr) {
                const int col = ci / cpr, chunk = ci % cpr;
                const float * xr = a.x + (int64_t) col * a.ldx + chunk * 256 + lane * 8;
                if ((((uintptr_t) xr) & 15) == 0) {
                    const float4 lo = *reinterpret_cast<const float4 *>(xr), hi = *reinterpret_cast<const float4 *>(xr + 4);
                    v[u][0] = lo.x; v[u][1] = lo.y; v[u][2] = lo.z; v[u][3] = lo.w; v[u][4] = hi.x; v[u][5] = hi.y; v[u][6] = hi.z; v[u][7] = hi.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[u][j] = xr[j];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < QU; ++u) {
            const int ci = c0 + u * NWARPS;
            if (ci < NCOLS * cpr) {
                const int col = ci / cpr, chunk = ci % cpr;
                const int e0 = chunk * 256 + lane * 8;
                int q[8]; float dval, sval;
                if (a.act_block == 256)                  quant_chunk256<B200_TYPE_Q8_K, B200_Q8_X86>(v[u], lane, q, dval, sval);
                else if (a.flavour == B200_Q8_X86)       quant_chunk256<B200_TYPE_Q8_0, B200_Q8_X86>(v[u], lane, q, dval, sval);
                else                                     quant_chunk256<B200_TYPE_Q8_0, B200_Q8_REF>(v[u], lane, q, dval, sval);
                int s8 = 0;
    #pragma unroll
                for (int j = 0; j < 8; ++j) s8 += (int) (int8_t) q[j];
                const int s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
                const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) (q[3] & 0xff) << 24);
                const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) (q[7] & 0xff) << 24);
                sts8(xs.qs + (uint32_t) (col * k + e0), w0, w1);
                if ((lane & 1) == 0) sts2(xs.bs + (uint32_t) (col * (k >> 4) + (e0 >> 4)) * 2, (uint16_t) (int16_t) s16);
                if (a.act_block == 256) { if (lane == 0) sts4(xs.d + (uint32_t) (col * cpr + chunk) * 4, __float_as_uint(dval)); }
                else if ((lane & 3) == 0) sts4(xs.d + (uint32_t) (col * (k >> 5) + (e0 >> 5)) * 4, __float_as_uint(dval));
                if (blockIdx.x == 0 && a.xq_out.qs) {
                    *reinterpret_cast<uint2 *>(a.xq_out.qs + col * k + e0) = make_uint2(w0, w1);
                    if ((lane & 1) == 0) a.xq_out.bs[col * (k >> 4) + (e0 >> 4)] = (int16_t) s16;
                    if (a.act_block == 256) { if (lane == 0) a.xq_out.d[col * cpr + chunk] = dval; }
                    else if ((lane & 3) == 0) a.xq_out.d[col * (k >> 5) + (e0 >> 5)] = dval;
                }
            }
        }
    }
    __syncthreads();
}


} // namespace

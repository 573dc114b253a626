// gemv.cu -- (1) decode GEMV: Y[n<=8][m] = W[m][k] . X[n][k] for quantised W streamed once from HBM.
//
// Reference counterparts: ggml_compute_forward_mul_mat + ggml_vec_dot_* on the CPU
// (ggml-cpu/ggml-cpu.c:1227-1427, ggml-cpu/quants.c) and mul_mat_vec_q on CUDA
// (ggml-cuda/mmvq.cu:141-335). This is neither: each warp owns R weight rows and walks k in steps
// of 1024 elements; per step every lane issues R*U independent 128-bit streaming loads (L1
// no-allocate) before any arithmetic so that >= 64 KB are in flight per SM, multiplies against
// the int8 activation planes (L1-resident, shared by all warps of the SM) with dp4a, and the
// warp reduces with shuffles at the end. HBM-bound by construction: algorithmic bytes per
// launch = m * row_size(type, k) (+ the activations, < 0.1 %).
//
// With B200_MM_PDL the kernel is launched with programmatic stream serialisation: weight loads of
// the first step are issued BEFORE griddepcontrol.wait, so the HBM pipe is already primed while
// the producer of X (previous kernel on the stream) drains.
#include "gemv_frag.cuh"
#include "actquant.cuh"
#include "gemv.h"

// fused row-split exchange: every thread of the CTA passes here exactly once, whichever way its warp leaves the kernel
__device__ __forceinline__ void gemv_finish(const GemvArgs & a) {
    if (a.peer.world > 1 && a.peer.signal) {
        __syncthreads();
        if (threadIdx.x == 0) peer_signal(a.peer);
    }
}

template <int TYPE, int NCOLS, int R, int U, int KV>
__global__ void __launch_bounds__(128) gemv_kernel(GemvArgs a) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int g = lane & 7;
    constexpr int BLK = Fmt<TYPE>::BLK;
    constexpr int BLK_PER_STEP = 1024 / BLK;           // 4 super-blocks or 32 blocks per warp step

    if (a.pdl) pdl_launch_dependents();

    // MUL_MAT_ID: blockIdx.y is one (token, slot) pair; the expert is read on the device
    GemvW W = a.W;
    const int8_t * xqs = a.xqs; const float * xd = a.xd; const int16_t * xbs = a.xbs;
    float * y = a.y;
    bool waited = !a.pdl;
    if (a.ids) {
        if (!waited) { pdl_wait(); waited = true; }   // ids come from the previous kernel too
        const int64_t t = blockIdx.y / a.n_used, s = blockIdx.y % a.n_used;
        const int e = a.ids[t * a.ids_row_stride + s];
        y += t * a.y_token_stride + s * a.ldy;
        if (e < 0 || e >= a.n_as) {   // the reference asserts on this; we write zeros
            const int64_t row = ((int64_t) blockIdx.x * 4 + warp) * R;
            if (lane < R && row + lane < a.m) y[row + lane] = 0.0f;
            return;
        }
        W.p0 += (int64_t) e * a.expert_stride; W.p1 += (int64_t) e * a.expert_stride;
        W.p2 += (int64_t) e * a.expert_stride; W.p3 += (int64_t) e * a.expert_stride;
        const int64_t xr = t * a.nb1 + (s % a.nb1);
        xqs += xr * a.k; xd += xr * (a.k / a.act_block); xbs += xr * (a.k >> 4);
    }

    const int64_t row0 = ((int64_t) blockIdx.x * 4 + warp) * R;
    if (row0 >= a.m) { if (KV == KV_PEER) gemv_finish(a); return; }

    float acc[R][NCOLS];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[r][c] = 0.0f;

    const int64_t nb = W.nb;
    const int64_t nsteps = (nb + BLK_PER_STEP - 1) / BLK_PER_STEP;
    const int64_t lane_blk = BLK == 256 ? (lane >> 3) : lane;

    for (int64_t s0 = 0; s0 < nsteps; s0 += U) {
        WFrag<TYPE> f[U][R];
        bool valid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t blk = (s0 + u) * BLK_PER_STEP + lane_blk;
            valid[u] = blk < nb;
            const int64_t bl = valid[u] ? blk : 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int64_t row = row0 + r < a.m ? row0 + r : a.m - 1;
                f[u][r].load(W, row, bl, g);
            }
        }
        if (!waited) { pdl_wait(); waited = true; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!valid[u]) continue;
            const int64_t blk = (s0 + u) * BLK_PER_STEP + lane_blk;
            XFrag x[NCOLS];
#pragma unroll
            for (int c = 0; c < NCOLS; ++c)
                xfrag_load<TYPE>(x[c], xqs + c * a.k, xd + c * (a.k / a.act_block), xbs + c * (a.k >> 4), blk, g, a.act_block);
#pragma unroll
            for (int r = 0; r < R; ++r) f[u][r].template dot<NCOLS>(x, acc[r], g);
        }
    }
    if (!waited) pdl_wait();

#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const float v = warp_sum(acc[r][c]);
            if (lane == 0 && row0 + r < a.m) {
                if (KV == KV_PEER) peer_store(a.peer, y + c * a.ldy + row0 + r, v);
                else y[c * a.ldy + row0 + r] = v;
            }
        }
    if (KV == KV_PEER) gemv_finish(a);
}

// ------------------------------------------------------------------------------------------------
template <int TYPE, int NCOLS, int R, int U>
static int launch_gemv(const GemvArgs & a, int64_t pairs, cudaStream_t s) {
    dim3 grid((unsigned) ((a.m + 4 * R - 1) / (4 * R)), (unsigned) pairs);
    if (a.peer.world > 1)  B200_CHECK(b200_launch(gemv_kernel<TYPE, NCOLS, R, U, KV_PEER>, grid, dim3(128), 0, s, a.pdl != 0, a));
    else                   B200_CHECK(b200_launch(gemv_kernel<TYPE, NCOLS, R, U, KV_PLAIN>, grid, dim3(128), 0, s, a.pdl != 0, a));
    return 0;
}

template <int TYPE>
static int dispatch_gemv(const GemvArgs & a, int ncols, int64_t pairs, int R, int U, cudaStream_t s) {
    if (ncols == 1) {
#define B200_RU(r, u) if (R == r && U == u) return launch_gemv<TYPE, 1, r, u>(a, pairs, s)
        B200_RU(1, 1); B200_RU(1, 2); B200_RU(1, 4);
        B200_RU(2, 1); B200_RU(2, 2); B200_RU(2, 4);
        B200_RU(4, 1); B200_RU(4, 2);
#undef B200_RU
        return launch_gemv<TYPE, 1, 2, 2>(a, pairs, s);
    }
    switch (ncols) {
        case 2: return launch_gemv<TYPE, 2, 2, 1>(a, pairs, s);
        case 3: return launch_gemv<TYPE, 3, 2, 1>(a, pairs, s);
        case 4: return launch_gemv<TYPE, 4, 2, 1>(a, pairs, s);
        case 5: return launch_gemv<TYPE, 5, 1, 1>(a, pairs, s);
        case 6: return launch_gemv<TYPE, 6, 1, 1>(a, pairs, s);
        case 7: return launch_gemv<TYPE, 7, 1, 1>(a, pairs, s);
        case 8: return launch_gemv<TYPE, 8, 1, 1>(a, pairs, s);
    }
    b200_set_error("gemv: ncols %d out of range", ncols);
    return -2;
}

// default (R, U): enough warps to cover 148 SMs x 16 warps, then as much ILP as the row count allows
static void gemv_pick(int type, int64_t m, int64_t k, int tune, int * R, int * U) {
    if (tune) { *R = tune & 0xff; *U = (tune >> 8) & 0xff; return; }
    const int64_t steps = (k + 1023) / 1024;
    int r = m >= 148 * 16 * 4 ? 4 : (m >= 148 * 16 * 2 ? 2 : 1);
    if (type == B200_TYPE_Q6_K || type == B200_TYPE_Q8_0) r = r > 2 ? 2 : r;   // fatter fragments
    int u = steps >= 4 ? (r >= 4 ? 2 : 4) : (steps >= 2 ? 2 : 1);
    if (r == 4 && u > 2) u = 2;
    *R = r; *U = u;
}

int b200_gemv_launch(int type, const GemvArgs & a, int ncols, int64_t pairs, int tune, cudaStream_t s) {
    int R, U;
    gemv_pick(type, a.m, a.k, tune, &R, &U);
    switch (type) {
        case B200_TYPE_Q4_K:   return dispatch_gemv<B200_TYPE_Q4_K>(a, ncols, pairs, R, U, s);
        case B200_TYPE_Q5_K:   return dispatch_gemv<B200_TYPE_Q5_K>(a, ncols, pairs, R, U, s);
        case B200_TYPE_Q6_K:   return dispatch_gemv<B200_TYPE_Q6_K>(a, ncols, pairs, R, U, s);
        case B200_TYPE_IQ4_XS: return dispatch_gemv<B200_TYPE_IQ4_XS>(a, ncols, pairs, R, U, s);
        case B200_TYPE_Q4_0:   return dispatch_gemv<B200_TYPE_Q4_0>(a, ncols, pairs, R, U, s);
        case B200_TYPE_Q8_0:   return dispatch_gemv<B200_TYPE_Q8_0>(a, ncols, pairs, R, U, s);
    }
    b200_set_error("gemv: unsupported weight type %d", type);
    return -2;
}

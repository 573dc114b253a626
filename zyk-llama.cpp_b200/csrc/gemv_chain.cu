// gemv_chain.cu -- (1) decode GEMV, chain version: a whole sequence of DEPENDENT MUL_MAT groups (every quantised matmul of
// one decode token: wq/wk/wv -> wo -> gate/up -> down -> ... -> output) in ONE persistent launch.
//
// Reference counterparts: the node loop of ggml_backend_sched_compute_splits -> graph_compute (ggml-backend.cpp:1443-1620) over
// ggml_compute_forward_mul_mat + ggml_vec_dot_* (ggml-cpu/ggml-cpu.c:1227-1427, ggml-cpu/quants.c:115-1181); on the CUDA side
// quantize_q8_1 + mul_mat_vec_q per node (ggml-cuda/quantize.cu:4-73, mmvq.cu:141-335).
//
// Why: one decode token is ~130 dependent GEMVs of 1.5-10 us each at HBM speed. As separate launches every one of them pays a
// fill (first DRAM round trip, X quantisation) and a drain (last rows) during which HBM idles, and the register-streaming
// kernels spend ~88 warp instructions per 512 weight bytes (issue bound). Here
//   * one CTA per SM stays resident for the whole chain. Weights do not depend on activations, so each half-CTA ("group" of
//     8 warps = one virtual consumer, VC) keeps a ring of NS bulk copies (cp.async.bulk global -> shared, mbarrier
//     complete_tx, L2 evict-first) in flight that runs AHEAD across node boundaries: while the grid waits for the X of node
//     i + 1, the weights of node i + 1 are already landing in shared memory;
//   * a node boundary is chosen per node from the pointers (NB_* below): tagged 64-bit {epoch, f32} words when X is exactly the
//     previous node's Y (no counter, no fence), a relaxed arrival counter when nobody in the chain writes X, a release /
//     acquire counter otherwise; then every CTA quantises the new X into its own shared memory (quant_chunk256: bit for bit
//     quantize_row_q8_K / q8_0 of the reference);
//   * shared-memory staging frees the lane <-> data mapping from coalescing: Q4_K is processed ONE 256-BLOCK PER LANE (the 144 B
//     block stride is bank-conflict free), so d/dmin, the 12 scale bytes and the min correction are decoded once per 256
//     weights instead of once per 32, and the min term is 4 dp2a against per-32 activation sums; Q6_K is one block per lane
//     as well (its QL / QH planes land through swizzling 2-D tensor copies so that the lanes hit different banks, unit_q6k) and
//     rebuilds the 6-bit values without shifts (nibble and 2-bit fields are masked in place and the power-of-two factor is
//     taken out of the int32 sum exactly); the other formats keep 8 lanes per block. The integer arithmetic per block is the CPU's, value for value; only the fp32
//     summation order differs (fixed, so results are deterministic).
// Work split: per node the rows of the group's matrices are cut by BYTES into one contiguous share per VC (host, at create
// time) and each share into items that fit a ring stage; an item is 1 (Q4_K) to 4 (Q6_K planes) bulk copies, or for Q6_K
// rows of a multiple of 8 units a few tensor copies (QL, QH) + 2 bulk copies (scales, d).
// Algorithmic bytes per launch = sum over all nodes of m * row_size(type, k); HBM bound.
#include "gemv_frag.cuh"
#include "actquant.cuh"
#include "sm100.cuh"

#include <cuda.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

namespace {
using namespace sm100;

constexpr int CH_GROUPS = 2;                 // virtual consumers per CTA
constexpr int CH_GWARPS = 8;                 // warps per group
constexpr int CH_GTHREADS = CH_GWARPS * 32;
constexpr int CH_CONSUMERS = CH_GROUPS * CH_GTHREADS;      // 16 consumer warps ...
constexpr int CH_THREADS = CH_CONSUMERS + 32;               // ... + one producer warp (one thread issues every bulk copy)
constexpr int CH_HDR = 512;                                 // mbarriers (full, empty) + per-stage item descriptors
constexpr int CH_NS = 3;                     // ring stages per group
constexpr int XU_BYTES = 336;                // one 256-element unit of quantised X in shared memory (bank-conflict-free stride)
constexpr int XU_BS16 = 256, XU_BS32 = 288, XU_D = 304;
constexpr int CH_SMEM_TOTAL = 227 * 1024;
constexpr int CH_MAX_COLS = 4;
constexpr int CH_T_Q6K_SW = 0x600;        // item type: Q6_K whose QL / QH planes arrive through swizzling tensor copies (one block per lane)

struct ChainMatD {
    const uint8_t * w;        // plane layout (canonical for Q4_K)
    float * y; long long ldy;
    int m, type, upr;         // rows, weight type, 256-element units per row
    int pad;
};
// how a node learns that its X is complete (decided at create time from the pointers):
enum { NB_NONE = 0,     // first node: stream order
       NB_FENCED = 1,   // X overlaps something an earlier node wrote: arrival counter with release / acquire fences
       NB_RELAXED = 2,  // X is written by nobody in the chain: the node is merely kept in stream order (counter, no fences)
       NB_TAGGED = 3,   // X is exactly the Y of a matrix of the previous node: the values arrive as 64-bit {epoch, f32} words,
                        // each one validates itself, no counter and no fence
       NB_RELAXED_LOCAL = 4 }; // row split only: as NB_RELAXED, but only THIS rank's CTAs are waited for (no data flows through the boundary and the
                        // boundary before it was a tagged one, i.e. every rank has already passed the node before the previous one: see chain_plan)
struct ChainNodeD {
    const float * x; long long ldx;
    int k, act_block, flavour;
    int mode;                             // NB_*
    const unsigned long long * xtag;      // NB_TAGGED: where the previous node publishes this X
    unsigned want;                        // counter value that means "every earlier signalling node is finished everywhere"
    int signal;                           // what the NEXT node needs from this one: 0 nothing, 1 relaxed arrival, 2 release arrival, 3 relaxed arrival on the rank-local counter
};
// one ring stage worth of rows of one matrix, self-contained (the producer thread reads nothing else)
struct ChainItemD {
    const uint8_t * src[4];   // plane pieces (nullptr / 0 bytes: unused)
    uint32_t bytes[4];
    float * y; long long ldy; // y already points at the item's first row
    int type, nrows;
    int segs, S;              // how the warps share the item (plan_item)
    unsigned long long * tag; // tagged copy of the item's rows for the next node (nullptr: nobody in the chain reads them next)
    long long pad;
};

struct ChainArgs {
    const ChainNodeD * nodes;
    const ChainItemD * items;
    const int * node_off;     // [vc][n_nodes + 1] -> index into items (items are VC-major, node-ordered)
    unsigned long long * counter;   // arrivals of the counter barriers, never reset: targets are relative to the launch number
    unsigned long long * done;      // row split: CTAs of all ranks that have finished the whole chain (never reset)
    unsigned long long * launches;  // launches of this chain so far (local): base of the counter targets and of the tags
    unsigned long long per_launch;  // arrivals one launch adds to `counter` (signalling nodes x CTAs x ranks)
    unsigned long long * counter_local;   // row split: arrivals of the rank-local barriers (only this rank adds)
    unsigned long long per_launch_local;
    int world, rank;                // row split over `world` GPUs (1: off): every Y row, tagged word and arrival goes to every rank,
    long long delta[B200_MAX_PEERS];//   local address + delta[g] = the same location of rank g's symmetric arena
    int n_nodes;
    int stage_bytes;          // bytes per ring stage
    int max_units;            // stride of the partial-sum buffers
    int x_bytes;              // shared memory reserved for the quantised X
    long long * timeline;     // B200_CHAIN_TIMELINE: clock64 samples per (CTA, node), see the kernel
    long long tag_stride;     // elements between the columns of a tagged vector
    int inflight;             // copies per group that may be in flight (issued, not landed): bounds the queueing delay the stream adds to the node barrier
    int debug;                // B200_CHAIN_DEBUG (timing experiments, results wrong): 1 no node wait, 2 no dot products, 4 no X quantiser, 8 no row sums
    int ns;                   // gemv_chainx_kernel: ring stages per CTA
    int lanes8;               // B200_CHAIN_LANES8=1: IQ4_XS / Q4_0 items on the older eight-lanes-per-block fragments (comparison runs)
};

// bytes per 256-element unit in plane p of the plane layout (repack.cu); 0 terminates
__host__ __device__ inline int plane_bytes(int type, int p) {
    switch (type) {
        case B200_TYPE_Q4_K:   return p == 0 ? 144 : 0;
        case B200_TYPE_Q5_K:   return p == 0 ? 176 : 0;
        case B200_TYPE_Q6_K:   return p == 0 ? 128 : p == 1 ? 64 : p == 2 ? 16 : p == 3 ? 2 : 0;
        case B200_TYPE_IQ4_XS: return p == 0 ? 128 : p == 1 ? 8 : 0;
        case B200_TYPE_Q4_0:   return p == 0 ? 128 : p == 1 ? 16 : 0;
        case B200_TYPE_Q8_0:   return p == 0 ? 256 : p == 1 ? 16 : 0;
    }
    return 0;
}
__host__ __device__ inline int unit_bytes(int type) { int s = 0; for (int p = 0; p < 4; ++p) s += plane_bytes(type, p); return s; }

__device__ __forceinline__ uint4 lds16(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ uint2 lds8(uint32_t a) { uint2 r; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a)); return r; }
__device__ __forceinline__ uint32_t lds4(uint32_t a) { uint32_t r; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a)); return r; }
__device__ __forceinline__ uint16_t lds2(uint32_t a) { uint16_t r; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(a)); return r; }
__device__ __forceinline__ int ldss1(uint32_t a) { int r; asm volatile("ld.shared.s8 %0, [%1];" : "=r"(r) : "r"(a)); return r; }
__device__ __forceinline__ float ldsf(uint32_t a) { float r; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(a)); return r; }
__device__ __forceinline__ void sts8(uint32_t a, uint32_t x, uint32_t y) { asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(a), "r"(x), "r"(y) : "memory"); }
__device__ __forceinline__ void sts2(uint32_t a, uint16_t x) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"(x) : "memory"); }
__device__ __forceinline__ void sts4(uint32_t a, uint32_t x) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(x) : "memory"); }
__device__ __forceinline__ void stsf(uint32_t a, float x) { asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(x) : "memory"); }

// X changes while the kernel runs (it is the Y of an earlier node): read it at L2, never through the non-coherent path
__device__ __forceinline__ float4 ldcg16(const float * p) {
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ float ldcg4(const float * p) { float r; asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory"); return r; }

// y and (when the next node reads these rows) the tagged copy; row split: into every rank's arena
struct YOut { float * y; long long ldy; unsigned long long * tag; long long tag_stride; unsigned epoch; int world; const long long * delta; };
__device__ __forceinline__ void put_y(const YOut & o, int c, int row, float v) {
    float * yp = o.y + c * o.ldy + row;
    const unsigned long long w = ((unsigned long long) o.epoch << 32) | (unsigned long long) __float_as_uint(v);
    if (o.world <= 1) {
        *yp = v;
        if (o.tag) asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" :: "l"(o.tag + c * o.tag_stride + row), "l"(w) : "memory");
    } else {
        for (int g = 0; g < o.world; ++g) {
            *reinterpret_cast<float *>(reinterpret_cast<char *>(yp) + o.delta[g]) = v;
            if (o.tag) asm volatile("st.relaxed.sys.global.b64 [%0], %1;" :: "l"(reinterpret_cast<char *>(o.tag + c * o.tag_stride + row) + o.delta[g]), "l"(w) : "memory");
        }
    }
}
__device__ __forceinline__ void group_bar(int grp) { asm volatile("bar.sync %0, %1;" :: "r"(grp + 1), "n"(CH_GTHREADS) : "memory"); }

__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p;
}
// 2-D tensor copy (box at (c0, c1)) global -> shared with an L2 policy; the tensor map lives in global memory (host-written)
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const void * tmap, int c0, int c1, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
                 :: "r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}

// position of 16-byte chunk `ch` (0..15) of a unit's int8 quants: the upper half is rotated by four chunks so that the eight
// lanes of a Q6_K block (chunks {0..3, 8..11} or {4..7, 12..15}) hit eight different bank groups
__host__ __device__ constexpr int xpos(int ch) { return ch < 8 ? ch : 8 + ((ch + 4) & 7); }

__device__ __forceinline__ int dot16_us(const uint4 & a, const uint4 & b) {
    int s = dp4a_us(a.x, b.x, 0); s = dp4a_us(a.y, b.y, s); s = dp4a_us(a.z, b.z, s); s = dp4a_us(a.w, b.w, s); return s;
}

// ---- activation quantiser into the shared-memory unit layout ---------------------------------------------------------
//   unit (col, u): [qs 256 B, chunks placed by xpos][bs16: 16 x i16][bs32: 8 x i16][d: 8 x f32 (Q8_K uses d[0])]
template <int NCOLS, int QU, int NW>
__device__ __forceinline__ void chain_quantise_x_t(const ChainNodeD & nd, uint32_t xs, int warp, int lane, long long tag_stride, unsigned epoch, bool sys) {
    const int upr = nd.k >> 8;
    const int total = NCOLS * upr;
    constexpr unsigned FULL = 0xffffffffu;
    for (int c0 = warp; c0 < total; c0 += NW * QU) {
        float v[QU][8];
#pragma unroll
        for (int q = 0; q < QU; ++q) {
            const int ci = c0 + q * NW;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[q][j] = 0.0f;
            if (ci < total) {
                const int col = NCOLS == 1 ? 0 : ci / upr, u = ci - col * upr;
                const float * xr = nd.x + (long long) col * nd.ldx + u * 256 + lane * 8;
                if (nd.mode == NB_TAGGED) {
                    // the previous node publishes this X as {epoch, value} words: spin until the lane's eight carry this launch's tag
                    const unsigned long long * tp = nd.xtag + (long long) col * tag_stride + u * 256 + lane * 8;
                    unsigned long long t0 = 0;
                    for (unsigned spin = 0;; ++spin) {
                        unsigned long long w[8];
#pragma unroll
                        for (int j = 0; j < 8; j += 2) {
                            if (sys) asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];" : "=l"(w[j]), "=l"(w[j + 1]) : "l"(tp + j) : "memory");
                            else     asm volatile("ld.relaxed.gpu.global.v2.b64 {%0, %1}, [%2];" : "=l"(w[j]), "=l"(w[j + 1]) : "l"(tp + j) : "memory");
                        }
                        bool ok = true;
#pragma unroll
                        for (int j = 0; j < 8; ++j) { ok = ok && (unsigned) (w[j] >> 32) == epoch; v[q][j] = __uint_as_float((unsigned) w[j]); }
                        if (__all_sync(0xffffffffu, ok)) break;
                        if ((spin & 1023u) == 1023u) {
                            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                            if (t0 == 0) t0 = t; else if (t - t0 > 4000000000ull) __trap();
                        }
                    }
                } else if ((((uintptr_t) xr) & 15) == 0) {
                    const float4 lo = ldcg16(xr), hi = ldcg16(xr + 4);
                    v[q][0] = lo.x; v[q][1] = lo.y; v[q][2] = lo.z; v[q][3] = lo.w; v[q][4] = hi.x; v[q][5] = hi.y; v[q][6] = hi.z; v[q][7] = hi.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[q][j] = ldcg4(xr + j);
                }
            }
        }
        int qq[QU][8]; float dval[QU];
        if (nd.act_block == 256) {
            // quantize_row_q8_K_ref (ggml-quants.c:2555-2591), the stages of the QU chunks interleaved by hand: the node boundary
            // waits for this. max |x| of the chunk by ONE redux on the bit patterns (order preserving for non-negative floats),
            // then the FIRST element that has it (strict '>' in the reference loop) keeps its sign
            uint32_t mb[QU]; float maxv[QU];
#pragma unroll
            for (int q = 0; q < QU; ++q) {
                float am = 0.0f;
#pragma unroll
                for (int j = 0; j < 8; ++j) am = fmaxf(am, fabsf(v[q][j]));
                mb[q] = __float_as_uint(am);
            }
#pragma unroll
            for (int q = 0; q < QU; ++q) mb[q] = __reduce_max_sync(FULL, mb[q]);
#pragma unroll
            for (int q = 0; q < QU; ++q) {
                float cand = 0.0f; bool have = false;
#pragma unroll
                for (int j = 7; j >= 0; --j) if (__float_as_uint(fabsf(v[q][j])) == mb[q]) { cand = v[q][j]; have = true; }
                const unsigned bal = __ballot_sync(FULL, have);
                maxv[q] = __shfl_sync(FULL, cand, (__ffs((int) bal) - 1) & 31);
            }
#pragma unroll
            for (int q = 0; q < QU; ++q) {
                if (mb[q] == 0u) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) qq[q][j] = 0;
                    dval[q] = 0.0f;
                } else {
                    const float iscale = __fdiv_rn(-127.0f, maxv[q]);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { const int t = __float2int_rn(__fmul_rn(iscale, v[q][j])); qq[q][j] = t > 127 ? 127 : t; }
                    dval[q] = __fdiv_rn(1.0f, iscale);
                }
            }
        } else {
#pragma unroll
            for (int q = 0; q < QU; ++q) {
                float sval;
                if (nd.flavour == B200_Q8_X86) quant_chunk256<B200_TYPE_Q8_0, B200_Q8_X86>(v[q], lane, qq[q], dval[q], sval);
                else                           quant_chunk256<B200_TYPE_Q8_0, B200_Q8_REF>(v[q], lane, qq[q], dval[q], sval);
            }
        }
#pragma unroll
        for (int q = 0; q < QU; ++q) {
            const int ci = c0 + q * NW;
            int s8 = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) s8 += (int) (int8_t) qq[q][j];
            const int s16 = s8 + __shfl_xor_sync(FULL, s8, 1);
            const int s32 = s16 + __shfl_xor_sync(FULL, s16, 2);
            if (ci < total) {
                const uint32_t w0 = (qq[q][0] & 0xff) | ((qq[q][1] & 0xff) << 8) | ((qq[q][2] & 0xff) << 16) | ((uint32_t) (qq[q][3] & 0xff) << 24);
                const uint32_t w1 = (qq[q][4] & 0xff) | ((qq[q][5] & 0xff) << 8) | ((qq[q][6] & 0xff) << 16) | ((uint32_t) (qq[q][7] & 0xff) << 24);
                const uint32_t ub = xs + (uint32_t) ci * XU_BYTES;
                const int ch = lane >> 1;
                const int pc = ch < 8 ? ch : 8 + ((ch + 4) & 7);
                sts8(ub + pc * 16 + (lane & 1) * 8, w0, w1);
                if ((lane & 1) == 0) sts2(ub + XU_BS16 + (lane >> 1) * 2, (uint16_t) (int16_t) s16);
                if ((lane & 3) == 0) sts2(ub + XU_BS32 + (lane >> 2) * 2, (uint16_t) (int16_t) s32);
                if (nd.act_block == 256) { if (lane == 0) stsf(ub + XU_D, dval[q]); }
                else if ((lane & 3) == 0) stsf(ub + XU_D + (lane >> 2) * 4, dval[q]);
            }
        }
    }
}

// chunks per warp that are quantised together: as many as the node needs, no dead slots (every CTA runs this once per node)
template <int NCOLS, int NW = CH_CONSUMERS / 32>
__device__ __forceinline__ void chain_quantise_x(const ChainNodeD & nd, uint32_t xs, int warp, int lane, long long tag_stride, unsigned epoch, bool sys) {
    const int total = NCOLS * (nd.k >> 8);
    if (total <= NW)          chain_quantise_x_t<NCOLS, 1, NW>(nd, xs, warp, lane, tag_stride, epoch, sys);
    else if (total <= 2 * NW) chain_quantise_x_t<NCOLS, 2, NW>(nd, xs, warp, lane, tag_stride, epoch, sys);
    else                      chain_quantise_x_t<NCOLS, 4, NW>(nd, xs, warp, lane, tag_stride, epoch, sys);
}

// ---- Q4_K: one 256-block per lane --------------------------------------------------------------------------------------
// block bytes (ggml-common.h:295-306): d, dmin (fp16), scales[12], qs[128]; arithmetic of ggml_vec_dot_q4_K_q8_K
// (ggml-cpu/quants.c:550-625) with get_scale_min_k4 (ggml-quants.c:703-711) done for the eight sub-blocks at once
template <int NCOLS>
__device__ __forceinline__ void unit_q4k(uint32_t wb, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint4 hdr = lds16(wb);
    const uint32_t sc03 = hdr.y & 0x3f3f3f3fu, mn03 = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc47 = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t mn47 = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
    int isum[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) isum[c] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint4 qa = lds16(wb + 16 + 32 * j), qb = lds16(wb + 32 + 32 * j);
        const uint4 la = and4(qa, 0x0f0f0f0fu), lb = and4(qb, 0x0f0f0f0fu);
        // high nibbles stay in place (16 x nibble as an unsigned byte); the factor leaves the int32 sum with an exact shift
        const uint4 ha = and4(qa, 0xf0f0f0f0u), hb = and4(qb, 0xf0f0f0f0u);
        const uint32_t scw = j < 2 ? sc03 : sc47;
        const int s0 = (int) ((scw >> (16 * (j & 1))) & 0xffu), s1 = (int) ((scw >> (16 * (j & 1) + 8)) & 0xffu);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
            const uint4 x0 = lds16(xc + xpos(4 * j) * 16), x1 = lds16(xc + xpos(4 * j + 1) * 16);
            const uint4 x2 = lds16(xc + xpos(4 * j + 2) * 16), x3 = lds16(xc + xpos(4 * j + 3) * 16);
            const int dlo = dot16(la, x0) + dot16(lb, x1);
            const int dhi = dot16_us(ha, x2) + dot16_us(hb, x3);
            isum[c] += s0 * dlo + s1 * (dhi >> 4);
        }
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
        const uint4 bs = lds16(xc + XU_BS32);              // 8 x i16: sums of 32 quants
        int msum = __dp2a_lo((int) bs.x, (int) mn03, 0);
        msum = __dp2a_hi((int) bs.y, (int) mn03, msum);
        msum = __dp2a_lo((int) bs.z, (int) mn47, msum);
        msum = __dp2a_hi((int) bs.w, (int) mn47, msum);
        const float dx = ldsf(xc + XU_D);
        acc[c] += dx * (dm.x * (float) isum[c] - dm.y * (float) msum);
    }
}

// ---- Q5_K: one 256-block per lane; the Q4_K scheme over 5-bit values (fifth bit: bit j of qh[l] for sub-block j) ----------
// block bytes (ggml-common.h:319-333): d, dmin, scales[12], qh[32], qs[128]; arithmetic of ggml_vec_dot_q5_K_q8_K
template <int NCOLS>
__device__ __forceinline__ void unit_q5k(uint32_t wb, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint4 hdr = lds16(wb), h0 = lds16(wb + 16), h1 = lds16(wb + 32);
    const uint32_t sc03 = hdr.y & 0x3f3f3f3fu, mn03 = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc47 = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t mn47 = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
    int isum[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) isum[c] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint4 qa = lds16(wb + 48 + 32 * j), qb = lds16(wb + 64 + 32 * j);
        const uint4 fa = and4(shr4(h0, 2 * j), 0x03030303u), fb = and4(shr4(h1, 2 * j), 0x03030303u);   // bit 0: sub-block 2j, bit 1: 2j + 1
        const uint4 la = make_uint4((qa.x & 0x0f0f0f0fu) | ((fa.x & 0x01010101u) << 4), (qa.y & 0x0f0f0f0fu) | ((fa.y & 0x01010101u) << 4),
                                    (qa.z & 0x0f0f0f0fu) | ((fa.z & 0x01010101u) << 4), (qa.w & 0x0f0f0f0fu) | ((fa.w & 0x01010101u) << 4));
        const uint4 lb = make_uint4((qb.x & 0x0f0f0f0fu) | ((fb.x & 0x01010101u) << 4), (qb.y & 0x0f0f0f0fu) | ((fb.y & 0x01010101u) << 4),
                                    (qb.z & 0x0f0f0f0fu) | ((fb.z & 0x01010101u) << 4), (qb.w & 0x0f0f0f0fu) | ((fb.w & 0x01010101u) << 4));
        const uint4 ha = make_uint4(((qa.x >> 4) & 0x0f0f0f0fu) | ((fa.x & 0x02020202u) << 3), ((qa.y >> 4) & 0x0f0f0f0fu) | ((fa.y & 0x02020202u) << 3),
                                    ((qa.z >> 4) & 0x0f0f0f0fu) | ((fa.z & 0x02020202u) << 3), ((qa.w >> 4) & 0x0f0f0f0fu) | ((fa.w & 0x02020202u) << 3));
        const uint4 hb = make_uint4(((qb.x >> 4) & 0x0f0f0f0fu) | ((fb.x & 0x02020202u) << 3), ((qb.y >> 4) & 0x0f0f0f0fu) | ((fb.y & 0x02020202u) << 3),
                                    ((qb.z >> 4) & 0x0f0f0f0fu) | ((fb.z & 0x02020202u) << 3), ((qb.w >> 4) & 0x0f0f0f0fu) | ((fb.w & 0x02020202u) << 3));
        const uint32_t scw = j < 2 ? sc03 : sc47;
        const int s0 = (int) ((scw >> (16 * (j & 1))) & 0xffu), s1 = (int) ((scw >> (16 * (j & 1) + 8)) & 0xffu);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
            const uint4 x0 = lds16(xc + xpos(4 * j) * 16), x1 = lds16(xc + xpos(4 * j + 1) * 16);
            const uint4 x2 = lds16(xc + xpos(4 * j + 2) * 16), x3 = lds16(xc + xpos(4 * j + 3) * 16);
            isum[c] += s0 * (dot16(la, x0) + dot16(lb, x1)) + s1 * (dot16(ha, x2) + dot16(hb, x3));
        }
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
        const uint4 bs = lds16(xc + XU_BS32);
        int msum = __dp2a_lo((int) bs.x, (int) mn03, 0);
        msum = __dp2a_hi((int) bs.y, (int) mn03, msum);
        msum = __dp2a_lo((int) bs.z, (int) mn47, msum);
        msum = __dp2a_hi((int) bs.w, (int) mn47, msum);
        const float dx = ldsf(xc + XU_D);
        acc[c] += dx * (dm.x * (float) isum[c] - dm.y * (float) msum);
    }
}

// ---- Q6_K: eight lanes per 256-block, planes QL 128 | QH 64 | SC 16 | D 2 ---------------------------------------------
// arithmetic of ggml_vec_dot_q6_K_q8_K (ggml-cpu/quants.c:705-760); lane g owns elements e0 + [0,16) and e0 + 64 + [0,16),
// e0 = 128 (g >> 2) + 32 ((g >> 1) & 1) + 16 (g & 1) (dequantize_row_q6_K, ggml-quants.c:1762-1792)
template <int NCOLS>
__device__ __forceinline__ void frag_q6k(uint32_t stage, int nunits, int u, int g, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint4 ql = lds16(stage + (uint32_t) (u * 128 + 16 * g));
    const uint4 qh = lds16(stage + (uint32_t) (nunits * 128 + u * 64 + 32 * (g >> 2) + 16 * (g & 1)));
    const int is = 8 * (g >> 2) + (g & 3);
    const uint32_t scb = stage + (uint32_t) (nunits * 192 + u * 16 + is);
    const int sc0 = ldss1(scb), sc1 = ldss1(scb + 4);
    const float dw = half_bits_to_float(lds2(stage + (uint32_t) (nunits * 208 + u * 2)));
    const int sh = g & 2;
    const uint32_t ma = 0x03030303u << sh, mb = ma << 4;
    const uint4 lo = and4(ql, 0x0f0f0f0fu), hi = and4(ql, 0xf0f0f0f0u);
    const uint4 ha = and4(qh, ma), hb = and4(qh, mb);
    const int ch0 = is, ch1 = is + 4;
    const int p0 = ch0 < 8 ? ch0 : 8 + ((ch0 + 4) & 7), p1 = ch1 < 8 ? ch1 : 8 + ((ch1 + 4) & 7);
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
        const uint4 x0 = lds16(xc + p0 * 16), x1 = lds16(xc + p1 * 16);
        const int bs0 = (int) (int16_t) lds2(xc + XU_BS16 + ch0 * 2), bs1 = (int) (int16_t) lds2(xc + XU_BS16 + ch1 * 2);
        const float dx = ldsf(xc + XU_D);
        // q6 = nibble | two bits << 4: the two-bit fields are masked in place (value h << sh resp. h << (sh + 4)) and the
        // power of two is taken out of the int32 sums exactly
        const int s0 = dot16(lo, x0) + (dot16(ha, x0) << (4 - sh)) - 32 * bs0;
        const int s1 = (dot16_us(hi, x1) >> 4) + (dot16_us(hb, x1) >> sh) - 32 * bs1;
        acc[c] += (dx * dw) * (float) (sc0 * s0 + sc1 * s1);
    }
}

// ---- Q6_K, one 256-block per lane (CH_T_Q6K_SW items) --------------------------------------------------------------------
// The producer lands QL and QH with 2-D tensor copies (SWIZZLE_128B / SWIZZLE_64B: 16-byte chunk c of unit u sits at chunk
// c ^ (u & 7) resp. c ^ ((u >> 1) & 3)), so 32 lanes reading the same chunk of 32 consecutive units hit 32 different bank
// groups. Stage: [QL 128 B x units][pad to 512][QH 64 B x units][SC 16 B x units][D 2 B x units]. Same integer arithmetic
// as frag_q6k; QH, the 16 scales, d and the activation scale are loaded once per block instead of once per 32 weights, and
// the -32 offset of the 6-bit values is one dp2a chain over the per-16 activation sums (sum_j sc_j * bs16_j).
__host__ __device__ inline int q6sw_qh0(int nunits) { return (nunits * 128 + 511) & ~511; }
__device__ __forceinline__ int sext8(uint32_t w, int byte) { return (int) (w << (24 - 8 * byte)) >> 24; }
template <int NCOLS>
__device__ __forceinline__ void unit_q6k(uint32_t stage, int nunits, int u, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint32_t qh0 = stage + (uint32_t) q6sw_qh0(nunits);
    const uint32_t qlp = (stage + (uint32_t) (u * 128)) ^ (uint32_t) ((u & 7) << 4);          // ^ (c << 4): chunk c
    const uint32_t qhp = (qh0 + (uint32_t) (u * 64)) ^ (uint32_t) (((u >> 1) & 3) << 4);
    const uint4 scw = lds16(qh0 + (uint32_t) (nunits * 64 + u * 16));
    const uint32_t sc[4] = {scw.x, scw.y, scw.z, scw.w};
    const float dw = half_bits_to_float(lds2(qh0 + (uint32_t) (nunits * 80 + u * 2)));
    int isum[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
        const uint4 b0 = lds16(xc + XU_BS16), b1 = lds16(xc + XU_BS16 + 16);            // 16 x i16: sums of 16 quants
        int m = __dp2a_lo((int) b0.x, (int) sc[0], 0);
        m = __dp2a_hi((int) b0.y, (int) sc[0], m);
        m = __dp2a_lo((int) b0.z, (int) sc[1], m);
        m = __dp2a_hi((int) b0.w, (int) sc[1], m);
        m = __dp2a_lo((int) b1.x, (int) sc[2], m);
        m = __dp2a_hi((int) b1.y, (int) sc[2], m);
        m = __dp2a_lo((int) b1.z, (int) sc[3], m);
        m = __dp2a_hi((int) b1.w, (int) sc[3], m);
        isum[c] = -32 * m;
    }
#pragma unroll
    for (int n = 0; n < 2; ++n) {
        const uint4 qhA = lds16(qhp ^ (uint32_t) ((2 * n) << 4)), qhB = lds16(qhp ^ (uint32_t) ((2 * n + 1) << 4));
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint4 ql = lds16(qlp ^ (uint32_t) ((4 * n + t) << 4));
            const uint4 qh = (t & 1) ? qhB : qhA;
            constexpr uint32_t M3 = 0x03030303u;
            const int sh = t & 2;
            const uint4 lo = and4(ql, 0x0f0f0f0fu), hi = and4(ql, 0xf0f0f0f0u);
            const uint4 ha = and4(qh, M3 << sh), hb = and4(qh, M3 << (sh + 4));
            const int is = 8 * n + t;
            const int sc0 = sext8(sc[is >> 2], is & 3), sc1 = sext8(sc[(is + 4) >> 2], is & 3);
            const int p0 = xpos(is), p1 = xpos(is + 4);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) {
                const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
                const uint4 x0 = lds16(xc + p0 * 16), x1 = lds16(xc + p1 * 16);
                const int s0 = dot16(lo, x0) + (dot16(ha, x0) << (4 - sh));
                const int s1 = (dot16_us(hi, x1) >> 4) + (dot16_us(hb, x1) >> sh);
                isum[c] += sc0 * s0 + sc1 * s1;
            }
        }
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const float dx = ldsf(xb + (uint32_t) (c * x_col_stride) + XU_D);
        acc[c] += (dx * dw) * (float) isum[c];
    }
}

// ---- IQ4_XS / Q4_0, one 256-element unit per lane ----------------------------------------------------------------------------
// The QS plane is 128 B per unit, so 32 lanes reading the same 16-byte chunk of 32 consecutive units would all hit the same four banks.
// Lane b therefore walks its eight 32-element sub-blocks in the order t ^ rho(b), rho(b) = (b & 7) ^ ((b >> 1) & 2): within every group
// of eight lanes the weight chunks AND the two activation chunks (placed by xpos, unit stride 336 B) fall into eight different 16-byte
// bank groups (exhaustive search over the permutations of 0..7; an additive rotation cannot satisfy both). Stage layout = the item's
// planes back to back, as the bulk copies land them.
__device__ __forceinline__ int lane_rho(int b) { return (b & 7) ^ ((b >> 1) & 2); }
__device__ __forceinline__ int xpos_rt(int ch) { return ch < 8 ? ch : 8 + ((ch + 4) & 7); }
// arithmetic of ggml_vec_dot_iq4_xs_q8_K (ggml-cpu/quants.c:1137-1181): per sub-block ls * sum(kvalues[nib] * q8), one d * d_x per super-block
template <int NCOLS>
__device__ __forceinline__ void unit_iq4xs(uint32_t stage, int nunits, int u, int b, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint2 hdr = lds8(stage + (uint32_t) (nunits * 128 + u * 8));          // d, scales_h | scales_l[4]
    const uint32_t sh = hdr.x >> 16;
    const uint32_t qs = stage + (uint32_t) (u * 128);
    const int rho = lane_rho(b);
    int isum[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) isum[c] = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int ib = t ^ rho;
        const uint4 q = lds16(qs + (uint32_t) (ib << 4));
        const int ls = (int) (((hdr.y >> (4 * ib)) & 0xfu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;
        const uint32_t p0 = (uint32_t) xpos_rt(2 * ib) << 4, p1 = (uint32_t) xpos_rt(2 * ib + 1) << 4;
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
            const uint4 x0 = lds16(xc + p0), x1 = lds16(xc + p1);
            int d = iq4_word_dot(q.x, x0.x, x1.x, 0);
            d = iq4_word_dot(q.y, x0.y, x1.y, d);
            d = iq4_word_dot(q.z, x0.z, x1.z, d);
            d = iq4_word_dot(q.w, x0.w, x1.w, d);
            isum[c] += ls * d;
        }
    }
    const float dw = half_bits_to_float((uint16_t) (hdr.x & 0xffffu));
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) acc[c] += (ldsf(xb + (uint32_t) (c * x_col_stride) + XU_D) * dw) * (float) isum[c];
}
// arithmetic of ggml_vec_dot_q4_0_q8_0 (ggml-cpu/quants.c:115-150): per 32-block (sum(nib * q8) - 8 * sum(q8)) * d_w * d_x
template <int NCOLS>
__device__ __forceinline__ void unit_q40(uint32_t stage, int nunits, int u, int b, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint32_t qs = stage + (uint32_t) (u * 128), dp = stage + (uint32_t) (nunits * 128 + u * 16);
    const int rho = lane_rho(b);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int ib = t ^ rho;
        const uint4 q = lds16(qs + (uint32_t) (ib << 4));
        const float dw = half_bits_to_float(lds2(dp + (uint32_t) (ib << 1)));
        const uint4 lo = and4(q, 0x0f0f0f0fu), hi = and4(shr4(q, 4), 0x0f0f0f0fu);
        const uint32_t p0 = (uint32_t) xpos_rt(2 * ib) << 4, p1 = (uint32_t) xpos_rt(2 * ib + 1) << 4;
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
            const uint4 x0 = lds16(xc + p0), x1 = lds16(xc + p1);
            const int bs = (int) (int16_t) lds2(xc + XU_BS32 + (uint32_t) (ib << 1));
            const int isum = dot16(lo, x0) + dot16(hi, x1) - 8 * bs;
            acc[c] += (ldsf(xc + XU_D + (uint32_t) (ib << 2)) * dw) * (float) isum;
        }
    }
}

// ---- Q8_0, one 256-element unit per lane: QS 256 B | D 16 B per unit. Lane b takes its sixteen 16-byte chunks in the order t ^ rho8(b): the weight chunks of
// eight neighbouring lanes always fall into eight different bank groups, the activation chunks collide two ways in half of the steps (no order does
// better with the xpos placement: exhaustive search). Chunks 2 i and 2 i + 1 are one 32-block: ggml_vec_dot_q8_0_q8_0 (ggml-cpu/quants.c:305-335).
__device__ __forceinline__ int lane_rho8(int b) { return (int) ((0x54672318u >> (4 * (b & 7))) & 0xfu); }
template <int NCOLS>
__device__ __forceinline__ void unit_q80(uint32_t stage, int nunits, int u, int b, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    const uint32_t qs = stage + (uint32_t) (u * 256), dp = stage + (uint32_t) (nunits * 256 + u * 16);
    const int rho = lane_rho8(b);
#pragma unroll
    for (int tp = 0; tp < 8; ++tp) {
        const int j0 = (2 * tp) ^ rho, j1 = j0 ^ 1, blk = j0 >> 1;
        const uint4 w0 = lds16(qs + (uint32_t) (j0 << 4)), w1 = lds16(qs + (uint32_t) (j1 << 4));
        const float dw = half_bits_to_float(lds2(dp + (uint32_t) (blk << 1)));
        const uint32_t p0 = (uint32_t) xpos_rt(j0) << 4, p1 = (uint32_t) xpos_rt(j1) << 4;
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
            const uint4 x0 = lds16(xc + p0), x1 = lds16(xc + p1);
            acc[c] += (ldsf(xc + XU_D + (uint32_t) (blk << 2)) * dw) * (float) (dot16(w0, x0) + dot16(w1, x1));
        }
    }
}

// ---- shapes the paths above do not take: eight lanes per 256 elements, fragments of gemv_frag.cuh filled from shared memory ----
template <int TYPE> __device__ __forceinline__ void wfrag_lds(WFrag<TYPE> & f, uint32_t stage, int nunits, int u, int g);
template <> __device__ __forceinline__ void wfrag_lds<B200_TYPE_IQ4_XS>(WFrag<B200_TYPE_IQ4_XS> & f, uint32_t stage, int nunits, int u, int g) {
    f.q = lds16(stage + (uint32_t) (u * 128 + 16 * g));
    f.hdr = lds8(stage + (uint32_t) (nunits * 128 + u * 8));
}
template <> __device__ __forceinline__ void wfrag_lds<B200_TYPE_Q4_0>(WFrag<B200_TYPE_Q4_0> & f, uint32_t stage, int nunits, int u, int g) {
    f.q = lds16(stage + (uint32_t) (u * 128 + 16 * g));
    f.d = lds2(stage + (uint32_t) (nunits * 128 + u * 16 + 2 * g));
}
template <> __device__ __forceinline__ void wfrag_lds<B200_TYPE_Q8_0>(WFrag<B200_TYPE_Q8_0> & f, uint32_t stage, int nunits, int u, int g) {
    f.q0 = lds16(stage + (uint32_t) (u * 256 + 32 * g));
    f.q1 = lds16(stage + (uint32_t) (u * 256 + 32 * g + 16));
    f.d = lds2(stage + (uint32_t) (nunits * 256 + u * 16 + 2 * g));
}
template <int TYPE, int NCOLS>
__device__ __forceinline__ void frag_legacy(uint32_t stage, int nunits, int u, int g, uint32_t xb, int x_col_stride, int act_block, float (&acc)[NCOLS]) {
    WFrag<TYPE> f;
    wfrag_lds<TYPE>(f, stage, nunits, u, g);
    XFrag x[NCOLS];
    const int ch0 = 2 * g, ch1 = 2 * g + 1;                // elements 32 g + [0,16) and + [16,32)
    const int p0 = ch0 < 8 ? ch0 : 8 + ((ch0 + 4) & 7), p1 = ch1 < 8 ? ch1 : 8 + ((ch1 + 4) & 7);
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const uint32_t xc = xb + (uint32_t) (c * x_col_stride);
        x[c].x0 = lds16(xc + p0 * 16); x[c].x1 = lds16(xc + p1 * 16);
        x[c].bs0 = (int) (int16_t) lds2(xc + XU_BS16 + ch0 * 2); x[c].bs1 = (int) (int16_t) lds2(xc + XU_BS16 + ch1 * 2);
        x[c].dx = ldsf(xc + XU_D + (act_block == 256 ? 0 : 4 * g));
    }
    f.template dot<NCOLS>(x, acc, g);
}

// How the eight warps of a group share one item (nrows rows of upr units each). A TASK is what one warp reduces on its own:
//   one block per lane (Q4_K): upr <= 32 -> a task is 32 / upr whole rows side by side (segmented shuffle reduction, Y stored
//     straight from the warp); upr > 32 -> a task is 32 consecutive units of one row;
//   eight lanes per block (the other formats): a task is a run of `S` units of one row, four units per warp pass.
// Rows cut into several tasks meet in a small partial buffer and are added in a fixed order after the group barrier.
struct ItemPlan { int segs, S; };
inline ItemPlan plan_item(bool lane_per_unit, int nrows, int upr) {       // host, at create time
    ItemPlan p;
    if (lane_per_unit) { p.S = 32; p.segs = upr <= 32 ? 1 : (upr + 31) >> 5; return p; }
    int want = (2 * CH_GWARPS + nrows - 1) / nrows;            // aim at >= 2 tasks per warp
    const int most = (upr + 3) >> 2;
    if (want > most) want = most;
    if (want < 1) want = 1;
    p.S = (((upr + want - 1) / want) + 3) & ~3;
    p.segs = (upr + p.S - 1) / p.S;
    return p;
}

template <int TYPE, int NCOLS>
__device__ __forceinline__ void item_lanes8(uint32_t stage, int nrows, int upr, uint32_t xs, int x_col_stride, int act_block,
                                            const ItemPlan pl, uint32_t part, int part_stride, const YOut & out, int first, int lane) {
    const int nunits = nrows * upr;
    const int j = lane >> 3, g = lane & 7;
    int row = 0, seg = first;
    for (int t = first; t < nrows * pl.segs; t += CH_GWARPS, seg += CH_GWARPS) {
        while (seg >= pl.segs) { seg -= pl.segs; ++row; }
        const int u_end = min(upr, (seg + 1) * pl.S);
        float acc[NCOLS];
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[c] = 0.0f;
        for (int b = seg * pl.S + j; b < u_end; b += 4) {
            const uint32_t xb = xs + (uint32_t) (b * XU_BYTES);
            if constexpr (TYPE == B200_TYPE_Q6_K) frag_q6k<NCOLS>(stage, nunits, row * upr + b, g, xb, x_col_stride, acc);
            else                                  frag_legacy<TYPE, NCOLS>(stage, nunits, row * upr + b, g, xb, x_col_stride, act_block, acc);
        }
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const float v = warp_sum(acc[c]);
            if (lane == 0) {
                if (pl.segs == 1) put_y(out, c, row, v);
                else stsf(part + (uint32_t) (c * part_stride + t) * 4, v);
            }
        }
    }
}

constexpr int LU_Q4K = 0, LU_Q6K = 1, LU_Q5K = 2, LU_IQ4XS = 3, LU_Q40 = 4, LU_Q80 = 5;       // block-per-lane formats
template <int Q6, int NCOLS>
__device__ __forceinline__ void lane_unit(uint32_t stage, int nunits, int u, int b, uint32_t xb, int x_col_stride, float (&acc)[NCOLS]) {
    if constexpr (Q6 == LU_Q6K)        unit_q6k<NCOLS>(stage, nunits, u, xb, x_col_stride, acc);
    else if constexpr (Q6 == LU_Q5K)   unit_q5k<NCOLS>(stage + (uint32_t) u * 176, xb, x_col_stride, acc);
    else if constexpr (Q6 == LU_IQ4XS) unit_iq4xs<NCOLS>(stage, nunits, u, b, xb, x_col_stride, acc);
    else if constexpr (Q6 == LU_Q40)   unit_q40<NCOLS>(stage, nunits, u, b, xb, x_col_stride, acc);
    else if constexpr (Q6 == LU_Q80)   unit_q80<NCOLS>(stage, nunits, u, b, xb, x_col_stride, acc);
    else                               unit_q4k<NCOLS>(stage + (uint32_t) u * 144, xb, x_col_stride, acc);
}
template <int NCOLS, int Q6 = LU_Q4K>
__device__ __forceinline__ void item_q4k(uint32_t stage, int nrows, int upr, uint32_t xs, int x_col_stride,
                                         const ItemPlan pl, uint32_t part, int part_stride, const YOut & out, int first, int lane,
                                         int rp, int sub, int b) {
    if (upr <= 32) {                                           // rp rows side by side in one warp: lane = (sub, b)
        for (int t = first; t * rp < nrows; t += CH_GWARPS) {
            const int row = t * rp + sub;
            const bool live = sub < rp && row < nrows;
            float acc[NCOLS];
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) acc[c] = 0.0f;
            if (live) lane_unit<Q6, NCOLS>(stage, nrows * upr, row * upr + b, b, xs + (uint32_t) (b * XU_BYTES), x_col_stride, acc);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) {
                float v = acc[c];
                for (int o = 1; o < upr; o <<= 1) {            // segmented: lanes of one row are consecutive, fixed order
                    const float w = __shfl_down_sync(0xffffffffu, v, o);
                    if (b + o < upr) v += w;
                }
                if (live && b == 0) put_y(out, c, row, v);
            }
        }
    } else {
        int row = 0, seg = first;
        for (int t = first; t < nrows * pl.segs; t += CH_GWARPS, seg += CH_GWARPS) {
            while (seg >= pl.segs) { seg -= pl.segs; ++row; }
            const int b = seg * 32 + lane;
            float acc[NCOLS];
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) acc[c] = 0.0f;
            if (b < upr) lane_unit<Q6, NCOLS>(stage, nrows * upr, row * upr + b, b, xs + (uint32_t) (b * XU_BYTES), x_col_stride, acc);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) {
                const float v = warp_sum(acc[c]);
                if (lane == 0) stsf(part + (uint32_t) (c * part_stride + t) * 4, v);
            }
        }
    }
}

__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void cons_bar() { asm volatile("bar.sync 3, %0;" :: "n"(CH_CONSUMERS) : "memory"); }

// KS = which item kinds are compiled in: 0 Q4_K / Q6_K one block per lane (what the 8B Q4_K_M file needs), 1 + Q5_K (70B) and Q6_K rows that are not a
// multiple of 8 units, 2 + Q4_0 / IQ4_XS, 3 + Q8_0. Code a chain never runs still costs it: with every item kind in one kernel the 8B token runs at 770
// tokens/s, 785 with Q4_K / Q5_K / Q6_K, 813 with KS = 0. (The effect is ptxas scheduling, not size as such: compiling the row-split and timeline branches
// out as well gave 760.)
template <int NCOLS, int KS>
__global__ void __launch_bounds__(CH_THREADS, 1) gemv_chain_kernel(const ChainArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NN = a.n_nodes;

    // shared memory: [full / empty mbarriers, item descriptors: CH_HDR][X units][partials: groups x 2 x NCOLS x max_units f32]
    //                [ring: groups x NS x stage_bytes]
    const uint32_t base = smem_u32(smem_raw);
    auto full_bar  = [&](int g, int st) { return base + (uint32_t) ((g * CH_NS + st) * 8); };
    auto empty_bar = [&](int g, int st) { return base + 64u + (uint32_t) ((g * CH_NS + st) * 8); };
    auto desc_at   = [&](int g, int st) { return base + 128u + (uint32_t) ((g * CH_NS + st) * 48); };
    const uint32_t xs = base + CH_HDR;
    const uint32_t part_bytes = (uint32_t) (NCOLS * a.max_units) * 4;
    const uint32_t ring0 = (xs + (uint32_t) a.x_bytes + CH_GROUPS * 2 * part_bytes + 1023u) & ~1023u;     // swizzle patterns repeat every 1024 B

    if (tid == 0) {
        for (int i = 0; i < CH_GROUPS * CH_NS; ++i) { mbar_init(base + i * 8, 1); mbar_init(base + 64 + i * 8, CH_GWARPS); }
        fence_barrier_init();
    }
    __syncthreads();

    // ---- producer: ONE thread keeps both rings full; the item list of a VC is node-ordered, so the copies run ahead of the
    //      node the consumers are working on for as long as there are free stages
    if (warp == CH_CONSUMERS / 32) {
        if (lane != 0) return;
        const uint64_t policy = policy_evict_first();
        int nxt[CH_GROUPS], end[CH_GROUPS], seq[CH_GROUPS], landed[CH_GROUPS];
        ChainItemD item[CH_GROUPS];
#pragma unroll
        for (int g = 0; g < CH_GROUPS; ++g) {
            const int * off = a.node_off + (long long) (blockIdx.x * CH_GROUPS + g) * (NN + 1);
            nxt[g] = off[0]; end[g] = off[NN]; seq[g] = 0; landed[g] = 0;
            if (nxt[g] < end[g]) item[g] = a.items[nxt[g]];
        }
        while (true) {
            bool pending = false;
#pragma unroll
            for (int g = 0; g < CH_GROUPS; ++g) {
                if (nxt[g] >= end[g]) continue;
                pending = true;
                const int st = seq[g] % CH_NS;
                while (landed[g] < seq[g] && mbar_test(full_bar(g, landed[g] % CH_NS), (uint32_t) ((landed[g] / CH_NS) & 1))) ++landed[g];
                if (seq[g] - landed[g] >= a.inflight) continue;
                if (!mbar_test(empty_bar(g, st), (uint32_t) (((seq[g] / CH_NS) & 1) ^ 1))) continue;
                const uint32_t dst = ring0 + (uint32_t) (g * CH_NS + st) * (uint32_t) a.stage_bytes, bar = full_bar(g, st), dsc = desc_at(g, st);
                const unsigned long long yp = (unsigned long long) item[g].y, ld = (unsigned long long) item[g].ldy;
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dsc), "r"((uint32_t) yp), "r"((uint32_t) (yp >> 32)), "r"((uint32_t) ld), "r"((uint32_t) (ld >> 32)) : "memory");
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dsc + 16), "r"((uint32_t) item[g].type), "r"((uint32_t) item[g].nrows), "r"((uint32_t) item[g].segs), "r"((uint32_t) item[g].S) : "memory");
                { const unsigned long long tg = (unsigned long long) item[g].tag; sts8(dsc + 32, (uint32_t) tg, (uint32_t) (tg >> 32)); }
                const uint32_t total = item[g].bytes[0] + item[g].bytes[1] + item[g].bytes[2] + item[g].bytes[3];
                mbar_arrive_expect_tx(bar, total);                       // release: the descriptor is visible to whoever sees the phase flip
                if (item[g].type == CH_T_Q6K_SW) {
                    // tensor copies of R rows (box = R x upr units x 128 / 64 bytes), single rows for the tail; SC and D as bulk copies
                    const int nr = item[g].nrows, u0 = (int) (item[g].pad & 0xffffffffll), R = (int) (item[g].pad >> 32);
                    const int upr_i = (int) (item[g].bytes[0] >> 7) / nr;
                    const uint32_t qh0 = dst + (uint32_t) q6sw_qh0(nr * upr_i);
                    int r = 0;
                    for (; r + R <= nr; r += R) {
                        tma_load_2d_hint(dst + (uint32_t) (r * upr_i * 128), item[g].src[0], 0, u0 + r * upr_i, bar, policy);
                        tma_load_2d_hint(qh0 + (uint32_t) (r * upr_i * 64), item[g].src[1], 0, u0 + r * upr_i, bar, policy);
                    }
                    for (; r < nr; ++r) {
                        tma_load_2d_hint(dst + (uint32_t) (r * upr_i * 128), item[g].src[0] + 2 * sizeof(CUtensorMap), 0, u0 + r * upr_i, bar, policy);
                        tma_load_2d_hint(qh0 + (uint32_t) (r * upr_i * 64), item[g].src[1] + 2 * sizeof(CUtensorMap), 0, u0 + r * upr_i, bar, policy);
                    }
                    bulk_g2s(qh0 + item[g].bytes[1], item[g].src[2], item[g].bytes[2], bar, policy);
                    bulk_g2s(qh0 + item[g].bytes[1] + item[g].bytes[2], item[g].src[3], item[g].bytes[3], bar, policy);
                } else {
                    uint32_t o = 0;
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        if (item[g].bytes[p]) bulk_g2s(dst + o, item[g].src[p], item[g].bytes[p], bar, policy);
                        o += item[g].bytes[p];
                    }
                }
                ++nxt[g]; ++seq[g];
                if (nxt[g] < end[g]) item[g] = a.items[nxt[g]];           // in flight while the ring drains
            }
            if (!pending) break;
        }
        return;
    }

    // ---- consumers ---------------------------------------------------------------------------------------------------------
    const int grp = tid / CH_GTHREADS, gtid = tid % CH_GTHREADS, gwarp = gtid >> 5;
    const uint32_t part0 = xs + (uint32_t) a.x_bytes + (uint32_t) grp * 2 * part_bytes;
    const int * off = a.node_off + (long long) (blockIdx.x * CH_GROUPS + grp) * (NN + 1);

    // B200_CHAIN_TIMELINE: per (CTA, node) eight clock64 values: node start, after the wait, after the X quantiser, node end;
    // then per group: cycles waiting for weight copies, cycles in the item loop
    long long * tl = a.timeline ? a.timeline + ((long long) blockIdx.x * NN) * 8 : nullptr;
    long long c_rel = 0;

    // launch number: bumped by CTA 0 when it leaves; nobody can still be reading it by then (see the end of the kernel)
    const unsigned long long launch = *a.launches;
    const unsigned epoch0 = (unsigned) (launch * (unsigned long long) (NN + 1));      // tags of this launch: epoch0 + 1 .. epoch0 + NN
    const unsigned long long arrivals0 = launch * a.per_launch;
    int pbuf = 0, seq = 0;
    ChainNodeD nd = a.nodes[0];
    int cnt = off[1] - off[0];
    for (int n = 0; n < NN; ++n) {
        ChainNodeD nd_next = nd; int cnt_next = 0;
        if (n + 1 < NN) { nd_next = a.nodes[n + 1]; cnt_next = off[n + 2] - off[n + 1]; }   // lands while this node runs
        if (tl && tid == 0) tl[n * 8 + 0] = clock64();
        if (nd.mode != NB_NONE && nd.mode != NB_TAGGED && tid == 0 && !(a.debug & 1)) {
            // every CTA has finished every earlier node (their arrivals are in). NB_FENCED: X overlaps what they wrote, so the
            // arrivals were releases and this is the acquire; NB_RELAXED: no data flows through this barrier
            const bool local = nd.mode == NB_RELAXED_LOCAL;
            const unsigned long long want = local ? launch * a.per_launch_local + (unsigned long long) nd.want : arrivals0 + (unsigned long long) nd.want * (unsigned) a.world;
            const unsigned long long * cp = local ? a.counter_local : a.counter;
            unsigned long long v, t0 = 0;
            for (unsigned spin = 0;; ++spin) {
                if (a.world > 1 && !local) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(cp) : "memory");
                else                       asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(cp) : "memory");
                if (v >= want) {
                    if (nd.mode == NB_FENCED) { if (a.world > 1) asm volatile("fence.acq_rel.sys;" ::: "memory"); else asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
                    break;
                }
                if ((spin & 4095u) == 4095u) {
                    unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    if (t0 == 0) t0 = t; else if (t - t0 > 4000000000ull) __trap();
                }
            }
        }
        cons_bar();
        if (tl && tid == 0) tl[n * 8 + 1] = clock64();
        if (!(a.debug & 4)) chain_quantise_x<NCOLS>(nd, xs, warp, lane, a.tag_stride, epoch0 + (unsigned) n, a.world > 1);   // tag of node n - 1's output
        cons_bar();
        if (tl && tid == 0) tl[n * 8 + 2] = clock64();

        const int upr = nd.k >> 8;
        const int x_col_stride = upr * XU_BYTES;
        const int q4_rp = upr <= 32 ? 32 / upr : 1, q4_sub = upr <= 32 ? lane / upr : 0, q4_b = lane - q4_sub * upr;
        long long c_wait = 0, c_loop = tl ? clock64() : 0;
        for (int i = 0; i < cnt; ++i, ++seq) {
            const int st = seq % CH_NS;
            const uint32_t stage = ring0 + (uint32_t) (grp * CH_NS + st) * (uint32_t) a.stage_bytes;
            const uint32_t part = part0 + (uint32_t) pbuf * part_bytes;
            const long long w0 = tl ? clock64() : 0;
            mbar_wait(full_bar(grp, st), (uint32_t) ((seq / CH_NS) & 1));
            if (tl) c_wait += clock64() - w0;
            const uint4 d0 = lds16(desc_at(grp, st));
            const uint4 d1 = lds16(desc_at(grp, st) + 16);
            const uint2 d2 = lds8(desc_at(grp, st) + 32);
            YOut out;
            out.y = (float *) (((unsigned long long) d0.y << 32) | d0.x);
            out.ldy = (long long) (((unsigned long long) d0.w << 32) | d0.z);
            out.tag = (unsigned long long *) (((unsigned long long) d2.y << 32) | d2.x);
            out.tag_stride = a.tag_stride; out.epoch = epoch0 + (unsigned) n + 1u;
            out.world = a.world; out.delta = a.delta;
            const int type = (int) d1.x, nrows = (int) d1.y;
            ItemPlan pl; pl.segs = (int) d1.z; pl.S = (int) d1.w;

            // phase A: every warp reduces its tasks (rows, or runs of units of a row)
            const int first = gwarp;
            if (a.debug & 2) {
            } else if (type == B200_TYPE_Q4_K) {
                item_q4k<NCOLS>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
            } else if (type == CH_T_Q6K_SW) {
                item_q4k<NCOLS, LU_Q6K>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
            } else if (KS >= 1 && type == B200_TYPE_Q5_K) {
                item_q4k<NCOLS, LU_Q5K>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
            } else if (KS >= 1 && type == B200_TYPE_Q6_K) {
                item_lanes8<B200_TYPE_Q6_K, NCOLS>(stage, nrows, upr, xs, x_col_stride, nd.act_block, pl, part, a.max_units, out, first, lane);
            } else if (KS < 2) {
            } else if (type == B200_TYPE_IQ4_XS) {
                if (a.lanes8) item_lanes8<B200_TYPE_IQ4_XS, NCOLS>(stage, nrows, upr, xs, x_col_stride, nd.act_block, pl, part, a.max_units, out, first, lane);
                else          item_q4k<NCOLS, LU_IQ4XS>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
            } else if (type == B200_TYPE_Q4_0) {
                if (a.lanes8) item_lanes8<B200_TYPE_Q4_0, NCOLS>(stage, nrows, upr, xs, x_col_stride, nd.act_block, pl, part, a.max_units, out, first, lane);
                else          item_q4k<NCOLS, LU_Q40>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
            } else {
                if constexpr (KS >= 3) {       // (its own instantiation: in the same kernel the unit-per-lane Q8_0 code cost the Q4_0 path 8 %)
                    if (a.lanes8) item_lanes8<B200_TYPE_Q8_0, NCOLS>(stage, nrows, upr, xs, x_col_stride, nd.act_block, pl, part, a.max_units, out, first, lane);
                    else          item_q4k<NCOLS, LU_Q80>(stage, nrows, upr, xs, x_col_stride, pl, part, a.max_units, out, first, lane, q4_rp, q4_sub, q4_b);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty_bar(grp, st));     // this warp is done with the stage (and its descriptor)

            // phase B (rows cut into several tasks only): add the pieces of a row in a fixed order
            if (pl.segs > 1) {
                group_bar(grp);
                if (!(a.debug & 8)) {
                    for (int rc = gtid; rc < nrows * NCOLS; rc += CH_GTHREADS) {
                        const int c = NCOLS == 1 ? 0 : rc / nrows, r = rc - c * nrows;
                        float v = 0.0f;
                        for (int sg = 0; sg < pl.segs; ++sg) v += ldsf(part + (uint32_t) (c * a.max_units + r * pl.segs + sg) * 4);
                        put_y(out, c, r, v);
                    }
                }
                pbuf ^= 1;
            }
        }
        if (tl && gtid == 0) { if (grp == 0) tl[n * 8 + 4] = c_wait; tl[n * 8 + 5 + 2 * grp] = clock64() - c_loop; }
        if (tl && tid == 0) c_rel = clock64();
        cons_bar();                            // every store of this CTA is issued; a release below publishes them (cumulativity)
        if (tid == 0 && nd.signal) {
            if (nd.signal == 3) asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter_local) : "memory");
            else if (a.world > 1) {             // row split: the arrival goes to every rank; a fenced boundary releases this CTA's peer stores first
                if (nd.signal == 2) asm volatile("fence.acq_rel.sys;" ::: "memory");
                for (int g = 0; g < a.world; ++g)
                    asm volatile("red.relaxed.sys.global.add.u64 [%0], 1;" :: "l"(reinterpret_cast<char *>(a.counter) + a.delta[g]) : "memory");
            } else if (nd.signal == 2) asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter) : "memory");
            else                       asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter) : "memory");
        }
        if (tl && tid == 0) { tl[n * 8 + 3] = clock64(); tl[n * 8 + 6] = tl[n * 8 + 3] - c_rel; }   // slot 6: cycles in the release (group 1's stall is dropped)
        nd = nd_next; cnt = cnt_next;
    }
    if (a.world > 1) {
        // row split: this rank's kernel may only complete when every rank's stores into its arena have been performed, so that
        // whatever follows on the stream sees the whole Y. Every CTA of every rank: system fence (covers the CTA's peer stores:
        // the barrier above + cumulativity), then one arrival on every rank's `done`; CTA 0 waits for all of them.
        if (tid == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            for (int g = 0; g < a.world; ++g)
                asm volatile("red.relaxed.sys.global.add.u64 [%0], 1;" :: "l"(reinterpret_cast<char *>(a.done) + a.delta[g]) : "memory");
            if (blockIdx.x == 0) {
                const unsigned long long want = (launch + 1ull) * (unsigned long long) gridDim.x * (unsigned) a.world;
                unsigned long long v, t0 = 0;
                for (unsigned spin = 0;; ++spin) {
                    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a.done) : "memory");
                    if (v >= want) break;
                    __nanosleep(100);
                    if ((spin & 1023u) == 1023u) {
                        unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                        if (t0 == 0) t0 = t; else if (t - t0 > 4000000000ull) __trap();
                    }
                }
            }
        }
    }
    // next launch, new tags and counter targets. CTA 0 gets here only after every CTA has passed the start of the last node
    // (a one-node chain uses neither), i.e. after everybody has read the old value
    if (blockIdx.x == 0 && tid == 0) *a.launches = launch + 1ull;
}

#include "gemv_chain_xreg.cuh"

// ---------------------------------------------------------------------------------------------------------------------------
struct Chain {
    int device = 0, sms = 0, ncols = 1, n_nodes = 0, n_launch_nodes = 0;
    ChainArgs args = {};
    void * d_blob = nullptr, * d_tags = nullptr;
    size_t smem = 0;
    bool cooperative = true;
    bool xreg = false;            // gemv_chainx_kernel (activations in registers) instead of gemv_chain_kernel
    int kset = 0;                 // item kinds the chain needs: gemv_chain_kernel<NCOLS, kset>
};

// consumer warps of gemv_chainx_kernel (8 or 12; B200_CHAIN_NW)
int cx_nw() { static const int v = [] { const char * e = getenv("B200_CHAIN_NW"); const int n = e ? atoi(e) : 8; return n == 12 ? 12 : 8; }(); return v; }
#define CX_NW cx_nw()

bool chain_type_ok(int t) { return t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K || t == B200_TYPE_IQ4_XS || t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0; }
int gcd_i(int a, int b) { while (b) { int t = a % b; a = b; b = t; } return a; }
// rows of one item must keep every plane piece a multiple of 16 bytes, in size and in source address
int row_granule(int type, int upr) {
    int g = 1;
    for (int p = 0; p < 4; ++p) {
        const int pb = plane_bytes(type, p);
        if (!pb) break;
        const int per_row = pb * upr;
        const int need = 16 / gcd_i(per_row, 16);
        g = g * need / gcd_i(g, need);
    }
    return g;
}

template <int NCOLS, int KS> cudaError_t chain_launch_t(const Chain & c, cudaStream_t s) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) c.sms); cfg.blockDim = dim3(CH_THREADS); cfg.dynamicSmemBytes = c.smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative; attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = c.cooperative ? 1 : 0;
    b200_count_launch();
    return cudaLaunchKernelEx(&cfg, gemv_chain_kernel<NCOLS, KS>, c.args);
}
cudaError_t chainx_launch(const Chain & c, cudaStream_t s) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) c.sms); cfg.blockDim = dim3((CX_NW + 1) * 32); cfg.dynamicSmemBytes = c.smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative; attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = c.cooperative ? 1 : 0;
    b200_count_launch();
    return CX_NW == 12 ? cudaLaunchKernelEx(&cfg, gemv_chainx_kernel<12>, c.args) : cudaLaunchKernelEx(&cfg, gemv_chainx_kernel<8>, c.args);
}

// [units][plane_bytes] uint8 plane of a Q6_K matrix; one box = the `upr` units of one weight row
typedef CUresult (*chain_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool chain_plane_map(CUtensorMap * tm, const void * base, uint32_t plane, uint64_t units, uint32_t upr, CUtensorMapSwizzle sw) {
    static chain_encode_fn fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (chain_encode_fn) p;
    }();
    if (!fn) return false;
    const cuuint64_t dims[2] = {plane, units};
    const cuuint64_t strides[1] = {plane};
    const cuuint32_t box[2] = {plane, upr};
    const cuuint32_t estr[2] = {1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ring stage: what is left after X, split over groups x stages; the partial buffers depend on the stage size
// (max_units = stage / 136 units of 4 B x 2 buffers x ncols per group), solved by stepping the stage size down. 0: no room
int chain_stage_bytes(int ncols, int64_t kmax, int total_smem, int * max_units) {
    const int64_t x_bytes = ((int64_t) ncols * (kmax >> 8) * XU_BYTES + 127) & ~(int64_t) 127;
    for (int st = 48 * 1024; st >= 4 * 1024; st -= 1024) {
        const int mu = st / 136 + 64;         // partial sums: at most max(rows, 2 x warps + rows) tasks per item
        const int64_t need = CH_HDR + x_bytes + (int64_t) CH_GROUPS * 2 * ncols * mu * 4 + 1024 + (int64_t) CH_GROUPS * CH_NS * st;
        if (need <= total_smem) { if (max_units) *max_units = mu; return st; }
    }
    return 0;
}
// gemv_chainx_kernel: header + two X buffers + per-stage partial sums + alignment slack + ns stages (multiples of 1 KB)
int chainx_stage_bytes(int64_t kmax, int total_smem, int ns) {
    const int64_t x_bytes = ((kmax >> 8) * XU_BYTES + 127) & ~(int64_t) 127;
    const int64_t left = (int64_t) total_smem - CX_HDR - 2 * x_bytes - (int64_t) ns * CX_PART_FLOATS * 4 - 1024;
    if (left <= 0) return 0;
    const int64_t st = left / ns / 1024 * 1024;
    return st >= 4096 ? (int) (st > 64 * 1024 ? 64 * 1024 : st) : 0;
}

const char * chain_check(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups, bool split = false) {
    if (!nodes || !group_sizes || n_groups < 1) return "no nodes";
    int idx = 0;
    const int64_t n0 = nodes[0].n;
    if (n0 < 1 || n0 > CH_MAX_COLS) return "n must be 1..4";
    for (int g = 0; g < n_groups; ++g) {
        if (group_sizes[g] < 1) return "empty group";
        const b200_mul_mat_args & a0 = nodes[idx];
        for (int i = 0; i < group_sizes[g]; ++i, ++idx) {
            const b200_mul_mat_args & a = nodes[idx];
            if (!a.w || !a.x || !a.y) return "null pointer";
            if (!chain_type_ok(a.type)) return "weight type";
            if (a.n != n0) return "all nodes of a chain need the same n";
            if (a.k <= 0 || a.k % 256 != 0 || a.k > 65536) return "k must be a multiple of 256 (<= 65536)";
            if (a.m < (split ? 0 : 1) || a.m > (1 << 30)) return "m out of range";      // a rank's slice of a small matrix may be empty
            if (a.x != a0.x || a.k != a0.k || a.x_row_stride != a0.x_row_stride || a.flavour != a0.flavour) return "nodes of a group must read the same X";
            if (b200_act_type(a.type) != b200_act_type(a0.type)) return "nodes of a group need the same activation format";
            if (a.layout != B200_LAYOUT_PLANES && a.type != B200_TYPE_Q4_K && a.type != B200_TYPE_Q5_K) return "plane layout required";
            if (a.w_row_stride != 0 && a.w_row_stride != (int64_t) b200_row_size(a.type, a.k)) return "packed rows required";
            if ((((uintptr_t) a.w) & 15) != 0) return "weights must be 16 B aligned";
            if (a.m % row_granule(a.type, (int) (a.k >> 8)) != 0) return "row count incompatible with 16-byte plane pieces";
            if (a.flags & B200_MM_X_PREQUANTISED) return "prequantised X";
        }
    }
    // every row granule must fit one ring stage next to the quantised activations of the widest node (n x k of 4 x 28672 does not)
    int64_t kmax = 0;
    for (int i = 0; i < idx; ++i) kmax = nodes[i].k > kmax ? nodes[i].k : kmax;
    const int stage = chain_stage_bytes((int) n0, kmax, CH_SMEM_TOTAL, nullptr);
    if (!stage) return "activations leave no room for the weight ring";
    for (int i = 0; i < idx; ++i) {
        const int upr = (int) (nodes[i].k >> 8);
        if ((int64_t) unit_bytes(nodes[i].type) * upr * row_granule(nodes[i].type, upr) > stage - 512) return "a weight row does not fit a ring stage next to the activations";
    }
    return nullptr;
}

// ---- the plan of a chain: pure host logic, no CUDA call (b200_chain_plan exposes it to the CPU tests) ------------------------------
struct Cut { int m; int64_t row; };          // where a share ends: (matrix of the group, row)
struct ChainPlan {
    bool xreg = false;                       // kernel: gemv_chainx_kernel (one ring, X in registers) or gemv_chain_kernel
    int groups = CH_GROUPS, ns = CH_NS, stage = 0, max_units = 0, x_bytes = 0, ncols = 1, VC = 0;
    size_t smem = 0;
    std::vector<int> mode, feed, first;      // per group: NB_*, matrix of the group whose Y the next group reads through tags (-1), first node
    std::vector<std::vector<Cut>> cuts;      // per group: end of share v = 0 .. VC - 1 (share v belongs to virtual consumer (v + 37 g) % VC)
    std::string error;
};

bool q6_swizzle_enabled() { return getenv("B200_CHAIN_Q6K_LANES8") == nullptr; }
bool q6_swizzled_shape(const b200_mul_mat_args & a) {
    const int64_t upr = a.k >> 8;
    return a.type == B200_TYPE_Q6_K && q6_swizzle_enabled() && upr % 8 == 0 && a.m * upr < (1ll << 31);
}

// shares of one group: the concatenated rows of its matrices cut by bytes into VC contiguous pieces, rounded to row granules
void chain_cut_group(const int64_t * m, const int * type, int cnt, int upr, int VC, std::vector<Cut> & out) {
    std::vector<int64_t> rb(cnt), cum(cnt + 1, 0);
    for (int i = 0; i < cnt; ++i) { rb[i] = (int64_t) unit_bytes(type[i]) * upr; cum[i + 1] = cum[i] + rb[i] * m[i]; }
    const int64_t T = cum[cnt];
    out.resize(VC);
    int pm = 0; int64_t prow = 0;
    for (int v = 0; v < VC; ++v) {
        int cm; int64_t crow;
        if (v == VC - 1) { cm = cnt - 1; crow = m[cnt - 1]; }
        else {
            const int64_t t = T * (v + 1) / VC;
            cm = 0;
            while (cm + 1 < cnt && t >= cum[cm + 1]) ++cm;
            const int gr = row_granule(type[cm], upr);
            int64_t r = ((t - cum[cm]) + rb[cm] * gr / 2) / (rb[cm] * gr) * gr;
            if (r > m[cm]) r = m[cm];
            crow = r;
        }
        if (cm < pm || (cm == pm && crow < prow)) { cm = pm; crow = prow; }
        out[v].m = cm; out[v].row = crow;
        pm = cm; prow = crow;
    }
}

// `sms` CTAs; split may be null. Decides the kernel, the ring geometry, the shares and how each node learns that its X is complete.
ChainPlan chain_plan(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups, int sms, int total_smem, const b200_chain_split * split) {
    ChainPlan P;
    const int world = split ? split->world : 1;
    P.ncols = (int) nodes[0].n;
    P.first.assign(n_groups + 1, 0);
    for (int g = 0; g < n_groups; ++g) P.first[g + 1] = P.first[g] + group_sizes[g];
    const int n_nodes = P.first[n_groups];
    int64_t kmax = 0;
    for (int i = 0; i < n_nodes; ++i) kmax = nodes[i].k > kmax ? nodes[i].k : kmax;

    // ---- kernel and ring geometry ------------------------------------------------------------------------------------------------
    // measured on B200 (8B Q4_K_M token): 685-732 tokens/s against 796 for gemv_chain_kernel -- with 8 warps the per-task latency chain, not
    // shared memory, bounds the consumers -- so the activations-in-registers kernel is opt-in: B200_CHAIN_XREG=1
    bool xreg = P.ncols == 1 && getenv("B200_CHAIN_XREG") && getenv("B200_CHAIN_XREG")[0] == '1';
    int ns = CX_MAX_NS < 6 ? CX_MAX_NS : 6;
    if (const char * e = getenv("B200_CHAIN_NS")) { const int v = atoi(e); if (v >= 2 && v <= CX_MAX_NS) ns = v; }
    int xstage = xreg ? chainx_stage_bytes(kmax, total_smem, ns) : 0;
    if (const char * e = getenv("B200_CHAIN_STAGE_KB")) { const int v = atoi(e) * 1024; if (v >= 4096 && v < xstage) xstage = v; }
    for (int i = 0; i < n_nodes && xreg; ++i) {
        const b200_mul_mat_args & a = nodes[i];
        const int upr = (int) (a.k >> 8);
        const int segs = upr <= 32 ? 1 : (upr + 31) >> 5;
        if (!(a.type == B200_TYPE_Q4_K || q6_swizzled_shape(a))) xreg = false;
        else if (CX_NW % segs != 0 || segs > CX_PART_FLOATS) xreg = false;
        else if ((int64_t) unit_bytes(a.type) * upr > xstage - 512) xreg = false;
    }
    P.xreg = xreg;
    if (xreg) {
        P.groups = 1; P.ns = ns; P.stage = xstage; P.max_units = 0;
        P.x_bytes = (int) (((kmax >> 8) * XU_BYTES + 127) & ~(int64_t) 127);
        P.smem = (size_t) CX_HDR + 2 * (size_t) P.x_bytes + (size_t) ns * CX_PART_FLOATS * 4 + 1024 + (size_t) ns * xstage;
    } else {
        P.groups = CH_GROUPS; P.ns = CH_NS;
        P.stage = chain_stage_bytes(P.ncols, kmax, total_smem, &P.max_units);
        if (const char * e = getenv("B200_CHAIN_STAGE_KB")) { const int v = atoi(e) * 1024; if (v >= 4096 && v < P.stage) P.stage = v; }
        P.x_bytes = (int) (((int64_t) P.ncols * (kmax >> 8) * XU_BYTES + 127) & ~127);
        P.smem = CH_HDR + (size_t) P.x_bytes + (size_t) CH_GROUPS * 2 * P.ncols * P.max_units * 4 + 1024 + (size_t) CH_GROUPS * CH_NS * P.stage;
        if (!P.stage) { P.error = "activations leave no room for the weight ring"; return P; }
    }
    P.VC = sms * P.groups;

    // ---- shares ------------------------------------------------------------------------------------------------------------------
    P.cuts.resize(n_groups);
    for (int g = 0; g < n_groups; ++g) {
        const int cnt = group_sizes[g];
        std::vector<int64_t> m(cnt); std::vector<int> type(cnt);
        for (int i = 0; i < cnt; ++i) { m[i] = nodes[P.first[g] + i].m; type[i] = nodes[P.first[g] + i].type; }
        chain_cut_group(m.data(), type.data(), cnt, (int) (nodes[P.first[g]].k >> 8), P.VC, P.cuts[g]);
    }

    // ---- how each node learns that its X is complete (NB_*): from the pointers -----------------------------------------------------
    auto lo_of = [&](int i) { return split ? split->row_lo[i] : (int64_t) 0; };          // first row of this rank's slice
    auto mf_of = [&](int i) { return split ? split->m_full[i] : nodes[i].m; };           // rows of the whole matrix
    // the whole Y of node i (every rank's rows: with the row split all of it is written into this rank's arena)
    auto y_lo = [&](int i) { return (const char *) (nodes[i].y - lo_of(i)); };
    auto y_hi = [&](int i) { return (const char *) (nodes[i].y - lo_of(i) + (nodes[i].n - 1) * nodes[i].y_row_stride + mf_of(i)); };
    auto x_lo = [&](int g) { return (const char *) nodes[P.first[g]].x; };
    auto x_hi = [&](int g) { const b200_mul_mat_args & a = nodes[P.first[g]]; return (const char *) (a.x + (a.n - 1) * a.x_row_stride + a.k); };
    const bool no_tags = getenv("B200_CHAIN_NO_TAGS") != nullptr, all_fenced = getenv("B200_CHAIN_ALL_FENCED") != nullptr;
    P.mode.assign(n_groups, NB_NONE); P.feed.assign(n_groups, -1);
    // does every CTA of every rank own rows of matrix `fm` of group g? (a CTA without any is not ordered by that matrix' tagged rows)
    auto every_cta_feeds = [&](int g, int fm) {
        const int cnt = group_sizes[g], upr = (int) (nodes[P.first[g]].k >> 8);
        for (int r = 0; r < world; ++r) {
            std::vector<int64_t> m(cnt); std::vector<int> type(cnt);
            for (int i = 0; i < cnt; ++i) {
                type[i] = nodes[P.first[g] + i].type;
                m[i] = nodes[P.first[g] + i].m;
                if (split && r != split->rank) { int64_t lo = 0, hi = 0; b200_row_split(mf_of(P.first[g] + i), world, r, &lo, &hi); m[i] = hi - lo; }
            }
            std::vector<Cut> cuts;
            if (split && r != split->rank) chain_cut_group(m.data(), type.data(), cnt, upr, P.VC, cuts);
            const std::vector<Cut> & cu = (split && r != split->rank) ? cuts : P.cuts[g];
            std::vector<char> has(sms, 0);
            Cut prev = {0, 0};
            for (int v = 0; v < P.VC; ++v) {
                const Cut end = cu[v];
                const int64_t lo = prev.m < fm ? 0 : (prev.m == fm ? prev.row : m[fm]);
                const int64_t hi = end.m > fm ? m[fm] : (end.m == fm ? end.row : 0);
                if (hi > lo) has[((v + g * 37) % P.VC) / P.groups] = 1;
                prev = end;
            }
            for (int c = 0; c < sms; ++c) if (!has[c]) return false;
        }
        return true;
    };
    for (int g = 1; g < n_groups; ++g) {
        const b200_mul_mat_args & a0 = nodes[P.first[g]];
        // read after write: X overlaps the Y of any earlier node
        bool hazard = false;
        for (int i = 0; i < P.first[g] && !hazard; ++i) hazard = x_lo(g) < y_hi(i) && y_lo(i) < x_hi(g);
        // write after read / write after write against the two previous groups: with a tagged boundary CTAs of neighbouring nodes coexist,
        // and a relaxed arrival does not order the stores of the two nodes
        bool waw_war = false;
        for (int h = g - 1; h >= 0 && h >= g - 2 && !waw_war; --h)
            for (int i = P.first[g]; i < P.first[g + 1] && !waw_war; ++i) {
                waw_war = y_lo(i) < x_hi(h) && x_lo(h) < y_hi(i);
                for (int j = P.first[h]; j < P.first[h + 1] && !waw_war; ++j) waw_war = y_lo(i) < y_hi(j) && y_lo(j) < y_hi(i);
            }
        P.mode[g] = hazard || waw_war || all_fenced ? NB_FENCED : NB_RELAXED;
        // X is exactly the Y of a matrix of the previous group, and every CTA owns rows of that matrix (so nobody can be more than one
        // node behind a CTA that has seen all of its tagged rows): the rows validate themselves, no counter and no fence
        for (int i = P.first[g - 1]; i < P.first[g] && !no_tags && !all_fenced && !waw_war; ++i)
            if (nodes[i].y - lo_of(i) == a0.x && mf_of(i) == a0.k && (a0.n == 1 || nodes[i].y_row_stride == a0.x_row_stride) &&
                every_cta_feeds(g - 1, i - P.first[g - 1])) { P.mode[g] = NB_TAGGED; P.feed[g - 1] = i - P.first[g - 1]; }
        // row split: a stream-order-only boundary need not wait for the other ranks when the boundary before it was a tagged one -- passing that one
        // already proved that every CTA of every rank has left the node before the previous one, so the ranks stay within one node of each other
        // and the three tag buffers (node g publishes into buffer g % 3) cannot be overwritten while a slower rank still reads them
        if (split && P.mode[g] == NB_RELAXED && g >= 2 && P.mode[g - 1] == NB_TAGGED && !getenv("B200_CHAIN_GLOBAL_BARRIERS")) P.mode[g] = NB_RELAXED_LOCAL;
    }
    return P;
}

} // namespace

extern "C" {

int b200_chain_supported(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups) {
    return chain_check(nodes, group_sizes, n_groups) == nullptr ? 1 : 0;
}

size_t b200_chain_scratch_bytes(int64_t max_m_full, int ncols) {
    const int64_t tmax = ((max_m_full > 256 ? max_m_full : 256) + 255) & ~(int64_t) 255;
    return 256 + (size_t) 3 * (size_t) (ncols > 0 ? ncols : 1) * (size_t) tmax * 8;     // counters + three tag buffers (node g publishes into buffer g % 3)
}

// host-only view of the plan a chain would get on a device with `sms` SMs (no CUDA call; pointers are only compared):
// modes[g] = NB_* (0 none, 1 fenced counter, 2 relaxed counter, 3 tagged words), *xreg = 1 when gemv_chainx_kernel takes the chain
int b200_chain_plan(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups, int sms, const b200_chain_split * split,
                    int32_t * modes, int32_t * xreg) {
    const char * why = chain_check(nodes, group_sizes, n_groups, split != nullptr);
    B200_REQUIRE(!why, "b200_chain_plan: %s", why);
    B200_REQUIRE(sms > 0 && modes, "b200_chain_plan: bad arguments");
    const ChainPlan P = chain_plan(nodes, group_sizes, n_groups, sms, CH_SMEM_TOTAL, split);
    B200_REQUIRE(P.error.empty(), "b200_chain_plan: %s", P.error.c_str());
    for (int g = 0; g < n_groups; ++g) modes[g] = P.mode[g];
    if (xreg) *xreg = P.xreg ? 1 : 0;
    return 0;
}

b200_chain_t b200_chain_create(int device, const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups) {
    return b200_chain_create_split(device, nodes, group_sizes, n_groups, nullptr);
}

b200_chain_t b200_chain_create_split(int device, const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups, const b200_chain_split * split) {
    const char * why = chain_check(nodes, group_sizes, n_groups, split != nullptr);
    if (why) { b200_set_error("b200_chain_create: %s", why); return nullptr; }
    const int world = split ? split->world : 1;
    const bool lanes8 = getenv("B200_CHAIN_LANES8") != nullptr;
    int kset = 0;
    { int total = 0; for (int g = 0; g < n_groups; ++g) total += group_sizes[g];
      for (int i = 0; i < total; ++i) {
          const int t = nodes[i].type;
          const int need = (t == B200_TYPE_Q4_K || (t == B200_TYPE_Q6_K && q6_swizzled_shape(nodes[i]))) ? 0 : (t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K) ? 1 : t == B200_TYPE_Q8_0 ? 3 : 2;
          kset = need > kset ? need : kset;
      } }
    if (split && (world < 2 || world > B200_MAX_PEERS || split->rank < 0 || split->rank >= world || !split->scratch || !split->row_lo || !split->m_full ||
                  split->peer_delta[split->rank] != 0 || (((uintptr_t) split->scratch) & 255) != 0)) {
        b200_set_error("b200_chain_create_split: bad descriptor (world %d rank %d)", world, split->rank); return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); b200_set_error("b200_chain_create: cudaSetDevice(%d) failed", device); return nullptr; }
    int sms = 0, smem_optin = 0, coop = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
    if (sms <= 0 || smem_optin < 200 * 1024) { cudaGetLastError(); b200_set_error("b200_chain_create: no usable sm_100 device"); return nullptr; }
    if (const char * e = getenv("B200_CHAIN_CTAS")) { const int v = atoi(e); if (v > 0 && v < sms) sms = v; }
    const int total_smem = smem_optin < CH_SMEM_TOTAL ? smem_optin : CH_SMEM_TOTAL;

    const ChainPlan P = chain_plan(nodes, group_sizes, n_groups, sms, total_smem, split);
    if (!P.error.empty()) { b200_set_error("b200_chain_create: %s", P.error.c_str()); return nullptr; }
    const int ncols = P.ncols, VC = P.VC, stage = P.stage, max_units = P.max_units, x_bytes = P.x_bytes;
    const std::vector<int> & mode = P.mode, & feed = P.feed, & first = P.first;

    std::vector<ChainNodeD> hn(n_groups);
    std::vector<ChainMatD> hm;       // host-side only: the items carry everything the kernel needs
    std::vector<CUtensorMap> maps;   // two per swizzled Q6_K matrix (QL, QH); copied into the blob, the items point at them
    std::vector<int> mat_map, mat_R; // per matrix: index of its first map or -1 (four: QL, QH for R rows per copy, QL, QH for one row), R

    int64_t tmax = 256;
    auto lo_of = [&](int i) { return split ? split->row_lo[i] : (int64_t) 0; };          // first row of this rank's slice
    auto mf_of = [&](int i) { return split ? split->m_full[i] : nodes[i].m; };           // rows of the whole matrix
    for (int i = 0; i < first[n_groups]; ++i) tmax = mf_of(i) > tmax ? mf_of(i) : tmax;
    tmax = (tmax + 255) & ~(int64_t) 255;
    unsigned long long * d_tags = nullptr;
    const size_t tag_bytes = (size_t) 3 * ncols * tmax * 8;
    if (split) {                               // tags and counters live in the symmetric arena: every rank stores into every rank's copy
        if (split->scratch_bytes < 256 + tag_bytes) { b200_set_error("b200_chain_create_split: scratch %zu < %zu bytes", split->scratch_bytes, 256 + tag_bytes); return nullptr; }
        d_tags = (unsigned long long *) ((char *) split->scratch + 256);
    } else if (cudaMalloc((void **) &d_tags, tag_bytes) != cudaSuccess || cudaMemset(d_tags, 0, tag_bytes) != cudaSuccess) {
        cudaGetLastError(); b200_set_error("b200_chain_create: cudaMalloc of %zu bytes failed", tag_bytes); return nullptr;
    }
    unsigned signalling_total = 0, signalling_local_total = 0;
    {
        unsigned signalling = 0, signalling_local = 0;      // arrivals so far on the all-ranks counter / on the rank-local counter
        for (int g = 0; g < n_groups; ++g) {
            hn[g].mode = mode[g];
            hn[g].xtag = mode[g] == NB_TAGGED ? d_tags + (size_t) ((g - 1) % 3) * ncols * tmax : nullptr;
            hn[g].want = (mode[g] == NB_RELAXED_LOCAL ? signalling_local : signalling) * (unsigned) sms;
            const int nm = g + 1 < n_groups ? mode[g + 1] : NB_NONE;
            hn[g].signal = nm == NB_FENCED ? 2 : nm == NB_RELAXED ? 1 : nm == NB_RELAXED_LOCAL ? 3 : 0;
            if (hn[g].signal == 3) ++signalling_local; else if (hn[g].signal) ++signalling;
        }
        signalling_total = signalling; signalling_local_total = signalling_local;
    }
    std::vector<std::vector<ChainItemD>> per_vc_node((size_t) VC * n_groups);
    std::vector<int> rot(VC, 0);     // gemv_chainx_kernel: tasks dealt to the warps so far, per virtual consumer (restarts with every node)
    int idx = 0;
    for (int g = 0; g < n_groups; ++g) {
        const b200_mul_mat_args & a0 = nodes[idx];
        hn[g].x = a0.x; hn[g].ldx = a0.x_row_stride; hn[g].k = (int) a0.k;
        hn[g].act_block = b200_act_type(a0.type) == B200_TYPE_Q8_K ? 256 : 32;
        hn[g].flavour = a0.flavour;
        const int upr = (int) (a0.k >> 8);
        const int cnt = group_sizes[g];
        const int mat_base = (int) hm.size();
        for (int i = 0; i < cnt; ++i) {
            const b200_mul_mat_args & a = nodes[idx + i];
            ChainMatD M; M.w = (const uint8_t *) a.w; M.y = a.y; M.ldy = a.y_row_stride; M.m = (int) a.m; M.type = a.type; M.upr = upr; M.pad = 0;
            hm.push_back(M);
            // Q6_K rows of a multiple of 8 units: QL / QH through swizzling tensor copies, one block per lane (unit_q6k)
            int map0 = -1, R = 1;
            if (q6_swizzled_shape(a) && a.m > 0) {
                const uint8_t * w = (const uint8_t *) a.w;
                // rows per tensor copy: as many as the box limit (256 units) and a ring stage allow, but at least two copies per full stage
                const int fit = (int) (((int64_t) stage - 512) / ((int64_t) 210 * upr));
                const int lim = 256 / upr < fit ? 256 / upr : fit;
                while (2 * R <= lim) R *= 2;
                int pf = 1; while (2 * pf <= fit) pf *= 2;
                if (R == pf && R > 2) R /= 2;
                CUtensorMap t4[4];
                if (chain_plane_map(&t4[0], w, 128, (uint64_t) a.m * upr, (uint32_t) (R * upr), CU_TENSOR_MAP_SWIZZLE_128B) &&
                    chain_plane_map(&t4[1], w + (int64_t) a.m * upr * 128, 64, (uint64_t) a.m * upr, (uint32_t) (R * upr), CU_TENSOR_MAP_SWIZZLE_64B) &&
                    chain_plane_map(&t4[2], w, 128, (uint64_t) a.m * upr, (uint32_t) upr, CU_TENSOR_MAP_SWIZZLE_128B) &&
                    chain_plane_map(&t4[3], w + (int64_t) a.m * upr * 128, 64, (uint64_t) a.m * upr, (uint32_t) upr, CU_TENSOR_MAP_SWIZZLE_64B)) {
                    map0 = (int) maps.size();
                    for (int t = 0; t < 4; ++t) maps.push_back(t4[t]);
                } else if (P.xreg) {
                    b200_set_error("b200_chain_create: cuTensorMapEncodeTiled failed for a Q6_K plane (m=%lld k=%lld)", (long long) a.m, (long long) a.k);
                    if (!split) cudaFree(d_tags);
                    return nullptr;
                }
            }
            if (a.type == B200_TYPE_Q6_K && map0 < 0 && kset < 1) kset = 1;      // this matrix takes the eight-lanes-per-block Q6_K items
            mat_R.push_back(R);
            mat_map.push_back(map0);
        }
        int pm = 0; int64_t prow = 0;
        // the VC order rotates from node to node so that rounding leftovers do not pile up on the same SMs
        for (int v = 0; v < VC; ++v) {
            const int cm = P.cuts[g][v].m; const int64_t crow = P.cuts[g][v].row;
            const int vc = (v + g * 37) % VC;
            std::vector<ChainItemD> & out = per_vc_node[(size_t) vc * n_groups + g];
            rot[vc] = 0;
            for (int mi = pm; mi <= cm; ++mi) {
                const int64_t r0 = mi == pm ? prow : 0, r1 = mi == cm ? crow : nodes[idx + mi].m;
                if (r1 <= r0) continue;
                const int type = nodes[idx + mi].type;
                const int gr = row_granule(type, upr);
                const int map0 = mat_map[mat_base + mi];
                int64_t max_rows = ((int64_t) stage - (map0 >= 0 ? 512 : 0)) / ((int64_t) unit_bytes(type) * upr) / gr * gr;
                const int xsegs = upr <= 32 ? 1 : (upr + 31) >> 5;
                if (P.xreg && xsegs > 1 && max_rows > CX_PART_FLOATS / xsegs) max_rows = CX_PART_FLOATS / xsegs;      // partial sums of an item: rows x segs floats
                if (max_rows < gr) { b200_set_error("b200_chain_create: a row granule of type %d, k=%lld does not fit a %d-byte ring stage", type, (long long) a0.k, stage); if (!split) cudaFree(d_tags); return nullptr; }
                const int64_t rows = r1 - r0;
                const int64_t nit = (rows + max_rows - 1) / max_rows;
                const int R = mat_R[mat_base + mi];
                const int64_t pg = map0 >= 0 && max_rows >= R ? (int64_t) R : (int64_t) gr;      // full items are whole tensor copies (R is a power of two, gr = 1 here)
                int64_t per = ((rows + nit - 1) / nit + pg - 1) / pg * pg;
                if (per > max_rows) per = max_rows / pg * pg;
                for (int64_t r = r0; r < r1; r += per) {
                    const ChainMatD & M = hm[mat_base + mi];
                    ChainItemD it = {};
                    it.nrows = (int) (r1 - r < per ? r1 - r : per); it.type = type;
                    it.y = M.y + r; it.ldy = M.ldy;
                    it.tag = feed[g] == mi ? d_tags + (size_t) (g % 3) * ncols * tmax + lo_of(idx + mi) + r : nullptr;
                    { const ItemPlan pl = plan_item(type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || map0 >= 0 || (!lanes8 && (type == B200_TYPE_IQ4_XS || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0)), it.nrows, upr); it.segs = pl.segs; it.S = pl.S; }
                    if (P.xreg) {
                        // `S` carries the rotation: which warp gets the item's first task, so that tasks go round the warps across items
                        if (it.segs == 1) { const int rp = 32 / upr; it.S = rot[vc] % CX_NW; rot[vc] = (rot[vc] + (it.nrows + rp - 1) / rp) % CX_NW; }
                        else { const int per_class = CX_NW / it.segs; it.S = rot[vc] % per_class; rot[vc] = (rot[vc] + it.nrows) % CX_NW; }
                    }
                    int64_t plane_off = 0;
                    for (int p = 0; p < 4; ++p) {
                        const int pb = plane_bytes(type, p);
                        if (!pb) break;
                        it.src[p] = M.w + plane_off + r * upr * pb;
                        it.bytes[p] = (uint32_t) ((int64_t) it.nrows * upr * pb);
                        plane_off += (int64_t) M.m * upr * pb;
                    }
                    if (map0 >= 0) {              // src[0], src[1]: offsets of the tensor maps for now (blob address not known yet)
                        it.type = CH_T_Q6K_SW; it.pad = (long long) (r * upr) | ((long long) R << 32);
                        it.src[0] = (const uint8_t *) (uintptr_t) ((size_t) map0 * sizeof(CUtensorMap));
                        it.src[1] = (const uint8_t *) (uintptr_t) ((size_t) (map0 + 1) * sizeof(CUtensorMap));
                    }
                    out.push_back(it);
                }
            }
            pm = cm; prow = crow;
        }
        idx += cnt;
    }
    std::vector<ChainItemD> hi;
    std::vector<int> hoff((size_t) VC * (n_groups + 1));
    for (int vc = 0; vc < VC; ++vc) {
        for (int g = 0; g < n_groups; ++g) {
            hoff[(size_t) vc * (n_groups + 1) + g] = (int) hi.size();
            const std::vector<ChainItemD> & v = per_vc_node[(size_t) vc * n_groups + g];
            hi.insert(hi.end(), v.begin(), v.end());
        }
        hoff[(size_t) vc * (n_groups + 1) + n_groups] = (int) hi.size();
    }

    auto al = [](size_t v) { return (v + 255) & ~(size_t) 255; };
    const size_t o_nodes = 256, o_items = o_nodes + al(hn.size() * sizeof(ChainNodeD)),
                 o_off = o_items + al(hi.size() * sizeof(ChainItemD) + 16), o_maps = o_off + al(hoff.size() * sizeof(int)),
                 total = o_maps + al(maps.size() * sizeof(CUtensorMap));
    std::vector<uint8_t> blob(total, 0);
    if (!maps.empty()) memcpy(blob.data() + o_maps, maps.data(), maps.size() * sizeof(CUtensorMap));
    memcpy(blob.data() + o_nodes, hn.data(), hn.size() * sizeof(ChainNodeD));
    if (!hi.empty()) memcpy(blob.data() + o_items, hi.data(), hi.size() * sizeof(ChainItemD));
    memcpy(blob.data() + o_off, hoff.data(), hoff.size() * sizeof(int));
    void * d = nullptr;
    if (cudaMalloc(&d, total) == cudaSuccess) {
        ChainItemD * bi = (ChainItemD *) (blob.data() + o_items);
        for (size_t t = 0; t < hi.size(); ++t)
            if (bi[t].type == CH_T_Q6K_SW) for (int p = 0; p < 2; ++p) bi[t].src[p] = (const uint8_t *) d + o_maps + (size_t) (uintptr_t) bi[t].src[p];
    }
    if (!d || cudaMemcpy(d, blob.data(), total, cudaMemcpyHostToDevice) != cudaSuccess) {
        b200_set_error("b200_chain_create: cudaMalloc/cudaMemcpy of %zu bytes failed: %s", total, cudaGetErrorString(cudaGetLastError()));
        if (d) cudaFree(d);
        if (!split) cudaFree(d_tags);
        return nullptr;
    }
    Chain * c = new Chain();
    c->d_tags = split ? nullptr : d_tags;
    c->device = device; c->sms = sms; c->ncols = ncols; c->n_nodes = n_groups; c->d_blob = d;
    c->xreg = P.xreg;
    c->kset = kset;
    uint8_t * db = (uint8_t *) d;
    // single GPU: counters in the chain's own blob; row split: in the symmetric arena (peers add to them)
    c->args.counter = split ? (unsigned long long *) split->scratch : (unsigned long long *) db;
    c->args.done = split ? (unsigned long long *) split->scratch + 8 : (unsigned long long *) db + 8;
    c->args.launches = (unsigned long long *) (db + 128);
    c->args.world = world; c->args.rank = split ? split->rank : 0;
    for (int g = 0; g < B200_MAX_PEERS; ++g) c->args.delta[g] = split && g < world ? split->peer_delta[g] : 0;
    c->args.per_launch = (unsigned long long) signalling_total * (unsigned) sms * (unsigned) world;
    c->args.counter_local = split ? (unsigned long long *) split->scratch + 16 : (unsigned long long *) db + 16;
    c->args.per_launch_local = (unsigned long long) signalling_local_total * (unsigned) sms;
    c->args.tag_stride = tmax;
    c->args.nodes = (const ChainNodeD *) (db + o_nodes);
    c->args.items = (const ChainItemD *) (db + o_items);
    c->args.node_off = (const int *) (db + o_off);
    if (getenv("B200_CHAIN_TIMELINE")) {
        void * t = nullptr;
        if (cudaMalloc(&t, (size_t) sms * n_groups * 64) == cudaSuccess) { cudaMemset(t, 0, (size_t) sms * n_groups * 64); c->args.timeline = (long long *) t; }
    }
    c->args.inflight = getenv("B200_CHAIN_INFLIGHT") ? atoi(getenv("B200_CHAIN_INFLIGHT")) : CH_NS;
    if (c->args.inflight < 1) c->args.inflight = 1;
    c->args.debug = getenv("B200_CHAIN_DEBUG") ? atoi(getenv("B200_CHAIN_DEBUG")) : 0;
    c->args.lanes8 = lanes8 ? 1 : 0;
    c->args.n_nodes = n_groups; c->args.stage_bytes = stage; c->args.max_units = max_units; c->args.x_bytes = x_bytes; c->args.ns = P.ns;
    c->smem = P.smem;
    c->cooperative = coop != 0 && !getenv("B200_CHAIN_NO_COOP");
    cudaError_t e = cudaSuccess;
    if (c->xreg) e = CX_NW == 12 ? cudaFuncSetAttribute(gemv_chainx_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem)
                                 : cudaFuncSetAttribute(gemv_chainx_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem);
    else switch (ncols) {
        case 1: e = kset == 3 ? cudaFuncSetAttribute(gemv_chain_kernel<1, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 2 ? cudaFuncSetAttribute(gemv_chain_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 1 ? cudaFuncSetAttribute(gemv_chain_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem)
                                  : cudaFuncSetAttribute(gemv_chain_kernel<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem); break;
        case 2: e = kset == 3 ? cudaFuncSetAttribute(gemv_chain_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 2 ? cudaFuncSetAttribute(gemv_chain_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 1 ? cudaFuncSetAttribute(gemv_chain_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem)
                                  : cudaFuncSetAttribute(gemv_chain_kernel<2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem); break;
        case 3: e = kset == 3 ? cudaFuncSetAttribute(gemv_chain_kernel<3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 2 ? cudaFuncSetAttribute(gemv_chain_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 1 ? cudaFuncSetAttribute(gemv_chain_kernel<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem)
                                  : cudaFuncSetAttribute(gemv_chain_kernel<3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem); break;
        default: e = kset == 3 ? cudaFuncSetAttribute(gemv_chain_kernel<4, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 2 ? cudaFuncSetAttribute(gemv_chain_kernel<4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem) :
                        kset == 1 ? cudaFuncSetAttribute(gemv_chain_kernel<4, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem)
                                  : cudaFuncSetAttribute(gemv_chain_kernel<4, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, total_smem); break;
    }
    if (e != cudaSuccess) {
        b200_set_error("b200_chain_create: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        cudaGetLastError(); cudaFree(d); if (!split) cudaFree(d_tags); delete c; return nullptr;
    }
    return (b200_chain_t) c;
}

int b200_chain_launch(b200_chain_t chain, b200_stream_t stream) {
    B200_REQUIRE(chain, "b200_chain_launch: null chain");
    Chain * c = (Chain *) chain;
    cudaStream_t s = (cudaStream_t) stream;
    cudaError_t e;
    if (c->xreg) e = chainx_launch(*c, s);
    else switch (c->ncols) {
        case 1: e = c->kset == 3 ? chain_launch_t<1, 3>(*c, s) : c->kset == 2 ? chain_launch_t<1, 2>(*c, s) : c->kset == 1 ? chain_launch_t<1, 1>(*c, s) : chain_launch_t<1, 0>(*c, s); break;
        case 2: e = c->kset == 3 ? chain_launch_t<2, 3>(*c, s) : c->kset == 2 ? chain_launch_t<2, 2>(*c, s) : c->kset == 1 ? chain_launch_t<2, 1>(*c, s) : chain_launch_t<2, 0>(*c, s); break;
        case 3: e = c->kset == 3 ? chain_launch_t<3, 3>(*c, s) : c->kset == 2 ? chain_launch_t<3, 2>(*c, s) : c->kset == 1 ? chain_launch_t<3, 1>(*c, s) : chain_launch_t<3, 0>(*c, s); break;
        default: e = c->kset == 3 ? chain_launch_t<4, 3>(*c, s) : c->kset == 2 ? chain_launch_t<4, 2>(*c, s) : c->kset == 1 ? chain_launch_t<4, 1>(*c, s) : chain_launch_t<4, 0>(*c, s); break;
    }
    B200_CHECK(e);
    return 0;
}

int b200_chain_info(b200_chain_t chain, int * ctas, int * stage_bytes, int * smem_bytes, int * n_nodes) {
    B200_REQUIRE(chain, "b200_chain_info: null chain");
    Chain * c = (Chain *) chain;
    if (ctas) *ctas = c->sms;
    if (stage_bytes) *stage_bytes = c->args.stage_bytes;
    if (smem_bytes) *smem_bytes = (int) c->smem;
    if (n_nodes) *n_nodes = c->n_nodes;
    return 0;
}

/* debugging aid (B200_CHAIN_TIMELINE=1 at create time): copies ctas x n_groups x 8 clock64 samples of the last launch */
int b200_chain_timeline(b200_chain_t chain, long long * host, size_t count) {
    B200_REQUIRE(chain && host, "b200_chain_timeline: null pointer");
    Chain * c = (Chain *) chain;
    B200_REQUIRE(c->args.timeline, "b200_chain_timeline: the chain was created without B200_CHAIN_TIMELINE");
    const size_t have = (size_t) c->sms * c->n_nodes * 8;
    B200_CHECK(cudaMemcpy(host, c->args.timeline, (count < have ? count : have) * 8, cudaMemcpyDeviceToHost));
    return 0;
}

int b200_chain_destroy(b200_chain_t chain) {
    if (!chain) return 0;
    Chain * c = (Chain *) chain;
    cudaSetDevice(c->device);
    cudaFree(c->d_blob);
    cudaFree(c->d_tags);
    if (c->args.timeline) cudaFree(c->args.timeline);
    delete c;
    return 0;
}

} // extern "C"

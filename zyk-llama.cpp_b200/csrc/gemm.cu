// gemm.cu -- (2) prefill GEMM: Y[n][m] = sum_k W[m][k] X[n][k] for n > 8, quantised W, on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), sm_100a only.
//
// Reference counterparts: ggml_compute_forward_mul_mat + ggml_gemm_q4_K_8x8_q8_K on the CPU
// (ggml-cpu/ggml-cpu.c:1227-1427, ggml-cpu/repack.cpp:1037-1124) and mul_mat_q (mma.sync int8) on
// CUDA (ggml-cuda/mmq.cuh:3459). This kernel is neither:
//
//   * persistent CTAs (one per SM, 640 threads) walk a unit list over the tiles of up to three weight matrices of one type that read
//     the same X (b200_gemm_launch_group); a tile is 256 weight rows x BN <= 256 activation rows on a CTA PAIR
//     (tcgen05.mma.cta_group::2, M = 256) or 128 x BN on one CTA; the K loop advances 64 elements a stage; the tiles of the last
//     partial wave are cut along k and meet in pre-zeroed Y with red.global.add;
//   * A (weights): one warp issues TMA copies of the RAW quantised planes (one 256-element super-block of 128 rows per raw stage,
//     SWIZZLE_128B / 64B so that the readers are bank-conflict free), 8 dequantiser warps turn them into the fp16 K-major
//     SWIZZLE_128B operand tile in shared memory (gemm_dequant.cuh) -- every weight byte is read once per N-tile and never exists
//     in HBM in expanded form;
//   * B (activations): a prologue kernel rescales each f32 row by a power of two and rounds it to fp16 (exact scaling, 11-bit
//     mantissa; no int8 step); TMA (cp.async.bulk.tensor, SWIZZLE_128B) stages 64 x BN boxes, each CTA of a pair half of it;
//   * one elected thread of the leader CTA issues tcgen05.mma.kind::f16 (K = 16) four times per stage, tcgen05.commit (multicast
//     to both CTAs) releases the stage; fp32 accumulation over the whole K lives in TMEM, TWO accumulators so that
//   * the epilogue of unit i (8 warps, two per TMEM lane quarter: tcgen05.ld, per-row power-of-two scale, coalesced 128-byte
//     stores) overlaps the MMAs of unit i + 1;
//   * programmatic dependent launch: weights are staged and dequantised before griddepcontrol.wait, only the activation TMA
//     and the Y stores come after it.
// Measured behaviour, ablations and what bounds the kernel: DESIGN.md section 5.
//
// Why fp16 operands instead of int8 x int8: the block formats carry a scale per 32 (or 16) weights
// and the CPU applies one activation scale per 32/256 elements, so an int8 MMA would need a TMEM
// round trip every K = 32..256 (epilogue bound), or weights folded with their sub-block scale, which
// needs 10 bits. Folding the scales into fp16 operands is exact for the integers, keeps a single
// epilogue and is closer to the f32 product than the CPU's own Q8 activation path.
#include "gemm.h"
#include "gemm_dequant.cuh"
#include "sm100.cuh"

#include <cuda.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

using namespace sm100;

namespace {

constexpr int BM = 128;            // weight rows per tile = UMMA M
constexpr int BK = 64;             // k-elements per stage = one 128-byte swizzle row of fp16
constexpr int A_STAGE = BM * BK * 2;
constexpr int DEQ_WARPS = 8;
constexpr int EPI_WARPS = 8;              // two warps per TMEM lane quarter: they take alternate 32-column blocks of the accumulator
constexpr int THREADS = 128 + DEQ_WARPS * 32 + EPI_WARPS * 32;   // warps 0..3: TMA X, MMA, TMEM, TMA W | 4..11: dequantisers | 12..19: epilogue
constexpr int MAX_STAGES = 8;
constexpr size_t SMEM_BUDGET = 224 * 1024;   // dynamic shared memory per CTA (227 KB usable minus the static barriers)

constexpr int GEMM_MAX_GROUP = 3;   // weight matrices of the same type and k that read the same X in ONE launch (wq/wk/wv, ffn_gate/ffn_up)

struct GemmArgs {
    GemvW W[GEMM_MAX_GROUP];
    int64_t m[GEMM_MAX_GROUP], n, k;
    float * y[GEMM_MAX_GROUP]; int64_t ldy[GEMM_MAX_GROUP];
    int nmat;
    int tile_begin[GEMM_MAX_GROUP + 1];   // the tiles of matrix i are tile_begin[i] .. tile_begin[i + 1] - 1 of the launch's tile list
    const float * xscale;       // per activation row: multiply the accumulator by this
    int bn;                     // activation rows per tile (multiple of 16, <= 256)
    int stages;
    int acc_cols;               // TMEM columns of one accumulator buffer: power of two >= max(32, bn); two buffers are allocated
    int debug;                  // B200_GEMM_DEBUG bit mask (timing experiments only: results are wrong when set)
    unsigned long long * trace; // B200_GEMM_TRACE: globaltimer samples of the leader CTAs, [pair][unit < 8][8] (debugging aid)
    int pdl;                    // launched with programmatic stream serialisation: weights are staged and dequantised before griddepcontrol.wait,
                                //   the activation loads and the Y stores come after it
    // work list of the persistent grid: units 0..full_tiles-1 are whole tiles (plain stores); the remaining tiles are cut
    // `ksplit` ways along k (units full_tiles + (tile - full_tiles) * ksplit + z) and meet in pre-zeroed Y with red.global.add
    int n_tiles;                // tiles along n; tile index = m_tile * n_tiles + n_tile (the n-tiles of one weight tile run side by side)
    int full_tiles, ksplit, total_units;
    // MUL_MAT_ID (MOE instantiations): the unit list is built on the device by moe_plan_kernel from the expert ids
    const int4 * moe_units;     // [*moe_total] {expert, first weight row of the tile (before the CTA's rank), first gathered X row, X rows that exist}
    const int * moe_total;      // number of units
    const int * moe_yoff;       // per gathered X row: element offset of its Y row (token * y_token_stride + slot * y_row_stride)
    long long expert_stride;    // bytes between the expert matrices (each one in the plane layout)
    PeerGather peer;            // row split (world > 1): the epilogue stores every Y value into every rank's arena, the last CTA announces the call
};

// ---- prologue: X f32 [n][k] (row stride ldx) -> fp16 [n][k] scaled by 2^-e per row, xscale[n] = 2^e ----
// VMAX: 128-bit loads per thread the one-pass path may hold (k <= 1024 VMAX); small VMAX = few registers = the whole launch resident at once
template <int VMAX>
__global__ void __launch_bounds__(256) act_half_kernel(const float * __restrict__ x, int64_t ldx, int64_t k, __half * __restrict__ xh,
                                                       float * __restrict__ xscale, bool pdl, const long long * __restrict__ src_off) {
    if (pdl) { pdl_launch_dependents(); pdl_wait(); }
    const int64_t row = blockIdx.x;
    // src_off (MUL_MAT_ID): output row `row` is the activation row at element offset src_off[row] (rows gathered by expert)
    const float * xr = src_off ? x + src_off[row] : x + row * ldx;
    __shared__ float red[8];
    // one pass when the row fits the registers of the CTA (k <= 16384, 16-byte aligned): 128-bit loads, the row is read once
    const bool vec = k <= 256 * 4 * VMAX && (k & 3) == 0 && (((uintptr_t) xr) & 15) == 0;
    float4 v[VMAX];
    float amax = 0.0f;
    if (vec) {
        const int nv = (int) (k >> 2);
#pragma unroll
        for (int i = 0; i < VMAX; ++i) {
            const int idx = threadIdx.x + i * 256;
            v[i] = idx < nv ? __ldg(reinterpret_cast<const float4 *>(xr) + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[i].x), fabsf(v[i].y)), fmaxf(fabsf(v[i].z), fabsf(v[i].w))));
        }
    } else {
        for (int64_t i = threadIdx.x; i < k; i += 256) amax = fmaxf(amax, fabsf(xr[i]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
    __syncthreads();
    amax = red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) amax = fmaxf(amax, red[i]);
    // amax * 2^-e in [2^14, 2^15): well inside fp16 range, 11 significant bits for the large elements
    int e = 0;
    if (amax > 0.0f && amax < 3.0e38f) { (void) frexpf(amax, &e); e -= 15; }
    const float down = ldexpf(1.0f, -e);
    if (vec) {
        const int nv = (int) (k >> 2);
        uint2 * out = reinterpret_cast<uint2 *>(xh + row * k);       // k % 4 == 0 and the workspace is 256 B aligned: 8-byte stores
#pragma unroll
        for (int i = 0; i < VMAX; ++i) {
            const int idx = threadIdx.x + i * 256;
            if (idx < nv) {
                const __half2 lo = __floats2half2_rn(v[i].x * down, v[i].y * down), hi = __floats2half2_rn(v[i].z * down, v[i].w * down);
                out[idx] = make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
            }
        }
    } else {
        for (int64_t i = threadIdx.x; i < k; i += 256) xh[row * k + i] = __float2half_rn(xr[i] * down);
    }
    if (threadIdx.x == 0) xscale[row] = ldexpf(1.0f, e);
}

// zeroes y[r][m_lo .. m) for r < n: the rows the k-split tiles accumulate into (a kernel, not a memset node, so that the
// programmatic launch edges between prologue, this and the GEMM survive stream capture)
__global__ void __launch_bounds__(256) zero_rows_kernel(float * __restrict__ y, int64_t ldy, int64_t m_lo, int64_t m, bool pdl) {
    if (pdl) { pdl_launch_dependents(); pdl_wait(); }
    float * yr = y + (int64_t) blockIdx.x * ldy;
    for (int64_t i = m_lo + threadIdx.x; i < m; i += 256) yr[i] = 0.0f;
}

// ---- raw weight staging: one stage = 128 rows x one 256-element super-block, in the plane layout ----------
// Each region is what one cp.async.bulk.tensor box writes: [128 rows][box bytes]; 128-byte boxes use
// SWIZZLE_128B and the 64-byte box SWIZZLE_64B so that 32 lanes reading 16 bytes of 32 consecutive rows
// hit distinct banks (gemm_dequant.cuh sw128/sw64); 144- and 16-byte rows are conflict-free as they are.
struct RawLoad { int map, smem_off, cstep, coff, bytes; };
template <int TYPE> struct Raw;
template <> struct Raw<B200_TYPE_Q4_K>   { static constexpr int BYTES = 18432, N = 1; static constexpr RawLoad L[3] = {{0, 0, 144, 0, 18432}, {}, {}}; };
template <> struct Raw<B200_TYPE_Q5_K>   { static constexpr int BYTES = 22528, N = 1; static constexpr RawLoad L[3] = {{0, 0, 176, 0, 22528}, {}, {}}; };
template <> struct Raw<B200_TYPE_Q6_K>   { static constexpr int BYTES = 26624, N = 3; static constexpr RawLoad L[3] = {{0, 0, 128, 0, 16384}, {1, 16384, 64, 0, 8192}, {2, 24576, 16, 0, 2048}}; };
template <> struct Raw<B200_TYPE_IQ4_XS> { static constexpr int BYTES = 16384, N = 1; static constexpr RawLoad L[3] = {{0, 0, 128, 0, 16384}, {}, {}}; };
template <> struct Raw<B200_TYPE_Q4_0>   { static constexpr int BYTES = 18432, N = 2; static constexpr RawLoad L[3] = {{0, 0, 128, 0, 16384}, {1, 16384, 16, 0, 2048}, {}}; };
template <> struct Raw<B200_TYPE_Q8_0>   { static constexpr int BYTES = 34816, N = 3; static constexpr RawLoad L[3] = {{0, 0, 256, 0, 16384}, {0, 16384, 256, 128, 16384}, {1, 32768, 16, 0, 2048}}; };
constexpr int RAW_STAGES = 2;
#ifndef B200_GEMM_PAIR_DEFAULT
#define B200_GEMM_PAIR_DEFAULT true
#endif

// ---- main kernel --------------------------------------------------------------------------------------
// Persistent: one CTA per SM walks the unit list (unit = tile, or k-slice of a tail tile) with stride gridDim.x.
// warp 0: TMA of the activation boxes | warp 1: MMA issuer | warp 2: TMEM allocator | warp 3: TMA of the raw weight
// planes | warps 4..11: dequantise raw -> fp16 operand tiles | warps 12..15: epilogue. Two accumulator buffers in
// TMEM: the epilogue of unit i (TMEM -> registers -> Y, 128 KB of stores) overlaps the MMAs of unit i + 1. All
// rings (operand stages, raw stages, accumulators) keep running across units.
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
struct StageCursor {     // ring position without a division per step
    int s; uint32_t ph; int stages;
    __device__ __forceinline__ StageCursor(int stages_, int start) : s(start % stages_), ph((uint32_t) (start / stages_) & 1u), stages(stages_) {}
    __device__ __forceinline__ void advance(int n) { s += n; if (s >= stages) { s -= stages; ph ^= 1u; } }   // n <= stages
};

// the weight-plane tensor maps of every matrix of the launch (kernel parameter, __grid_constant__: TMA reads them in place)
struct WMapSet { CUtensorMap m[GEMM_MAX_GROUP][3]; };

struct Unit { int64_t m0; int n0, sb_begin, num_sb, mat, n_left; bool split; };
template <bool PAIR, bool MOE = false>
__device__ __forceinline__ Unit unit_decode(const GemmArgs & a, int u, uint32_t rank) {
    if (MOE) {
        const int4 d = __ldg(a.moe_units + u);
        Unit r;
        r.mat = d.x; r.m0 = (int64_t) d.y + (int64_t) rank * BM; r.n0 = d.z; r.n_left = d.w;
        r.sb_begin = 0; r.num_sb = (int) (a.k / 256); r.split = false;
        return r;
    }
    int tile = u, z = 0, ks = 1;
    if (u >= a.full_tiles) { const int v = u - a.full_tiles; tile = a.full_tiles + v / a.ksplit; z = v % a.ksplit; ks = a.ksplit; }
    const int total_sb = (int) (a.k / 256);
    Unit r;
    r.mat = (a.nmat > 1 && tile >= a.tile_begin[1]) + (a.nmat > 2 && tile >= a.tile_begin[2]);
    tile -= a.tile_begin[r.mat];
    r.m0 = (int64_t) (tile / a.n_tiles) * (PAIR ? 2 * BM : BM) + (int64_t) rank * BM;
    r.n0 = (tile % a.n_tiles) * a.bn;
    r.sb_begin = (int) ((int64_t) total_sb * z / ks);
    r.num_sb = (int) ((int64_t) total_sb * (z + 1) / ks) - r.sb_begin;
    r.split = ks > 1;
    r.n_left = (int) (a.n - r.n0 < a.bn ? a.n - r.n0 : a.bn);
    return r;
}

// 3-D tiled load (MUL_MAT_ID weight planes: [expert][row][bytes])
__device__ __forceinline__ void tma_load_3d(uint32_t smem_dst, const void * tmap, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 :: "r"(smem_dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

template <int TYPE, bool PAIR, bool MOE = false>
__device__ __forceinline__ void gemm_body(const CUtensorMap & tmap_x, const WMapSet & wmaps, const GemmArgs & a) {
    // MOE: the number of units comes from the plan kernel that ran before this one (plain stream order, no programmatic launch)
    const int total_units = MOE ? *a.moe_total : a.total_units;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_full[MAX_STAGES], bar_empty[MAX_STAGES], raw_full[RAW_STAGES], raw_empty[RAW_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // PAIR: the two CTAs of a cluster (one TPC) share one cta_group::2 tile of 256 weight rows: each dequantises its own 128
    // rows and stages HALF of the activation box; the leader issues the MMAs for both and its commits release both CTAs.
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    const int unit0 = PAIR ? (int) (blockIdx.x >> 1) : (int) blockIdx.x, unit_step = PAIR ? (int) (gridDim.x >> 1) : (int) gridDim.x;
    const int stages = a.stages, bn = a.bn;
    const int bn_cta = PAIR ? bn / 2 : bn;                             // activation rows this CTA stages
    const uint32_t b_stage = (uint32_t) bn_cta * 128u;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;      // swizzled tiles need 1024 B alignment
    const uint32_t smem_w = smem0;
    const uint32_t smem_a = smem0 + (uint32_t) (RAW_STAGES * Raw<TYPE>::BYTES), smem_b = smem_a + (uint32_t) stages * A_STAGE;

    if (a.pdl) pdl_launch_dependents();      // the next kernel may start its own prologue; it waits for this grid before touching memory
    if (warp == 0 && lane == 0) prefetch_tensormap(&tmap_x);
    if (warp == 3 && lane == 0) {
        for (int i = 0; i < a.nmat; ++i) {
            prefetch_tensormap(&wmaps.m[i][0]);
            if (Raw<TYPE>::N > 1) prefetch_tensormap(&wmaps.m[i][1]);
            if (TYPE == B200_TYPE_Q6_K) prefetch_tensormap(&wmaps.m[i][2]);
        }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(smem_u32(&bar_full[s]), 1 + (PAIR ? 2 : 1) * (DEQ_WARPS / 2));   // TMA expect_tx arrive + 4 producer warps (of each CTA)
            mbar_init(smem_u32(&bar_empty[s]), 1);                  // tcgen05.commit
        }
        for (int s = 0; s < RAW_STAGES; ++s) {
            mbar_init(smem_u32(&raw_full[s]), 1);                   // expect_tx arrive
            mbar_init(smem_u32(&raw_empty[s]), DEQ_WARPS);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&acc_full[s]), 1);                   // tcgen05.commit after the last MMA of a unit
            mbar_init(smem_u32(&acc_empty[s]), (PAIR ? 2 : 1) * EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        if (PAIR) { tmem_alloc_pair(smem_u32(&tmem_base_smem), (uint32_t) (2 * a.acc_cols)); tmem_relinquish_pair(); }
        else      { tmem_alloc(smem_u32(&tmem_base_smem), (uint32_t) (2 * a.acc_cols)); tmem_relinquish(); }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();   // both CTAs' barriers exist before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_smem;

    if (warp == 0) {
        // ===== TMA producer: activation boxes (64 k x bn rows) =====
        if (lane == 0) {
            if (a.pdl) pdl_wait();               // X (fp16 copy) is written by the kernel before this one
            StageCursor c(stages, 0);
            for (int u = unit0; u < total_units; u += unit_step) {
                const Unit w = unit_decode<PAIR, MOE>(a, u, rank);
                for (int kb = 0; kb < 4 * w.num_sb; ++kb, c.advance(1)) {
                    mbar_wait(smem_u32(&bar_empty[c.s]), c.ph ^ 1u);
                    if (a.debug & 1) { if (leader) mbar_arrive(smem_u32(&bar_full[c.s])); continue; }
                    if (leader) mbar_arrive_expect_tx(smem_u32(&bar_full[c.s]), (PAIR ? 2u : 1u) * b_stage);
                    if (PAIR) tma_load_2d_pair(smem_b + (uint32_t) c.s * b_stage, &tmap_x, (4 * w.sb_begin + kb) * BK, w.n0 + (int) rank * bn_cta, smem_u32(&bar_full[c.s]));
                    else      tma_load_2d(smem_b + (uint32_t) c.s * b_stage, &tmap_x, (4 * w.sb_begin + kb) * BK, w.n0, smem_u32(&bar_full[c.s]));
                }
            }
        }
    } else if (warp == 3) {
        // ===== TMA producer: raw quantised planes, one super-block (256 k) of 128 rows per stage =====
        if (lane == 0) {
            StageCursor c(RAW_STAGES, 0);
            for (int u = unit0; u < total_units; u += unit_step) {
                const Unit w = unit_decode<PAIR, MOE>(a, u, rank);
                for (int sb = 0; sb < w.num_sb; ++sb, c.advance(1)) {
                    mbar_wait(smem_u32(&raw_empty[c.s]), c.ph ^ 1u);
                    if (a.debug & 4) { mbar_arrive(smem_u32(&raw_full[c.s])); continue; }
                    mbar_arrive_expect_tx(smem_u32(&raw_full[c.s]), (uint32_t) Raw<TYPE>::BYTES);
#pragma unroll
                    for (int i = 0; i < Raw<TYPE>::N; ++i) {
                        constexpr RawLoad L0 = Raw<TYPE>::L[0], L1 = Raw<TYPE>::L[1], L2 = Raw<TYPE>::L[2];
                        const RawLoad L = i == 0 ? L0 : i == 1 ? L1 : L2;
                        const CUtensorMap * tm = &wmaps.m[MOE ? 0 : w.mat][L.map];
                        if (MOE) tma_load_3d(smem_w + (uint32_t) (c.s * Raw<TYPE>::BYTES + L.smem_off), tm, (w.sb_begin + sb) * L.cstep + L.coff, (int) w.m0, w.mat,
                                             smem_u32(&raw_full[c.s]));
                        else     tma_load_2d(smem_w + (uint32_t) (c.s * Raw<TYPE>::BYTES + L.smem_off), tm, (w.sb_begin + sb) * L.cstep + L.coff, (int) w.m0,
                                             smem_u32(&raw_full[c.s]));
                    }
                }
            }
        }
    } else if (warp == 1 && leader) {
        // ===== MMA issuer (leader CTA only when paired) =====
        const uint32_t idesc = umma_idesc_f16(PAIR ? 2 * BM : BM, bn);
        StageCursor c(stages, 0);
        int i = 0;
        for (int u = unit0; u < total_units; u += unit_step, ++i) {
            const Unit w = unit_decode<PAIR, MOE>(a, u, rank);
            const int num_kb = 4 * w.num_sb;
            const int buf = i & 1;
            const uint32_t tmem_acc = tmem_d + (uint32_t) (buf * a.acc_cols);
            unsigned long long * tr = a.trace && lane == 0 && i < 8 ? a.trace + ((size_t) (blockIdx.x >> (PAIR ? 1 : 0)) * 8 + i) * 8 : nullptr;
            if (tr) tr[0] = gtime();
            mbar_wait(smem_u32(&acc_empty[buf]), (uint32_t) ((i >> 1) & 1) ^ 1u);     // the epilogue has drained this buffer
            tc_fence_after();
            if (tr) tr[1] = gtime();
            for (int kb = 0; kb < num_kb; ++kb, c.advance(1)) {
                mbar_wait(smem_u32(&bar_full[c.s]), c.ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(smem_a + (uint32_t) c.s * A_STAGE);
                    const uint64_t bdesc = umma_desc_kmajor_sw128(smem_b + (uint32_t) c.s * b_stage);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {   // 16 fp16 = 32 bytes = 2 descriptor units along K inside the swizzle atom
                        if (PAIR) mma_f16_ss_pair(tmem_acc, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t) ((kb | k) != 0));
                        else      mma_f16_ss(tmem_acc, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t) ((kb | k) != 0));
                    }
                    if (PAIR) {
                        mma_commit_pair(smem_u32(&bar_empty[c.s]), 3);
                        if (kb == num_kb - 1) mma_commit_pair(smem_u32(&acc_full[buf]), 3);
                    } else {
                        mma_commit(smem_u32(&bar_empty[c.s]));
                        if (kb == num_kb - 1) mma_commit(smem_u32(&acc_full[buf]));
                    }
                }
                __syncwarp();
                if (tr && kb == 0) tr[2] = gtime();
            }
            if (tr) tr[3] = gtime();
        }
    } else if (warp >= 4 && warp < 4 + DEQ_WARPS) {
        // ===== weight producers: thread = tile row; the two halves of the group alternate 64-element k-blocks =====
        const int t = threadIdx.x - 128;
        const int half = t >> 7, r = t & 127;
        StageCursor c(stages, half), rc(RAW_STAGES, 0);
        for (int u = unit0; u < total_units; u += unit_step) {
            const Unit w = unit_decode<PAIR, MOE>(a, u, rank);
            const int64_t m_rows = a.m[MOE ? 0 : w.mat];
            const int64_t row = w.m0 + r < m_rows ? w.m0 + r : m_rows - 1;
            GemvW W = a.W[MOE ? 0 : w.mat];
            if (MOE) { const long long eo = (long long) w.mat * a.expert_stride; W.p0 += eo; W.p1 += eo; W.p2 += eo; W.p3 += eo; }
            uint2 aux_next = Deq<TYPE>::aux_load(W, row, w.sb_begin);
            for (int sb = 0; sb < w.num_sb; ++sb, rc.advance(1)) {
                const uint2 aux = aux_next;
                if (sb + 1 < w.num_sb) aux_next = Deq<TYPE>::aux_load(W, row, w.sb_begin + sb + 1);
                const uint32_t raw = smem_w + (uint32_t) (rc.s * Raw<TYPE>::BYTES);
                mbar_wait(smem_u32(&raw_full[rc.s]), rc.ph);
#pragma unroll
                for (int jj = 0; jj < 2; ++jj, c.advance(2)) {
                    const int j = 2 * jj + half;
                    Deq<TYPE> cur;
                    cur.load_smem(raw, r, j, aux);
                    mbar_wait(smem_u32(&bar_empty[c.s]), c.ph ^ 1u);
                    if (!(a.debug & 2)) cur.store(smem_a + (uint32_t) c.s * A_STAGE + (uint32_t) r * 128u, r, j);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        if (PAIR && !leader) mbar_arrive_cluster(smem_u32(&bar_full[c.s]), 0);
                        else mbar_arrive(smem_u32(&bar_full[c.s]));
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&raw_empty[rc.s]));
            }
        }
    } else if (warp >= 4 + DEQ_WARPS) {
        // ===== epilogue: TMEM -> registers -> Y (thread = weight row, 16 activation rows per load) =====
        const int q = warp & 3;                                  // TMEM lane quarter this warp may touch
        const int eh = (warp - 4 - DEQ_WARPS) >> 2;              // which of the quarter's two warps: column blocks eh, eh + 2, ...
        int i = 0;
        if (a.pdl) pdl_wait();                   // Y may still be read by earlier kernels; xscale comes from the prologue kernel
        for (int u = unit0; u < total_units; u += unit_step, ++i) {
            const Unit w = unit_decode<PAIR, MOE>(a, u, rank);
            const int buf = i & 1;
            const uint32_t tmem_acc = tmem_d + (uint32_t) (buf * a.acc_cols) + ((uint32_t) (32 * q) << 16);
            const int64_t mrow = w.m0 + 32 * q + lane;
            // lane l carries the scales of activation rows 32 * c8 + l of the tile: all loads are issued before the wait for the
            // accumulator, so their latency hides behind the MMAs of the unit (one load per iteration, consumed by the next one,
            // made the epilogue a chain of 8 L2 round trips: ~7 us per unit, exposed at the end of every launch)
            const int n_left = w.n_left;                                              // activation rows of this tile that exist (>= 1)
            float xsv[8];
            int yov[MOE ? 8 : 1];                                                     // MOE: where the Y row of each gathered X row lives
#pragma unroll
            for (int c8 = 0; c8 < 8; ++c8) {
                const int nn = c8 * 32 + lane;
                xsv[c8] = c8 * 32 < bn ? __ldg(a.xscale + w.n0 + (nn < n_left ? nn : n_left - 1)) : 0.0f;
                if (MOE) yov[c8] = c8 * 32 < bn ? __ldg(a.moe_yoff + w.n0 + (nn < n_left ? nn : n_left - 1)) : 0;
            }
            unsigned long long * tr = a.trace && leader && q == 0 && eh == 0 && lane == 0 && i < 8 ? a.trace + ((size_t) (blockIdx.x >> (PAIR ? 1 : 0)) * 8 + i) * 8 : nullptr;
            if (tr) tr[4] = gtime();
            mbar_wait(smem_u32(&acc_full[buf]), (uint32_t) ((i >> 1) & 1));
            tc_fence_after();
            if (tr) tr[5] = gtime();
            const bool row_ok = mrow < a.m[MOE ? 0 : w.mat] && !(a.debug & 16);
            const int64_t ldy = MOE ? 0 : a.ldy[w.mat];
            float * yp = MOE ? a.y[0] + (row_ok ? mrow : 0) : a.y[w.mat] + (int64_t) (w.n0 + 32 * eh) * ldy + (row_ok ? mrow : 0);
            for (int cc = eh; cc * 32 < bn; cc += EPI_WARPS / 4, yp += (EPI_WARPS / 4) * 32 * ldy) {     // columns past bn (inside acc_cols) are read and dropped
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_acc + (uint32_t) (cc * 32), v);
                float xs_lane = xsv[0];
                int yo_lane = MOE ? yov[0] : 0;
#pragma unroll
                for (int c8 = 1; c8 < 8; ++c8) if (cc == c8) { xs_lane = xsv[c8]; if (MOE) yo_lane = yov[c8]; }
                tmem_ld_wait();
                if (MOE) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float val = __uint_as_float(v[j]) * __shfl_sync(0xffffffffu, xs_lane, j);
                        const int off = __shfl_sync(0xffffffffu, yo_lane, j);
                        if (row_ok && cc * 32 + j < n_left) yp[off] = val;
                    }
                } else if (!w.split && a.peer.world > 1) {
                    // row split: the value goes to the same place of every rank's arena (NVLink peer stores straight from the epilogue: the
                    // exchange overlaps the MMAs of the next unit instead of following the GEMM as a copy kernel)
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float val = __uint_as_float(v[j]) * __shfl_sync(0xffffffffu, xs_lane, j);
                        if (row_ok && cc * 32 + j < n_left) peer_store(a.peer, yp + (int64_t) j * ldy, val);
                    }
                } else if (!w.split) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float val = __uint_as_float(v[j]) * __shfl_sync(0xffffffffu, xs_lane, j);
                        if (row_ok && cc * 32 + j < n_left) yp[(int64_t) j * ldy] = val;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float val = __uint_as_float(v[j]) * __shfl_sync(0xffffffffu, xs_lane, j);
                        if (row_ok && cc * 32 + j < n_left) atomicAdd(yp + (int64_t) j * ldy, val);      // result unused -> RED.E.ADD.F32
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (tr) tr[6] = gtime();
            if (lane == 0) {
                if (PAIR && !leader) mbar_arrive_cluster(smem_u32(&acc_empty[buf]), 0);
                else mbar_arrive(smem_u32(&acc_empty[buf]));
            }
        }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();   // nobody leaves while the peer may still touch its shared memory / TMEM
    if (!MOE && a.peer.world > 1 && a.peer.signal && threadIdx.x == 0) peer_signal(a.peer);   // every peer store of this CTA precedes the barrier above
    if (warp == 2) {
        if (PAIR) tmem_dealloc_pair(tmem_d, (uint32_t) (2 * a.acc_cols)); else tmem_dealloc(tmem_d, (uint32_t) (2 * a.acc_cols));
    }
}

template <int TYPE>
__global__ void __launch_bounds__(THREADS, 1) gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ WMapSet wmaps, const GemmArgs a) {
    gemm_body<TYPE, false>(tmap_x, wmaps, a);
}

template <int TYPE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ WMapSet wmaps, const GemmArgs a) {
    gemm_body<TYPE, true>(tmap_x, wmaps, a);
}

template <int TYPE>
__global__ void __launch_bounds__(THREADS, 1) gemm_moe_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ WMapSet wmaps, const GemmArgs a) {
    gemm_body<TYPE, false, true>(tmap_x, wmaps, a);
}
template <int TYPE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm_moe_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ WMapSet wmaps, const GemmArgs a) {
    gemm_body<TYPE, true, true>(tmap_x, wmaps, a);
}

// ---- MUL_MAT_ID plan: sort the (token, slot) pairs by expert and lay out the unit list, all on the device ----------------------
// (what mm_ids_helper does for the reference's CUDA backend, ggml-cuda/mmid.cu:28-117, and the row grouping of the CPU driver,
// ggml-cpu/ggml-cpu.c:1589-1602; here the ids never visit the host). ONE CTA: counts per expert in shared memory, an exclusive scan,
// then every valid pair takes the next free row of its expert's range of the gathered activation matrix. The order of the rows
// inside an expert's range depends on the atomics, the VALUE of every output row does not (a row is one column of the MMA).
// Pairs whose id is out of range are left out; moe_zero_invalid_kernel zeroes their Y rows (as the GEMV path does).
struct MoePlanArgs {
    const int32_t * ids; long long ids_row_stride; int n_tokens, n_used, n_as, nb1;
    long long b_row_stride, b_token_stride, y_row_stride, y_token_stride;
    int bn, tile_m, m_tiles, max_units;
    int * row_begin;           // [n_as + 1] first gathered row of every expert
    int * total;               // number of units
    int4 * units;              // [max_units]
    int * yoff;                // [pairs]
    long long * src_off;       // [pairs]
};
__global__ void __launch_bounds__(1024) moe_plan_kernel(const MoePlanArgs p) {
    extern __shared__ int sm[];           // cnt[n_as], begin[n_as + 1], ubeg[n_as + 1]
    int * cnt = sm, * begin = sm + p.n_as, * ubeg = begin + p.n_as + 1;
    const int pairs = p.n_tokens * p.n_used;
    for (int e = threadIdx.x; e < p.n_as; e += blockDim.x) cnt[e] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < pairs; i += blockDim.x) {
        const int t = i / p.n_used, sl = i - t * p.n_used;
        const int e = p.ids[(long long) t * p.ids_row_stride + sl];
        if (e >= 0 && e < p.n_as) atomicAdd(&cnt[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int r = 0, u = 0;
        for (int e = 0; e < p.n_as; ++e) {
            begin[e] = r; ubeg[e] = u;
            r += cnt[e];
            u += p.m_tiles * ((cnt[e] + p.bn - 1) / p.bn);
        }
        begin[p.n_as] = r; ubeg[p.n_as] = u;
        *p.total = u < p.max_units ? u : p.max_units;
    }
    __syncthreads();
    for (int e = threadIdx.x; e <= p.n_as; e += blockDim.x) p.row_begin[e] = begin[e];
    // unit list: expert-major, then weight tile, then the n-tiles of that weight tile side by side (they share the weight tile in L2)
    for (int e = threadIdx.x; e < p.n_as; e += blockDim.x) {
        const int nt = (cnt[e] + p.bn - 1) / p.bn;
        int u = ubeg[e];
        for (int mt = 0; mt < p.m_tiles; ++mt)
            for (int j = 0; j < nt; ++j, ++u)
                if (u < p.max_units) {
                    const int left = cnt[e] - j * p.bn;
                    p.units[u] = make_int4(e, mt * p.tile_m, begin[e] + j * p.bn, left < p.bn ? left : p.bn);
                }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < p.n_as; e += blockDim.x) cnt[e] = 0;          // now the fill cursors
    __syncthreads();
    for (int i = threadIdx.x; i < pairs; i += blockDim.x) {
        const int t = i / p.n_used, sl = i - t * p.n_used;
        const int e = p.ids[(long long) t * p.ids_row_stride + sl];
        if (e < 0 || e >= p.n_as) continue;
        const int pos = begin[e] + atomicAdd(&cnt[e], 1);
        p.yoff[pos] = (int) ((long long) t * p.y_token_stride + (long long) sl * p.y_row_stride);
        p.src_off[pos] = (long long) t * p.b_token_stride + (p.nb1 == 1 ? 0ll : (long long) sl * p.b_row_stride);
    }
    // rows past the valid pairs (ids out of range): the gather still converts a row for them (nobody reads the result): row 0
    for (int pos = begin[p.n_as] + (int) threadIdx.x; pos < pairs; pos += blockDim.x) { p.yoff[pos] = 0; p.src_off[pos] = 0; }
}
// one CTA per (token, slot) pair: zero the Y row of a pair whose expert id is out of range
__global__ void __launch_bounds__(256) moe_zero_invalid_kernel(const int32_t * __restrict__ ids, long long ids_row_stride, int n_used, int n_as,
                                                               float * __restrict__ y, long long y_row_stride, long long y_token_stride, long long m) {
    const int i = blockIdx.x, t = i / n_used, sl = i - t * n_used;
    const int e = ids[(long long) t * ids_row_stride + sl];
    if (e >= 0 && e < n_as) return;
    float * yr = y + (long long) t * y_token_stride + (long long) sl * y_row_stride;
    for (long long j = threadIdx.x; j < m; j += 256) yr[j] = 0.0f;
}

// ---- host side ------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_fn) p;
    }();
    return fn;
}

size_t align256(size_t v) { return (v + 255) & ~(size_t) 255; }

struct WMaps { CUtensorMap m[3]; };

// [rows][inner_bytes] uint8 plane, boxes of box_inner bytes x 128 rows; rows past `rows` are zero-filled by the TMA unit
bool make_plane_map(CUtensorMap * tm, const void * base, uint64_t inner_bytes, uint64_t rows, uint32_t box_inner, CUtensorMapSwizzle sw) {
    const cuuint64_t dims[2] = {inner_bytes, rows};
    const cuuint64_t strides[1] = {inner_bytes};
    const cuuint32_t box[2] = {box_inner, (cuuint32_t) BM};
    const cuuint32_t estr[2] = {1, 1};
    return get_encode_tiled()(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool make_weight_maps(int type, const GemvW & W, int64_t m, WMaps & w) {
    const uint64_t nb = (uint64_t) W.nb, rows = (uint64_t) m;
    bool ok = true;
    switch (type) {
        case B200_TYPE_Q4_K:
            ok = make_plane_map(&w.m[0], W.p0, nb * 144, rows, 144, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q5_K:
            ok = make_plane_map(&w.m[0], W.p0, nb * 176, rows, 176, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q6_K:
            ok = make_plane_map(&w.m[0], W.p0, nb * 128, rows, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map(&w.m[1], W.p1, nb * 64, rows, 64, CU_TENSOR_MAP_SWIZZLE_64B) &&
                 make_plane_map(&w.m[2], W.p2, nb * 16, rows, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            break;
        case B200_TYPE_IQ4_XS:
            ok = make_plane_map(&w.m[0], W.p0, nb * 128, rows, 128, CU_TENSOR_MAP_SWIZZLE_128B);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q4_0:
            ok = make_plane_map(&w.m[0], W.p0, nb * 16, rows, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map(&w.m[1], W.p1, nb * 2, rows, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[2] = w.m[0]; break;
        case B200_TYPE_Q8_0:
            ok = make_plane_map(&w.m[0], W.p0, nb * 32, rows, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map(&w.m[1], W.p1, nb * 2, rows, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[2] = w.m[0]; break;
        default: ok = false;
    }
    return ok;
}

template <int TYPE> constexpr int raw_ring_bytes() { return RAW_STAGES * Raw<TYPE>::BYTES; }
int raw_ring_bytes_of(int type) {
    switch (type) {
        case B200_TYPE_Q4_K:   return raw_ring_bytes<B200_TYPE_Q4_K>();
        case B200_TYPE_Q5_K:   return raw_ring_bytes<B200_TYPE_Q5_K>();
        case B200_TYPE_Q6_K:   return raw_ring_bytes<B200_TYPE_Q6_K>();
        case B200_TYPE_Q8_0:   return raw_ring_bytes<B200_TYPE_Q8_0>();
        case B200_TYPE_Q4_0:   return raw_ring_bytes<B200_TYPE_Q4_0>();
        case B200_TYPE_IQ4_XS: return raw_ring_bytes<B200_TYPE_IQ4_XS>();
    }
    return 0;
}

template <int TYPE>
int launch_gemm(const CUtensorMap & tm, const WMapSet & w, const GemmArgs & g, dim3 grid, size_t smem, cudaStream_t s, bool pair) {
    static unsigned configured = 0;
    if (b200_first_on_device(configured)) {
        B200_CHECK(cudaFuncSetAttribute(gemm_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SMEM_BUDGET + 1024)));
        B200_CHECK(cudaFuncSetAttribute(gemm_pair_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SMEM_BUDGET + 1024)));
    }
    if (pair) B200_CHECK(b200_launch(gemm_pair_kernel<TYPE>, grid, dim3(THREADS), smem, s, g.pdl != 0, tm, w, g));
    else      B200_CHECK(b200_launch(gemm_kernel<TYPE>, grid, dim3(THREADS), smem, s, g.pdl != 0, tm, w, g));
    return 0;
}

} // namespace

size_t b200_gemm_workspace(int type, int64_t n, int64_t k) {
    if (!b200_type_is_weight(type) || k % 256 != 0) return 0;
    return align256((size_t) n * k * sizeof(__half)) + align256((size_t) n * sizeof(float));
}

bool b200_gemm_wants(int type, int64_t m, int64_t n, int64_t k, bool force) {
    if (!b200_type_is_weight(type) || k % 256 != 0 || m <= 0) return false;
    if (k * 2 > (int64_t) 1 << 31 || n > (int64_t) 1 << 30) return false;
    return force ? n >= 1 : n > 8;
}

int b200_gemm_launch(const b200_mul_mat_args * a, cudaStream_t s) { return b200_gemm_launch_group(a, 1, s); }

// ---- MUL_MAT_ID on the tensor cores: expert gather + one grouped GEMM launch ------------------------------------------------------
namespace {
struct MoeShape { int bn, tile_m, m_tiles, max_units; bool pair; int64_t pairs; };
MoeShape moe_shape(int64_t m, int64_t n_as, int64_t pairs) {
    MoeShape h;
    h.pairs = pairs;
    h.pair = m > BM;
    h.tile_m = h.pair ? 2 * BM : BM;
    h.m_tiles = (int) ((m + h.tile_m - 1) / h.tile_m);
    // rows per n-tile: the expected rows per active expert plus four standard deviations of a uniform routing, so that most experts are ONE
    // tile (an n-tile more costs another pass over the expert's weights); multiple of 32, at most 256
    const double act = (double) (pairs < n_as ? pairs : n_as), avg = (double) pairs / act;
    int bn = (int) (avg + 4.0 * sqrt(avg) + 1.0);
    bn = (bn + 31) & ~31;
    h.bn = bn < 32 ? 32 : bn > 256 ? 256 : bn;
    const int64_t nt = pairs / h.bn + (pairs < n_as ? pairs : n_as);       // sum over experts of ceil(cnt / bn) <= pairs / bn + active experts
    const int64_t mu = nt * h.m_tiles;
    h.max_units = (int) (mu < (1 << 22) ? mu : (1 << 22));
    return h;
}
struct MoeCarve { __half * xh; float * xscale; int * row_begin; int * total; int4 * units; int * yoff; long long * src_off; size_t bytes; };
MoeCarve moe_carve(void * ws, int64_t k, int64_t n_as, const MoeShape & h) {
    MoeCarve c;
    uint8_t * b = (uint8_t *) ws;
    size_t o = 0;
    c.xh = (__half *) (b + o);           o += align256((size_t) (h.pairs + 256) * k * sizeof(__half));      // + one tile of slack rows: boxes may run past the last row
    c.xscale = (float *) (b + o);        o += align256((size_t) h.pairs * sizeof(float));
    c.row_begin = (int *) (b + o);       o += align256((size_t) (n_as + 1) * sizeof(int));
    c.total = (int *) (b + o);           o += 256;
    c.units = (int4 *) (b + o);          o += align256((size_t) h.max_units * sizeof(int4));
    c.yoff = (int *) (b + o);            o += align256((size_t) h.pairs * sizeof(int));
    c.src_off = (long long *) (b + o);   o += align256((size_t) h.pairs * sizeof(long long));
    c.bytes = o;
    return c;
}
template <int TYPE>
int launch_gemm_moe(const CUtensorMap & tm, const WMapSet & w, const GemmArgs & g, dim3 grid, size_t smem, cudaStream_t s, bool pair) {
    static unsigned configured = 0;
    if (b200_first_on_device(configured)) {
        B200_CHECK(cudaFuncSetAttribute(gemm_moe_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SMEM_BUDGET + 1024)));
        B200_CHECK(cudaFuncSetAttribute(gemm_moe_pair_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SMEM_BUDGET + 1024)));
    }
    if (pair) B200_CHECK(b200_launch(gemm_moe_pair_kernel<TYPE>, grid, dim3(THREADS), smem, s, false, tm, w, g));
    else      B200_CHECK(b200_launch(gemm_moe_kernel<TYPE>, grid, dim3(THREADS), smem, s, false, tm, w, g));
    return 0;
}
// [expert][rows][inner_bytes] uint8 plane, boxes of box_inner bytes x 128 rows x 1 expert
bool make_plane_map3(CUtensorMap * tm, const void * base, uint64_t inner_bytes, uint64_t rows, uint64_t experts, uint64_t expert_stride, uint32_t box_inner, CUtensorMapSwizzle sw) {
    const cuuint64_t dims[3] = {inner_bytes, rows, experts};
    const cuuint64_t strides[2] = {inner_bytes, expert_stride};
    const cuuint32_t box[3] = {box_inner, (cuuint32_t) BM, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return get_encode_tiled()(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
bool make_weight_maps3(int type, const GemvW & W, int64_t m, int64_t e, int64_t es, WMaps & w) {
    const uint64_t nb = (uint64_t) W.nb, rows = (uint64_t) m, E = (uint64_t) e, S = (uint64_t) es;
    bool ok = true;
    switch (type) {
        case B200_TYPE_Q4_K:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 144, rows, E, S, 144, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q5_K:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 176, rows, E, S, 176, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q6_K:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 128, rows, E, S, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map3(&w.m[1], W.p1, nb * 64, rows, E, S, 64, CU_TENSOR_MAP_SWIZZLE_64B) &&
                 make_plane_map3(&w.m[2], W.p2, nb * 16, rows, E, S, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            break;
        case B200_TYPE_IQ4_XS:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 128, rows, E, S, 128, CU_TENSOR_MAP_SWIZZLE_128B);
            w.m[1] = w.m[0]; w.m[2] = w.m[0]; break;
        case B200_TYPE_Q4_0:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 16, rows, E, S, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map3(&w.m[1], W.p1, nb * 2, rows, E, S, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[2] = w.m[0]; break;
        case B200_TYPE_Q8_0:
            ok = make_plane_map3(&w.m[0], W.p0, nb * 32, rows, E, S, 128, CU_TENSOR_MAP_SWIZZLE_128B) &&
                 make_plane_map3(&w.m[1], W.p1, nb * 2, rows, E, S, 16, CU_TENSOR_MAP_SWIZZLE_NONE);
            w.m[2] = w.m[0]; break;
        default: ok = false;
    }
    return ok;
}
} // namespace

size_t b200_gemm_moe_workspace(int type, int64_t m, int64_t k, int64_t n_as, int64_t pairs) {
    if (!b200_type_is_weight(type) || k % 256 != 0 || pairs <= 0 || n_as <= 0) return 0;
    const MoeShape h = moe_shape(m, n_as, pairs);
    return moe_carve(nullptr, k, n_as, h).bytes + 256;
}

// is the grouped GEMM the better path? (fast = plane layout, 16-byte aligned expert stride). A pair costs the GEMV one pass over an expert
// matrix, the GEMM costs every ACTIVE expert one dequantising pass that is several times dearer per weight (B200: 8 us vs 72 us for
// 4096 x 14336): take the GEMM when there are at least eight pairs per expert that can be active
bool b200_gemm_moe_wants(int type, int64_t m, int64_t k, int64_t n_as, int64_t n_used, int64_t n_tokens, int64_t y_span, bool force) {
    if (!b200_type_is_weight(type) || k % 256 != 0 || m <= 0 || n_as <= 0 || n_as > 2048) return false;
    const int64_t pairs = n_tokens * n_used;
    if (pairs <= 0 || pairs >= ((int64_t) 1 << 24) || y_span >= ((int64_t) 1 << 31) || k * 2 > (int64_t) 1 << 31) return false;
    if (force) return true;
    const int64_t act = pairs < n_as ? pairs : n_as;
    return pairs > 8 && pairs >= 8 * act;
}

int b200_gemm_moe_launch(const b200_mul_mat_id_args * a, cudaStream_t s) {
    encode_tiled_fn encode = get_encode_tiled();
    B200_REQUIRE(encode != nullptr, "gemm: cuTensorMapEncodeTiled is not available from this driver");
    const int64_t pairs = a->n_tokens * a->n_used, k = a->k;
    const MoeShape h = moe_shape(a->m, a->n_as, pairs);
    const MoeCarve c = moe_carve(a->workspace, k, a->n_as, h);
    B200_REQUIRE(a->workspace && a->workspace_size >= c.bytes, "mul_mat_id (GEMM): workspace %zu < %zu bytes", a->workspace_size, c.bytes);

    // 1. plan: pairs sorted by expert, unit list, scatter / gather offsets
    MoePlanArgs p = {};
    p.ids = a->ids; p.ids_row_stride = a->ids_row_stride; p.n_tokens = (int) a->n_tokens; p.n_used = (int) a->n_used; p.n_as = (int) a->n_as; p.nb1 = (int) a->nb1;
    p.b_row_stride = a->b_row_stride; p.b_token_stride = a->b_token_stride; p.y_row_stride = a->y_row_stride; p.y_token_stride = a->y_token_stride;
    p.bn = h.bn; p.tile_m = h.tile_m; p.m_tiles = h.m_tiles; p.max_units = h.max_units;
    p.row_begin = c.row_begin; p.total = c.total; p.units = c.units; p.yoff = c.yoff; p.src_off = c.src_off;
    B200_CHECK(b200_launch(moe_plan_kernel, dim3(1), dim3(1024), (size_t) (3 * a->n_as + 2) * sizeof(int), s, false, p));
    B200_CHECK(b200_launch(moe_zero_invalid_kernel, dim3((unsigned) pairs), dim3(256), 0, s, false, a->ids, (long long) a->ids_row_stride, (int) a->n_used, (int) a->n_as,
                           a->y, (long long) a->y_row_stride, (long long) a->y_token_stride, (long long) a->m));
    // 2. gathered fp16 activations: row p of xh is the activation row of the p-th sorted pair. Rows past the valid pairs (invalid ids) keep
    //    whatever the workspace held: they are only ever read into tile columns the epilogue drops
    if (k <= 4096)      B200_CHECK(b200_launch(act_half_kernel<4>, dim3((unsigned) pairs), dim3(256), 0, s, false, a->b, (int64_t) 0, k, c.xh, c.xscale, false, (const long long *) c.src_off));
    else if (k <= 8192) B200_CHECK(b200_launch(act_half_kernel<8>, dim3((unsigned) pairs), dim3(256), 0, s, false, a->b, (int64_t) 0, k, c.xh, c.xscale, false, (const long long *) c.src_off));
    else                B200_CHECK(b200_launch(act_half_kernel<16>, dim3((unsigned) pairs), dim3(256), 0, s, false, a->b, (int64_t) 0, k, c.xh, c.xscale, false, (const long long *) c.src_off));

    // 3. one GEMM launch over (expert, weight tile, n-tile)
    const int bn = h.bn, bn_cta = h.pair ? bn / 2 : bn;
    GemmArgs g = {};
    WMapSet wm;
    g.nmat = 1; g.n = pairs; g.k = k;
    g.W[0] = gemv_w_make(a->type, a->as, a->m, k);
    g.m[0] = a->m; g.y[0] = a->y; g.ldy[0] = 0;
    for (int i = 1; i < GEMM_MAX_GROUP; ++i) { g.W[i] = g.W[0]; g.m[i] = g.m[0]; g.y[i] = g.y[0]; g.ldy[i] = 0; }
    WMaps one;
    B200_REQUIRE(make_weight_maps3(a->type, g.W[0], a->m, a->n_as, a->as_expert_stride, one),
                 "mul_mat_id (GEMM): cuTensorMapEncodeTiled failed for the expert planes (type %d m=%lld k=%lld experts=%lld stride=%lld)", a->type,
                 (long long) a->m, (long long) k, (long long) a->n_as, (long long) a->as_expert_stride);
    for (int i = 0; i < GEMM_MAX_GROUP; ++i) for (int j = 0; j < 3; ++j) wm.m[i][j] = one.m[j];
    g.xscale = c.xscale; g.bn = bn; g.n_tiles = 1;
    g.acc_cols = bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256;
    const size_t per_stage = (size_t) A_STAGE + (size_t) bn_cta * 128;
    const size_t raw_ring = (size_t) raw_ring_bytes_of(a->type);
    const int stages = (int) ((SMEM_BUDGET - raw_ring) / per_stage);
    g.stages = stages > MAX_STAGES ? MAX_STAGES : stages;
    const size_t smem = raw_ring + (size_t) g.stages * per_stage + 1024;
    g.full_tiles = h.max_units; g.ksplit = 1; g.total_units = h.max_units;
    g.moe_units = c.units; g.moe_total = c.total; g.moe_yoff = c.yoff; g.expert_stride = a->as_expert_stride;
    if (const char * e = getenv("B200_GEMM_DEBUG")) g.debug = atoi(e);

    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t) k, (cuuint64_t) (pairs + 256)};        // the slack rows exist in the workspace: boxes never leave it
    const cuuint64_t strides[1] = {(cuuint64_t) k * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t) BK, (cuuint32_t) bn_cta};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, c.xh, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(cr == CUDA_SUCCESS, "mul_mat_id (GEMM): cuTensorMapEncodeTiled failed (%d) for pairs=%lld k=%lld bn=%d", (int) cr, (long long) pairs, (long long) k, bn);

    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    const int workers = h.pair ? sms / 2 : sms;
    const unsigned nw = (unsigned) (h.max_units < workers ? h.max_units : workers);
    dim3 grid(h.pair ? 2 * nw : nw);
    switch (a->type) {
        case B200_TYPE_Q4_K:   return launch_gemm_moe<B200_TYPE_Q4_K>(tm, wm, g, grid, smem, s, h.pair);
        case B200_TYPE_Q5_K:   return launch_gemm_moe<B200_TYPE_Q5_K>(tm, wm, g, grid, smem, s, h.pair);
        case B200_TYPE_Q6_K:   return launch_gemm_moe<B200_TYPE_Q6_K>(tm, wm, g, grid, smem, s, h.pair);
        case B200_TYPE_Q8_0:   return launch_gemm_moe<B200_TYPE_Q8_0>(tm, wm, g, grid, smem, s, h.pair);
        case B200_TYPE_Q4_0:   return launch_gemm_moe<B200_TYPE_Q4_0>(tm, wm, g, grid, smem, s, h.pair);
        case B200_TYPE_IQ4_XS: return launch_gemm_moe<B200_TYPE_IQ4_XS>(tm, wm, g, grid, smem, s, h.pair);
    }
    b200_set_error("mul_mat_id (GEMM): unsupported weight type %d", a->type);
    return -2;
}

// can these nodes share ONE GEMM launch? same weight type, same X (pointer, n, k, stride) and workspace
bool b200_gemm_groupable(const b200_mul_mat_args * args, int count) {
    if (count < 1 || count > GEMM_MAX_GROUP) return false;
    for (int i = 1; i < count; ++i)
        if (args[i].type != args[0].type || args[i].x != args[0].x || args[i].n != args[0].n || args[i].k != args[0].k ||
            args[i].x_row_stride != args[0].x_row_stride || args[i].workspace != args[0].workspace) return false;
    return true;
}

// `count` <= GEMM_MAX_GROUP weight matrices of the same type against the same X: the tiles of all of them form one work list for
// the persistent grid (wq/wk/wv and ffn_gate/ffn_up: 7 -> 4 launches per layer, and the 2 x 112 tiles of gate + up fill 74 CTA
// pairs in 3.03 waves instead of 2 x 2). args[0].flags decides whether X still has to be converted.
int b200_gemm_launch_group(const b200_mul_mat_args * args, int count, cudaStream_t s) { return b200_gemm_launch_group_ev(args, count, s, nullptr); }
int b200_gemm_launch_peer(const b200_mul_mat_args * a, cudaStream_t s, const PeerGather * peer) { return b200_gemm_launch_group_ev(a, 1, s, nullptr, peer); }

int b200_gemm_launch_group_ev(const b200_mul_mat_args * args, int count, cudaStream_t s, cudaEvent_t after_prologue, const PeerGather * peer) {
    encode_tiled_fn encode = get_encode_tiled();
    B200_REQUIRE(encode != nullptr, "gemm: cuTensorMapEncodeTiled is not available from this driver");
    B200_REQUIRE(b200_gemm_groupable(args, count), "gemm: the %d nodes of the group do not share type / X / workspace", count);
    const b200_mul_mat_args * a = &args[0];
    const int64_t n = a->n, k = a->k;
    const bool pdl = (a->flags & B200_MM_PDL) != 0;
    int64_t m_max = 0;
    for (int i = 0; i < count; ++i) m_max = args[i].m > m_max ? args[i].m : m_max;

    // workspace: fp16 activations [n][k] + per-row scale
    __half * xh = (__half *) a->workspace;
    float * xscale = (float *) ((uint8_t *) a->workspace + align256((size_t) n * k * sizeof(__half)));
    if (!(a->flags & B200_MM_X_PREQUANTISED))
    {
        if (k <= 4096)      B200_CHECK(b200_launch(act_half_kernel<4>, dim3((unsigned) n), dim3(256), 0, s, pdl, a->x, a->x_row_stride, k, xh, xscale, pdl, (const long long *) nullptr));
        else if (k <= 8192) B200_CHECK(b200_launch(act_half_kernel<8>, dim3((unsigned) n), dim3(256), 0, s, pdl, a->x, a->x_row_stride, k, xh, xscale, pdl, (const long long *) nullptr));
        else                B200_CHECK(b200_launch(act_half_kernel<16>, dim3((unsigned) n), dim3(256), 0, s, pdl, a->x, a->x_row_stride, k, xh, xscale, pdl, (const long long *) nullptr));
    }
    if (after_prologue) B200_CHECK(cudaEventRecord(after_prologue, s));

    // N tiling: equal tiles of at most 256 rows, multiple of 16
    const int n_tiles = (int) ((n + 255) / 256);
    int bn = (int) ((n + n_tiles - 1) / n_tiles);
    // CTA pairs (cta_group::2, 256 weight rows per tile) when there is more than one 128-row tile; B200_GEMM_PAIR=0/1 overrides
    static const int pair_env = [] { const char * e = getenv("B200_GEMM_PAIR"); return e ? atoi(e) : -1; }();
    const bool pair = (pair_env < 0 ? B200_GEMM_PAIR_DEFAULT : pair_env != 0) && m_max > BM;
    bn = pair ? (bn + 31) & ~31 : (bn + 15) & ~15;     // each CTA of a pair stages bn / 2 rows, a multiple of 16
    const int bn_cta = pair ? bn / 2 : bn;
    const int tile_m = pair ? 2 * BM : BM;
    GemmArgs g = {};
    WMapSet wm;
    g.nmat = count;
    g.n = n; g.k = k;
    int64_t tiles = 0;
    for (int i = 0; i < count; ++i) {
        g.W[i] = gemv_w_make(args[i].type, args[i].w, args[i].m, k);
        g.m[i] = args[i].m; g.y[i] = args[i].y; g.ldy[i] = args[i].y_row_stride;
        g.tile_begin[i] = (int) tiles;
        tiles += (args[i].m + tile_m - 1) / tile_m * n_tiles;
        WMaps one;
        B200_REQUIRE(make_weight_maps(args[i].type, g.W[i], args[i].m, one), "gemm: cuTensorMapEncodeTiled failed for the weight planes (type %d m=%lld k=%lld)",
                     args[i].type, (long long) args[i].m, (long long) k);
        for (int j = 0; j < 3; ++j) wm.m[i][j] = one.m[j];
    }
    for (int i = count; i <= GEMM_MAX_GROUP; ++i) g.tile_begin[i] = (int) tiles;
    for (int i = count; i < GEMM_MAX_GROUP; ++i) { g.W[i] = g.W[0]; g.m[i] = g.m[0]; g.y[i] = g.y[0]; g.ldy[i] = g.ldy[0]; for (int j = 0; j < 3; ++j) wm.m[i][j] = wm.m[0][j]; }
    B200_REQUIRE(tiles < ((int64_t) 1 << 26), "gemm: too many tiles (%lld)", (long long) tiles);
    g.xscale = xscale;
    g.bn = bn;
    g.n_tiles = n_tiles;
    g.acc_cols = bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256;
    const size_t per_stage = (size_t) A_STAGE + (size_t) bn_cta * 128;
    const size_t raw_ring = (size_t) raw_ring_bytes_of(a->type);
    int stages = (int) ((SMEM_BUDGET - raw_ring) / per_stage);
    g.stages = stages > MAX_STAGES ? MAX_STAGES : stages;
    const size_t smem = raw_ring + (size_t) g.stages * per_stage + 1024;

    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t) k, (cuuint64_t) n};
    const cuuint64_t strides[1] = {(cuuint64_t) k * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t) BK, (cuuint32_t) bn_cta};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xh, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(cr == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed (%d) for n=%lld k=%lld bn=%d", (int) cr, (long long) n, (long long) k, bn);

    // work list: whole waves of full tiles, then the remaining tiles cut along k so that the last wave also fills the SMs
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    const int workers = pair ? sms / 2 : sms;             // CTAs, or CTA pairs
    const int total_sb = (int) (k / 256);
    int64_t full = tiles / workers * workers;
    int64_t rest = tiles - full;
    int ks = 1;
    if (peer && peer->world > 1) { g.peer = *peer; rest = 0; full = tiles; }      // row split: no k-split tiles (they meet through atomics in local memory only)
    while (rest > 0 && ks < 16 && rest * ks * 2 <= workers && total_sb / (ks * 2) >= 2) ks *= 2;
    if (const char * e = getenv("B200_GEMM_KSPLIT")) { const int v = atoi(e); if (v >= 1 && v <= 16 && total_sb >= v) { ks = v; full = 0; rest = tiles; } }
    if (ks == 1) { full = tiles; rest = 0; }
    g.full_tiles = (int) full; g.ksplit = ks; g.total_units = (int) (full + rest * ks);
    if (const char * e = getenv("B200_GEMM_DEBUG")) g.debug = atoi(e);
    static unsigned long long * trace_buf = [] {
        unsigned long long * p = nullptr;
        if (getenv("B200_GEMM_TRACE") && cudaMallocManaged((void **) &p, 148 * 8 * 8 * 8) == cudaSuccess) {
            memset(p, 0, 148 * 8 * 8 * 8);
            static unsigned long long * keep; keep = p;
            atexit([] {
                cudaDeviceSynchronize();
                unsigned long long t0 = ~0ull;
                for (int i = 0; i < 148 * 64; ++i) if (keep[i] && keep[i] < t0) t0 = keep[i];
                for (int pr = 0; pr < 74; pr += 24) for (int u = 0; u < 8; ++u) {
                    const unsigned long long * r = keep + ((size_t) pr * 8 + u) * 8;
                    if (!r[0]) continue;
                    fprintf(stderr, "gemm trace pair %2d unit %d: mma start %7.2f  acc_empty %7.2f  first stage %7.2f  issued %7.2f | epi wait-from %7.2f  acc_full %7.2f  done %7.2f us\n", pr, u,
                            (r[0] - t0) * 1e-3, (r[1] - t0) * 1e-3, (r[2] - t0) * 1e-3, (r[3] - t0) * 1e-3, (r[4] - t0) * 1e-3, (r[5] - t0) * 1e-3, (r[6] - t0) * 1e-3);
                }
            });
        }
        return p;
    }();
    g.trace = trace_buf;
    static const bool pdl_env = [] { const char * e = getenv("B200_GEMM_PDL"); return !e || atoi(e) != 0; }();
    g.pdl = pdl && pdl_env ? 1 : 0;
    if (rest > 0) {     // zero the weight rows the split tiles accumulate into (full tiles of the same rows overwrite later, in stream order)
        for (int i = 0; i < count; ++i) {
            const int64_t first = full > g.tile_begin[i] ? full - g.tile_begin[i] : 0;       // first split tile of this matrix (local index)
            const int64_t m_lo = first / n_tiles * tile_m;
            if (m_lo >= args[i].m) continue;
            B200_CHECK(b200_launch(zero_rows_kernel, dim3((unsigned) n), dim3(256), 0, s, pdl, args[i].y, args[i].y_row_stride, m_lo, args[i].m, pdl));
        }
    }
    const unsigned nw = (unsigned) (g.total_units < workers ? g.total_units : workers);
    dim3 grid(pair ? 2 * nw : nw);
    switch (a->type) {
        case B200_TYPE_Q4_K:   return launch_gemm<B200_TYPE_Q4_K>(tm, wm, g, grid, smem, s, pair);
        case B200_TYPE_Q5_K:   return launch_gemm<B200_TYPE_Q5_K>(tm, wm, g, grid, smem, s, pair);
        case B200_TYPE_Q6_K:   return launch_gemm<B200_TYPE_Q6_K>(tm, wm, g, grid, smem, s, pair);
        case B200_TYPE_Q8_0:   return launch_gemm<B200_TYPE_Q8_0>(tm, wm, g, grid, smem, s, pair);
        case B200_TYPE_Q4_0:   return launch_gemm<B200_TYPE_Q4_0>(tm, wm, g, grid, smem, s, pair);
        case B200_TYPE_IQ4_XS: return launch_gemm<B200_TYPE_IQ4_XS>(tm, wm, g, grid, smem, s, pair);
    }
    b200_set_error("gemm: unsupported weight type %d", a->type);
    return -2;
}

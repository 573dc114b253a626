// gemm.cu -- (2) prefill GEMM: Y[n][m] = sum_k W[m][k] X[n][k] for n > 8, quantised W, on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), sm_100a only.
//
// Reference counterparts: ggml_compute_forward_mul_mat + ggml_gemm_q4_K_8x8_q8_K on the CPU
// (ggml-cpu/ggml-cpu.c:1227-1427, ggml-cpu/repack.cpp:1037-1124) and mul_mat_q (mma.sync int8) on
// CUDA (ggml-cuda/mmq.cuh:3459). This kernel is neither:
//
//   * a tile is 128 weight rows x BN <= 256 activation rows; the K loop advances 64 elements a step;
//   * A (weights): 8 producer warps stream the quantised blocks straight from HBM (plane layout of
//     repack.cu, 128-bit loads), dequantise in registers to fp16 and write the K-major
//     SWIZZLE_128B operand tile into shared memory (gemm_dequant.cuh) -- every weight byte is read
//     once per N-tile and never exists in HBM in expanded form;
//   * B (activations): a prologue kernel rescales each f32 row by a power of two and rounds it to fp16
//     (exact scaling, 11-bit mantissa; no int8 step), TMA (cp.async.bulk.tensor, SWIZZLE_128B) stages
//     64 x BN boxes into shared memory;
//   * one elected thread issues tcgen05.mma.kind::f16 (M = 128, N = BN, K = 16) four times per stage,
//     tcgen05.commit releases the stage; fp32 accumulation over the whole K lives in TMEM;
//   * epilogue: the producer warps read the accumulator with tcgen05.ld, multiply by the per-row
//     power-of-two scale and store Y with coalesced 128-byte rows.
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
#include <stdlib.h>

using namespace sm100;

namespace {

constexpr int BM = 128;            // weight rows per tile = UMMA M
constexpr int BK = 64;             // k-elements per stage = one 128-byte swizzle row of fp16
constexpr int A_STAGE = BM * BK * 2;
constexpr int DEQ_WARPS = 8;
constexpr int THREADS = 128 + DEQ_WARPS * 32;   // warp 0: TMA, warp 1: MMA, warp 2: TMEM, warp 3: spare, 4..11: producers/epilogue
constexpr int MAX_STAGES = 8;
constexpr size_t SMEM_BUDGET = 200 * 1024;

struct GemmArgs {
    GemvW W;
    int64_t m, n, k;
    float * y; int64_t ldy;
    const float * xscale;       // per activation row: multiply the accumulator by this
    int bn;                     // activation rows per tile (multiple of 16, <= 256)
    int stages;
    int tmem_cols;              // power of two >= max(32, bn)
    int ksplit;                 // gridDim.z CTAs share a tile along k; partial tiles meet in Y with red.global.add (Y pre-zeroed)
};

// ---- prologue: X f32 [n][k] (row stride ldx) -> fp16 [n][k] scaled by 2^-e per row, xscale[n] = 2^e ----
__global__ void __launch_bounds__(256) act_half_kernel(const float * __restrict__ x, int64_t ldx, int64_t k, __half * __restrict__ xh,
                                                       float * __restrict__ xscale, bool pdl) {
    if (pdl) { pdl_launch_dependents(); pdl_wait(); }
    const int64_t row = blockIdx.x;
    const float * xr = x + row * ldx;
    __shared__ float red[8];
    float amax = 0.0f;
    for (int64_t i = threadIdx.x; i < k; i += 256) amax = fmaxf(amax, fabsf(xr[i]));
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
    for (int64_t i = threadIdx.x; i < k; i += 256) xh[row * k + i] = __float2half_rn(xr[i] * down);
    if (threadIdx.x == 0) xscale[row] = ldexpf(1.0f, e);
}

// ---- main kernel --------------------------------------------------------------------------------------
template <int TYPE>
__global__ void __launch_bounds__(THREADS, 1) gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const GemmArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_full[MAX_STAGES], bar_empty[MAX_STAGES], bar_acc;
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int stages = a.stages, bn = a.bn;
    const uint32_t b_stage = (uint32_t) bn * 128u;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;      // SWIZZLE_128B tiles need 1024 B alignment
    const uint32_t smem_a = smem0, smem_b = smem0 + (uint32_t) stages * A_STAGE;
    const int total_kb = (int) (a.k / BK);
    const int kb_begin = (int) ((int64_t) total_kb * blockIdx.z / a.ksplit), kb_end = (int) ((int64_t) total_kb * (blockIdx.z + 1) / a.ksplit);
    const int num_kb = kb_end - kb_begin;   // k-blocks of this CTA; stage/phase bookkeeping runs on the local index
    const int64_t m0 = (int64_t) blockIdx.x * BM;
    const int n0 = blockIdx.y * bn;

    if (warp == 0 && lane == 0) prefetch_tensormap(&tmap_x);
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(smem_u32(&bar_full[s]), 1 + DEQ_WARPS / 2);   // TMA expect_tx arrive + the 4 producer warps of one half
            mbar_init(smem_u32(&bar_empty[s]), 1);                  // tcgen05.commit
        }
        mbar_init(smem_u32(&bar_acc), 1);
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc(smem_u32(&tmem_base_smem), (uint32_t) a.tmem_cols); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_smem;

    if (warp == 0) {
        // ===== TMA producer: activation boxes (64 k x bn rows) =====
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % stages; const uint32_t ph = (uint32_t) (kb / stages) & 1u;
                mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1u);
                mbar_arrive_expect_tx(smem_u32(&bar_full[s]), b_stage);
                tma_load_2d(smem_b + (uint32_t) s * b_stage, &tmap_x, (kb_begin + kb) * BK, n0, smem_u32(&bar_full[s]));
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const uint32_t idesc = umma_idesc_f16(BM, bn);
        for (int kb = 0; kb < num_kb; ++kb) {
            const int s = kb % stages; const uint32_t ph = (uint32_t) (kb / stages) & 1u;
            mbar_wait(smem_u32(&bar_full[s]), ph);
            tc_fence_after();
            if (lane == 0) {
                const uint64_t adesc = umma_desc_kmajor_sw128(smem_a + (uint32_t) s * A_STAGE);
                const uint64_t bdesc = umma_desc_kmajor_sw128(smem_b + (uint32_t) s * b_stage);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)     // 16 fp16 = 32 bytes = 2 descriptor units along K inside the swizzle atom
                    mma_f16_ss(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t) ((kb | k) != 0));
                mma_commit(smem_u32(&bar_empty[s]));
                if (kb == num_kb - 1) mma_commit(smem_u32(&bar_acc));
            }
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ===== weight producers: dequantise 128 rows x 64 k per stage; two halves alternate k-blocks =====
        const int t = threadIdx.x - 128;
        const int half = t >> 7, r = t & 127;
        const int64_t row = m0 + r < a.m ? m0 + r : a.m - 1;
        Deq<TYPE> cur, nxt;
        if (half < num_kb) nxt.load(a.W, row, kb_begin + half);
        for (int kb = half; kb < num_kb; kb += 2) {
            cur = nxt;
            if (kb + 2 < num_kb) nxt.load(a.W, row, kb_begin + kb + 2);
            const int s = kb % stages; const uint32_t ph = (uint32_t) (kb / stages) & 1u;
            mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1u);
            cur.store(smem_a + (uint32_t) s * A_STAGE + (uint32_t) r * 128u, r, kb_begin + kb);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&bar_full[s]));
        }
        // ===== epilogue: TMEM -> registers -> Y (thread = weight row, 16 activation rows per load) =====
        mbar_wait(smem_u32(&bar_acc), 0);
        tc_fence_after();
        const int q = warp & 3;                                  // TMEM lane quarter this warp may touch
        const int64_t mrow = m0 + 32 * q + lane;
        for (int c = (warp - 4) >> 2; c * 16 < bn; c += 2) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(tmem_d + ((uint32_t) (32 * q) << 16) + (uint32_t) (c * 16), v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int64_t n = (int64_t) n0 + c * 16 + j;
                if (n < a.n && mrow < a.m) {
                    const float val = __uint_as_float(v[j]) * __ldg(a.xscale + n);
                    if (a.ksplit == 1) a.y[n * a.ldy + mrow] = val;
                    else atomicAdd(a.y + n * a.ldy + mrow, val);      // result unused -> RED.E.ADD.F32
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_d, (uint32_t) a.tmem_cols);
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

template <int TYPE>
int launch_gemm(const CUtensorMap & tm, const GemmArgs & g, dim3 grid, size_t smem, cudaStream_t s, bool pdl) {
    static bool configured = false;
    if (!configured) {
        B200_CHECK(cudaFuncSetAttribute(gemm_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SMEM_BUDGET + 1024)));
        configured = true;
    }
    (void) pdl;
    B200_CHECK(b200_launch(gemm_kernel<TYPE>, grid, dim3(THREADS), smem, s, false, tm, g));
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

int b200_gemm_launch(const b200_mul_mat_args * a, cudaStream_t s) {
    encode_tiled_fn encode = get_encode_tiled();
    B200_REQUIRE(encode != nullptr, "gemm: cuTensorMapEncodeTiled is not available from this driver");
    const int64_t n = a->n, k = a->k, m = a->m;
    const bool pdl = (a->flags & B200_MM_PDL) != 0;

    // workspace: fp16 activations [n][k] + per-row scale
    __half * xh = (__half *) a->workspace;
    float * xscale = (float *) ((uint8_t *) a->workspace + align256((size_t) n * k * sizeof(__half)));
    if (!(a->flags & B200_MM_X_PREQUANTISED))
        B200_CHECK(b200_launch(act_half_kernel, dim3((unsigned) n), dim3(256), 0, s, pdl, a->x, a->x_row_stride, k, xh, xscale, pdl));

    // N tiling: equal tiles of at most 256 rows, multiple of 16
    const int n_tiles = (int) ((n + 255) / 256);
    int bn = (int) ((n + n_tiles - 1) / n_tiles);
    bn = (bn + 15) & ~15;
    GemmArgs g = {};
    g.W = gemv_w_make(a->type, a->w, m, k);
    g.m = m; g.n = n; g.k = k;
    g.y = a->y; g.ldy = a->y_row_stride;
    g.xscale = xscale;
    g.bn = bn;
    g.tmem_cols = bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256;
    const size_t per_stage = (size_t) A_STAGE + (size_t) bn * 128;
    int stages = (int) ((SMEM_BUDGET - 1024) / per_stage);
    g.stages = stages > MAX_STAGES ? MAX_STAGES : stages;
    const size_t smem = (size_t) g.stages * per_stage + 1024;

    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t) k, (cuuint64_t) n};
    const cuuint64_t strides[1] = {(cuuint64_t) k * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t) BK, (cuuint32_t) bn};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xh, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(cr == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed (%d) for n=%lld k=%lld bn=%d", (int) cr, (long long) n, (long long) k, bn);

    // fill the machine: when the tile grid is much smaller than the SM count, split k over gridDim.z
    const int64_t tiles = ((m + BM - 1) / BM) * n_tiles;
    int ks = 1;
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    while (ks < 8 && tiles * ks * 2 <= sms && (k / BK) / (ks * 2) >= 4) ks *= 2;
    if (const char * e = getenv("B200_GEMM_KSPLIT")) { const int v = atoi(e); if (v >= 1 && v <= 16 && k / BK >= v) ks = v; }
    g.ksplit = ks;
    if (ks > 1) B200_CHECK(cudaMemset2DAsync(a->y, (size_t) a->y_row_stride * sizeof(float), 0, (size_t) m * sizeof(float), (size_t) n, s));
    dim3 grid((unsigned) ((m + BM - 1) / BM), (unsigned) n_tiles, (unsigned) ks);
    switch (a->type) {
        case B200_TYPE_Q4_K:   return launch_gemm<B200_TYPE_Q4_K>(tm, g, grid, smem, s, pdl);
        case B200_TYPE_Q6_K:   return launch_gemm<B200_TYPE_Q6_K>(tm, g, grid, smem, s, pdl);
        case B200_TYPE_Q8_0:   return launch_gemm<B200_TYPE_Q8_0>(tm, g, grid, smem, s, pdl);
        case B200_TYPE_Q4_0:   return launch_gemm<B200_TYPE_Q4_0>(tm, g, grid, smem, s, pdl);
        case B200_TYPE_IQ4_XS: return launch_gemm<B200_TYPE_IQ4_XS>(tm, g, grid, smem, s, pdl);
    }
    b200_set_error("gemm: unsupported weight type %d", a->type);
    return -2;
}

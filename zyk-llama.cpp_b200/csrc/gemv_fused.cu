// gemv_fused.cu -- (1) decode GEMV, fused version: Y[n<=4][m] = W[m][k] . X[n][k] for one or several
// weight matrices that share X (wq/wk/wv, ffn_gate/ffn_up), activation quantiser included, ONE launch.
//
// Reference counterparts: ggml_compute_forward_mul_mat + ggml_vec_dot_* (ggml-cpu/ggml-cpu.c:1227-1427,
// ggml-cpu/quants.c) and quantize_q8_1 + mul_mat_vec_q (ggml-cuda/quantize.cu:4-73, mmvq.cu:141-335).
//
//   * persistent grid (SMs x CTAs-per-SM), 8 warps per CTA; the row pairs of all matrices of the group form
//     one index space that the warps of the whole grid walk interleaved (warp w takes pairs w, w + W, ...),
//     so neighbouring warps stream neighbouring rows and there is no tail wave per matrix;
//   * before griddepcontrol.wait every warp prefetches its first row pairs into L2 (weights do not depend on
//     the previous kernel); after the wait the CTA quantises X into shared memory itself -- the arithmetic
//     is quant_chunk256 (actquant.cuh), bit for bit the reference's quantize_row_q8_K / q8_0 -- so there is
//     no separate quantiser launch and the int8 activations are read from shared memory, not L1/L2;
//   * per row pair and 1024-element step every lane issues its 128-bit streaming loads (L1 no-allocate) for
//     both rows and U steps before any arithmetic, then runs the integer block arithmetic of the CPU
//     vec_dots (gemv_frag.cuh) with the activation fragment shared by the two rows; warp-shuffle reduction.
// Algorithmic bytes per launch = sum over the group of m * row_size(type, k); HBM bound.
#include "gemv_fused_common.cuh"

namespace {

constexpr int FWARPS = 8;
constexpr int FTHREADS = FWARPS * 32;

// rows-in-flight per lane: U steps x R rows of 128-bit loads; the thin fragments afford U = 4 (a whole k = 4096 row)
template <int TYPE> struct Unroll { static constexpr int U = (TYPE == B200_TYPE_Q6_K || TYPE == B200_TYPE_Q8_0) ? 2 : 4; };

// one warp, one row pair of matrix M: stream the two rows, U steps of loads in flight. `mid` runs after the loads of
// the first batch have been issued and before anything is computed (the first pair of a warp passes the
// wait-for-previous-kernel + activation quantiser there, so that phase overlaps the HBM latency of the weights).
template <int TYPE, int NCOLS, int KV, typename Mid>
__device__ __forceinline__ void row_pair(const FusedMat & M, int64_t row0, const XS & xs, int lane, const PeerGather & peer, Mid mid) {
    constexpr int U = Unroll<TYPE>::U;
    constexpr int BLK = Fmt<TYPE>::BLK;
    constexpr int BPS = 1024 / BLK;
    const int g = lane & 7;
    const int lane_blk = BLK == 256 ? (lane >> 3) : lane;
    const int nb = (int) M.W.nb;
    const int nsteps = (nb + BPS - 1) / BPS;
    const int s_begin = 0, s_end = nsteps;
    float acc[R][NCOLS];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[r][c] = 0.0f;
    bool first = true;
    for (int s0 = s_begin; s0 < s_end || first; s0 += U) {
        WFrag<TYPE> f[U][R];
        bool valid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int blk = (s0 + u) * BPS + lane_blk;
            valid[u] = blk < nb && s0 + u < s_end;
            const int bl = valid[u] ? blk : 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int64_t row = row0 + r < M.m ? row0 + r : M.m - 1;
                f[u][r].load(M.W, row, bl, g);
            }
        }
        if (first) { mid(); first = false; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!valid[u]) continue;
            const int blk = (s0 + u) * BPS + lane_blk;
            XFrag x[NCOLS];
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) xfrag_lds<TYPE>(x[c], xs, c, blk, g);
#pragma unroll
            for (int r = 0; r < R; ++r) f[u][r].template dot<NCOLS>(x, acc[r], g);
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            const float v = warp_sum(acc[r][c]);
            if (lane == 0 && row0 + r < M.m) {
                if (KV == KV_PEER) peer_store(peer, M.y + c * M.ldy + row0 + r, v);
                else M.y[c * M.ldy + row0 + r] = v;
            }
        }
}

template <int NCOLS, int KV, typename Mid>
__device__ __forceinline__ void row_pair_any(const FusedMat & M, int64_t row0, const XS & xs, int lane, const PeerGather & peer, Mid mid) {
    switch (M.type) {
        case B200_TYPE_Q4_K:   row_pair<B200_TYPE_Q4_K, NCOLS, KV>(M, row0, xs, lane, peer, mid); break;
        case B200_TYPE_Q6_K:   row_pair<B200_TYPE_Q6_K, NCOLS, KV>(M, row0, xs, lane, peer, mid); break;
        case B200_TYPE_IQ4_XS: row_pair<B200_TYPE_IQ4_XS, NCOLS, KV>(M, row0, xs, lane, peer, mid); break;
        case B200_TYPE_Q4_0:   row_pair<B200_TYPE_Q4_0, NCOLS, KV>(M, row0, xs, lane, peer, mid); break;
        default:               row_pair<B200_TYPE_Q8_0, NCOLS, KV>(M, row0, xs, lane, peer, mid); break;
    }
}

template <int NCOLS, int KV>
__global__ void __launch_bounds__(FTHREADS, 2) gemv_fused_kernel(const FusedArgs a) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k = (int) a.k;
    XS xs;
    xs.qs = smem_u32(smem_raw);                                        // int8 [NCOLS][k]
    xs.d = xs.qs + (uint32_t) (NCOLS * k);                             // f32  [NCOLS][k / act_block]
    xs.bs = xs.d + (uint32_t) (NCOLS * (k / a.act_block)) * 4;         // i16  [NCOLS][k / 16]
    xs.k = k; xs.act_block = a.act_block;

    const int64_t gw = (int64_t) blockIdx.x * FWARPS + warp, GW = (int64_t) gridDim.x * FWARPS;
    if (a.pdl) pdl_launch_dependents();

    // the weight loads of the warp's first row pair are issued first (they do not depend on the previous kernel);
    // waiting for the previous kernel and quantising X happens while they are in flight
    auto prologue = [&]() {
        if (a.pdl) pdl_wait();
        if (KV == KV_PEER && a.peer.wait) {          // X is complete once every rank has finished the previous call
            if (lane == 0) peer_wait(a.peer);
            __syncwarp();
        }
        quantise_x<NCOLS, FWARPS>(a, xs, warp, lane);
    };
    auto nothing = []() {};
    int g = 0;
    if (gw < a.total_pairs) {
        while (gw >= a.mat[g].pair_begin + (a.mat[g].m + R - 1) / R) ++g;
        row_pair_any<NCOLS, KV>(a.mat[g], (gw - a.mat[g].pair_begin) * R, xs, lane, a.peer, prologue);
        for (int64_t p = gw + GW; p < a.total_pairs; p += GW) {
            while (p >= a.mat[g].pair_begin + (a.mat[g].m + R - 1) / R) ++g;
            row_pair_any<NCOLS, KV>(a.mat[g], (p - a.mat[g].pair_begin) * R, xs, lane, a.peer, nothing);
        }
    } else {
        prologue();
    }
    if (KV == KV_PEER && a.peer.signal) {            // slices are in every rank's arena: the last CTA announces it
        __syncthreads();                             // CTA barrier, then ONE system fence per CTA (inside peer_signal) covers all warps' stores
        if (threadIdx.x == 0) peer_signal(a.peer);
    }
}

template <int NCOLS>
int launch_fused(const FusedArgs & a, int grid, size_t smem, cudaStream_t s) {
    static unsigned configured = 0;
    if (b200_first_on_device(configured)) {
        B200_CHECK(cudaFuncSetAttribute(gemv_fused_kernel<NCOLS, KV_PLAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        B200_CHECK(cudaFuncSetAttribute(gemv_fused_kernel<NCOLS, KV_PEER>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    }
    if (a.peer.world > 1) B200_CHECK(b200_launch(gemv_fused_kernel<NCOLS, KV_PEER>, dim3((unsigned) grid), dim3(FTHREADS), smem, s, a.pdl != 0, a));
    else                  B200_CHECK(b200_launch(gemv_fused_kernel<NCOLS, KV_PLAIN>, dim3((unsigned) grid), dim3(FTHREADS), smem, s, a.pdl != 0, a));
    return 0;
}

} // namespace

// Can this group take the fused kernel? All matrices: plane layout (or canonical Q4_K), packed, 16 B aligned,
// k % 256 == 0, same activation format, quantised X fits in shared memory.
bool b200_gemv_stream_wants(const GemvGroup & g, bool force) {
    if (g.count < 1 || g.count > B200_MAX_GROUP || g.ncols < 1 || g.ncols > 4) return false;
    if (g.k % 256 != 0 || g.k > 65536) return false;
    const int act = b200_act_type(g.type[0]);
    for (int i = 0; i < g.count; ++i) {
        if (b200_act_type(g.type[i]) != act || g.m[i] <= 0) return false;
        if (g.type[i] == B200_TYPE_Q5_K) return false;   // register GEMV only: keeps the fused kernel's code size where it was
        if ((((uintptr_t) g.w[i]) & 15) != 0) return false;
    }
    const int B = act == B200_TYPE_Q8_K ? 256 : 32;
    const size_t xbytes = (size_t) g.ncols * (g.k + g.k / B * 4 + g.k / 16 * 2);
    // every CTA quantises the whole of X: past 8192 activation elements the separate quantiser kernel + register GEMV wins
    // (measured on the llama-3-8B chain: 586 -> 622 tokens/s when the k = 14336 down projections take the two-kernel path)
    static const int64_t max_x = [] { const char * e = getenv("B200_FUSED_MAX_X"); return e ? (int64_t) atoll(e) : (int64_t) 8192; }();
    if (!force && (int64_t) g.ncols * g.k > max_x) return false;
    return xbytes <= 96 * 1024;
}

int b200_gemv_stream_launch(const GemvGroup & g, cudaStream_t s) {
    FusedArgs a = {};
    a.nmat = g.count;
    int64_t pairs = 0;
    for (int i = 0; i < g.count; ++i) {
        a.mat[i].W = gemv_w_make(g.type[i], g.w[i], g.m[i], g.k);
        a.mat[i].type = g.type[i];
        a.mat[i].m = g.m[i];
        a.mat[i].y = g.y[i];
        a.mat[i].ldy = g.ldy[i];
        a.mat[i].pair_begin = pairs;
        pairs += (g.m[i] + R - 1) / R;
    }
    a.total_pairs = pairs;
    a.k = g.k; a.x = g.x; a.ldx = g.ldx;
    a.flavour = g.flavour;
    a.act_block = b200_act_type(g.type[0]) == B200_TYPE_Q8_K ? 256 : 32;
    a.pdl = g.pdl;
    a.xq_out = g.xq_out;
    a.peer = g.peer;

    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    const size_t smem = (size_t) g.ncols * (g.k + g.k / a.act_block * 4 + g.k / 16 * 2) + 64;
    int per_sm = 2;
    if (const char * e = getenv("B200_FUSED_CTAS_PER_SM")) { const int v = atoi(e); if (v > 0) per_sm = v; }
    int64_t grid = (int64_t) sms * per_sm;
    // (sharing a row pair between 2..8 warps for long-k / few-row shapes was measured slower on B200 and removed: DESIGN.md)
    const int64_t need = (pairs + FWARPS - 1) / FWARPS;
    if (grid > need) grid = need;
    switch (g.ncols) {
        case 1: return launch_fused<1>(a, (int) grid, smem, s);
        case 2: return launch_fused<2>(a, (int) grid, smem, s);
        case 3: return launch_fused<3>(a, (int) grid, smem, s);
        case 4: return launch_fused<4>(a, (int) grid, smem, s);
    }
    b200_set_error("gemv_fused: ncols %d out of range", g.ncols);
    return -2;
}

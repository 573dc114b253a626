// mul_mat.cu -- C-ABI entry points b200_mul_mat / b200_mul_mat_id: quantise X, pick the kernel.
// Replaces ggml_compute_forward_mul_mat / _mul_mat_id (ggml-cpu/ggml-cpu.c:1227-1427, 1500-1676).
#include "actquant.cuh"
#include "gemv.h"
#include "gemm.h"
#include <string.h>
#include <stdlib.h>

static bool planes_fast_path(int type, int layout, int64_t m, int64_t k, const void * w, int64_t w_row_stride) {
    const bool packed = w_row_stride == 0 || w_row_stride == (int64_t) b200_row_size(type, k);
    if (!packed || (k % 256) != 0 || (((uintptr_t) w) & 15) != 0) return false;
    if (layout == B200_LAYOUT_PLANES) return true;
    return type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K;   // canonical Q4_K / Q5_K are their own plane layout
    (void) m;
}

extern "C" {

size_t b200_mul_mat_workspace(int type, int64_t n, int64_t k) {
    const int act = b200_act_type(type);
    if (act < 0) return 0;
    size_t q = xq_planes_bytes(act, n, k);
    size_t g = b200_gemm_workspace(type, n, k);
    return (q > g ? q : g) + 256;
}

// one tiny kernel each for the two-kernel paths (quantiser + GEMV, fp16 prologue + GEMM) of the fused row-split exchange
static __global__ void peer_wait_kernel(const PeerGather p) { if (threadIdx.x == 0) peer_wait(p); }
static __global__ void peer_signal_kernel(const PeerGather p) { if (threadIdx.x == 0) peer_signal(p); }
// prefill: the GEMM keeps its epilogue local (k-split tiles meet with red.global.add); this kernel then copies the finished slice
// y[n][0..m) (row stride ldy) into every other rank's arena with 128-bit stores and, when asked, announces the end of the call
static __global__ void __launch_bounds__(256) peer_scatter_kernel(const float * __restrict__ y, int64_t ldy, int64_t m, int64_t n, const PeerGather p) {
    const bool vec = (m % 4 == 0) && (ldy % 4 == 0) && ((((uintptr_t) y) & 15) == 0);
    const int64_t per_row = vec ? m / 4 : m, total = per_row * n;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = i / per_row, c = i % per_row;
        if (vec) {
            const float4 * src = reinterpret_cast<const float4 *>(y + r * ldy) + c;
            const float4 v = *src;
            for (int g = 0; g < p.world; ++g)
                if (g != p.rank) *reinterpret_cast<float4 *>(reinterpret_cast<char *>(const_cast<float4 *>(src)) + p.delta[g]) = v;
        } else {
            const float * src = y + r * ldy + c;
            const float v = *src;
            for (int g = 0; g < p.world; ++g)
                if (g != p.rank) *reinterpret_cast<float *>(reinterpret_cast<char *>(const_cast<float *>(src)) + p.delta[g]) = v;
        }
    }
    if (p.signal) {
        __syncthreads();
        if (threadIdx.x == 0) peer_signal(p);
    }
}

// `peer` (may be null): fused row-split exchange for this node; peer->wait / peer->signal say whether this node's first
// kernel waits for the earlier calls and whether its last kernel announces the end of the call
static int mul_mat_impl(const b200_mul_mat_args * a, cudaStream_t s, const PeerGather * peer, const ActGlu * glu = nullptr) {
    B200_REQUIRE(a && a->w && a->x && a->y, "b200_mul_mat: null pointer");
    B200_REQUIRE(b200_type_is_weight(a->type), "b200_mul_mat: unsupported weight type %d", a->type);
    const int be = b200_block_elems(a->type);
    B200_REQUIRE(a->k > 0 && a->k % be == 0, "b200_mul_mat: k=%lld not a multiple of the block size %d", (long long) a->k, be);
    if (a->m == 0 || a->n == 0) return 0;
    B200_REQUIRE(a->layout == B200_LAYOUT_CANONICAL || b200_repack_supported(a->type, a->m, a->k),
                 "b200_mul_mat: plane layout is not defined for type %d m=%lld k=%lld", a->type, (long long) a->m, (long long) a->k);
    B200_REQUIRE(a->layout == B200_LAYOUT_CANONICAL || a->w_row_stride == 0 || a->w_row_stride == (int64_t) b200_row_size(a->type, a->k),
                 "b200_mul_mat: plane layout needs packed rows");
    const size_t need = b200_mul_mat_workspace(a->type, a->n, a->k);
    B200_REQUIRE(a->workspace && a->workspace_size >= need, "b200_mul_mat: workspace %zu < %zu bytes", a->workspace_size, need);
    B200_REQUIRE((((uintptr_t) a->workspace) & 255) == 0, "b200_mul_mat: workspace must be 256 B aligned");

    const bool pdl = (a->flags & B200_MM_PDL) != 0;
    const bool fast = planes_fast_path(a->type, a->layout, a->m, a->k, a->w, a->w_row_stride);

    // (2) prefill: tensor-core GEMM on plane layout
    if (!glu && a->algo != 1 && fast && b200_gemm_wants(a->type, a->m, a->n, a->k, a->algo == 2)) {
        if (peer && peer->wait) B200_CHECK(b200_launch(peer_wait_kernel, dim3(1), dim3(32), 0, s, false, *peer));
        if (!peer) return b200_gemm_launch(a, s);
        // row split. Default: the GEMM keeps its epilogue local (k-split tiles meet through atomics in local memory) and a copy kernel sends the slice to
        // the peers. B200_GEMM_PEER_EPILOGUE=1: the epilogue itself stores into every rank's arena and the last CTA announces the call -- measured
        // SLOWER at N = 2 (pp512 37.9 k vs 53 k tokens/s): a rank's slices are 4 - 56 tiles for 74 CTA pairs, and without the k split they leave
        // most SMs idle; it pays only once the same-X nodes of a rank share one launch
        static const bool fused_epi = [] { const char * e = getenv("B200_GEMM_PEER_EPILOGUE"); return e && e[0] == '1'; }();
        if (fused_epi) return b200_gemm_launch_peer(a, s, peer);
        const int rc = b200_gemm_launch(a, s);
        if (rc) return rc;
        const int64_t work = (a->m % 4 == 0 ? a->m / 4 : a->m) * a->n;
        const unsigned blocks = (unsigned) (work / 256 + 1 < 148 * 8 ? work / 256 + 1 : 148 * 8);
        B200_CHECK(b200_launch(peer_scatter_kernel, dim3(blocks), dim3(256), 0, s, false, (const float *) a->y, a->y_row_stride, a->m, a->n, *peer));
        return 0;
    }
    B200_REQUIRE(a->algo != 2, "b200_mul_mat: GEMM forced but shape/layout is not eligible (type %d m=%lld n=%lld k=%lld)",
                 a->type, (long long) a->m, (long long) a->n, (long long) a->k);

    const int act = b200_act_type(a->type);
    const XqPlanes pl = xq_planes_carve(a->workspace, act, a->n, a->k);

    // (1) decode, streaming kernel: one launch, activation quantiser fused (quantised X also left in the workspace)
    if (!glu && fast && a->n <= 4 && a->algo != 1 && !(a->flags & B200_MM_X_PREQUANTISED)) {
        GemvGroup g = {};
        g.count = 1; g.type[0] = a->type; g.w[0] = a->w; g.m[0] = a->m; g.y[0] = a->y;
        g.ldy[0] = a->y_row_stride;
        g.k = a->k; g.ncols = (int) a->n; g.x = a->x; g.ldx = a->x_row_stride;
        g.flavour = a->flavour; g.pdl = pdl; g.xq_out = pl;
        if (peer) g.peer = *peer;
        if (b200_gemv_stream_wants(g, a->algo == 3)) return b200_gemv_stream_launch(g, s);
    }
    B200_REQUIRE(a->algo != 3, "b200_mul_mat: streaming GEMV forced but the call is not eligible (type %d m=%lld n=%lld k=%lld)",
                 a->type, (long long) a->m, (long long) a->n, (long long) a->k);

    // (3) + (1): quantise activations into the workspace planes, then register GEMV in column groups of <= 8
    if (peer && peer->wait) B200_CHECK(b200_launch(peer_wait_kernel, dim3(1), dim3(32), 0, s, false, *peer));
    if (!(a->flags & B200_MM_X_PREQUANTISED)) {
        int rc = b200_act_quant_planes(act, a->flavour, a->x, a->x_row_stride, pl, a->n, a->k, s, pdl, glu);
        if (rc) return rc;
    }
    B200_REQUIRE(!peer || fast, "b200_mul_mat_gather: the row-split exchange needs the plane layout (type %d m=%lld k=%lld)", a->type,
                 (long long) a->m, (long long) a->k);
    for (int64_t c0 = 0; c0 < a->n; c0 += 8) {
        const int nc = (int) (a->n - c0 < 8 ? a->n - c0 : 8);
        int rc;
        if (fast) {
            GemvArgs g = {};
            g.W = gemv_w_make(a->type, a->w, a->m, a->k);
            g.m = a->m; g.k = a->k;
            g.xqs = pl.qs + c0 * a->k; g.xd = pl.d + c0 * (a->k / pl.block); g.xbs = pl.bs + c0 * (a->k >> 4);
            g.act_block = pl.block;
            g.y = a->y + c0 * a->y_row_stride; g.ldy = a->y_row_stride;
            g.pdl = pdl;
            if (peer) { g.peer = *peer; g.peer.wait = 0; g.peer.signal = peer->signal && c0 + 8 >= a->n; }
            rc = b200_gemv_launch(a->type, g, nc, 1, a->reserved, s);
        } else {
            GemvGenericArgs g = {};
            g.w = (const uint8_t *) a->w;
            g.w_row_stride = a->w_row_stride ? a->w_row_stride : (int64_t) b200_row_size(a->type, a->k);
            g.m = a->m; g.k = a->k;
            g.xqs = pl.qs + c0 * a->k; g.xd = pl.d + c0 * (a->k / pl.block); g.xbs = pl.bs + c0 * (a->k >> 4);
            g.act_block = pl.block;
            g.y = a->y + c0 * a->y_row_stride; g.ldy = a->y_row_stride; g.ncols = nc;
            g.pdl = pdl;
            B200_REQUIRE(a->layout == B200_LAYOUT_CANONICAL, "b200_mul_mat: plane layout with unsupported stride/alignment");
            rc = b200_gemv_generic_launch(a->type, g, 1, s);
        }
        if (rc) return rc;
    }
    return 0;
}

int b200_mul_mat(const b200_mul_mat_args * a, b200_stream_t stream) { return mul_mat_impl(a, (cudaStream_t) stream, nullptr); }

// Y = W * glu(x, x2): the activation quantiser gates on the fly (act_quant_kernel), then the register GEMV. Decode shapes only (n <= 8).
int b200_mul_mat_glu(const b200_mul_mat_args * a, const float * x2, int64_t x2_row_stride, int32_t glu_op, int32_t swapped, b200_stream_t stream) {
    B200_REQUIRE(a && x2, "b200_mul_mat_glu: null pointer");
    B200_REQUIRE(glu_op >= 0 && glu_op <= 2, "b200_mul_mat_glu: op %d (0 reglu, 1 geglu, 2 swiglu)", glu_op);
    B200_REQUIRE(a->n >= 1 && a->n <= 8 && !(a->flags & B200_MM_X_PREQUANTISED) && a->algo != 2 && a->algo != 3, "b200_mul_mat_glu: decode shapes (n <= 8) on the quantiser + GEMV path only");
    const ActGlu g = {x2, x2_row_stride, (int) glu_op, (int) swapped};
    return mul_mat_impl(a, (cudaStream_t) stream, nullptr, &g);
}

// do `count` nodes that read the same X all take the GEMM, or all take the quantiser + register GEMV with the same activation
// format? Then the first call converts X into the common workspace and the others reuse it (B200_MM_X_PREQUANTISED).
static bool group_shares_x(const b200_mul_mat_args * args, int count) {
    const b200_mul_mat_args & a0 = args[0];
    if (!a0.workspace || count < 2) return false;
    int path0 = -1;
    for (int i = 0; i < count; ++i) {
        const b200_mul_mat_args & a = args[i];
        if (!(a.x == a0.x && a.n == a0.n && a.k == a0.k && a.x_row_stride == a0.x_row_stride && a.workspace == a0.workspace &&
              a.flavour == a0.flavour && a.algo == 0 && b200_type_is_weight(a.type) && !(a.flags & B200_MM_X_PREQUANTISED))) return false;
        const bool fast = planes_fast_path(a.type, a.layout, a.m, a.k, a.w, a.w_row_stride);
        int path = 2;                                                         // quantiser + register / generic GEMV
        if (fast && b200_gemm_wants(a.type, a.m, a.n, a.k, false)) path = 0;  // tensor-core GEMM
        else if (fast && a.n <= 4) {
            GemvGroup g = {};
            g.count = 1; g.type[0] = a.type; g.w[0] = a.w; g.m[0] = a.m; g.k = a.k; g.ncols = (int) a.n;
            if (b200_gemv_stream_wants(g)) path = 1;                          // fused kernel: quantises for itself
        }
        if (path == 1 || b200_act_type(a.type) != b200_act_type(a0.type)) return false;
        if (path0 < 0) path0 = path; else if (path != path0) return false;
    }
    return true;
}

// several MUL_MAT nodes that read the same X (Q/K/V, gate/up): one streaming launch when eligible
// side stream + fork / join events of the calling thread's current device (created on first use; B200_GEMM_SIDE=0: off)
struct SideStream { cudaStream_t s; cudaEvent_t ready, done; };
static SideStream * side_stream() {
    static const bool on = [] { const char * e = getenv("B200_GEMM_SIDE"); return !(e && e[0] == '0'); }();
    if (!on) return nullptr;
    static thread_local SideStream per_dev[16] = {};
    static thread_local bool have[16] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
    if (!have[dev]) {
        SideStream t = {};
        if (cudaStreamCreateWithFlags(&t.s, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&t.ready, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&t.done, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        per_dev[dev] = t; have[dev] = true;
    }
    return &per_dev[dev];
}

int b200_mul_mat_group(const b200_mul_mat_args * args, int count, b200_stream_t stream) {
    B200_REQUIRE(args && count >= 1, "b200_mul_mat_group: no arguments");
    const b200_mul_mat_args & a0 = args[0];
    bool same = count <= B200_MAX_GROUP && a0.n <= 4 && a0.workspace != nullptr;
    for (int i = 0; i < count && same; ++i) {
        const b200_mul_mat_args & a = args[i];
        same = a.w && a.x && a.y && a.x == a0.x && a.n == a0.n && a.k == a0.k && a.x_row_stride == a0.x_row_stride &&
               a.flavour == a0.flavour && a.algo == 0 && b200_type_is_weight(a.type) &&
               a.m > 0 && a.n > 0 && a.k % b200_block_elems(a.type) == 0 && !(a.flags & B200_MM_X_PREQUANTISED) &&
               planes_fast_path(a.type, a.layout, a.m, a.k, a.w, a.w_row_stride);
    }
    if (same) {
        GemvGroup g = {};
        g.count = count;
        for (int i = 0; i < count; ++i) { g.type[i] = args[i].type; g.w[i] = args[i].w; g.m[i] = args[i].m; g.y[i] = args[i].y; g.ldy[i] = args[i].y_row_stride; }
        g.k = a0.k; g.ncols = (int) a0.n; g.x = a0.x; g.ldx = a0.x_row_stride;
        g.flavour = a0.flavour; g.pdl = (a0.flags & B200_MM_PDL) != 0;
        const int act = b200_act_type(a0.type);
        if (a0.workspace_size >= xq_planes_bytes(act, a0.n, a0.k) && (((uintptr_t) a0.workspace) & 255) == 0)
            g.xq_out = xq_planes_carve(a0.workspace, act, a0.n, a0.k);
        if (b200_gemv_stream_wants(g)) return b200_gemv_stream_launch(g, (cudaStream_t) stream);
    }
    // prefill: nodes of the same weight type share ONE tensor-core GEMM launch (one tile list over up to 3 matrices), and all of
    // them the fp16 copy of X
    if (group_shares_x(args, count)) {
        bool all_gemm = true;
        for (int i = 0; i < count && all_gemm; ++i)
            all_gemm = planes_fast_path(args[i].type, args[i].layout, args[i].m, args[i].k, args[i].w, args[i].w_row_stride) &&
                       b200_gemm_wants(args[i].type, args[i].m, args[i].n, args[i].k, false) && args[i].m > 0 &&
                       args[i].workspace_size >= b200_mul_mat_workspace(args[i].type, args[i].n, args[i].k) && (((uintptr_t) args[i].workspace) & 255) == 0;
        static const bool group_gemm = [] { const char * e = getenv("B200_GEMM_GROUP"); return !(e && e[0] == '0'); }();
        if (all_gemm && group_gemm) {
            bool done[16] = {false};
            bool first = true, forked = false;
            B200_REQUIRE(count <= 16, "b200_mul_mat_group: at most 16 nodes per call");
            // the buckets after the first only need the fp16 X of the first one's prologue: they run on a side stream next to the
            // first bucket's GEMM (wq + wk fill 80 of the 148 SMs, the Q6_K wv of every other layer takes the rest) and join at the end
            SideStream * side = side_stream();
            cudaStream_t main_s = (cudaStream_t) stream;
            for (int i = 0; i < count; ++i) {
                if (done[i]) continue;
                b200_mul_mat_args bucket[3];
                int nb = 0, left = 0;
                for (int j = i; j < count && nb < 3; ++j)
                    if (!done[j] && args[j].type == args[i].type) { bucket[nb++] = args[j]; done[j] = true; }
                for (int j = 0; j < count; ++j) left += done[j] ? 0 : 1;
                int rc;
                if (first) {
                    const bool fork = side && left > 0;
                    rc = b200_gemm_launch_group_ev(bucket, nb, main_s, fork ? side->ready : nullptr);
                    if (!rc && fork) { B200_CHECK(cudaStreamWaitEvent(side->s, side->ready, 0)); forked = true; }
                } else {
                    bucket[0].flags |= B200_MM_X_PREQUANTISED;
                    rc = b200_gemm_launch_group(bucket, nb, forked ? side->s : main_s);
                }
                first = false;
                if (rc) return rc;
            }
            if (forked) {
                B200_CHECK(cudaEventRecord(side->done, side->s));
                B200_CHECK(cudaStreamWaitEvent(main_s, side->done, 0));
            }
            return 0;
        }
    }
    // not groupable into one launch: one call per node; the nodes share the converted X (fp16 copy for the GEMM, int8 planes for
    // the register GEMV) when they all take the same of those two paths
    const bool share_x = group_shares_x(args, count);
    for (int i = 0; i < count; ++i) {
        b200_mul_mat_args a = args[i];
        if (share_x && i > 0) a.flags |= B200_MM_X_PREQUANTISED;
        int rc = mul_mat_impl(&a, (cudaStream_t) stream, nullptr);
        if (rc) return rc;
    }
    return 0;
}


// B200_PEER_DEBUG (timing experiments only, results wrong): 1 = nobody waits, 2 = nobody announces
static int peer_debug_flags() { static const int v = [] { const char * e = getenv("B200_PEER_DEBUG"); return e ? atoi(e) : 0; }(); return v; }

// row slices of `count` nodes that read the same X + the exchange fused into the kernels (include/b200_mm.h)
int b200_mul_mat_gather(const b200_mul_mat_args * args, int count, const b200_row_gather * gt, b200_stream_t stream) {
    cudaStream_t s = (cudaStream_t) stream;
    B200_REQUIRE(args && count >= 1 && gt, "b200_mul_mat_gather: no arguments");
    B200_REQUIRE(gt->world >= 2 && gt->world <= B200_MAX_PEERS && gt->rank >= 0 && gt->rank < gt->world && gt->sync && gt->step &&
                 gt->seq >= 0 && gt->seq < gt->calls_per_step && gt->peer_delta[gt->rank] == 0,
                 "b200_mul_mat_gather: bad descriptor (world %d rank %d seq %lld/%lld)", gt->world, gt->rank, (long long) gt->seq, (long long) gt->calls_per_step);
    PeerGather p = {};
    p.world = gt->world; p.rank = gt->rank;
    for (int g = 0; g < gt->world; ++g) p.delta[g] = gt->peer_delta[g];
    p.sync = (unsigned long long *) gt->sync; p.step = (const unsigned long long *) gt->step;
    p.seq = (unsigned long long) gt->seq; p.calls = (unsigned long long) gt->calls_per_step;

    const int peer_debug = peer_debug_flags();
    const int dbg_wait = (peer_debug & 1) ? 0 : 1, dbg_signal = (peer_debug & 2) ? 0 : 1;
    // ranks whose slice of a matrix is empty still take part in the exchange
    b200_mul_mat_args live[B200_MAX_GROUP];
    int nlive = 0;
    B200_REQUIRE(count <= B200_MAX_GROUP, "b200_mul_mat_gather: at most %d nodes per call", B200_MAX_GROUP);
    for (int i = 0; i < count; ++i) if (args[i].m > 0 && args[i].n > 0) live[nlive++] = args[i];
    if (nlive == 0) {
        p.signal = 1;
        B200_CHECK(b200_launch(peer_signal_kernel, dim3(1), dim3(32), 0, s, false, p));
        return 0;
    }
    const b200_mul_mat_args & a0 = live[0];
    bool same = a0.n <= 4 && a0.workspace != nullptr;
    for (int i = 0; i < nlive && same; ++i) {
        const b200_mul_mat_args & a = live[i];
        same = a.w && a.x && a.y && a.x == a0.x && a.n == a0.n && a.k == a0.k && a.x_row_stride == a0.x_row_stride &&
               a.flavour == a0.flavour && a.algo == 0 && b200_type_is_weight(a.type) && a.k % b200_block_elems(a.type) == 0 &&
               !(a.flags & B200_MM_X_PREQUANTISED) && planes_fast_path(a.type, a.layout, a.m, a.k, a.w, a.w_row_stride);
    }
    if (same) {
        GemvGroup g = {};
        g.count = nlive;
        for (int i = 0; i < nlive; ++i) { g.type[i] = live[i].type; g.w[i] = live[i].w; g.m[i] = live[i].m; g.y[i] = live[i].y; g.ldy[i] = live[i].y_row_stride; }
        g.k = a0.k; g.ncols = (int) a0.n; g.x = a0.x; g.ldx = a0.x_row_stride;
        g.flavour = a0.flavour; g.pdl = (a0.flags & B200_MM_PDL) != 0;
        g.peer = p; g.peer.wait = dbg_wait; g.peer.signal = dbg_signal;
        if (b200_gemv_stream_wants(g)) return b200_gemv_stream_launch(g, s);
    }
    // one node at a time: the first waits, the last announces; the nodes share the converted X where they can
    const bool share_x = group_shares_x(live, nlive);
    for (int i = 0; i < nlive; ++i) {
        b200_mul_mat_args a = live[i];
        if (share_x && i > 0) a.flags |= B200_MM_X_PREQUANTISED;
        PeerGather pi = p;
        pi.wait = (i == 0) && dbg_wait; pi.signal = (i == nlive - 1) && dbg_signal;
        int rc = mul_mat_impl(&a, s, &pi);
        if (rc) return rc;
    }
    return 0;
}

int b200_row_gather_wait(const b200_row_gather * gt, b200_stream_t stream) {
    B200_REQUIRE(gt && gt->world >= 2 && gt->world <= B200_MAX_PEERS && gt->sync && gt->step && gt->seq >= 0 && gt->seq <= gt->calls_per_step,
                 "b200_row_gather_wait: bad descriptor");
    if (peer_debug_flags()) return 0;
    PeerGather p = {};
    p.world = gt->world; p.rank = gt->rank;
    p.sync = (unsigned long long *) gt->sync; p.step = (const unsigned long long *) gt->step;
    p.seq = (unsigned long long) gt->seq; p.calls = (unsigned long long) gt->calls_per_step;
    B200_CHECK(b200_launch(peer_wait_kernel, dim3(1), dim3(32), 0, (cudaStream_t) stream, false, p));
    return 0;
}

static __global__ void step_kernel(unsigned long long * step) { if (threadIdx.x == 0) *step += 1ull; }
int b200_row_gather_step(uint64_t * step, b200_stream_t stream) {
    B200_REQUIRE(step, "b200_row_gather_step: null pointer");
    B200_CHECK(b200_launch(step_kernel, dim3(1), dim3(32), 0, (cudaStream_t) stream, false, (unsigned long long *) step));
    return 0;
}

int b200_ipc_export(const void * dev_ptr, void * handle64) {
    B200_REQUIRE(dev_ptr && handle64, "b200_ipc_export: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    B200_CHECK(cudaIpcGetMemHandle((cudaIpcMemHandle_t *) handle64, const_cast<void *>(dev_ptr)));
    return 0;
}
int b200_ipc_import(const void * handle64, void ** dev_ptr) {
    B200_REQUIRE(dev_ptr && handle64, "b200_ipc_import: null pointer");
    cudaIpcMemHandle_t h; memcpy(&h, handle64, sizeof h);
    B200_CHECK(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int b200_ipc_close(void * dev_ptr) {
    B200_REQUIRE(dev_ptr, "b200_ipc_close: null pointer");
    B200_CHECK(cudaIpcCloseMemHandle(dev_ptr));
    return 0;
}

size_t b200_mul_mat_id_workspace(int type, int64_t k, int64_t n_used, int64_t nb1, int64_t n_tokens) {
    const int act = b200_act_type(type);
    if (act < 0) return 0;
    (void) n_used;
    return xq_planes_bytes(act, n_tokens * nb1, k) + 256;
}

size_t b200_mul_mat_id_workspace_ex(const b200_mul_mat_id_args * a) {
    if (!a) return 0;
    const size_t v = b200_mul_mat_id_workspace(a->type, a->k, a->n_used, a->nb1, a->n_tokens);
    const size_t g = b200_gemm_moe_workspace(a->type, a->m, a->k, a->n_as, a->n_tokens * a->n_used);
    return v > g ? v : g;
}

static int mul_mat_id_impl(const b200_mul_mat_id_args * a, b200_stream_t stream, const ActGlu * glu);
int b200_mul_mat_id(const b200_mul_mat_id_args * a, b200_stream_t stream) { return mul_mat_id_impl(a, stream, nullptr); }
// Y = W_expert * glu(b, b2) per (token, slot): the per-pair GEMV path with the gate folded into the activation quantiser (decode: few pairs)
int b200_mul_mat_id_glu(const b200_mul_mat_id_args * a, const float * b2, int64_t b2_row_stride, int32_t glu_op, int32_t swapped, b200_stream_t stream) {
    B200_REQUIRE(a && b2, "b200_mul_mat_id_glu: null pointer");
    B200_REQUIRE(glu_op >= 0 && glu_op <= 2, "b200_mul_mat_id_glu: op %d (0 reglu, 1 geglu, 2 swiglu)", glu_op);
    B200_REQUIRE(a->n_tokens * a->n_used <= 8 && a->algo != 2 && !(a->flags & B200_MM_X_PREQUANTISED) && a->b_token_stride == a->nb1 * a->b_row_stride,
                 "b200_mul_mat_id_glu: decode shapes (at most 8 pairs, b rows packed token by token) on the per-pair GEMV path only");
    const ActGlu g = {b2, b2_row_stride, (int) glu_op, (int) swapped};
    return mul_mat_id_impl(a, stream, &g);
}
static int mul_mat_id_impl(const b200_mul_mat_id_args * a, b200_stream_t stream, const ActGlu * glu) {
    cudaStream_t s = (cudaStream_t) stream;
    B200_REQUIRE(a && a->as && a->b && a->ids && a->y, "b200_mul_mat_id: null pointer");
    B200_REQUIRE(b200_type_is_weight(a->type), "b200_mul_mat_id: unsupported weight type %d", a->type);
    const int be = b200_block_elems(a->type);
    B200_REQUIRE(a->k > 0 && a->k % be == 0, "b200_mul_mat_id: k=%lld not a multiple of %d", (long long) a->k, be);
    B200_REQUIRE(a->nb1 == 1 || a->nb1 == a->n_used, "b200_mul_mat_id: b has %lld rows per token, want 1 or n_used=%lld",
                 (long long) a->nb1, (long long) a->n_used);
    if (a->m == 0 || a->n_tokens == 0 || a->n_used == 0) return 0;
    const size_t need = b200_mul_mat_id_workspace(a->type, a->k, a->n_used, a->nb1, a->n_tokens);
    B200_REQUIRE(a->workspace && a->workspace_size >= need, "b200_mul_mat_id: workspace %zu < %zu bytes", a->workspace_size, need);
    B200_REQUIRE((((uintptr_t) a->workspace) & 255) == 0, "b200_mul_mat_id: workspace must be 256 B aligned");
    const bool pdl = (a->flags & B200_MM_PDL) != 0;
    const int act = b200_act_type(a->type);
    const int64_t rows = a->n_tokens * a->nb1;
    // (4) many pairs per expert: sort the pairs by expert on the device, gather their activation rows, ONE tensor-core GEMM launch
    //     over (expert, tile), Y rows scattered back (gemm.cu: b200_gemm_moe_launch)
    {
        const bool fast_id = planes_fast_path(a->type, a->layout, a->m, a->k, a->as, 0) && (a->as_expert_stride % 16) == 0;
        const int64_t y_span = (a->n_tokens - 1) * a->y_token_stride + (a->n_used - 1) * a->y_row_stride + a->m;
        const bool strides_ok = a->y_token_stride >= 0 && a->y_row_stride >= 0 && a->b_row_stride >= 0 && a->b_token_stride >= 0;
        const bool want = !glu && a->algo != 1 && fast_id && strides_ok && b200_gemm_moe_wants(a->type, a->m, a->k, a->n_as, a->n_used, a->n_tokens, y_span, a->algo == 2);
        const bool room = a->workspace_size >= b200_gemm_moe_workspace(a->type, a->m, a->k, a->n_as, a->n_tokens * a->n_used);
        B200_REQUIRE(a->algo != 2 || (want && room), "b200_mul_mat_id: GEMM forced but the call is not eligible (layout / strides / workspace %zu bytes, see b200_mul_mat_id_workspace_ex)",
                     a->workspace_size);
        if (want && room) return b200_gemm_moe_launch(a, s);
    }
    const XqPlanes pl = xq_planes_carve(a->workspace, act, rows, a->k);

    // quantise every b row once: token-major, nb1 rows per token (b strides may be arbitrary). B200_MM_X_PREQUANTISED: the previous call left exactly
    // these rows in the workspace (ffn_gate_exps / ffn_up_exps read the same activations)
    if (a->flags & B200_MM_X_PREQUANTISED) {
    } else if (a->b_token_stride == a->nb1 * a->b_row_stride) {
        int rc = b200_act_quant_planes(act, a->flavour, a->b, a->b_row_stride, pl, rows, a->k, s, pdl, glu);
        if (rc) return rc;
    } else {
        for (int64_t t = 0; t < a->n_tokens; ++t) {
            XqPlanes pt = pl;
            pt.qs += t * a->nb1 * a->k; pt.d += t * a->nb1 * (a->k / pl.block); pt.bs += t * a->nb1 * (a->k >> 4);
            int rc = b200_act_quant_planes(act, a->flavour, a->b + t * a->b_token_stride, a->b_row_stride, pt, a->nb1, a->k, s, pdl);
            if (rc) return rc;
        }
    }
    const int64_t pairs = a->n_tokens * a->n_used;
    const bool fast = planes_fast_path(a->type, a->layout, a->m, a->k, a->as, 0) && (a->as_expert_stride % 16) == 0;
    B200_REQUIRE(pairs <= 65535, "b200_mul_mat_id: %lld (token, slot) pairs exceed one launch; split the batch", (long long) pairs);
    if (fast) {
        GemvArgs g = {};
        g.W = gemv_w_make(a->type, a->as, a->m, a->k);
        g.m = a->m; g.k = a->k;
        g.xqs = pl.qs; g.xd = pl.d; g.xbs = pl.bs; g.act_block = pl.block;
        g.y = a->y; g.ldy = a->y_row_stride;
        g.ids = a->ids; g.ids_row_stride = a->ids_row_stride; g.n_used = a->n_used; g.nb1 = a->nb1; g.n_as = a->n_as;
        g.expert_stride = a->as_expert_stride; g.y_token_stride = a->y_token_stride;
        g.pdl = pdl;
        return b200_gemv_launch(a->type, g, 1, pairs, a->reserved, s);
    }
    B200_REQUIRE(a->layout == B200_LAYOUT_CANONICAL, "b200_mul_mat_id: plane layout with unsupported stride/alignment");
    GemvGenericArgs g = {};
    g.w = (const uint8_t *) a->as; g.w_row_stride = (int64_t) b200_row_size(a->type, a->k);
    g.m = a->m; g.k = a->k;
    g.xqs = pl.qs; g.xd = pl.d; g.xbs = pl.bs; g.act_block = pl.block;
    g.y = a->y; g.ldy = a->y_row_stride; g.ncols = 1;
    g.ids = a->ids; g.ids_row_stride = a->ids_row_stride; g.n_used = a->n_used; g.nb1 = a->nb1; g.n_as = a->n_as;
    g.expert_stride = a->as_expert_stride; g.y_token_stride = a->y_token_stride;
    g.pdl = pdl;
    return b200_gemv_generic_launch(a->type, g, pairs, s);
}

} // extern "C"

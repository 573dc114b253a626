// mul_mat.cu -- C-ABI entry points b200_mul_mat / b200_mul_mat_id: quantise X, pick the kernel.
// Replaces ggml_compute_forward_mul_mat / _mul_mat_id (ggml-cpu/ggml-cpu.c:1227-1427, 1500-1676).
#include "actquant.cuh"
#include "gemv.h"
#include "gemm.h"

static bool planes_fast_path(int type, int layout, int64_t m, int64_t k, const void * w, int64_t w_row_stride) {
    const bool packed = w_row_stride == 0 || w_row_stride == (int64_t) b200_row_size(type, k);
    if (!packed || (k % 256) != 0 || (((uintptr_t) w) & 15) != 0) return false;
    if (layout == B200_LAYOUT_PLANES) return true;
    return type == B200_TYPE_Q4_K;   // canonical Q4_K is its own plane layout
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

int b200_mul_mat(const b200_mul_mat_args * a, b200_stream_t stream) {
    cudaStream_t s = (cudaStream_t) stream;
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
    if (a->algo != 1 && fast && b200_gemm_wants(a->type, a->m, a->n, a->k, a->algo == 2)) {
        return b200_gemm_launch(a, s);
    }
    B200_REQUIRE(a->algo != 2, "b200_mul_mat: GEMM forced but shape/layout is not eligible (type %d m=%lld n=%lld k=%lld)",
                 a->type, (long long) a->m, (long long) a->n, (long long) a->k);

    const int act = b200_act_type(a->type);
    const XqPlanes pl = xq_planes_carve(a->workspace, act, a->n, a->k);

    // (1) decode, streaming kernel: one launch, activation quantiser fused (quantised X also left in the workspace)
    if (fast && a->n <= 4 && a->algo != 1 && !(a->flags & B200_MM_X_PREQUANTISED)) {
        GemvGroup g = {};
        g.count = 1; g.type[0] = a->type; g.w[0] = a->w; g.m[0] = a->m; g.y[0] = a->y;
        g.ldy[0] = a->y_row_stride;
        g.k = a->k; g.ncols = (int) a->n; g.x = a->x; g.ldx = a->x_row_stride;
        g.flavour = a->flavour; g.pdl = pdl; g.xq_out = pl;
        if (b200_gemv_stream_wants(g)) return b200_gemv_stream_launch(g, s);
    }
    B200_REQUIRE(a->algo != 3, "b200_mul_mat: streaming GEMV forced but the call is not eligible (type %d m=%lld n=%lld k=%lld)",
                 a->type, (long long) a->m, (long long) a->n, (long long) a->k);

    // (3) + (1): quantise activations into the workspace planes, then register GEMV in column groups of <= 8
    if (!(a->flags & B200_MM_X_PREQUANTISED)) {
        int rc = b200_act_quant_planes(act, a->flavour, a->x, a->x_row_stride, pl, a->n, a->k, s, pdl);
        if (rc) return rc;
    }
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

// several MUL_MAT nodes that read the same X (Q/K/V, gate/up): one streaming launch when eligible
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
    // not groupable into one launch: one call per node. Prefill nodes that all take the GEMM share the fp16 copy of X.
    bool share_x = a0.n > 8 && a0.workspace != nullptr;
    for (int i = 0; i < count && share_x; ++i) {
        const b200_mul_mat_args & a = args[i];
        share_x = a.x == a0.x && a.n == a0.n && a.k == a0.k && a.x_row_stride == a0.x_row_stride && a.workspace == a0.workspace &&
                  a.algo == 0 && b200_type_is_weight(a.type) && planes_fast_path(a.type, a.layout, a.m, a.k, a.w, a.w_row_stride) &&
                  b200_gemm_wants(a.type, a.m, a.n, a.k, false);
    }
    for (int i = 0; i < count; ++i) {
        b200_mul_mat_args a = args[i];
        if (share_x && i > 0) a.flags |= B200_MM_X_PREQUANTISED;
        int rc = b200_mul_mat(&a, stream);
        if (rc) return rc;
    }
    return 0;
}

size_t b200_mul_mat_id_workspace(int type, int64_t k, int64_t n_used, int64_t nb1, int64_t n_tokens) {
    const int act = b200_act_type(type);
    if (act < 0) return 0;
    (void) n_used;
    return xq_planes_bytes(act, n_tokens * nb1, k) + 256;
}

int b200_mul_mat_id(const b200_mul_mat_id_args * a, b200_stream_t stream) {
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
    const XqPlanes pl = xq_planes_carve(a->workspace, act, rows, a->k);

    // quantise every b row once: token-major, nb1 rows per token (b strides may be arbitrary)
    if (a->b_token_stride == a->nb1 * a->b_row_stride) {
        int rc = b200_act_quant_planes(act, a->flavour, a->b, a->b_row_stride, pl, rows, a->k, s, pdl);
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

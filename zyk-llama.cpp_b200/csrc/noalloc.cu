// noalloc.cu -- b200_mul_mat_noalloc: the decode MUL_MAT (activation quantiser + register GEMV, kernels 3 + 1b of DESIGN.md)
// for callers that may not allocate: code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// No counterpart in the reference. b200_mul_mat takes the scratch memory of
// the quantised activations (int8 planes + scales + block sums) from the caller's `workspace`. Here that scratch is a
// file-scope device array of B200_NOALLOC_SCRATCH_BYTES: every GPU that loads the library holds it until exit, every call on
// a device shares it, and it keeps the last call's contents. The quantiser overwrites every byte the GEMV reads, so nothing
// has to start at zero. Two calls that overlap on the GPU (different streams, no dependency) race on the array.
//
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: its address is asked for once per
// device (that device current) with cudaGetSymbolAddress and handed to the kernels as an ordinary pointer. The entry point
// calls no cudaMalloc / cudaFree / stream or event creation, neither directly nor through the launchers it uses
// (b200_act_quant_planes, b200_gemv_launch, b200_gemv_generic_launch: kernel launches only).
#include "actquant.cuh"
#include "gemv.h"

#include <mutex>

#define B200_NOALLOC_SCRATCH_BYTES (1u << 20)     /* 1 MiB: e.g. n = 8 rows of k = 65536 need 0.79 MiB */
#define B200_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// device address of the array on the CURRENT device, resolved once per device
static int noalloc_scratch(void ** out) {
    static std::mutex mu;
    static void * ptr[B200_NOALLOC_MAX_DEVICES] = {nullptr};
    int dev = -1;
    B200_CHECK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < B200_NOALLOC_MAX_DEVICES, "b200_mul_mat_noalloc: device %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void * p = nullptr;
        B200_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        B200_REQUIRE(p != nullptr && (((uintptr_t) p) & 255) == 0, "b200_mul_mat_noalloc: scratch symbol resolved to %p", p);
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return 0;
}

extern "C" {

size_t b200_noalloc_scratch_max(void) { return B200_NOALLOC_SCRATCH_BYTES; }

// bytes of the shared array a call with this shape uses (0: type unknown)
size_t b200_mul_mat_noalloc_scratch(int type, int64_t n, int64_t k) {
    const int act = b200_act_type(type);
    return act < 0 ? 0 : xq_planes_bytes(act, n, k);
}

int b200_mul_mat_noalloc(const b200_mul_mat_args * a, b200_stream_t stream) {
    cudaStream_t s = (cudaStream_t) stream;
    B200_REQUIRE(a && a->w && a->x && a->y, "b200_mul_mat_noalloc: null pointer");
    B200_REQUIRE(b200_type_is_weight(a->type), "b200_mul_mat_noalloc: unsupported weight type %d", a->type);
    const int be = b200_block_elems(a->type);
    B200_REQUIRE(a->k > 0 && a->k % be == 0, "b200_mul_mat_noalloc: k=%lld not a multiple of the block size %d", (long long) a->k, be);
    B200_REQUIRE(a->m >= 0 && a->n >= 0, "b200_mul_mat_noalloc: negative shape");
    if (a->m == 0 || a->n == 0) return 0;
    B200_REQUIRE(a->layout == B200_LAYOUT_CANONICAL || b200_repack_supported(a->type, a->m, a->k),
                 "b200_mul_mat_noalloc: plane layout is not defined for type %d m=%lld k=%lld", a->type, (long long) a->m, (long long) a->k);
    B200_REQUIRE(!(a->flags & B200_MM_X_PREQUANTISED), "b200_mul_mat_noalloc: the shared scratch holds no earlier call's X");
    const int act = b200_act_type(a->type);
    const size_t need = xq_planes_bytes(act, a->n, a->k);
    B200_REQUIRE(need <= B200_NOALLOC_SCRATCH_BYTES,
                 "b200_mul_mat_noalloc: n=%lld x k=%lld needs %zu bytes of scratch, the fixed array holds %u (use b200_mul_mat with a workspace)",
                 (long long) a->n, (long long) a->k, need, (unsigned) B200_NOALLOC_SCRATCH_BYTES);
    void * scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;

    const bool pdl = (a->flags & B200_MM_PDL) != 0;
    const bool packed = a->w_row_stride == 0 || a->w_row_stride == (int64_t) b200_row_size(a->type, a->k);
    const bool fast = packed && (a->k % 256) == 0 && (((uintptr_t) a->w) & 15) == 0 &&
                      (a->layout == B200_LAYOUT_PLANES || a->type == B200_TYPE_Q4_K || a->type == B200_TYPE_Q5_K);
    B200_REQUIRE(fast || a->layout == B200_LAYOUT_CANONICAL, "b200_mul_mat_noalloc: plane layout with unsupported stride/alignment");

    const XqPlanes pl = xq_planes_carve(scratch, act, a->n, a->k);
    if (int rc = b200_act_quant_planes(act, a->flavour, a->x, a->x_row_stride, pl, a->n, a->k, s, pdl)) return rc;
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
            rc = b200_gemv_generic_launch(a->type, g, 1, s);
        }
        if (rc) return rc;
    }
    return 0;
}

} // extern "C"

// gemv.h -- internal interface of the decode GEMV kernels (gemv.cu, gemv_generic.cu).
#pragma once
#include "gemv_frag.cuh"
#include "actquant.cuh"

struct GemvArgs {
    GemvW W;                 // plane pointers of ONE [m][k] matrix (expert 0 for MUL_MAT_ID)
    int64_t m, k;
    // quantised activations, plane layout (actquant.cuh); column c is `c` rows further
    const int8_t * xqs; const float * xd; const int16_t * xbs;
    int act_block;           // 32 (Q8_0 scales) or 256 (Q8_K scales)
    float * y; int64_t ldy;  // y[c * ldy + row]
    // MUL_MAT_ID (ids == nullptr for plain MUL_MAT)
    const int32_t * ids; int64_t ids_row_stride; int64_t n_used, nb1, n_as, expert_stride, y_token_stride;
    int pdl;
    PeerGather peer;         // fused row-split exchange: peer stores + arrival (the activation quantiser does the wait)
};

// plane-layout fast path (also canonical packed Q4_K). ncols in 1..8. `pairs` = grid.y
// (n_tokens * n_used for MUL_MAT_ID, else 1). tune: 0 = auto, else R | U << 8.
int b200_gemv_launch(int type, const GemvArgs & a, int ncols, int64_t pairs, int tune, cudaStream_t s);

// canonical-layout fallback for shapes/strides the plane path does not take
struct GemvGenericArgs {
    const uint8_t * w; int64_t w_row_stride;   // bytes
    int64_t m, k;
    const int8_t * xqs; const float * xd; const int16_t * xbs; int act_block;
    float * y; int64_t ldy; int ncols;
    const int32_t * ids; int64_t ids_row_stride; int64_t n_used, nb1, n_as, expert_stride, y_token_stride;
    int pdl;
};
int b200_gemv_generic_launch(int type, const GemvGenericArgs & a, int64_t pairs, cudaStream_t s);

// streaming GEMV (gemv_stream.cu): up to B200_MAX_GROUP matrices that share the same f32 X (ncols <= 4 rows),
// activation quantiser fused. xq_out (optional, qs == nullptr to skip) receives the quantised X planes.
#define B200_MAX_GROUP 4
struct GemvGroup {
    int count;
    int type[B200_MAX_GROUP];
    const void * w[B200_MAX_GROUP];
    int64_t m[B200_MAX_GROUP];
    float * y[B200_MAX_GROUP];
    int64_t k; int ncols;
    int64_t ldy[B200_MAX_GROUP];
    const float * x; int64_t ldx;
    int flavour, pdl;
    XqPlanes xq_out;
    PeerGather peer;         // fused row-split exchange (world <= 1: off)
};
bool b200_gemv_stream_wants(const GemvGroup & g, bool force = false);   // force: caller set algo == 3
int  b200_gemv_stream_launch(const GemvGroup & g, cudaStream_t s);

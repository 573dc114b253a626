// quantize.cu -- (3) vectorised activation quantiser: host launchers + C-ABI entry point.
#include "actquant.cuh"

template <int QT, int FL, bool CANON>
static int launch_act_quant(const float * x, int64_t ldx, uint8_t * y, const XqPlanes & pl, int64_t nrows, int64_t k,
                            cudaStream_t s, bool pdl, const ActGlu & glu) {
    if (nrows == 0 || k == 0) return 0;
    const int64_t nchunks = (k + 255) / 256;
    dim3 grid((unsigned) ((nchunks + 3) / 4), (unsigned) (nrows < 32768 ? nrows : 32768));
    B200_CHECK(b200_launch(act_quant_kernel<QT, FL, CANON>, grid, dim3(128), 0, s, pdl, x, ldx, nrows, k, y, pl, pdl, glu));
    return 0;
}

template <bool CANON>
static int dispatch_act_quant(int qt, int flavour, const float * x, int64_t ldx, uint8_t * y, const XqPlanes & pl,
                              int64_t nrows, int64_t k, cudaStream_t s, bool pdl, const ActGlu * glu_in = nullptr) {
    ActGlu glu = {nullptr, 0, -1, 0};
    if (glu_in) glu = *glu_in;
    const bool x86 = flavour != B200_Q8_REF;
    switch (qt) {
        case B200_TYPE_Q8_K:
            B200_REQUIRE(k % 256 == 0, "quantize: Q8_K needs k %% 256 == 0 (k=%lld)", (long long) k);
            return launch_act_quant<B200_TYPE_Q8_K, B200_Q8_X86, CANON>(x, ldx, y, pl, nrows, k, s, pdl, glu);
        case B200_TYPE_Q8_0:
            B200_REQUIRE(k % 32 == 0, "quantize: Q8_0 needs k %% 32 == 0 (k=%lld)", (long long) k);
            return x86 ? launch_act_quant<B200_TYPE_Q8_0, B200_Q8_X86, CANON>(x, ldx, y, pl, nrows, k, s, pdl, glu)
                       : launch_act_quant<B200_TYPE_Q8_0, B200_Q8_REF, CANON>(x, ldx, y, pl, nrows, k, s, pdl, glu);
        case B200_TYPE_Q8_1:
            B200_REQUIRE(CANON, "quantize: Q8_1 is only produced as canonical blocks");
            B200_REQUIRE(k % 32 == 0, "quantize: Q8_1 needs k %% 32 == 0 (k=%lld)", (long long) k);
            return x86 ? launch_act_quant<B200_TYPE_Q8_1, B200_Q8_X86, CANON>(x, ldx, y, pl, nrows, k, s, pdl, glu)
                       : launch_act_quant<B200_TYPE_Q8_1, B200_Q8_REF, CANON>(x, ldx, y, pl, nrows, k, s, pdl, glu);
    }
    b200_set_error("quantize: unsupported activation type %d", qt);
    return -2;
}

int b200_act_quant_canonical(int q8_type, int flavour, const float * x, int64_t ldx, void * y, int64_t nrows, int64_t k, cudaStream_t s) {
    XqPlanes none = {};
    return dispatch_act_quant<true>(q8_type, flavour, x, ldx, (uint8_t *) y, none, nrows, k, s, false);
}

int b200_act_quant_planes(int act_type, int flavour, const float * x, int64_t ldx, const XqPlanes & pl, int64_t nrows, int64_t k,
                          cudaStream_t s, bool pdl, const ActGlu * glu) {
    return dispatch_act_quant<false>(act_type, flavour, x, ldx, nullptr, pl, nrows, k, s, pdl, glu);
}

extern "C" int b200_quantize_rows(int q8_type, int flavour, const float * x, int64_t x_row_stride, void * y, int64_t nrows,
                                  int64_t k, b200_stream_t stream) {
    if (nrows == 0 || k == 0) return 0;   // empty input: nothing to do, pointers may be null
    B200_REQUIRE(x && y, "b200_quantize_rows: null pointer");
    return b200_act_quant_canonical(q8_type, flavour, x, x_row_stride, y, nrows, k, (cudaStream_t) stream);
}

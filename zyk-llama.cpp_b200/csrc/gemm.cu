// gemm.cu -- (2) prefill GEMM. Placeholder until the tcgen05 kernel lands: declines every shape,
// so b200_mul_mat runs the GEMV in column groups of 8 (correct, not fast).
#include "gemm.h"

size_t b200_gemm_workspace(int, int64_t, int64_t) { return 0; }
bool b200_gemm_wants(int, int64_t, int64_t, int64_t, bool) { return false; }
int b200_gemm_launch(const b200_mul_mat_args *, cudaStream_t) {
    b200_set_error("gemm: not built");
    return -2;
}

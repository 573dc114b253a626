/* oracle/ggml_oracle.h -- TEST INFRASTRUCTURE ONLY (see ggml_oracle.c). */
#ifndef GGML_ORACLE_H
#define GGML_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ggml type ids (ggml/include/ggml.h:389-430) */
enum {
    ORC_F32 = 0, ORC_Q4_0 = 2, ORC_Q8_0 = 8, ORC_Q8_1 = 9, ORC_Q4_K = 12, ORC_Q5_K = 13, ORC_Q6_K = 14,
    ORC_Q8_K = 15, ORC_IQ4_XS = 23
};

/* activation-quantiser flavour */
enum { ORC_Q8_REF = 0 /* ggml-quants.c *_ref */, ORC_Q8_X86 = 1 /* arch/x86/quants.c AVX2 */ };

float    orc_fp16_to_fp32(uint16_t h);
uint16_t orc_fp32_to_fp16(float f);

int      orc_block_elems(int type);             /* 32 or 256, 0 if unknown           */
size_t   orc_block_bytes(int type);
size_t   orc_row_size(int type, int64_t k);     /* bytes of one row of k elements    */
int      orc_vec_dot_type(int type);            /* activation format the CPU pairs   */

void orc_dequantize_row(int type, const void * src, float * dst, int64_t k);

void orc_quantize_row_q8_0(const float * x, void * y, int64_t k, int flavour);
void orc_quantize_row_q8_1(const float * x, void * y, int64_t k, int flavour);
void orc_quantize_row_q8_K(const float * x, void * y, int64_t k);
void orc_quantize_row(int q8_type, const float * x, void * y, int64_t k, int flavour);

/* one output element: weight row (type) . quantised activation row (orc_vec_dot_type(type)) */
float orc_vec_dot(int type, int64_t k, const void * w_row, const void * xq_row);

/* Y[n][m] = sum_k W[m][k] X[n][k]; w = m canonical rows, x = f32 [n][k], y = f32 [n][m] */
void orc_mul_mat(int type, int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y, int flavour);

/* MUL_MAT_ID (ggml.c:3217-3243): as[n_as][m][k], b[n_tokens][nb1][k] (nb1 == n_used or 1),
 * ids[n_tokens][n_used] -> y[n_tokens][n_used][m] */
void orc_mul_mat_id(int type, int64_t m, int64_t k, int64_t n_as, int64_t n_used, int64_t nb1, int64_t n_tokens,
                    const void * as, const float * b, const int32_t * ids, float * y, int flavour);

/* plain f32 product against the dequantised weights, accumulated in double (the "true" answer the
 * NMSE tolerance of test-backend-ops is ultimately about) */
void orc_mul_mat_f64(int type, int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y);

/* ---- load-time repack: the B200 plane layout (DESIGN.md "HBM layout") and its inverse ---- */
int  orc_b200_repack_supported(int type, int64_t m, int64_t k);
void orc_b200_repack(int type, int64_t m, int64_t k, const void * canonical, void * planes);
void orc_b200_unrepack(int type, int64_t m, int64_t k, const void * planes, void * canonical);

/* ---- the reference CPU's own load-time repacks / interleaved activation formats ---- */
/* block_q4_Kx8 interleave of 8 rows (repack.cpp:1501-1575, 1652-1681); m % 8 == 0, k % 256 == 0 */
int  orc_cpu_repack_q4_K_8x8(int64_t m, int64_t k, const void * canonical, void * out);
/* fork ZYK transposed Q4_0 layout (repack.cpp:1746-1796) */
int  orc_cpu_repack_q4_0_zyk(int64_t m, int64_t k, const void * canonical, void * out);
/* fork activation format block_q8_2_x4 (repack.h:88-101, repack.cpp:2324-2462) */
void orc_quantize_row_q8_2_x4(const float * x, void * y, int64_t k);
/* CPU GEMM activation format block_q8_Kx4: 4 rows x k f32 (row stride k) -> k / 256 blocks of 1168 bytes
 * (repack.h:55-61, ggml_quantize_mat_q8_K_4x8_generic repack.cpp:180-229) */
void orc_quantize_mat_q8_K_4x8(const float * x, void * y, int64_t k);
/* the fork's ZYK Q4_0 matmul end to end: q8_2_x4 activations, unsigned nibbles with the -8 * s bias, fma per block in block
 * order (arch/x86/repack.cpp:186-243, 460-508); w canonical block_q4_0 rows, x f32 [n][k], y f32 [n][m] */
void orc_mul_mat_q4_0_zyk(int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y);

#ifdef __cplusplus
}
#endif
#endif

// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" shim over the UNMODIFIED reference's public API (ggml.h / ggml-cpu.h /
// ggml-backend.h), compiled into oracle/_ref/<variant>/libref_harness.so together with the
// reference's own libggml-{base,cpu}.so (see oracle/Makefile). It contains no arithmetic of its
// own: every number it returns was produced by reference code. Used for
//   * pinning oracle/ggml_oracle.c against the reference (tests/test_oracle_vs_ref.py),
//   * generating the committed golden vectors (tests/golden/make_golden.py),
//   * timing the reference CPU path as bench.py's cpu_baseline / --impl reference arm.
// Nothing on the product path may link or load this file.
#include "ggml.h"
#include "ggml-cpu.h"
#include "ggml-backend.h"
#include "ggml-alloc.h"

#include <algorithm>
#include <chrono>
#include <vector>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

extern "C" {

int ref_type_block_size(int type) { return (int) ggml_blck_size((ggml_type) type); }
size_t ref_row_size(int type, int64_t k) { return ggml_row_size((ggml_type) type, k); }
const char * ref_type_name(int type) { return ggml_type_name((ggml_type) type); }
int ref_vec_dot_type(int type) { return (int) ggml_get_type_traits_cpu((ggml_type) type)->vec_dot_type; }

// weight quantiser used by llama-quantize (ggml_quantize_chunk, ggml/include/ggml.h:2680)
size_t ref_quantize_weights(int type, const float * src, void * dst, int64_t nrows, int64_t k) {
    ggml_quantize_init((ggml_type) type);
    return ggml_quantize_chunk((ggml_type) type, src, dst, 0, nrows, k, nullptr);
}

// to_float of the base type traits (ggml-quants.c dequantize_row_*)
int ref_dequantize(int type, const void * src, float * dst, int64_t n) {
    auto fn = ggml_get_type_traits((ggml_type) type)->to_float;
    if (!fn) return 0;
    fn(src, dst, n);
    return 1;
}

// activation quantiser the CPU backend really runs (arch/x86/quants.c SIMD variant)
void ref_from_float_cpu(int type, const float * x, void * y, int64_t k) {
    ggml_cpu_init();
    ggml_get_type_traits_cpu((ggml_type) type)->from_float(x, y, k);
}

// scalar *_ref activation quantiser (ggml-quants.c)
// returns 0 when the base traits have no from_float_ref (e.g. Q8_K: only the CPU backend quantises to it)
int ref_from_float_ref(int type, const float * x, void * y, int64_t k) {
    auto fn = ggml_get_type_traits((ggml_type) type)->from_float_ref;
    if (!fn) return 0;
    fn(x, y, k);
    return 1;
}

// the CPU backend's vec_dot for weight type `type` against its vec_dot_type activations
void ref_vec_dot(int type, int n, float * s, const void * w, const void * xq) {
    ggml_cpu_init();
    ggml_get_type_traits_cpu((ggml_type) type)->vec_dot(n, s, 0, w, 0, xq, 0, 1);
}

static ggml_backend_buffer_type_t find_cpu_extra_buft(const char * name_substr) {
    ggml_backend_dev_t dev = ggml_backend_dev_by_type(GGML_BACKEND_DEVICE_TYPE_CPU);
    if (!dev) return nullptr;
    auto fn = (ggml_backend_dev_get_extra_bufts_t)
        ggml_backend_reg_get_proc_address(ggml_backend_dev_backend_reg(dev), "ggml_backend_dev_get_extra_bufts");
    if (!fn) return nullptr;
    ggml_backend_buffer_type_t * bufts = fn(dev);
    for (; bufts && *bufts; ++bufts) {
        if (strstr(ggml_backend_buft_name(*bufts), name_substr)) return *bufts;
    }
    return nullptr;
}

// Y[n][m] = W[m][k] . X[n][k] through ggml_mul_mat on the reference CPU backend.
//   w      : quantised weights, canonical ggml bytes, m rows
//   repack : 0 = plain CPU buffer (what test-backend-ops compares against, generic vec_dot path)
//            1 = weights placed in the CPU_REPACK extra buffer type (the fork's ZYK / 8x8 kernels);
//                returns -2 if that buffer type refuses the tensor
//   iters  : number of timed graph computes after one warm-up; *secs = mean seconds per compute
int ref_mul_mat(int type, int64_t m, int64_t n, int64_t k, const void * w, const float * x, float * y,
                int n_threads, int repack, int iters, double * secs) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);

    ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), nullptr, true };
    ggml_context * ctx_w = ggml_init(ip);
    ggml_context * ctx   = ggml_init(ip);
    ggml_tensor * W = ggml_new_tensor_2d(ctx_w, (ggml_type) type, k, m);
    ggml_tensor * X = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, k, n);
    ggml_tensor * Y = ggml_mul_mat(ctx, W, X);

    int rc = 0;
    ggml_backend_buffer_t buf_w = nullptr;
    if (repack) {
        ggml_backend_buffer_type_t buft = find_cpu_extra_buft("REPACK");
        if (buft) buf_w = ggml_backend_alloc_ctx_tensors_from_buft(ctx_w, buft);
        if (!buf_w) rc = -2;
        // init_tensor attaches the repack traits only if a repack layout exists for this type/shape
        if (buf_w && !W->extra) rc = -2;
    } else {
        buf_w = ggml_backend_alloc_ctx_tensors(ctx_w, cpu);
    }
    ggml_backend_buffer_t buf = nullptr;
    if (rc == 0) {
        if (repack) ggml_backend_buffer_set_usage(buf_w, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        buf = ggml_backend_alloc_ctx_tensors(ctx, cpu);
        ggml_backend_tensor_set(W, w, 0, ggml_nbytes(W));
        ggml_backend_tensor_set(X, x, 0, ggml_nbytes(X));
        ggml_cgraph * gf = ggml_new_graph(ctx);
        ggml_build_forward_expand(gf, Y);
        if (ggml_backend_graph_compute(cpu, gf) != GGML_STATUS_SUCCESS) rc = -3;
        if (rc == 0 && iters > 0) {
            // median of the per-iteration times: one descheduled worker thread stalls ggml's spin barrier for tens of
            // milliseconds on a shared host, which a mean would charge to the reference
            std::vector<double> dt((size_t) iters);
            for (int i = 0; i < iters; ++i) {
                auto t0 = std::chrono::steady_clock::now();
                ggml_backend_graph_compute(cpu, gf);
                dt[(size_t) i] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            }
            std::sort(dt.begin(), dt.end());
            if (secs) *secs = dt[dt.size() / 2];
        }
        if (rc == 0) ggml_backend_tensor_get(Y, y, 0, ggml_nbytes(Y));
    }
    if (buf) ggml_backend_buffer_free(buf);
    if (buf_w) ggml_backend_buffer_free(buf_w);
    ggml_free(ctx); ggml_free(ctx_w);
    ggml_backend_free(cpu);
    return rc;
}

// MUL_MAT_ID: as[k,m,n_as], b[k, nb1 (n_used or 1), n_tokens], ids i32[n_used, n_tokens] -> y[m, n_used, n_tokens]
int ref_mul_mat_id(int type, int64_t m, int64_t k, int64_t n_as, int64_t n_used, int64_t nb1, int64_t n_tokens,
                   const void * as, const float * b, const int32_t * ids, float * y, int n_threads) {
    ggml_backend_t cpu = ggml_backend_cpu_init();
    if (!cpu) return -1;
    ggml_backend_cpu_set_n_threads(cpu, n_threads);
    ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_tensor * AS  = ggml_new_tensor_3d(ctx, (ggml_type) type, k, m, n_as);
    ggml_tensor * B   = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, k, nb1, n_tokens);
    ggml_tensor * IDS = ggml_new_tensor_2d(ctx, GGML_TYPE_I32, n_used, n_tokens);
    ggml_tensor * Y   = ggml_mul_mat_id(ctx, AS, B, IDS);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, cpu);
    ggml_backend_tensor_set(AS, as, 0, ggml_nbytes(AS));
    ggml_backend_tensor_set(B, b, 0, ggml_nbytes(B));
    ggml_backend_tensor_set(IDS, ids, 0, ggml_nbytes(IDS));
    ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, Y);
    int rc = ggml_backend_graph_compute(cpu, gf) == GGML_STATUS_SUCCESS ? 0 : -3;
    if (rc == 0) ggml_backend_tensor_get(Y, y, 0, ggml_nbytes(Y));
    ggml_backend_buffer_free(buf);
    ggml_free(ctx);
    ggml_backend_free(cpu);
    return rc;
}

// Bytes of a weight matrix after the reference's load-time repack (CPU_REPACK buffer set_tensor,
// ggml/src/ggml-cpu/repack.cpp:3141-3151). The buffer is host memory, so tensor->data is readable.
// Returns number of bytes written to `out` (== ggml_nbytes) or <0 when the type has no repack here.
int64_t ref_repack_dump(int type, int64_t m, int64_t k, const void * w, void * out) {
    ggml_backend_buffer_type_t buft = find_cpu_extra_buft("REPACK");
    if (!buft) return -1;
    ggml_init_params ip = { ggml_tensor_overhead() * 4, nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_tensor * W = ggml_new_tensor_2d(ctx, (ggml_type) type, k, m);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, buft);
    int64_t rc = -2;
    if (buf) {
        ggml_backend_buffer_set_usage(buf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        if (W->extra) { // init_tensor attached repack traits => this type/shape is repacked
            ggml_backend_tensor_set(W, w, 0, ggml_nbytes(W));
            memcpy(out, W->data, ggml_nbytes(W));
            rc = (int64_t) ggml_nbytes(W);
        }
        ggml_backend_buffer_free(buf);
    }
    ggml_free(ctx);
    return rc;
}

} // extern "C"

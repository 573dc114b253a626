// glue_ops.cu -- the non-matmul ops of a llama-arch graph, so that llama.cpp / llama-bench keep a whole
// layer on the device between the quantised MUL_MATs instead of bouncing to the CPU backend after every one of them:
// GET_ROWS, RMS_NORM, ADD / MUL (broadcast), ROPE (normal / neox), SET_ROWS (KV cache writes), CPY / CONT, SOFT_MAX (scale + mask),
// GLU (swiglu / geglu / reglu) and the f16 x f32 MUL_MAT of attention over the KV cache.
//
// Semantics follow the reference CPU backend, op by op (ggml/src/ggml-cpu/ops.cpp; cited at each entry point); the plugin claims
// exactly the variants implemented here (supports_op) and the reference's test-backend-ops compares them with its CPU backend.
// All kernels are HBM / L2 bound elementwise or row kernels: one thread per element or one CTA per row, no tensor cores (the
// attention matmuls of a 512-token ubatch are the exception that would merit them).
#include <stdlib.h>
#include "common.cuh"

#include <math.h>

namespace {

struct T4 { const char * data; long long ne[4], nb[4]; int type; };
struct T4w { char * data; long long ne[4], nb[4]; int type; };

T4 t4(const b200_tensor * t) { T4 r; r.data = (const char *) t->data; r.type = t->type; for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; } return r; }
T4w t4w(const b200_tensor * t) { T4w r; r.data = (char *) t->data; r.type = t->type; for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; } return r; }
long long nelem(const b200_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
bool contiguous_f(const b200_tensor * t, long long esz) {
    return t->nb[0] == esz && t->nb[1] == t->nb[0] * t->ne[0] && t->nb[2] == t->nb[1] * t->ne[1] && t->nb[3] == t->nb[2] * t->ne[2];
}
// B200_GLUE_PDL=0: launch the glue kernels without the programmatic-dependent-launch attribute
bool glue_pdl() { static const bool on = [] { const char * e = getenv("B200_GLUE_PDL"); return !(e && e[0] == '0'); }(); return on; }
unsigned blocks_for(long long n, int per) { const long long b = (n + per - 1) / per; return (unsigned) (b < 1 ? 1 : b > 0x7fffffffll ? 0x7fffffffll : b); }

// element index -> 4-D coordinates; I = unsigned (32-bit divisions, ~5x cheaper) whenever the tensor has fewer than 2^31 elements
template <typename I>
__device__ __forceinline__ void unravel(I i, const long long * ne, I & i0, I & i1, I & i2, I & i3) {
    const I n0 = (I) ne[0], n1 = (I) ne[1], n2 = (I) ne[2];
    i0 = i % n0; I r = i / n0; i1 = r % n1; r /= n1; i2 = r % n2; i3 = r / n2;
}
#define B200_LAUNCH_IDX(kernel, total, ...) \
    do { if ((total) < (1ll << 31)) B200_CHECK(b200_launch(kernel<unsigned>, dim3(blocks_for(total, 256)), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), __VA_ARGS__)); \
         else                       B200_CHECK(b200_launch(kernel<long long>, dim3(blocks_for(total, 256)), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), __VA_ARGS__)); } while (0)

// ---- dequantise ONE element of a canonical ggml row (dequantize_row_*, ggml-quants.c:307-325, 401-415, 1352-1375, 1762-1792, 2530-2551) ----
__device__ __constant__ signed char kvalues_iq4nl_d[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};

__device__ __forceinline__ float deq_one(int type, const uint8_t * row, long long e) {
    switch (type) {
        case B200_TYPE_F32: return reinterpret_cast<const float *>(row)[e];
        case B200_TYPE_F16: return __half2float(reinterpret_cast<const __half *>(row)[e]);
        case B200_TYPE_Q4_0: {
            const uint8_t * b = row + (e >> 5) * 18; const int j = (int) (e & 31);
            const float d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b));
            const int q = j < 16 ? (b[2 + j] & 0x0f) : (b[2 + j - 16] >> 4);
            return d * (float) (q - 8);
        }
        case B200_TYPE_Q8_0: {
            const uint8_t * b = row + (e >> 5) * 34;
            return half_bits_to_float(*reinterpret_cast<const uint16_t *>(b)) * (float) (int8_t) b[2 + (e & 31)];
        }
        case B200_TYPE_Q4_K: {
            const uint8_t * b = row + (e >> 8) * 144; const int r = (int) (e & 255);
            const int g = r >> 6, hi = (r >> 5) & 1, l = r & 31, is = 2 * g + hi;
            const uint8_t * sc = b + 4;
            int s, m;
            if (is < 4) { s = sc[is] & 63; m = sc[is + 4] & 63; }
            else { s = (sc[is + 4] & 0x0f) | ((sc[is - 4] >> 6) << 4); m = (sc[is + 4] >> 4) | ((sc[is] >> 6) << 4); }
            const float d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b)), dmin = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b + 2));
            const uint8_t qb = b[16 + 32 * g + l];
            return d * (float) s * (float) (hi ? (qb >> 4) : (qb & 0x0f)) - dmin * (float) m;
        }
        case B200_TYPE_Q5_K: {
            const uint8_t * b = row + (e >> 8) * 176; const int r = (int) (e & 255);
            const int g = r >> 6, hi = (r >> 5) & 1, l = r & 31, is = 2 * g + hi;
            const uint8_t * sc = b + 4;
            int s, m;
            if (is < 4) { s = sc[is] & 63; m = sc[is + 4] & 63; }
            else { s = (sc[is + 4] & 0x0f) | ((sc[is - 4] >> 6) << 4); m = (sc[is + 4] >> 4) | ((sc[is] >> 6) << 4); }
            const float d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b)), dmin = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b + 2));
            const uint8_t qb = b[48 + 32 * g + l];
            const int v = (hi ? (qb >> 4) : (qb & 0x0f)) | (((b[16 + l] >> is) & 1) << 4);
            return d * (float) s * (float) v - dmin * (float) m;
        }
        case B200_TYPE_Q6_K: {
            const uint8_t * b = row + (e >> 8) * 210; const int r = (int) (e & 255);
            const int n = r >> 7, rr = r & 127, quarter = rr >> 5, l = rr & 31;
            const uint8_t qlb = b[64 * n + l + ((quarter & 1) ? 32 : 0)];
            const int nib = quarter < 2 ? (qlb & 0x0f) : (qlb >> 4);
            const int h = (b[128 + 32 * n + l] >> (2 * quarter)) & 3;
            const int scv = (int) (int8_t) b[192 + 8 * n + (l >> 4) + 2 * quarter];
            const float d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b + 208));
            return d * (float) scv * (float) ((nib | (h << 4)) - 32);
        }
        case B200_TYPE_IQ4_XS: {
            const uint8_t * b = row + (e >> 8) * 136; const int r = (int) (e & 255);
            const int ib = r >> 5, j = r & 31;
            const float d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(b));
            const unsigned sh = *reinterpret_cast<const uint16_t *>(b + 2);
            const int ls = (int) (((b[4 + (ib >> 1)] >> (4 * (ib & 1))) & 0x0f) | (((sh >> (2 * ib)) & 3u) << 4));
            const uint8_t qb = b[8 + 16 * ib + (j & 15)];
            return d * (float) (ls - 32) * (float) kvalues_iq4nl_d[j < 16 ? (qb & 0x0f) : (qb >> 4)];
        }
    }
    return 0.0f;
}

// ---- GET_ROWS (ggml_compute_forward_get_rows, ggml-cpu/ops.cpp): dst[:, i10, i11, i12] = row src1[i10, i11, i12] of matrix (i11, i12) of src0 ----
template <typename I>
__global__ void __launch_bounds__(256) get_rows_kernel(const T4 a, const T4 ids, const T4w d) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const I total = (I) (d.ne[0] * d.ne[1] * d.ne[2] * d.ne[3]);
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I e, i10, i11, i12;
        unravel<I>(i, d.ne, e, i10, i11, i12);
        const int r = *reinterpret_cast<const int *>(ids.data + (long long) i10 * ids.nb[0] + (long long) i11 * ids.nb[1] + (long long) i12 * ids.nb[2]);
        const uint8_t * row = reinterpret_cast<const uint8_t *>(a.data + (long long) r * a.nb[1] + (long long) i11 * a.nb[2] + (long long) i12 * a.nb[3]);
        *reinterpret_cast<float *>(d.data + (long long) e * d.nb[0] + (long long) i10 * d.nb[1] + (long long) i11 * d.nb[2] + (long long) i12 * d.nb[3]) = deq_one(a.type, row, (long long) e);
    }
}

// ---- RMS_NORM (ggml_compute_forward_rms_norm_f32): y = x / sqrt(mean(x^2) + eps), one CTA per row ----
__global__ void __launch_bounds__(256) rms_norm_kernel(const T4 a, const T4w d, float eps) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const long long row = blockIdx.x;
    const long long i1 = row % a.ne[1], i2 = (row / a.ne[1]) % a.ne[2], i3 = row / (a.ne[1] * a.ne[2]);
    const float * x = reinterpret_cast<const float *>(a.data + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float * y = reinterpret_cast<float *>(d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    float s = 0.0f;
    for (long long i = threadIdx.x; i < a.ne[0]; i += 256) { const float v = x[i]; s += v * v; }
    __shared__ float red[8];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i];
    const float scale = 1.0f / sqrtf(s / (float) a.ne[0] + eps);
    for (long long i = threadIdx.x; i < a.ne[0]; i += 256) y[i] = x[i] * scale;
}

// ---- fused [ADD ->] RMS_NORM -> MUL(weight): sum = a + b (stored when `sum` is given), y = sum / sqrt(mean(sum^2) + eps) * w ----
// (what the llama graph does at every residual: ffn_inp = attn_out + inp; norm = rms_norm(ffn_inp); ffn_norm = norm * ffn_norm.weight;
//  the reference's CUDA backend fuses the same pattern, ggml-cuda.cu ggml_cuda_can_fuse / rms_norm_mul_f32)
// one CTA per row, 1024 threads for a few rows (decode) and 256 for many (a prefill ubatch: more rows in flight per SM, eight loads per thread): the row is read once with 128-bit loads (two per thread: rows of up to 8192 elements stay in registers), longer or
// unaligned rows take the scalar loop. w has ne[0] == row length (checked by the caller), so it is indexed without a modulo.
template <int NT>
__global__ void __launch_bounds__(NT) add_rms_norm_mul_kernel(const T4 a, const T4 b, int has_b, const T4w sum, int has_sum, const T4 w, const T4w d, float eps) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const unsigned row = blockIdx.x;
    const unsigned n1 = (unsigned) a.ne[1], n2 = (unsigned) a.ne[2];
    const unsigned i1 = row % n1, i2 = (row / n1) % n2, i3 = row / (n1 * n2);
    const float * x = reinterpret_cast<const float *>(a.data + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
    const float * xb = has_b ? reinterpret_cast<const float *>(b.data + (long long) (i1 % (unsigned) b.ne[1]) * b.nb[1] + (long long) (i2 % (unsigned) b.ne[2]) * b.nb[2] + (long long) (i3 % (unsigned) b.ne[3]) * b.nb[3]) : nullptr;
    float * so = has_sum ? reinterpret_cast<float *>(sum.data + (long long) i1 * sum.nb[1] + (long long) i2 * sum.nb[2] + (long long) i3 * sum.nb[3]) : nullptr;
    const float * wr = reinterpret_cast<const float *>(w.data + (long long) (i1 % (unsigned) w.ne[1]) * w.nb[1] + (long long) (i2 % (unsigned) w.ne[2]) * w.nb[2] + (long long) (i3 % (unsigned) w.ne[3]) * w.nb[3]);
    float * y = reinterpret_cast<float *>(d.data + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]);
    const int n = (int) a.ne[0];
    __shared__ float red[32];
    const bool vec = n <= 8192 && (n & 3) == 0 && ((((uintptr_t) x) | ((uintptr_t) y) | ((uintptr_t) wr) | (xb ? (uintptr_t) xb : 0) | (so ? (uintptr_t) so : 0)) & 15) == 0;
    constexpr int NV = 8192 / (4 * NT);                   // 128-bit pieces per thread: 2 (1024 threads) or 8 (256 threads)
    float4 v[NV];
    float s = 0.0f;
    if (vec) {
#pragma unroll
        for (int t = 0; t < NV; ++t) {
            const int i4 = threadIdx.x + t * NT;
            float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i4 * 4 < n) {
                e = reinterpret_cast<const float4 *>(x)[i4];
                if (xb) { const float4 f = reinterpret_cast<const float4 *>(xb)[i4]; e.x += f.x; e.y += f.y; e.z += f.z; e.w += f.w; }
                if (so) reinterpret_cast<float4 *>(so)[i4] = e;
            }
            v[t] = e; s += e.x * e.x + e.y * e.y + e.z * e.z + e.w * e.w;
        }
    } else {
        for (int i = threadIdx.x; i < n; i += NT) { float e = x[i]; if (xb) e += xb[i]; if (so) so[i] = e; s += e * e; }
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    s = (threadIdx.x & 31) < NT / 32 ? red[threadIdx.x & 31] : 0.0f;
    s = warp_sum(s);
    const float scale = 1.0f / sqrtf(s / (float) n + eps);
    if (vec) {
#pragma unroll
        for (int t = 0; t < NV; ++t) {
            const int i4 = threadIdx.x + t * NT;
            if (i4 * 4 < n) {
                const float4 ww = reinterpret_cast<const float4 *>(wr)[i4];
                reinterpret_cast<float4 *>(y)[i4] = make_float4(v[t].x * scale * ww.x, v[t].y * scale * ww.y, v[t].z * scale * ww.z, v[t].w * scale * ww.w);
            }
        }
    } else {
        // (so holds the sums when it exists; otherwise recompute: y may alias x only element by element)
        for (int i = threadIdx.x; i < n; i += NT) { float e = so ? so[i] : (xb ? x[i] + xb[i] : x[i]); y[i] = e * scale * wr[i]; }
    }
}

// ---- fused decode attention: softmax(scale * K q + mask) V for a few query tokens, one CTA per (head, token) --------------------------------
// replaces MUL_MAT(K view f16, Q) -> SOFT_MAX(scale, mask) -> MUL_MAT(V^T view f16, P) -> PERMUTE -> CONT of the llama graph
// (src/llama-graph.cpp build_attn_mha, the path without flash attention). K: [head_dim, n_kv, n_head_kv] rows contiguous; V^T: [n_kv, head_dim,
// n_head_kv] rows contiguous; q: [head_dim, n_tokens, n_head]; mask [n_kv, n_tokens(+pad)] f32 / f16 or none; out: [head_dim, n_head, n_tokens] contiguous.
// Latency is what matters here (32 CTAs, a few KB each): every phase issues all its loads before it uses them.
//   scores: four lanes per KV row (head_dim / 4 elements each as 128-bit loads), 64 rows per CTA pass;
//   output: thread (d, part) walks its part of row d of V^T with 128-bit loads, the parts meet in shared memory.
__global__ void __launch_bounds__(256) attn_decode_kernel(const T4 K, const T4 q, const T4 m, int has_mask, const T4 V, float * __restrict__ out, float scale) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    extern __shared__ float sc[];                 // n_kv scores, then head_dim q values, then 256 partial outputs
    __shared__ float red[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, t = blockIdx.y;
    const int n_kv = (int) K.ne[1], hd = (int) K.ne[0], n_head = (int) q.ne[2];
    const int hk = h / (n_head / (int) K.ne[2]);
    float * qs = sc + n_kv, * part = qs + hd;
    const float * qv = reinterpret_cast<const float *>(q.data + (long long) t * q.nb[1] + (long long) h * q.nb[2]);
    for (int e = threadIdx.x; e < hd; e += 256) qs[e] = qv[e];
    __syncthreads();
    const char * mp = has_mask ? m.data + (long long) t * m.nb[1] + (long long) (h % (int) m.ne[2]) * m.nb[2] : nullptr;
    const char * kb = K.data + (long long) hk * K.nb[2];
    // phase 1: scores. lane quad g = threadIdx.x / 4 owns rows g, g + 64, ...; sub = lane & 3 owns elements [sub * hd / 4, (sub + 1) * hd / 4)
    const int g = threadIdx.x >> 2, sub = threadIdx.x & 3, per = hd >> 2;      // per is a multiple of 8 (checked by the caller)
    float mx = -INFINITY;
    for (int j0 = 0; j0 < n_kv; j0 += 64) {
        const int j = j0 + g;
        float a = 0.0f;
        if (j < n_kv) {
            const uint4 * kr = reinterpret_cast<const uint4 *>(kb + (long long) j * K.nb[1]) + sub * (per >> 3);
            for (int c = 0; c < (per >> 3); ++c) {
                const uint4 kk = kr[c];
                const float * qq = qs + sub * per + c * 8;
                const float2 k0 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.x)), k1 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.y));
                const float2 k2 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.z)), k3 = __half22float2(*reinterpret_cast<const __half2 *>(&kk.w));
                a += k0.x * qq[0] + k0.y * qq[1] + k1.x * qq[2] + k1.y * qq[3] + k2.x * qq[4] + k2.y * qq[5] + k3.x * qq[6] + k3.y * qq[7];
            }
        }
        a += __shfl_xor_sync(0xffffffffu, a, 1);
        a += __shfl_xor_sync(0xffffffffu, a, 2);
        if (j < n_kv) {
            a *= scale;
            if (mp) a += m.type == B200_TYPE_F16 ? __half2float(reinterpret_cast<const __half *>(mp)[j]) : reinterpret_cast<const float *>(mp)[j];
            if (sub == 0) sc[j] = a;
            mx = fmaxf(mx, a);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
    __syncthreads();
    // phase 2: exponentials and their sum
    float s = 0.0f;
    for (int j = threadIdx.x; j < n_kv; j += 256) { const float e = expf(sc[j] - mx); sc[j] = e; s += e; }
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i];
    const float inv = 1.0f / s;
    // phase 3: out[d] = sum_j p[j] V^T[d][j]. Aligned rows of a multiple of 8 entries: a warp takes four rows of V^T at a time and its lanes walk
    // them in consecutive 16-byte pieces (coalesced: one 512-byte run per row and step), four independent loads in flight per lane
    const char * vbase = V.data + (long long) hk * V.nb[2];
    if ((n_kv & 7) == 0 && (hd & 3) == 0 && (((uintptr_t) vbase | (uintptr_t) V.nb[1]) & 15) == 0) {
        __syncthreads();                                    // every sc[j] holds its exponential
        for (int d0 = warp * 4; d0 < hd; d0 += 32) {
            float acc4[4] = {0.f, 0.f, 0.f, 0.f};
            for (int j = lane * 8; j < n_kv; j += 256) {
                const float4 p0 = *reinterpret_cast<const float4 *>(sc + j), p1 = *reinterpret_cast<const float4 *>(sc + j + 4);
                uint4 vv[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) vv[r] = *reinterpret_cast<const uint4 *>(vbase + (long long) (d0 + r) * V.nb[1] + j * 2);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float2 v0 = __half22float2(*reinterpret_cast<const __half2 *>(&vv[r].x)), v1 = __half22float2(*reinterpret_cast<const __half2 *>(&vv[r].y));
                    const float2 v2 = __half22float2(*reinterpret_cast<const __half2 *>(&vv[r].z)), v3 = __half22float2(*reinterpret_cast<const __half2 *>(&vv[r].w));
                    acc4[r] += v0.x * p0.x + v0.y * p0.y + v1.x * p0.z + v1.y * p0.w + v2.x * p1.x + v2.y * p1.y + v3.x * p1.z + v3.y * p1.w;
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const float v = warp_sum(acc4[r]);
                if (lane == 0) out[((long long) t * n_head + h) * hd + d0 + r] = v * inv;
            }
        }
        return;
    }
    // any other shape: `parts` threads share a row of V^T
    const int parts = hd <= 256 ? 256 / hd : 1;
    const int d = threadIdx.x % hd, pt = threadIdx.x / hd;
    float acc = 0.0f;
    if (threadIdx.x < hd * parts) {
        const int chunk = (((n_kv + parts - 1) / parts) + 7) & ~7;
        const int j_lo = pt * chunk, j_hi = min(n_kv, j_lo + chunk);
        const __half * vr = reinterpret_cast<const __half *>(V.data + (long long) d * V.nb[1] + (long long) hk * V.nb[2]);
        int j = j_lo;
        if ((((uintptr_t) vr) & 15) == 0) {
            for (; j + 8 <= j_hi; j += 8) {
                const uint4 vv = *reinterpret_cast<const uint4 *>(vr + j);
                const float2 v0 = __half22float2(*reinterpret_cast<const __half2 *>(&vv.x)), v1 = __half22float2(*reinterpret_cast<const __half2 *>(&vv.y));
                const float2 v2 = __half22float2(*reinterpret_cast<const __half2 *>(&vv.z)), v3 = __half22float2(*reinterpret_cast<const __half2 *>(&vv.w));
                acc += v0.x * sc[j] + v0.y * sc[j + 1] + v1.x * sc[j + 2] + v1.y * sc[j + 3] + v2.x * sc[j + 4] + v2.y * sc[j + 5] + v3.x * sc[j + 6] + v3.y * sc[j + 7];
            }
        }
        for (; j < j_hi; ++j) acc += __half2float(vr[j]) * sc[j];
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x < hd) {
        float r = 0.0f;
        for (int p = 0; p < parts; ++p) r += part[p * hd + threadIdx.x];
        out[((long long) t * n_head + h) * hd + threadIdx.x] = r * inv;
    }
    // head_dim > 256: the rows past 256 are handled by the same threads in further rounds
    for (int d2 = threadIdx.x + 256; d2 < hd && parts == 1; d2 += 256) {
        const __half * vr = reinterpret_cast<const __half *>(V.data + (long long) d2 * V.nb[1] + (long long) hk * V.nb[2]);
        float r = 0.0f;
        for (int j = 0; j < n_kv; ++j) r += __half2float(vr[j]) * sc[j];
        out[((long long) t * n_head + h) * hd + d2] = r * inv;
    }
}

// ---- ADD / MUL / DIV with ggml broadcasting of src1 (ggml_compute_forward_add / _mul / _div: src1 dims divide src0 dims) ----
template <int OP, typename I>
__global__ void __launch_bounds__(256) binary_kernel(const T4 a, const T4 b, const T4w d) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const I total = (I) (d.ne[0] * d.ne[1] * d.ne[2] * d.ne[3]);
    const I b0 = (I) b.ne[0], b1 = (I) b.ne[1], b2 = (I) b.ne[2], b3 = (I) b.ne[3];
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I i0, i1, i2, i3;
        unravel<I>(i, d.ne, i0, i1, i2, i3);
        const float x = *reinterpret_cast<const float *>(a.data + (long long) i0 * a.nb[0] + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
        const I j0 = b0 == (I) d.ne[0] ? i0 : i0 % b0, j1 = b1 == 1 ? 0 : i1 % b1, j2 = b2 == 1 ? 0 : i2 % b2, j3 = b3 == 1 ? 0 : i3 % b3;
        const float y = *reinterpret_cast<const float *>(b.data + (long long) j0 * b.nb[0] + (long long) j1 * b.nb[1] + (long long) j2 * b.nb[2] + (long long) j3 * b.nb[3]);
        *reinterpret_cast<float *>(d.data + (long long) i0 * d.nb[0] + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]) = OP == 0 ? x + y : OP == 1 ? x * y : x / y;
    }
}
// ---- the router of a mixture-of-experts layer (build_moe_ffn, src/llama-graph.cpp): SCALE, CLAMP, SUM_ROWS, ARGSORT --------------------------
// x * scale + bias (ggml_compute_forward_scale) and min(max(x, lo), hi) (ggml_compute_forward_clamp), f32, any strides
template <int OP, typename I>
__global__ void __launch_bounds__(256) unary_kernel(const T4 a, const T4w d, float p0, float p1) {
    pdl_launch_dependents(); pdl_wait();
    const I total = (I) (d.ne[0] * d.ne[1] * d.ne[2] * d.ne[3]);
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I i0, i1, i2, i3;
        unravel<I>(i, d.ne, i0, i1, i2, i3);
        const float x = *reinterpret_cast<const float *>(a.data + (long long) i0 * a.nb[0] + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
        const float y = OP == 0 ? x * p0 + p1 : fminf(fmaxf(x, p0), p1);
        *reinterpret_cast<float *>(d.data + (long long) i0 * d.nb[0] + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]) = y;
    }
}
// ggml_compute_forward_sum_rows: dst[0, i1, i2, i3] = sum over i0; one warp per row
__global__ void __launch_bounds__(128) sum_rows_kernel(const T4 a, const T4w d, long long rows) {
    pdl_launch_dependents(); pdl_wait();
    const int lane = threadIdx.x & 31;
    const long long row = (long long) blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const long long i1 = row % a.ne[1], i2 = (row / a.ne[1]) % a.ne[2], i3 = row / (a.ne[1] * a.ne[2]);
    const char * x = a.data + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3];
    float s = 0.0f;
    for (long long i = lane; i < a.ne[0]; i += 32) s += *reinterpret_cast<const float *>(x + i * a.nb[0]);
    s = warp_sum(s);
    if (lane == 0) *reinterpret_cast<float *>(d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = s;
}
// ggml_compute_forward_argsort_f32: dst row = the indices that sort the src row (order 0 ascending, 1 descending). One CTA per row, bitonic
// network over the next power of two in shared memory (rows of up to 1024 elements: expert scores); padding sorts last in either order.
__global__ void __launch_bounds__(256) argsort_kernel(const T4 a, const T4w d, int order, int npad) {
    pdl_launch_dependents(); pdl_wait();
    __shared__ float key[1024];
    __shared__ int idx[1024];
    const long long row = blockIdx.x;
    const long long i1 = row % a.ne[1], i2 = (row / a.ne[1]) % a.ne[2], i3 = row / (a.ne[1] * a.ne[2]);
    const char * x = a.data + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3];
    const int n = (int) a.ne[0];
    for (int i = threadIdx.x; i < npad; i += 256) { idx[i] = i; key[i] = i < n ? *reinterpret_cast<const float *>(x + (long long) i * a.nb[0]) : 0.0f; }
    __syncthreads();
    // "p before q" in the requested order; a padded slot (index >= n) goes after every real one
    auto before = [&](int p, int q) {
        const int ip = idx[p], iq = idx[q];
        if (ip >= n) return false;
        if (iq >= n) return true;
        return order == 0 ? key[p] < key[q] : key[p] > key[q];
    };
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < npad; i += 256) {
                const int l = i ^ j;
                if (l > i) {
                    const bool up = (i & k) == 0;
                    const bool swap = up ? before(l, i) : before(i, l);
                    if (swap) { const float tk = key[i]; key[i] = key[l]; key[l] = tk; const int ti = idx[i]; idx[i] = idx[l]; idx[l] = ti; }
                }
            }
            __syncthreads();
        }
    }
    int * out = reinterpret_cast<int *>(d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    for (int i = threadIdx.x; i < n; i += 256) out[i] = idx[i];
}

// ---- ROPE, modes normal (0) and neox (2), ext_factor == 0 (ggml_compute_forward_rope_flt + ggml_rope_cache_init + rope_yarn, ops.cpp:5524-5790) ----
struct RopeP { int n_dims, mode; float freq_base, freq_scale, attn_factor; };
// one thread = one rotation pair position (ip, token) for a chunk of `hc` heads: the angle (powf, sincosf) is computed once per chunk. hc = 8 for a
// prefill ubatch (the angle arithmetic dominated: 11 -> 5 us per 512 x 32 x 128 tensor), 1 for a few tokens (decode wants the threads, not the reuse)
inline int rope_hc(const b200_tensor * src) { return src->ne[2] * src->ne[3] >= 32 ? 8 : 1; }      // template instantiations: 1 and 8
__device__ __forceinline__ void rope_angle(const RopeP & p, float theta_scale, int pos, int ip, const float * ff, float & sn, float & cs) {
    const float theta = (float) pos * powf(theta_scale, (float) ip) / (ff ? ff[ip] : 1.0f) * p.freq_scale;
    sincosf(theta, &sn, &cs);
    sn *= p.attn_factor; cs *= p.attn_factor;
}
template <typename I, int ROPE_HC>
__global__ void __launch_bounds__(256) rope_kernel(const T4 a, const int * __restrict__ pos, const float * __restrict__ ff, const T4w d, const RopeP p) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    if constexpr (ROPE_HC == 1) {      // one thread per pair and head (decode: the kernel is all latency, keep it short)
        const long long hne[4] = {a.ne[0] / 2, a.ne[1], a.ne[2], a.ne[3]};
        const I total = (I) (hne[0] * hne[1] * hne[2] * hne[3]);
        const float theta_scale = powf(p.freq_base, -2.0f / (float) p.n_dims);
        for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
            I ip, i1, i2, i3;
            unravel<I>(i, hne, ip, i1, i2, i3);
            const char * src = a.data + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3];
            char * dst = d.data + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3];
            if (2 * (int) ip >= p.n_dims) {                // the dimensions past n_dims are copied
                const int e = 2 * (int) ip;
                reinterpret_cast<float *>(dst)[e] = reinterpret_cast<const float *>(src)[e];
                reinterpret_cast<float *>(dst)[e + 1] = reinterpret_cast<const float *>(src)[e + 1];
                continue;
            }
            const float theta = (float) pos[i2] * powf(theta_scale, (float) ip) / (ff ? ff[ip] : 1.0f) * p.freq_scale;
            float sn, cs;
            sincosf(theta, &sn, &cs);
            sn *= p.attn_factor; cs *= p.attn_factor;
            const int e0 = p.mode == 2 ? (int) ip : 2 * (int) ip, e1 = p.mode == 2 ? (int) ip + p.n_dims / 2 : 2 * (int) ip + 1;
            const float x0 = reinterpret_cast<const float *>(src)[e0], x1 = reinterpret_cast<const float *>(src)[e1];
            reinterpret_cast<float *>(dst)[e0] = x0 * cs - x1 * sn;
            reinterpret_cast<float *>(dst)[e1] = x0 * sn + x1 * cs;
        }
    } else {
        const long long hch = (a.ne[1] + ROPE_HC - 1) / ROPE_HC;
        const long long hne[4] = {a.ne[0] / 2, hch, a.ne[2], a.ne[3]};
        const I total = (I) (hne[0] * hne[1] * hne[2] * hne[3]);
        const float theta_scale = powf(p.freq_base, -2.0f / (float) p.n_dims);
        for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
            I ip, ic, i2, i3;
            unravel<I>(i, hne, ip, ic, i2, i3);
            const bool rot = 2 * (int) ip < p.n_dims;           // the dimensions past n_dims are copied
            float sn = 0.0f, cs = 1.0f;
            if (rot) rope_angle(p, theta_scale, pos[i2], (int) ip, ff, sn, cs);
            const int e0 = rot && p.mode == 2 ? (int) ip : 2 * (int) ip, e1 = rot && p.mode == 2 ? (int) ip + p.n_dims / 2 : 2 * (int) ip + 1;
            const long long h1 = (long long) ic * ROPE_HC + ROPE_HC < a.ne[1] ? (long long) ic * ROPE_HC + ROPE_HC : a.ne[1];
            for (long long i1 = (long long) ic * ROPE_HC; i1 < h1; ++i1) {
                const float * src = reinterpret_cast<const float *>(a.data + i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
                float * dst = reinterpret_cast<float *>(d.data + i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]);
                const float x0 = src[e0], x1 = src[e1];
                dst[e0] = rot ? x0 * cs - x1 * sn : x0;
                dst[e1] = rot ? x0 * sn + x1 * cs : x1;
            }
        }
    }
}

__device__ __forceinline__ void store_as(char * p, int type, float v) {
    if (type == B200_TYPE_F16) *reinterpret_cast<__half *>(p) = __float2half_rn(v); else *reinterpret_cast<float *>(p) = v;
}
__device__ __forceinline__ float load_as(const char * p, int type) {
    return type == B200_TYPE_F16 ? __half2float(*reinterpret_cast<const __half *>(p)) : *reinterpret_cast<const float *>(p);
}

// ---- SET_ROWS (ggml_compute_forward_set_rows_f32): dst row src1[i01, i02 % ne11, i03 % ne12] of matrix (i02, i03) = src0 row (i01, i02, i03) ----
template <typename I>
__global__ void __launch_bounds__(256) set_rows_kernel(const T4 a, const T4 idx, const T4w d) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const I total = (I) (a.ne[0] * a.ne[1] * a.ne[2] * a.ne[3]);
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I i0, i1, i2, i3;
        unravel<I>(i, a.ne, i0, i1, i2, i3);
        const char * ip = idx.data + (long long) i1 * idx.nb[0] + (long long) (i2 % (I) idx.ne[1]) * idx.nb[1] + (long long) (i3 % (I) idx.ne[2]) * idx.nb[2];
        const long long r = idx.type == B200_TYPE_I32 ? (long long) *reinterpret_cast<const int *>(ip) : *reinterpret_cast<const long long *>(ip);
        const float v = *reinterpret_cast<const float *>(a.data + (long long) i0 * a.nb[0] + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
        store_as(d.data + (long long) i0 * d.nb[0] + r * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3], d.type, v);
    }
}

// ---- decode / prefill KV-cache prologue in ONE launch: ROPE(k) whose result goes to its own tensor AND, converted, into the K cache rows
// (the SET_ROWS that follows it in build_attn, src/llama-graph.cpp), plus optionally the independent SET_ROWS of V in the tail blocks.
// `sa` = shape of the SET_ROWS source (a reshaped, contiguous view of the ROPE result: element L of one is element L of the other).
template <typename I, int ROPE_HC>
__global__ void __launch_bounds__(256) rope_set_rows_kernel(const T4 a, const int * __restrict__ pos, const float * __restrict__ ff, const T4w d, const RopeP p,
                                                           const T4 sa, const T4 idx, const T4w cache, unsigned kblocks,
                                                           const T4 va, const T4 vidx, const T4w vcache) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    if constexpr (ROPE_HC == 1) {      // one thread per pair and head (decode: the kernel is all latency, keep it short)
        if (blockIdx.x >= kblocks) {                              // SET_ROWS of V: the body of set_rows_kernel on the tail blocks
            const I total = (I) (va.ne[0] * va.ne[1] * va.ne[2] * va.ne[3]);
            for (I i = (I) (blockIdx.x - kblocks) * 256 + threadIdx.x; i < total; i += (I) (gridDim.x - kblocks) * 256) {
                I i0, i1, i2, i3;
                unravel<I>(i, va.ne, i0, i1, i2, i3);
                const char * ip = vidx.data + (long long) i1 * vidx.nb[0] + (long long) (i2 % (I) vidx.ne[1]) * vidx.nb[1] + (long long) (i3 % (I) vidx.ne[2]) * vidx.nb[2];
                const long long r = vidx.type == B200_TYPE_I32 ? (long long) *reinterpret_cast<const int *>(ip) : *reinterpret_cast<const long long *>(ip);
                const float v = *reinterpret_cast<const float *>(va.data + (long long) i0 * va.nb[0] + (long long) i1 * va.nb[1] + (long long) i2 * va.nb[2] + (long long) i3 * va.nb[3]);
                store_as(vcache.data + (long long) i0 * vcache.nb[0] + r * vcache.nb[1] + (long long) i2 * vcache.nb[2] + (long long) i3 * vcache.nb[3], vcache.type, v);
            }
            return;
        }
        const long long hne[4] = {a.ne[0] / 2, a.ne[1], a.ne[2], a.ne[3]};
        const I total = (I) (hne[0] * hne[1] * hne[2] * hne[3]);
        const float theta_scale = powf(p.freq_base, -2.0f / (float) p.n_dims);
        for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) kblocks * 256) {
            I ip, i1, i2, i3;
            unravel<I>(i, hne, ip, i1, i2, i3);
            const char * src = a.data + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3];
            int e0, e1; float y0, y1;
            if (2 * (int) ip >= p.n_dims) {
                e0 = 2 * (int) ip; e1 = e0 + 1;
                y0 = reinterpret_cast<const float *>(src)[e0]; y1 = reinterpret_cast<const float *>(src)[e1];
            } else {
                const float theta = (float) pos[i2] * powf(theta_scale, (float) ip) / (ff ? ff[ip] : 1.0f) * p.freq_scale;
                float sn, cs;
                sincosf(theta, &sn, &cs);
                sn *= p.attn_factor; cs *= p.attn_factor;
                e0 = p.mode == 2 ? (int) ip : 2 * (int) ip; e1 = p.mode == 2 ? (int) ip + p.n_dims / 2 : 2 * (int) ip + 1;
                const float x0 = reinterpret_cast<const float *>(src)[e0], x1 = reinterpret_cast<const float *>(src)[e1];
                y0 = x0 * cs - x1 * sn; y1 = x0 * sn + x1 * cs;
            }
            // d is contiguous (host check): linear element index of (e, i1, i2, i3)
            const I base = (I) (((long long) i3 * d.ne[2] + i2) * d.ne[1] + i1) * (I) d.ne[0];
            float * dp = reinterpret_cast<float *>(d.data) + base;
            dp[e0] = y0; dp[e1] = y1;
    #pragma unroll
            for (int h = 0; h < 2; ++h) {
                const I L = base + (I) (h ? e1 : e0);
                I j0, j1, j2, j3;
                unravel<I>(L, sa.ne, j0, j1, j2, j3);
                const char * xp = idx.data + (long long) j1 * idx.nb[0] + (long long) (j2 % (I) idx.ne[1]) * idx.nb[1] + (long long) (j3 % (I) idx.ne[2]) * idx.nb[2];
                const long long r = idx.type == B200_TYPE_I32 ? (long long) *reinterpret_cast<const int *>(xp) : *reinterpret_cast<const long long *>(xp);
                store_as(cache.data + (long long) j0 * cache.nb[0] + r * cache.nb[1] + (long long) j2 * cache.nb[2] + (long long) j3 * cache.nb[3], cache.type, h ? y1 : y0);
            }
        }
    } else {
        if (blockIdx.x >= kblocks) {                              // SET_ROWS of V: the body of set_rows_kernel on the tail blocks
            const I total = (I) (va.ne[0] * va.ne[1] * va.ne[2] * va.ne[3]);
            for (I i = (I) (blockIdx.x - kblocks) * 256 + threadIdx.x; i < total; i += (I) (gridDim.x - kblocks) * 256) {
                I i0, i1, i2, i3;
                unravel<I>(i, va.ne, i0, i1, i2, i3);
                const char * ip = vidx.data + (long long) i1 * vidx.nb[0] + (long long) (i2 % (I) vidx.ne[1]) * vidx.nb[1] + (long long) (i3 % (I) vidx.ne[2]) * vidx.nb[2];
                const long long r = vidx.type == B200_TYPE_I32 ? (long long) *reinterpret_cast<const int *>(ip) : *reinterpret_cast<const long long *>(ip);
                const float v = *reinterpret_cast<const float *>(va.data + (long long) i0 * va.nb[0] + (long long) i1 * va.nb[1] + (long long) i2 * va.nb[2] + (long long) i3 * va.nb[3]);
                store_as(vcache.data + (long long) i0 * vcache.nb[0] + r * vcache.nb[1] + (long long) i2 * vcache.nb[2] + (long long) i3 * vcache.nb[3], vcache.type, v);
            }
            return;
        }
        const long long hch = (a.ne[1] + ROPE_HC - 1) / ROPE_HC;
        const long long hne[4] = {a.ne[0] / 2, hch, a.ne[2], a.ne[3]};
        const I total = (I) (hne[0] * hne[1] * hne[2] * hne[3]);
        const float theta_scale = powf(p.freq_base, -2.0f / (float) p.n_dims);
        for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) kblocks * 256) {
            I ip, ic, i2, i3;
            unravel<I>(i, hne, ip, ic, i2, i3);
            const bool rot = 2 * (int) ip < p.n_dims;
            float sn = 0.0f, cs = 1.0f;
            if (rot) rope_angle(p, theta_scale, pos[i2], (int) ip, ff, sn, cs);
            const int e0 = rot && p.mode == 2 ? (int) ip : 2 * (int) ip, e1 = rot && p.mode == 2 ? (int) ip + p.n_dims / 2 : 2 * (int) ip + 1;
            const long long h1 = (long long) ic * ROPE_HC + ROPE_HC < a.ne[1] ? (long long) ic * ROPE_HC + ROPE_HC : a.ne[1];
            for (long long i1 = (long long) ic * ROPE_HC; i1 < h1; ++i1) {
                const float * src = reinterpret_cast<const float *>(a.data + i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
                const float x0 = src[e0], x1 = src[e1];
                const float y0 = rot ? x0 * cs - x1 * sn : x0, y1 = rot ? x0 * sn + x1 * cs : x1;
                // d is contiguous (host check): linear element index of (e, i1, i2, i3)
                const I base = (I) ((((long long) i3 * d.ne[2] + i2) * d.ne[1] + i1) * d.ne[0]);
                float * dp = reinterpret_cast<float *>(d.data) + base;
                dp[e0] = y0; dp[e1] = y1;
    #pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const I L = base + (I) (h ? e1 : e0);
                    I j0, j1, j2, j3;
                    unravel<I>(L, sa.ne, j0, j1, j2, j3);
                    const char * xp = idx.data + (long long) j1 * idx.nb[0] + (long long) (j2 % (I) idx.ne[1]) * idx.nb[1] + (long long) (j3 % (I) idx.ne[2]) * idx.nb[2];
                    const long long r = idx.type == B200_TYPE_I32 ? (long long) *reinterpret_cast<const int *>(xp) : *reinterpret_cast<const long long *>(xp);
                    store_as(cache.data + (long long) j0 * cache.nb[0] + r * cache.nb[1] + (long long) j2 * cache.nb[2] + (long long) j3 * cache.nb[3], cache.type, h ? y1 : y0);
                }
            }
        }
    }
}

// ---- CPY / CONT / DUP between f32 and f16 (ggml_compute_forward_dup): element i of src (row-major over its shape) -> element i of dst ----
template <typename I>
__global__ void __launch_bounds__(256) cpy_kernel(const T4 a, const T4w d) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const I total = (I) (d.ne[0] * d.ne[1] * d.ne[2] * d.ne[3]);
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I s0, s1, s2, s3, d0, d1, d2, d3;
        unravel<I>(i, a.ne, s0, s1, s2, s3);
        unravel<I>(i, d.ne, d0, d1, d2, d3);
        store_as(d.data + (long long) d0 * d.nb[0] + (long long) d1 * d.nb[1] + (long long) d2 * d.nb[2] + (long long) d3 * d.nb[3], d.type,
                 load_as(a.data + (long long) s0 * a.nb[0] + (long long) s1 * a.nb[1] + (long long) s2 * a.nb[2] + (long long) s3 * a.nb[3], a.type));
    }
}

// ---- SOFT_MAX with scale and an optional f32 / f16 mask, max_bias == 0, no sinks (ggml_compute_forward_soft_max_f32, ops.cpp:5160-5280) ----
__global__ void __launch_bounds__(256) soft_max_kernel(const T4 a, const T4 m, int has_mask, const T4w d, float scale) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const long long row = blockIdx.x;
    const long long i1 = row % a.ne[1], i2 = (row / a.ne[1]) % a.ne[2], i3 = row / (a.ne[1] * a.ne[2]);
    const float * x = reinterpret_cast<const float *>(a.data + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float * y = reinterpret_cast<float *>(d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const char * mp = has_mask ? m.data + i1 * m.nb[1] + (i2 % m.ne[2]) * m.nb[2] + (i3 % m.ne[3]) * m.nb[3] : nullptr;
    const long long n = a.ne[0];
    __shared__ float red[8];
    auto val = [&](long long i) { float v = x[i] * scale; if (mp) v += m.type == B200_TYPE_F16 ? __half2float(reinterpret_cast<const __half *>(mp)[i]) : reinterpret_cast<const float *>(mp)[i]; return v; };
    float mx = -INFINITY;
    for (long long i = threadIdx.x; i < n; i += 256) mx = fmaxf(mx, val(i));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    mx = red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
    __syncthreads();
    float s = 0.0f;
    for (long long i = threadIdx.x; i < n; i += 256) { const float e = expf(val(i) - mx); y[i] = e; s += e; }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i];
    const float inv = 1.0f / s;
    for (long long i = threadIdx.x; i < n; i += 256) y[i] *= inv;
}

// rows of up to 1024 elements (attention scores over the KV cache): one WARP per row, the row lives in registers between the passes.
// NJ = elements per lane (8 / 16 / 32 for rows of up to 256 / 512 / 1024); 32-bit index arithmetic (the host checks the row count)
template <int NJ>
__global__ void __launch_bounds__(128) soft_max_warp_kernel(const T4 a, const T4 m, int has_mask, const T4w d, float scale, unsigned rows) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const int lane = threadIdx.x & 31;
    const unsigned row = blockIdx.x * 4u + (threadIdx.x >> 5);
    if (row >= rows) return;
    const unsigned ne1 = (unsigned) a.ne[1], ne2 = (unsigned) a.ne[2];
    const unsigned i1 = row % ne1, r12 = row / ne1, i2 = r12 % ne2, i3 = r12 / ne2;
    const float * x = reinterpret_cast<const float *>(a.data + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
    float * y = reinterpret_cast<float *>(d.data + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]);
    const char * mp = has_mask ? m.data + (long long) i1 * m.nb[1] + (long long) (i2 % (unsigned) m.ne[2]) * m.nb[2] + (long long) (i3 % (unsigned) m.ne[3]) * m.nb[3] : nullptr;
    const int n = (int) a.ne[0];
    const bool mh = m.type == B200_TYPE_F16;
    float v[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) { const int i = lane + 32 * j; v[j] = i < n ? x[i] : -INFINITY; }
    if (mp) {
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int i = lane + 32 * j;
            const float mk = i < n ? (mh ? __half2float(reinterpret_cast<const __half *>(mp)[i]) : reinterpret_cast<const float *>(mp)[i]) : 0.0f;
            v[j] = v[j] * scale + mk;
        }
    } else {
#pragma unroll
        for (int j = 0; j < NJ; ++j) v[j] *= scale;
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NJ; ++j) mx = fmaxf(mx, v[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) { v[j] = lane + 32 * j < n ? expf(v[j] - mx) : 0.0f; s += v[j]; }
    s = warp_sum(s);
    const float inv = 1.0f / s;
#pragma unroll
    for (int j = 0; j < NJ; ++j) { const int i = lane + 32 * j; if (i < n) y[i] = v[j] * inv; }
}

// ---- GLU: reglu 0, geglu 1, swiglu 2 (enum ggml_glu_op, ggml.h; ggml_compute_forward_glu, ops.cpp): dst = act(x) * g ----
__device__ __forceinline__ float glu_act(int op, float x) {
    if (op == 0) return x > 0.0f ? x : 0.0f;
    if (op == 1) return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x)));
    return x / (1.0f + expf(-x));
}
template <typename I>
__global__ void __launch_bounds__(256) glu_kernel(const T4 a, const T4 b, int split, const T4w d, int op, int swapped) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const I nc = (I) d.ne[0], total = (I) (d.ne[0] * d.ne[1] * d.ne[2] * d.ne[3]);
    for (I i = (I) blockIdx.x * 256 + threadIdx.x; i < total; i += (I) gridDim.x * 256) {
        I i0, i1, i2, i3;
        unravel<I>(i, d.ne, i0, i1, i2, i3);
        const float * ar = reinterpret_cast<const float *>(a.data + (long long) i1 * a.nb[1] + (long long) i2 * a.nb[2] + (long long) i3 * a.nb[3]);
        float x, g;
        if (split) {
            const float * br = reinterpret_cast<const float *>(b.data + (long long) i1 * b.nb[1] + (long long) i2 * b.nb[2] + (long long) i3 * b.nb[3]);
            x = ar[i0]; g = br[i0];
        } else {
            x = ar[swapped ? i0 + nc : i0]; g = ar[swapped ? i0 : i0 + nc];
        }
        *reinterpret_cast<float *>(d.data + (long long) i0 * d.nb[0] + (long long) i1 * d.nb[1] + (long long) i2 * d.nb[2] + (long long) i3 * d.nb[3]) = glu_act(op, x) * g;
    }
}

// 2-D, split operands, rows of a multiple of four floats, everything 16-byte aligned (the SwiGLU of a prefill ubatch): float4 per thread
__global__ void __launch_bounds__(256) glu_vec4_kernel(const T4 a, const T4 b, const T4w d, int op, int swapped, long long total4) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const long long q = d.ne[0] >> 2;
    for (long long i = (long long) blockIdx.x * 256 + threadIdx.x; i < total4; i += (long long) gridDim.x * 256) {
        const long long row = i / q, c = (i - row * q) * 16;
        float4 x = *reinterpret_cast<const float4 *>(a.data + row * a.nb[1] + c), g = *reinterpret_cast<const float4 *>(b.data + row * b.nb[1] + c);
        if (swapped) { const float4 t = x; x = g; g = t; }
        float4 r;
        r.x = glu_act(op, x.x) * g.x; r.y = glu_act(op, x.y) * g.y; r.z = glu_act(op, x.z) * g.z; r.w = glu_act(op, x.w) * g.w;
        *reinterpret_cast<float4 *>(d.data + row * d.nb[1] + c) = r;
    }
}

// ---- MUL_MAT with f16 / f32 src0 rows and f32 src1 (attention over the KV cache): dst[m, n, i2, i3] = sum_k a[k, m, i2 / r2, i3 / r3] b[k, n, i2, i3] ----
// few activation rows: one warp per (src0 row, batch), every lane strides over k, all n columns at once (n <= 8)
template <int NCOLS, bool AH>
__global__ void __launch_bounds__(128) mm_rows_kernel(const T4 a, const T4 b, const T4w d, long long r2, long long r3) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    const int lane = threadIdx.x & 31;
    const long long row = (long long) blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= a.ne[1]) return;
    const long long i2 = blockIdx.y % d.ne[2], i3 = blockIdx.y / d.ne[2];
    const char * ar = a.data + row * a.nb[1] + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3];
    const char * bb = b.data + i2 * b.nb[2] + i3 * b.nb[3];
    float acc[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) acc[c] = 0.0f;
    for (long long k = lane; k < a.ne[0]; k += 32) {
        const float w = AH ? __half2float(reinterpret_cast<const __half *>(ar)[k]) : reinterpret_cast<const float *>(ar)[k];
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[c] += w * reinterpret_cast<const float *>(bb + c * b.nb[1])[k];
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const float v = warp_sum(acc[c]);
        if (lane == 0) *reinterpret_cast<float *>(d.data + row * d.nb[0] + c * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = v;
    }
}
// few src0 rows with a long k (the router of a mixture-of-experts layer: 4096 -> 8): one CTA of 256 threads per row instead of one warp
template <int NCOLS, bool AH>
__global__ void __launch_bounds__(256) mm_rowblock_kernel(const T4 a, const T4 b, const T4w d, long long r2, long long r3) {
    pdl_launch_dependents(); pdl_wait();
    __shared__ float red[NCOLS][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long row = blockIdx.x;
    const long long i2 = blockIdx.y % d.ne[2], i3 = blockIdx.y / d.ne[2];
    const char * ar = a.data + row * a.nb[1] + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3];
    const char * bb = b.data + i2 * b.nb[2] + i3 * b.nb[3];
    float acc[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) acc[c] = 0.0f;
    for (long long k = threadIdx.x; k < a.ne[0]; k += 256) {
        const float w = AH ? __half2float(reinterpret_cast<const __half *>(ar)[k]) : reinterpret_cast<const float *>(ar)[k];
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[c] += w * reinterpret_cast<const float *>(bb + c * b.nb[1])[k];
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) { const float v = warp_sum(acc[c]); if (lane == 0) red[c][warp] = v; }
    __syncthreads();
    if (threadIdx.x < NCOLS) {
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += red[threadIdx.x][w];
        *reinterpret_cast<float *>(d.data + row * d.nb[0] + threadIdx.x * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = v;
    }
}
// many activation rows: 64 x 64 output tile per CTA of 256 threads (4 x 4 per thread), k in steps of 16 through shared memory
template <bool AH>
__global__ void __launch_bounds__(256) mm_tile_kernel(const T4 a, const T4 b, const T4w d, long long r2, long long r3) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    __shared__ float As[16][64 + 4], Bs[16][64 + 4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const long long m0 = (long long) blockIdx.x * 64, n0 = (long long) blockIdx.y * 64;
    const long long i2 = blockIdx.z % d.ne[2], i3 = blockIdx.z / d.ne[2];
    const char * ab = a.data + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3];
    const char * bb = b.data + i2 * b.nb[2] + i3 * b.nb[3];
    const long long K = a.ne[0], M = a.ne[1], N = b.ne[1];
    float acc[4][4] = {};
    for (long long k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {               // 64 rows x 16 k = 1024 values per operand, 4 per thread
            const int idx = threadIdx.x + t * 256, r = idx >> 4, kk = idx & 15;
            const long long k = k0 + kk;
            float va = 0.0f, vb = 0.0f;
            if (k < K && m0 + r < M) va = AH ? __half2float(reinterpret_cast<const __half *>(ab + (m0 + r) * a.nb[1])[k]) : reinterpret_cast<const float *>(ab + (m0 + r) * a.nb[1])[k];
            if (k < K && n0 + r < N) vb = reinterpret_cast<const float *>(bb + (n0 + r) * b.nb[1])[k];
            As[kk][r] = va; Bs[kk][r] = vb;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { av[i] = As[kk][tx * 4 + i]; bv[i] = Bs[kk][ty * 4 + i]; }
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[j][i] += av[i] * bv[j];
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const long long m = m0 + tx * 4 + i, n = n0 + ty * 4 + j;
            if (m < M && n < N) *reinterpret_cast<float *>(d.data + m * d.nb[0] + n * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = acc[j][i];
        }
}

// f16 rows x many f32 rows on the tensor cores (the K^T Q and V P products of prefill attention: K = 128 / n_kv, 512 query tokens, 32 heads):
// 64 x 64 output tile per CTA of four warps, k in steps of 32; the activations are rounded to f16 on their way into shared memory (what the
// reference CPU does for an F16 src0: vec_dot_type F16, ggml-cpu.c:217-235) and products accumulate in f32 (mma.sync.m16n8k16; these
// are small batched matmuls between the quantised GEMMs, the warp-level MMA is enough to take them off the critical path).
// Needs K % 32 == 0 and 16-byte aligned rows of both operands (host check); everything else stays on mm_tile_kernel.
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__global__ void __launch_bounds__(128) mm_mma_kernel(const T4 a, const T4 b, const T4w d, long long r2, long long r3) {
    pdl_launch_dependents(); pdl_wait();      // programmatic dependent launch (no-ops without the launch attribute)
    constexpr int BK = 32, LD = BK + 8;                       // row stride 40 halves = 20 words: the fragment loads below are conflict-free
    __shared__ __align__(16) __half As[64][LD], Bs[64][LD];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const long long m0 = (long long) blockIdx.x * 64, n0 = (long long) blockIdx.y * 64;
    const long long i2 = blockIdx.z % d.ne[2], i3 = blockIdx.z / d.ne[2];
    const char * ab = a.data + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3];
    const char * bb = b.data + i2 * b.nb[2] + i3 * b.nb[3];
    const long long K = a.ne[0], M = a.ne[1], N = b.ne[1];
    // loads: thread -> row tid / 2, 16 consecutive k of the 32 (tid & 1)
    const int lr = tid >> 1, lk = (tid & 1) * 16;
    const bool a_ok = m0 + lr < M, b_ok = n0 + lr < N;
    const char * ap = ab + (a_ok ? m0 + lr : 0) * a.nb[1] + lk * 2;
    const char * bp = bb + (b_ok ? n0 + lr : 0) * b.nb[1] + lk * 4;
    uint4 ra[2]; float4 rb[4];
    auto fetch = [&](long long k0) {
        if (a_ok) { ra[0] = *reinterpret_cast<const uint4 *>(ap + k0 * 2); ra[1] = *reinterpret_cast<const uint4 *>(ap + k0 * 2 + 16); }
        else { ra[0] = make_uint4(0, 0, 0, 0); ra[1] = ra[0]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) rb[i] = b_ok ? *reinterpret_cast<const float4 *>(bp + k0 * 4 + 16 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    auto stash = [&]() {
        *reinterpret_cast<uint4 *>(&As[lr][lk]) = ra[0];
        *reinterpret_cast<uint4 *>(&As[lr][lk + 8]) = ra[1];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __half2 h0 = __floats2half2_rn(rb[i].x, rb[i].y), h1 = __floats2half2_rn(rb[i].z, rb[i].w);
            *reinterpret_cast<uint2 *>(&Bs[lr][lk + 4 * i]) = make_uint2(*reinterpret_cast<const uint32_t *>(&h0), *reinterpret_cast<const uint32_t *>(&h1));
        }
    };
    // warp tile: 32 (m) x 32 (n) = 2 x 4 MMAs per k16
    const int wm = (warp & 1) * 32, wn = (warp >> 1) * 32;
    float acc[2][4][4] = {};
    fetch(0);
    for (long long k0 = 0; k0 < K; k0 += BK) {
        __syncthreads();                                        // the previous step's fragment loads are done
        stash();
        __syncthreads();
        if (k0 + BK < K) fetch(k0 + BK);                        // in flight during the MMAs
#pragma unroll
        for (int kk = 0; kk < BK; kk += 16) {
            uint32_t fa[2][4], fb[4][2];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const __half * r0 = &As[wm + 16 * i + g][kk + 2 * t], * r1 = &As[wm + 16 * i + g + 8][kk + 2 * t];
                fa[i][0] = *reinterpret_cast<const uint32_t *>(r0);     fa[i][1] = *reinterpret_cast<const uint32_t *>(r1);
                fa[i][2] = *reinterpret_cast<const uint32_t *>(r0 + 8); fa[i][3] = *reinterpret_cast<const uint32_t *>(r1 + 8);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const __half * r = &Bs[wn + 8 * j + g][kk + 2 * t];
                fb[j][0] = *reinterpret_cast<const uint32_t *>(r); fb[j][1] = *reinterpret_cast<const uint32_t *>(r + 8);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma16816(acc[i][j], fa[i], fb[j]);
        }
    }
    char * db = d.data + i2 * d.nb[2] + i3 * d.nb[3];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const long long m = m0 + wm + 16 * i + g + 8 * (e >> 1), n = n0 + wn + 8 * j + 2 * t + (e & 1);
                if (m < M && n < N) *reinterpret_cast<float *>(db + m * 4 + n * d.nb[1]) = acc[i][j][e];
            }
}

// ---- prefill attention for the MUL_MAT(K, q) -> SOFT_MAX -> MUL_MAT(V^T, p) -> PERMUTE -> CONT run of build_attn (no FLASH_ATTN_EXT in the graph):
// one pass over the KV cache with an online softmax, both products on the tensor cores (mma.sync.m16n8k16, f16 operands, f32 accumulation;
// q and the probabilities are rounded to f16 exactly where the reference CPU rounds them: src1 of an F16 matmul). The score matrix never
// reaches memory. CTA = 64 query tokens of one head (four warps x 16 rows), KV tiles of 64; a tile whose mask is -inf for all 64 x 64
// entries (the upper triangle of a causal ubatch) is skipped before its K / V are loaded. HD = head dimension (64 or 128).
template <int HD>
__global__ void __launch_bounds__(128) attn_prefill_kernel(const T4 k, const T4 q, const T4 m, int has_mask, const T4 v, float * __restrict__ dst, float scale) {
    pdl_launch_dependents(); pdl_wait();
    constexpr int LDK = HD + 8, LDV = 64 + 8;                  // row strides in halves: conflict-free fragment loads (see mm_mma_kernel)
    __shared__ __align__(16) __half Ks[64][LDK];               // K tile [kv][d]; first used to stage the q tile [token][d]
    __shared__ __align__(16) __half Vs[HD][LDV];               // V^T tile [d][kv]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int n_kv = (int) k.ne[1], n_tok = (int) q.ne[1], n_head = (int) q.ne[2];
    const int head = blockIdx.y, kvh = head / (n_head / (int) k.ne[2]);
    const int tok0 = blockIdx.x * 64;
    const char * kb = k.data + (long long) kvh * k.nb[2];
    const char * vb = v.data + (long long) kvh * v.nb[2];
    const char * qb = q.data + (long long) head * q.nb[2];

    // q tile -> f16 fragments (A operand of the score product), kept in registers for the whole pass
    for (int i = tid; i < 64 * (HD / 4); i += 128) {
        const int r = i / (HD / 4), c = (i - r * (HD / 4)) * 4;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tok0 + r < n_tok) x = *reinterpret_cast<const float4 *>(qb + (long long) (tok0 + r) * q.nb[1] + c * 4);
        const __half2 h0 = __floats2half2_rn(x.x, x.y), h1 = __floats2half2_rn(x.z, x.w);
        *reinterpret_cast<uint2 *>(&Ks[r][c]) = make_uint2(*reinterpret_cast<const uint32_t *>(&h0), *reinterpret_cast<const uint32_t *>(&h1));
    }
    __syncthreads();
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
        const __half * r0 = &Ks[warp * 16 + g][ks * 16 + 2 * t], * r1 = &Ks[warp * 16 + g + 8][ks * 16 + 2 * t];
        qf[ks][0] = *reinterpret_cast<const uint32_t *>(r0);     qf[ks][1] = *reinterpret_cast<const uint32_t *>(r1);
        qf[ks][2] = *reinterpret_cast<const uint32_t *>(r0 + 8); qf[ks][3] = *reinterpret_cast<const uint32_t *>(r1 + 8);
    }
    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = 0.f; o[j][1] = 0.f; o[j][2] = 0.f; o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int tokr[2] = {tok0 + warp * 16 + g, tok0 + warp * 16 + g + 8};
    const bool mh = m.type == B200_TYPE_F16;

    for (int kv0 = 0; kv0 < n_kv; kv0 += 64) {
        // mask values of this thread's 2 x 16 score entries (columns 8 j + 2 t + {0, 1})
        float mk[2][16];
        int finite = 0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool live = tokr[h] < n_tok;
            const char * mp = has_mask && live ? m.data + (long long) tokr[h] * m.nb[1] : nullptr;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float a = 0.f, b = 0.f;
                if (mp) {
                    const int c = kv0 + 8 * j + 2 * t;
                    if (mh) { const __half2 hv = *reinterpret_cast<const __half2 *>(mp + c * 2); a = __low2float(hv); b = __high2float(hv); }
                    else    { const float2 fv = *reinterpret_cast<const float2 *>(mp + c * 4); a = fv.x; b = fv.y; }
                }
                mk[h][2 * j] = a; mk[h][2 * j + 1] = b;
                finite |= (live && (a > -INFINITY || b > -INFINITY)) ? 1 : 0;
            }
        }
        __syncthreads();                                        // the previous tile's fragment loads are done (and the q staging, first time round)
        if (!__syncthreads_or(finite)) continue;                // nothing of this tile is visible to any of the 64 tokens
        for (int i = tid; i < 64 * (HD / 8); i += 128) {        // K tile: 64 rows x HD halves
            const int r = i / (HD / 8), c = (i - r * (HD / 8)) * 8;
            *reinterpret_cast<uint4 *>(&Ks[r][c]) = *reinterpret_cast<const uint4 *>(kb + (long long) (kv0 + r) * k.nb[1] + c * 2);
        }
        for (int i = tid; i < HD * 8; i += 128) {               // V^T tile: HD rows x 64 halves
            const int r = i >> 3, c = (i & 7) * 8;
            *reinterpret_cast<uint4 *>(&Vs[r][c]) = *reinterpret_cast<const uint4 *>(vb + (long long) r * v.nb[1] + (long long) (kv0 + c) * 2);
        }
        __syncthreads();
        // scores: 16 rows x 64 kv per warp
        float sc[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) { sc[j][0] = 0.f; sc[j][1] = 0.f; sc[j][2] = 0.f; sc[j][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const __half * r = &Ks[8 * j + g][ks * 16 + 2 * t];
                const uint32_t fb[2] = {*reinterpret_cast<const uint32_t *>(r), *reinterpret_cast<const uint32_t *>(r + 8)};
                mma16816(sc[j], qf[ks], fb);
            }
        // online softmax over the tile (rows g and g + 8 of the warp; a row lives in the four lanes of a quad)
        uint32_t pf[4][4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                sc[j][2 * h] = sc[j][2 * h] * scale + mk[h][2 * j];
                sc[j][2 * h + 1] = sc[j][2 * h + 1] * scale + mk[h][2 * j + 1];
                mx = fmaxf(mx, fmaxf(sc[j][2 * h], sc[j][2 * h + 1]));
            }
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            const float mnew = fmaxf(mrow[h], mx);
            const float base = mnew == -INFINITY ? 0.f : mnew;          // a row that has seen nothing but -inf so far: keep everything at zero
            const float corr = expf(mrow[h] - base);                   // mrow = -inf -> 0
            float sum = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float p0 = expf(sc[j][2 * h] - base), p1 = expf(sc[j][2 * h + 1] - base);
                sc[j][2 * h] = p0; sc[j][2 * h + 1] = p1;
                sum += p0 + p1;
            }
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            lrow[h] = lrow[h] * corr + sum;
            mrow[h] = mnew;
#pragma unroll
            for (int j = 0; j < HD / 8; ++j) { o[j][2 * h] *= corr; o[j][2 * h + 1] *= corr; }
        }
        // probabilities as the A operand of the second product: two adjacent 8-column score tiles are one 16-wide k step
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            const __half2 a0 = __floats2half2_rn(sc[2 * ks][0], sc[2 * ks][1]), a1 = __floats2half2_rn(sc[2 * ks][2], sc[2 * ks][3]);
            const __half2 a2 = __floats2half2_rn(sc[2 * ks + 1][0], sc[2 * ks + 1][1]), a3 = __floats2half2_rn(sc[2 * ks + 1][2], sc[2 * ks + 1][3]);
            pf[ks][0] = *reinterpret_cast<const uint32_t *>(&a0); pf[ks][1] = *reinterpret_cast<const uint32_t *>(&a1);
            pf[ks][2] = *reinterpret_cast<const uint32_t *>(&a2); pf[ks][3] = *reinterpret_cast<const uint32_t *>(&a3);
        }
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
            for (int j = 0; j < HD / 8; ++j) {
                const __half * r = &Vs[8 * j + g][ks * 16 + 2 * t];
                const uint32_t fb[2] = {*reinterpret_cast<const uint32_t *>(r), *reinterpret_cast<const uint32_t *>(r + 8)};
                mma16816(o[j], pf[ks], fb);
            }
    }
    // dst [head_dim, n_head, n_tokens] contiguous
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (tokr[h] >= n_tok) continue;
        const float inv = 1.0f / lrow[h];
        float * dp = dst + ((long long) tokr[h] * n_head + head) * HD;
#pragma unroll
        for (int j = 0; j < HD / 8; ++j) *reinterpret_cast<float2 *>(dp + 8 * j + 2 * t) = make_float2(o[j][2 * h] * inv, o[j][2 * h + 1] * inv);
    }
}

// ---- the router of build_moe_ffn as ONE launch (softmax gating, weights renormalised over the used experts): SOFT_MAX -> ARGSORT (descending) ->
// GET_ROWS of the probabilities at the first n_used indices -> SUM_ROWS -> CLAMP -> DIV. One warp per token, lane e = expert e (n_expert <= 32);
// every intermediate tensor of the six nodes is written, so the rest of the graph reads them as if the nodes had run one by one.
__global__ void __launch_bounds__(128) moe_route_kernel(const float * __restrict__ logits, float * __restrict__ probs, int * __restrict__ order, float * __restrict__ picked,
                                                        float * __restrict__ sum, float * __restrict__ weights, int n_expert, int n_used, int n_tok, float lo, float hi) {
    pdl_launch_dependents(); pdl_wait();
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (t >= n_tok) return;
    const bool live = lane < n_expert;
    const float x = live ? logits[(long long) t * n_expert + lane] : -INFINITY;
    float mx = x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float e = live ? expf(x - mx) : 0.0f;
    const float p = e * (1.0f / warp_sum(e));
    if (live && probs) probs[(long long) t * n_expert + lane] = p;
    int rank = 0;                                            // position of this expert in descending order (ties: lower index first)
    for (int j = 0; j < n_expert; ++j) { const float pj = __shfl_sync(0xffffffffu, p, j); rank += (pj > p || (pj == p && j < lane)) ? 1 : 0; }
    if (live) order[(long long) t * n_expert + rank] = lane;
    const bool used = live && rank < n_used;
    if (used && picked) picked[(long long) t * n_used + rank] = p;
    float s = warp_sum(used ? p : 0.0f);
    s = fminf(fmaxf(s, lo), hi);
    if (lane == 0 && sum) sum[t] = s;
    if (used) weights[(long long) t * n_used + rank] = p / s;
}

} // namespace

extern "C" {

int b200_op_get_rows(const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(src0 && ids && dst && src0->data && ids->data && dst->data, "b200_op_get_rows: null tensor");
    B200_REQUIRE(ids->type == B200_TYPE_I32 && dst->type == B200_TYPE_F32, "b200_op_get_rows: ids must be i32 and dst f32");
    B200_REQUIRE(src0->type == B200_TYPE_F32 || src0->type == B200_TYPE_F16 || b200_type_is_weight(src0->type), "b200_op_get_rows: unsupported source type %d", src0->type);
    const long long n = nelem(dst);
    if (n == 0) return 0;
    B200_LAUNCH_IDX(get_rows_kernel, n, t4(src0), t4(ids), t4w(dst));
    return 0;
}

int b200_op_rms_norm(const b200_tensor * src, const b200_tensor * dst, float eps, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_rms_norm: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && src->nb[0] == 4 && dst->nb[0] == 4, "b200_op_rms_norm: f32 rows with unit stride only");
    const long long rows = src->ne[1] * src->ne[2] * src->ne[3];
    if (rows == 0 || src->ne[0] == 0) return 0;
    B200_CHECK(b200_launch(rms_norm_kernel, dim3((unsigned) rows), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), t4(src), t4w(dst), eps));
    return 0;
}

int b200_op_add_rms_norm_mul(const b200_tensor * a, const b200_tensor * b, const b200_tensor * sum, const b200_tensor * w, const b200_tensor * dst, float eps, b200_stream_t stream) {
    B200_REQUIRE(a && w && dst && a->data && w->data && dst->data && (!b || b->data) && (!sum || sum->data), "b200_op_add_rms_norm_mul: null tensor");
    B200_REQUIRE(a->type == B200_TYPE_F32 && w->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && a->nb[0] == 4 && w->nb[0] == 4 && dst->nb[0] == 4 &&
                 (!b || (b->type == B200_TYPE_F32 && b->nb[0] == 4 && b->ne[0] == a->ne[0])) && (!sum || (sum->type == B200_TYPE_F32 && sum->nb[0] == 4)),
                 "b200_op_add_rms_norm_mul: f32 rows with unit stride only");
    B200_REQUIRE(w->ne[0] > 0 && a->ne[0] % w->ne[0] == 0, "b200_op_add_rms_norm_mul: weight does not broadcast over the rows");
    const long long rows = a->ne[1] * a->ne[2] * a->ne[3];
    if (rows == 0 || a->ne[0] == 0) return 0;
    B200_REQUIRE(w->ne[0] == a->ne[0], "b200_op_add_rms_norm_mul: the weight must have the row length");
    if (rows >= 128)
        B200_CHECK(b200_launch(add_rms_norm_mul_kernel<256>, dim3((unsigned) rows), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), t4(a), b ? t4(b) : t4(a), b ? 1 : 0,
                               sum ? t4w(sum) : t4w(dst), sum ? 1 : 0, t4(w), t4w(dst), eps));
    else
        B200_CHECK(b200_launch(add_rms_norm_mul_kernel<1024>, dim3((unsigned) rows), dim3(1024), 0, (cudaStream_t) stream, glue_pdl(), t4(a), b ? t4(b) : t4(a), b ? 1 : 0,
                               sum ? t4w(sum) : t4w(dst), sum ? 1 : 0, t4(w), t4w(dst), eps));
    return 0;
}

int b200_op_attn_decode(const b200_tensor * k, const b200_tensor * q, const b200_tensor * mask, const b200_tensor * v, const b200_tensor * dst, float scale, b200_stream_t stream) {
    B200_REQUIRE(k && q && v && dst && k->data && q->data && v->data && dst->data && (!mask || mask->data), "b200_op_attn_decode: null tensor");
    B200_REQUIRE(k->type == B200_TYPE_F16 && v->type == B200_TYPE_F16 && q->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && k->nb[0] == 2 && v->nb[0] == 2 && q->nb[0] == 4,
                 "b200_op_attn_decode: f16 K / V rows, f32 q");
    const long long hd = k->ne[0], n_kv = k->ne[1], n_head = q->ne[2], n_tok = q->ne[1];
    B200_REQUIRE(q->ne[0] == hd && v->ne[0] == n_kv && v->ne[1] == hd && v->ne[2] == k->ne[2] && k->ne[2] > 0 && n_head % k->ne[2] == 0 && k->ne[3] == 1 && q->ne[3] == 1 &&
                 hd % 32 == 0 && (k->nb[1] % 16) == 0 && (k->nb[2] % 16) == 0 && ((uintptr_t) k->data % 16) == 0 && (v->nb[1] % 2) == 0 && (v->nb[2] % 2) == 0, "b200_op_attn_decode: shapes");
    B200_REQUIRE(n_kv >= 1 && n_kv <= 11000 && n_tok >= 1 && n_tok <= 65535, "b200_op_attn_decode: n_kv %lld / n_tokens %lld out of range", n_kv, n_tok);
    B200_REQUIRE(!mask || ((mask->type == B200_TYPE_F32 || mask->type == B200_TYPE_F16) && mask->ne[0] >= n_kv && mask->ne[1] >= n_tok), "b200_op_attn_decode: bad mask");
    B200_REQUIRE(contiguous_f(dst, 4) && nelem(dst) == hd * n_head * n_tok, "b200_op_attn_decode: dst must be the contiguous [head_dim, n_head, n_tokens] result");
    B200_CHECK(b200_launch(attn_decode_kernel, dim3((unsigned) n_head, (unsigned) n_tok), dim3(256), (size_t) (n_kv + hd + 256) * sizeof(float), (cudaStream_t) stream, glue_pdl(),
                           t4(k), t4(q), mask ? t4(mask) : t4(q), mask ? 1 : 0, t4(v), (float *) dst->data, scale));
    return 0;
}

int b200_op_attn_prefill(const b200_tensor * k, const b200_tensor * q, const b200_tensor * mask, const b200_tensor * v, const b200_tensor * dst, float scale, b200_stream_t stream) {
    B200_REQUIRE(k && q && v && dst && k->data && q->data && v->data && dst->data && (!mask || mask->data), "b200_op_attn_prefill: null tensor");
    B200_REQUIRE(k->type == B200_TYPE_F16 && v->type == B200_TYPE_F16 && q->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && k->nb[0] == 2 && v->nb[0] == 2 && q->nb[0] == 4,
                 "b200_op_attn_prefill: f16 K / V rows, f32 q");
    const long long hd = k->ne[0], n_kv = k->ne[1], n_head = q->ne[2], n_tok = q->ne[1];
    auto al16 = [](const b200_tensor * t) { return ((uintptr_t) t->data % 16) == 0 && t->nb[1] % 16 == 0 && t->nb[2] % 16 == 0; };
    B200_REQUIRE((hd == 64 || hd == 128) && q->ne[0] == hd && v->ne[0] == n_kv && v->ne[1] == hd && v->ne[2] == k->ne[2] && k->ne[2] > 0 && n_head % k->ne[2] == 0 &&
                 k->ne[3] == 1 && q->ne[3] == 1 && v->ne[3] == 1 && n_kv >= 64 && n_kv % 64 == 0 && al16(k) && al16(v) && al16(q), "b200_op_attn_prefill: shapes / alignment");
    B200_REQUIRE(n_tok >= 1 && n_head >= 1 && n_head <= 65535, "b200_op_attn_prefill: n_tokens %lld / n_head %lld out of range", n_tok, n_head);
    B200_REQUIRE(!mask || ((mask->type == B200_TYPE_F32 || mask->type == B200_TYPE_F16) && mask->ne[0] >= n_kv && mask->ne[1] >= n_tok && mask->ne[2] == 1 && mask->ne[3] == 1 &&
                           mask->nb[0] == (mask->type == B200_TYPE_F16 ? 2 : 4) && ((uintptr_t) mask->data % 8) == 0 && mask->nb[1] % 8 == 0), "b200_op_attn_prefill: bad mask");
    B200_REQUIRE(contiguous_f(dst, 4) && nelem(dst) == hd * n_head * n_tok && ((uintptr_t) dst->data % 8) == 0, "b200_op_attn_prefill: dst must be the contiguous [head_dim, n_head, n_tokens] result");
    const dim3 grid(blocks_for(n_tok, 64), (unsigned) n_head);
    if (hd == 128) B200_CHECK(b200_launch(attn_prefill_kernel<128>, grid, dim3(128), 0, (cudaStream_t) stream, glue_pdl(), t4(k), t4(q), mask ? t4(mask) : t4(q), mask ? 1 : 0, t4(v), (float *) dst->data, scale));
    else           B200_CHECK(b200_launch(attn_prefill_kernel<64>, grid, dim3(128), 0, (cudaStream_t) stream, glue_pdl(), t4(k), t4(q), mask ? t4(mask) : t4(q), mask ? 1 : 0, t4(v), (float *) dst->data, scale));
    return 0;
}

int b200_op_moe_route(const b200_tensor * logits, const b200_tensor * probs, const b200_tensor * order, const b200_tensor * picked, const b200_tensor * sum,
                      const b200_tensor * weights, int n_used, float clamp_lo, float clamp_hi, b200_stream_t stream) {
    B200_REQUIRE(logits && order && weights && logits->data && order->data && weights->data && (!probs || probs->data) && (!picked || picked->data) && (!sum || sum->data), "b200_op_moe_route: null tensor");
    const long long n_expert = logits->ne[0], n_tok = logits->ne[1];
    B200_REQUIRE(logits->type == B200_TYPE_F32 && (!probs || probs->type == B200_TYPE_F32) && order->type == B200_TYPE_I32 && (!picked || picked->type == B200_TYPE_F32) &&
                 (!sum || sum->type == B200_TYPE_F32) && weights->type == B200_TYPE_F32, "b200_op_moe_route: f32 tensors, i32 order");
    B200_REQUIRE(n_expert >= 1 && n_expert <= 32 && n_used >= 1 && n_used <= n_expert && n_tok >= 1 && n_tok < (1ll << 30) && logits->ne[2] == 1 && logits->ne[3] == 1,
                 "b200_op_moe_route: %lld experts / %d used / %lld tokens", n_expert, n_used, n_tok);
    B200_REQUIRE(contiguous_f(logits, 4) && (!probs || (contiguous_f(probs, 4) && nelem(probs) == n_expert * n_tok)) && contiguous_f(order, 4) &&
                 (!picked || (contiguous_f(picked, 4) && nelem(picked) == n_used * n_tok)) && (!sum || (contiguous_f(sum, 4) && nelem(sum) == n_tok)) && contiguous_f(weights, 4) &&
                 nelem(order) == n_expert * n_tok && nelem(weights) == n_used * n_tok,
                 "b200_op_moe_route: contiguous tensors of the shapes of the six nodes");
    B200_CHECK(b200_launch(moe_route_kernel, dim3(blocks_for(n_tok, 4)), dim3(128), 0, (cudaStream_t) stream, glue_pdl(), (const float *) logits->data, probs ? (float *) probs->data : nullptr,
                           (int *) order->data, picked ? (float *) picked->data : nullptr, sum ? (float *) sum->data : nullptr, (float *) weights->data, (int) n_expert, n_used, (int) n_tok, clamp_lo, clamp_hi));
    return 0;
}

int b200_op_binary(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(a && b && dst && a->data && b->data && dst->data, "b200_op_binary: null tensor");
    B200_REQUIRE(a->type == B200_TYPE_F32 && b->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32, "b200_op_binary: f32 only");
    B200_REQUIRE(op >= 0 && op <= 2, "b200_op_binary: op %d (0 add, 1 mul, 2 div)", op);
    for (int i = 0; i < 4; ++i) B200_REQUIRE(b->ne[i] > 0 && dst->ne[i] % b->ne[i] == 0 && a->ne[i] == dst->ne[i], "b200_op_binary: src1 does not broadcast over src0");
    const long long n = nelem(dst);
    if (n == 0) return 0;
    const bool small = n < (1ll << 31);
    const dim3 grid(blocks_for(n, 256));
    cudaStream_t st = (cudaStream_t) stream;
    if (op == 0) { if (small) B200_CHECK(b200_launch(binary_kernel<0, unsigned>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst)));
                   else       B200_CHECK(b200_launch(binary_kernel<0, long long>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst))); }
    else if (op == 1) { if (small) B200_CHECK(b200_launch(binary_kernel<1, unsigned>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst)));
                        else       B200_CHECK(b200_launch(binary_kernel<1, long long>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst))); }
    else              { if (small) B200_CHECK(b200_launch(binary_kernel<2, unsigned>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst)));
                        else       B200_CHECK(b200_launch(binary_kernel<2, long long>, grid, dim3(256), 0, st, glue_pdl(), t4(a), t4(b), t4w(dst))); }
    return 0;
}

int b200_op_unary(int op, const b200_tensor * src, const b200_tensor * dst, float p0, float p1, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_unary: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32, "b200_op_unary: f32 only");
    B200_REQUIRE(op == 0 || op == 1, "b200_op_unary: op %d (0 scale: x * p0 + p1, 1 clamp to [p0, p1])", op);
    for (int i = 0; i < 4; ++i) B200_REQUIRE(src->ne[i] == dst->ne[i], "b200_op_unary: shapes differ");
    const long long n = nelem(dst);
    if (n == 0) return 0;
    const bool small = n < (1ll << 31);
    const dim3 grid(blocks_for(n, 256));
    cudaStream_t st = (cudaStream_t) stream;
    if (op == 0) { if (small) B200_CHECK(b200_launch(unary_kernel<0, unsigned>, grid, dim3(256), 0, st, glue_pdl(), t4(src), t4w(dst), p0, p1));
                   else       B200_CHECK(b200_launch(unary_kernel<0, long long>, grid, dim3(256), 0, st, glue_pdl(), t4(src), t4w(dst), p0, p1)); }
    else         { if (small) B200_CHECK(b200_launch(unary_kernel<1, unsigned>, grid, dim3(256), 0, st, glue_pdl(), t4(src), t4w(dst), p0, p1));
                   else       B200_CHECK(b200_launch(unary_kernel<1, long long>, grid, dim3(256), 0, st, glue_pdl(), t4(src), t4w(dst), p0, p1)); }
    return 0;
}

int b200_op_sum_rows(const b200_tensor * src, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_sum_rows: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && dst->ne[0] == 1 && dst->ne[1] == src->ne[1] && dst->ne[2] == src->ne[2] && dst->ne[3] == src->ne[3],
                 "b200_op_sum_rows: f32 [k, ...] -> f32 [1, ...]");
    const long long rows = src->ne[1] * src->ne[2] * src->ne[3];
    if (rows == 0) return 0;
    B200_CHECK(b200_launch(sum_rows_kernel, dim3(blocks_for(rows, 4)), dim3(128), 0, (cudaStream_t) stream, glue_pdl(), t4(src), t4w(dst), rows));
    return 0;
}

int b200_op_argsort(const b200_tensor * src, const b200_tensor * dst, int order, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_argsort: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_I32 && dst->nb[0] == 4, "b200_op_argsort: f32 -> i32 indices");
    B200_REQUIRE(order == 0 || order == 1, "b200_op_argsort: order %d (0 ascending, 1 descending)", order);
    B200_REQUIRE(src->ne[0] >= 1 && src->ne[0] <= 1024, "b200_op_argsort: rows of 1..1024 elements (got %lld)", (long long) src->ne[0]);
    for (int i = 0; i < 4; ++i) B200_REQUIRE(src->ne[i] == dst->ne[i], "b200_op_argsort: shapes differ");
    const long long rows = src->ne[1] * src->ne[2] * src->ne[3];
    if (rows == 0) return 0;
    B200_REQUIRE(rows < (1ll << 31), "b200_op_argsort: too many rows");
    int npad = 1;
    while (npad < src->ne[0]) npad <<= 1;
    B200_CHECK(b200_launch(argsort_kernel, dim3((unsigned) rows), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), t4(src), t4w(dst), order, npad));
    return 0;
}

int b200_op_rope(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst, const b200_rope_params * p, b200_stream_t stream) {
    B200_REQUIRE(src && pos && dst && p && src->data && pos->data && dst->data, "b200_op_rope: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && pos->type == B200_TYPE_I32 && src->nb[0] == 4 && dst->nb[0] == 4, "b200_op_rope: f32 rows, i32 positions");
    B200_REQUIRE((p->mode == 0 || p->mode == 2) && p->ext_factor == 0.0f && p->n_dims > 0 && p->n_dims % 2 == 0 && p->n_dims <= src->ne[0] && src->ne[0] % 2 == 0,
                 "b200_op_rope: mode %d / ext_factor %g / n_dims %d not supported", p->mode, (double) p->ext_factor, p->n_dims);
    B200_REQUIRE(pos->ne[0] >= src->ne[2], "b200_op_rope: %lld positions for %lld tokens", (long long) pos->ne[0], (long long) src->ne[2]);
    if (nelem(src) == 0) return 0;
    const int ROPE_HC = rope_hc(src);
    const long long n = (src->ne[0] / 2) * ((src->ne[1] + ROPE_HC - 1) / ROPE_HC) * src->ne[2] * src->ne[3];      // one thread per (pair, chunk of heads, token)
    RopeP rp; rp.n_dims = p->n_dims; rp.mode = p->mode; rp.freq_base = p->freq_base; rp.freq_scale = p->freq_scale; rp.attn_factor = p->attn_factor;
    const float * ffp = freq_factors ? (const float *) freq_factors->data : (const float *) nullptr;
    const dim3 grid(blocks_for(n, 256));
    cudaStream_t st = (cudaStream_t) stream;
    if (nelem(src) < (1ll << 31)) {
        if (ROPE_HC == 1) B200_CHECK(b200_launch(rope_kernel<unsigned, 1>, grid, dim3(256), 0, st, glue_pdl(), t4(src), (const int *) pos->data, ffp, t4w(dst), rp));
        else              B200_CHECK(b200_launch(rope_kernel<unsigned, 8>, grid, dim3(256), 0, st, glue_pdl(), t4(src), (const int *) pos->data, ffp, t4w(dst), rp));
    } else {
        if (ROPE_HC == 1) B200_CHECK(b200_launch(rope_kernel<long long, 1>, grid, dim3(256), 0, st, glue_pdl(), t4(src), (const int *) pos->data, ffp, t4w(dst), rp));
        else              B200_CHECK(b200_launch(rope_kernel<long long, 8>, grid, dim3(256), 0, st, glue_pdl(), t4(src), (const int *) pos->data, ffp, t4w(dst), rp));
    }
    return 0;
}

int b200_op_rope_set_rows(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst, const b200_rope_params * p,
                          const b200_tensor * sr_src, const b200_tensor * sr_idx, const b200_tensor * sr_dst,
                          const b200_tensor * v_src, const b200_tensor * v_idx, const b200_tensor * v_dst, b200_stream_t stream) {
    B200_REQUIRE(src && pos && dst && p && sr_src && sr_idx && sr_dst && src->data && pos->data && dst->data && sr_idx->data && sr_dst->data, "b200_op_rope_set_rows: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && pos->type == B200_TYPE_I32 && src->nb[0] == 4, "b200_op_rope_set_rows: f32 rows, i32 positions");
    B200_REQUIRE((p->mode == 0 || p->mode == 2) && p->ext_factor == 0.0f && p->n_dims > 0 && p->n_dims % 2 == 0 && p->n_dims <= src->ne[0] && src->ne[0] % 2 == 0,
                 "b200_op_rope_set_rows: mode %d / ext_factor %g / n_dims %d not supported", p->mode, (double) p->ext_factor, p->n_dims);
    B200_REQUIRE(pos->ne[0] >= src->ne[2], "b200_op_rope_set_rows: %lld positions for %lld tokens", (long long) pos->ne[0], (long long) src->ne[2]);
    B200_REQUIRE(contiguous_f(dst, 4) && sr_src->data == dst->data && sr_src->type == B200_TYPE_F32 && contiguous_f(sr_src, 4) && nelem(sr_src) == nelem(dst),
                 "b200_op_rope_set_rows: the SET_ROWS source must be a contiguous reshape of the ROPE result");
    B200_REQUIRE((sr_dst->type == B200_TYPE_F32 || sr_dst->type == B200_TYPE_F16) && (sr_idx->type == B200_TYPE_I64 || sr_idx->type == B200_TYPE_I32), "b200_op_rope_set_rows: cache f32 / f16, indices i64 / i32");
    if (v_src) {
        B200_REQUIRE(v_idx && v_dst && v_src->data && v_idx->data && v_dst->data && v_src->type == B200_TYPE_F32 && (v_dst->type == B200_TYPE_F32 || v_dst->type == B200_TYPE_F16) &&
                     (v_idx->type == B200_TYPE_I64 || v_idx->type == B200_TYPE_I32), "b200_op_rope_set_rows: bad second SET_ROWS");
    }
    const long long nv = v_src ? nelem(v_src) : 0;
    if (nelem(src) == 0) return v_src ? b200_op_set_rows(v_src, v_idx, v_dst, stream) : 0;
    const int ROPE_HC = rope_hc(src);
    const long long n = (src->ne[0] / 2) * ((src->ne[1] + ROPE_HC - 1) / ROPE_HC) * src->ne[2] * src->ne[3];
    RopeP rp; rp.n_dims = p->n_dims; rp.mode = p->mode; rp.freq_base = p->freq_base; rp.freq_scale = p->freq_scale; rp.attn_factor = p->attn_factor;
    const unsigned kb = blocks_for(n, 256), vb = nv ? blocks_for(nv, 256) : 0;
    const float * ffp = freq_factors ? (const float *) freq_factors->data : (const float *) nullptr;
    const T4 tv = v_src ? t4(v_src) : t4(src), tvi = v_src ? t4(v_idx) : t4(sr_idx); const T4w tvc = v_src ? t4w(v_dst) : t4w(sr_dst);
#define B200_RSR(I, HC) B200_CHECK(b200_launch(rope_set_rows_kernel<I, HC>, dim3(kb + vb), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), t4(src), (const int *) pos->data, ffp, \
                                               t4w(dst), rp, t4(sr_src), t4(sr_idx), t4w(sr_dst), kb, tv, tvi, tvc))
    if (nelem(dst) < (1ll << 31) && nv < (1ll << 31)) { if (ROPE_HC == 1) B200_RSR(unsigned, 1); else B200_RSR(unsigned, 8); }
    else                                              { if (ROPE_HC == 1) B200_RSR(long long, 1); else B200_RSR(long long, 8); }
#undef B200_RSR
    return 0;
}

int b200_op_set_rows(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(src && idx && dst && src->data && idx->data && dst->data, "b200_op_set_rows: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && (dst->type == B200_TYPE_F32 || dst->type == B200_TYPE_F16) && (idx->type == B200_TYPE_I64 || idx->type == B200_TYPE_I32),
                 "b200_op_set_rows: f32 -> f32 / f16 with i64 / i32 indices only");
    const long long n = nelem(src);
    if (n == 0) return 0;
    B200_LAUNCH_IDX(set_rows_kernel, n, t4(src), t4(idx), t4w(dst));
    return 0;
}

int b200_op_cpy(const b200_tensor * src, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_cpy: null tensor");
    B200_REQUIRE((src->type == B200_TYPE_F32 || src->type == B200_TYPE_F16) && (dst->type == B200_TYPE_F32 || dst->type == B200_TYPE_F16), "b200_op_cpy: f32 / f16 only");
    B200_REQUIRE(nelem(src) == nelem(dst), "b200_op_cpy: element counts differ");
    const long long n = nelem(dst);
    if (n == 0) return 0;
    if (src->type == dst->type && contiguous_f(src, src->type == B200_TYPE_F32 ? 4 : 2) && contiguous_f(dst, dst->type == B200_TYPE_F32 ? 4 : 2)) {
        B200_CHECK(cudaMemcpyAsync(dst->data, src->data, (size_t) n * (src->type == B200_TYPE_F32 ? 4 : 2), cudaMemcpyDeviceToDevice, (cudaStream_t) stream));
        return 0;
    }
    B200_LAUNCH_IDX(cpy_kernel, n, t4(src), t4w(dst));
    return 0;
}

int b200_op_soft_max(const b200_tensor * src, const b200_tensor * mask, const b200_tensor * dst, float scale, b200_stream_t stream) {
    B200_REQUIRE(src && dst && src->data && dst->data, "b200_op_soft_max: null tensor");
    B200_REQUIRE(src->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && src->nb[0] == 4 && dst->nb[0] == 4, "b200_op_soft_max: f32 rows with unit stride only");
    B200_REQUIRE(!mask || ((mask->type == B200_TYPE_F32 || mask->type == B200_TYPE_F16) && mask->data && mask->ne[0] >= src->ne[0] && mask->ne[1] >= src->ne[1]), "b200_op_soft_max: bad mask");
    const long long rows = src->ne[1] * src->ne[2] * src->ne[3];
    if (rows == 0 || src->ne[0] == 0) return 0;
    T4 m = mask ? t4(mask) : t4(src);
    if (src->ne[0] <= 1024 && rows >= 512 && rows < (1ll << 31)) {        // enough rows to fill the machine with one warp each
        const dim3 grid(blocks_for(rows, 4));
        const T4 ts = t4(src); const T4w td = t4w(dst);
        if (src->ne[0] <= 256)      B200_CHECK(b200_launch(soft_max_warp_kernel<8>, grid, dim3(128), 0, (cudaStream_t) stream, glue_pdl(), ts, m, mask ? 1 : 0, td, scale, (unsigned) rows));
        else if (src->ne[0] <= 512) B200_CHECK(b200_launch(soft_max_warp_kernel<16>, grid, dim3(128), 0, (cudaStream_t) stream, glue_pdl(), ts, m, mask ? 1 : 0, td, scale, (unsigned) rows));
        else                        B200_CHECK(b200_launch(soft_max_warp_kernel<32>, grid, dim3(128), 0, (cudaStream_t) stream, glue_pdl(), ts, m, mask ? 1 : 0, td, scale, (unsigned) rows));
        return 0;
    }
    B200_CHECK(b200_launch(soft_max_kernel, dim3((unsigned) rows), dim3(256), 0, (cudaStream_t) stream, glue_pdl(), t4(src), m, mask ? 1 : 0, t4w(dst), scale));
    return 0;
}

int b200_op_glu(const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, int glu_op, int swapped, b200_stream_t stream) {
    B200_REQUIRE(a && dst && a->data && dst->data, "b200_op_glu: null tensor");
    B200_REQUIRE(a->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && (!b || b->type == B200_TYPE_F32) && a->nb[0] == 4 && (!b || b->nb[0] == 4), "b200_op_glu: f32 rows with unit stride only");
    B200_REQUIRE(glu_op >= 0 && glu_op <= 2, "b200_op_glu: op %d (0 reglu, 1 geglu, 2 swiglu)", glu_op);
    B200_REQUIRE(b ? (a->ne[0] == dst->ne[0] && b->ne[0] == dst->ne[0]) : a->ne[0] / 2 == dst->ne[0], "b200_op_glu: shapes");     // (an odd row length drops its last element, as the reference does)
    const long long n = nelem(dst);
    if (n == 0) return 0;
    auto al16 = [](const b200_tensor * t) { return ((uintptr_t) t->data % 16) == 0 && t->nb[1] % 16 == 0; };
    if (b && dst->ne[2] == 1 && dst->ne[3] == 1 && dst->ne[0] % 4 == 0 && dst->nb[0] == 4 && al16(a) && al16(b) && al16(dst) && n >= (1 << 16)) {
        const long long total4 = n / 4;
        const long long blocks = (total4 + 255) / 256;
        B200_CHECK(b200_launch(glu_vec4_kernel, dim3((unsigned) (blocks > 148 * 16 ? 148 * 16 : blocks)), dim3(256), 0, (cudaStream_t) stream, glue_pdl(),
                               t4(a), t4(b), t4w(dst), glu_op, swapped, total4));
        return 0;
    }
    B200_LAUNCH_IDX(glu_kernel, n, t4(a), b ? t4(b) : t4(a), b ? 1 : 0, t4w(dst), glu_op, swapped);
    return 0;
}

int b200_op_mul_mat_f(const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, b200_stream_t stream) {
    B200_REQUIRE(a && b && dst && a->data && b->data && dst->data, "b200_op_mul_mat_f: null tensor");
    B200_REQUIRE((a->type == B200_TYPE_F16 || a->type == B200_TYPE_F32) && b->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32, "b200_op_mul_mat_f: f16 / f32 x f32 -> f32 only");
    B200_REQUIRE(a->nb[0] == (a->type == B200_TYPE_F16 ? 2 : 4) && b->nb[0] == 4 && dst->nb[0] == 4, "b200_op_mul_mat_f: rows must have unit stride");
    B200_REQUIRE(a->ne[0] == b->ne[0] && dst->ne[0] == a->ne[1] && dst->ne[1] == b->ne[1] && dst->ne[2] == b->ne[2] && dst->ne[3] == b->ne[3] &&
                 a->ne[2] > 0 && a->ne[3] > 0 && b->ne[2] % a->ne[2] == 0 && b->ne[3] % a->ne[3] == 0, "b200_op_mul_mat_f: shapes");
    if (nelem(dst) == 0) return 0;
    const long long r2 = b->ne[2] / a->ne[2], r3 = b->ne[3] / a->ne[3], batch = dst->ne[2] * dst->ne[3], n = b->ne[1];
    B200_REQUIRE(batch <= 65535, "b200_op_mul_mat_f: %lld batches exceed one launch", batch);
    cudaStream_t s = (cudaStream_t) stream;
    const bool ah = a->type == B200_TYPE_F16;
    if (n <= 8 && a->ne[1] <= 64 && a->ne[0] >= 1024) {
        dim3 grid((unsigned) a->ne[1], (unsigned) batch);
#define B200_MMB(nc) case nc: if (ah) B200_CHECK(b200_launch(mm_rowblock_kernel<nc, true>, grid, dim3(256), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3)); \
                              else    B200_CHECK(b200_launch(mm_rowblock_kernel<nc, false>, grid, dim3(256), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3)); break
        switch ((int) n) { B200_MMB(1); B200_MMB(2); B200_MMB(3); B200_MMB(4); B200_MMB(5); B200_MMB(6); B200_MMB(7); B200_MMB(8); }
#undef B200_MMB
        return 0;
    }
    if (n <= 8) {
        dim3 grid(blocks_for(a->ne[1], 4), (unsigned) batch);
#define B200_MMR(nc) case nc: if (ah) B200_CHECK(b200_launch(mm_rows_kernel<nc, true>, grid, dim3(128), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3)); \
                              else    B200_CHECK(b200_launch(mm_rows_kernel<nc, false>, grid, dim3(128), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3)); break
        switch ((int) n) { B200_MMR(1); B200_MMR(2); B200_MMR(3); B200_MMR(4); B200_MMR(5); B200_MMR(6); B200_MMR(7); B200_MMR(8); }
#undef B200_MMR
        return 0;
    }
    dim3 grid(blocks_for(a->ne[1], 64), blocks_for(n, 64), (unsigned) batch);
    static const bool mma_on = [] { const char * e = getenv("B200_ATTN_MMA"); return !(e && e[0] == '0'); }();
    auto aligned16 = [](const b200_tensor * t) { return ((uintptr_t) t->data % 16) == 0 && t->nb[1] % 16 == 0 && t->nb[2] % 16 == 0 && t->nb[3] % 16 == 0; };
    if (ah && mma_on && a->ne[0] % 32 == 0 && aligned16(a) && aligned16(b)) {
        B200_CHECK(b200_launch(mm_mma_kernel, grid, dim3(128), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3));
        return 0;
    }
    if (ah) B200_CHECK(b200_launch(mm_tile_kernel<true>, grid, dim3(256), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3));
    else    B200_CHECK(b200_launch(mm_tile_kernel<false>, grid, dim3(256), 0, s, glue_pdl(), t4(a), t4(b), t4w(dst), r2, r3));
    return 0;
}

} // extern "C"

// repack.cu -- (5) load-time repack of canonical ggml blocks into the B200 plane layout, and the
// exact inverse used by get_tensor. The reference's counterpart is the CPU_REPACK buffer's
// set_tensor hook (ggml/src/ggml-cpu/repack.cpp:3141-3151); the layout itself is our own:
//
//   A matrix of m rows x nb blocks keeps its byte size. Block fields of one kind are gathered
//   into a plane [m][nb][field bytes]; planes are concatenated in this order:
//     Q4_0  : QS 16 | D 2                    (canonical block: d @0, qs @2)
//     Q8_0  : QS 32 | D 2                    (d @0, qs @2)
//     Q4_K  : identity -- 144 B blocks are already 16 B aligned (hdr 16 B + 128 B nibbles)
//     Q6_K  : QL 128 | QH 64 | SC 16 | D 2   (ql @0, qh @128, scales @192, d @208)
//     IQ4_XS: QS 128 | HDR 8                 (d, scales_h, scales_l[4] @0, qs @8)
//   With k % 256 == 0 every plane base and plane row is a multiple of 16 B, so the GEMV/GEMM
//   kernels use 128-bit loads only. Pure byte shuffle => bit-exact round trip.
#include "planes.cuh"

// one thread moves one unit (min(len,16) bytes, as 2-byte pieces: canonical blocks are only
// 2 B aligned) between canonical block b / field p and plane p.
template <bool INVERSE>
__global__ void __launch_bounds__(256) repack_kernel(const uint8_t * __restrict__ src, uint8_t * __restrict__ dst,
                                                     PlaneSpec ps, int64_t nblk, int64_t units_total) {
    for (int64_t u = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; u < units_total; u += (int64_t) gridDim.x * blockDim.x) {
        // locate plane
        int p = 0; int64_t rem = u;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i < ps.n) {
                const int64_t cnt = nblk * (ps.len[i] / ps.unit[i]);
                if (p == i && rem >= cnt) { rem -= cnt; p = i + 1; }
            }
        }
        const int len = ps.len[p], unit = ps.unit[p], upb = len / unit;
        const int64_t b = rem / upb; const int w = (int) (rem % upb);
        const int64_t canon = b * ps.block_bytes + ps.off[p] + w * unit;
        const int64_t plane = ps.base_blocks[p] * nblk + b * len + w * unit;
        if (!INVERSE) {
            const uint16_t * s = reinterpret_cast<const uint16_t *>(src + canon);
            uint16_t t[8];
            for (int i = 0; i < unit / 2; ++i) t[i] = s[i];
            if (unit == 16) *reinterpret_cast<uint4 *>(dst + plane) = *reinterpret_cast<uint4 *>(t);
            else { uint16_t * d = reinterpret_cast<uint16_t *>(dst + plane); for (int i = 0; i < unit / 2; ++i) d[i] = t[i]; }
        } else {
            uint16_t t[8];
            if (unit == 16) *reinterpret_cast<uint4 *>(t) = *reinterpret_cast<const uint4 *>(src + plane);
            else { const uint16_t * s = reinterpret_cast<const uint16_t *>(src + plane); for (int i = 0; i < unit / 2; ++i) t[i] = s[i]; }
            uint16_t * d = reinterpret_cast<uint16_t *>(dst + canon);
            for (int i = 0; i < unit / 2; ++i) d[i] = t[i];
        }
    }
}

static int run_repack(int type, int64_t m, int64_t k, const void * src, void * dst, cudaStream_t s, bool inverse) {
    B200_REQUIRE(b200_repack_supported(type, m, k), "repack: unsupported type %d / shape m=%lld k=%lld", type, (long long) m, (long long) k);
    B200_REQUIRE(src && dst && src != dst, "repack: needs distinct non-null src/dst");
    B200_REQUIRE((((uintptr_t) src | (uintptr_t) dst) & 15) == 0, "repack: pointers must be 16 B aligned");
    PlaneSpec ps;
    plane_spec(type, &ps);
    const int64_t nblk = m * (k / b200_block_elems(type));
    if (ps.n == 1) { // identity layout
        B200_CHECK(cudaMemcpyAsync(dst, src, (size_t) nblk * ps.block_bytes, cudaMemcpyDeviceToDevice, s));
        return 0;
    }
    int64_t units = 0;
    for (int i = 0; i < ps.n; ++i) units += nblk * (ps.len[i] / ps.unit[i]);
    const int64_t blocks = (units + 255) / 256;
    dim3 grid((unsigned) (blocks < 148 * 32 ? blocks : 148 * 32));
    if (!inverse) B200_CHECK(b200_launch(repack_kernel<false>, grid, dim3(256), 0, s, false, (const uint8_t *) src, (uint8_t *) dst, ps, nblk, units));
    else          B200_CHECK(b200_launch(repack_kernel<true>, grid, dim3(256), 0, s, false, (const uint8_t *) src, (uint8_t *) dst, ps, nblk, units));
    return 0;
}

extern "C" {

int b200_repack_supported(int type, int64_t m, int64_t k) {
    return b200_type_is_weight(type) && m > 0 && k > 0 && (k % 256) == 0;
}
int b200_repack(int type, int64_t m, int64_t k, const void * canonical, void * planes, b200_stream_t stream) {
    return run_repack(type, m, k, canonical, planes, (cudaStream_t) stream, false);
}
int b200_unrepack(int type, int64_t m, int64_t k, const void * planes, void * canonical, b200_stream_t stream) {
    return run_repack(type, m, k, planes, canonical, (cudaStream_t) stream, true);
}

} // extern "C"

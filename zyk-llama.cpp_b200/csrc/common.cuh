// common.cuh -- shared host/device helpers for the B200 quantised-matmul kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200_mm.h"

// ---- error plumbing: no exceptions cross the C ABI, every entry point returns a status ---------
void b200_set_error(const char * fmt, ...);
void b200_count_launch(void);

#define B200_CHECK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t err__ = (expr);                                                               \
        if (err__ != cudaSuccess) {                                                               \
            b200_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(err__)); \
            return -1;                                                                            \
        }                                                                                         \
    } while (0)

#define B200_REQUIRE(cond, ...)                                                                   \
    do {                                                                                          \
        if (!(cond)) {                                                                            \
            b200_set_error(__VA_ARGS__);                                                          \
            return -2;                                                                            \
        }                                                                                         \
    } while (0)

// every kernel launch of this library goes through here (launch counter + PDL attribute)
template <typename... KArgs, typename... Args>
static inline cudaError_t b200_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                      cudaStream_t stream, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    b200_count_launch();
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

#ifdef __CUDACC__

static constexpr int WARP = 32;

// 128-bit streaming load: weights are read exactly once per launch, keep them out of L1
__device__ __forceinline__ uint4 ld_stream16(const void * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ld_stream8(const void * p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ uint16_t ld_stream2(const void * p) {
    uint16_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return r;
}
// 128-bit cached load (activations: re-read by every warp of the SM, want them in L1)
__device__ __forceinline__ uint4 ld_cached16(const void * p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }

__device__ __forceinline__ void prefetch_l2(const void * p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// programmatic dependent launch (PDL): let the next kernel start, wait for the previous one
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float half_bits_to_float(uint16_t h) { return __half2float(__ushort_as_half(h)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// byte `i` (0..15, runtime) of a 16-byte register quad
__device__ __forceinline__ uint32_t byte_of(const uint4 & v, int i) {
    const uint32_t w = (i < 8) ? ((i < 4) ? v.x : v.y) : ((i < 12) ? v.z : v.w);
    return (w >> (8 * (i & 3))) & 0xffu;
}

#endif // __CUDACC__

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
// "first launch of this kernel on the CURRENT device": function attributes (dynamic shared memory size) are per device, and one process may
// drive several GPUs (the plugin's row-split buffer type). `mask` is a function-local static of the caller.
static inline bool b200_first_on_device(unsigned & mask) {
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned bit = 1u << (dev & 31);
    if (mask & bit) return false;
    mask |= bit;
    return true;
}

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
// kernel variants: the row-split exchange costs registers in the streaming loops (Q6_K register GEMV: 72 -> 80, one CTA less
// per SM), so it is compiled in only where it is used
enum { KV_PLAIN = 0, KV_PEER = 1 };
// 128-bit cached load (activations: re-read by every warp of the SM, want them in L1)
__device__ __forceinline__ uint4 ld_cached16(const void * p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }

__device__ __forceinline__ void prefetch_l2(const void * p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// programmatic dependent launch (PDL): let the next kernel start, wait for the previous one
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }


// ---- fused row-split exchange (include/b200_mm.h: b200_row_gather) ---------------------------------------
struct PeerGather {
    int world, rank;                    // world <= 1: off
    long long delta[8];
    unsigned long long * sync;          // [0] arrivals (peers add), [1] CTAs of the signalling kernel that are done
    const unsigned long long * step;
    unsigned long long seq, calls;
    int wait, signal;                   // this launch waits for the earlier calls / announces the end of this call
};

// one thread: block until every rank has finished every call before (step, seq). Traps after ~10 s instead of hanging the GPU.
__device__ __forceinline__ void peer_wait(const PeerGather & p) {
    const unsigned long long want = ((*p.step - 1ull) * p.calls + p.seq) * (unsigned long long) p.world;
    unsigned long long t0 = 0, v;
    for (unsigned spin = 0;; ++spin) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p.sync) : "memory");
        if (v >= want) break;
        __nanosleep(40);
        if ((spin & 1023u) == 1023u) {
            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t; else if (t - t0 > 10000000000ull) __trap();
        }
    }
}
// the value goes to the same location of every rank's arena (delta[rank] == 0)
__device__ __forceinline__ void peer_store(const PeerGather & p, float * addr, float v) {
    for (int g = 0; g < p.world; ++g) *reinterpret_cast<float *>(reinterpret_cast<char *>(addr) + p.delta[g]) = v;
}
// ONE thread per CTA, after a CTA barrier that follows all of the CTA's peer stores (barrier + this thread's system fence order them):
// the last CTA of the grid announces the end of the call to every rank
__device__ __forceinline__ void peer_signal(const PeerGather & p) {
    // release at GPU scope on the local CTA counter (cheap), one system-scope release by the last CTA: release sequences are
    // cumulative, so every CTA's peer stores are ordered before the arrival the other ranks acquire at system scope
    const unsigned long long n = (unsigned long long) gridDim.x * gridDim.y * gridDim.z;
    unsigned long long old;
    asm volatile("atom.acq_rel.gpu.global.add.u64 %0, [%1], 1;" : "=l"(old) : "l"(p.sync + 1) : "memory");
    if (old == n - 1) {
        p.sync[1] = 0;                  // the next signalling kernel on this stream starts from zero (it runs after this one)
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        for (int g = 0; g < p.world; ++g) {
            unsigned long long * dst = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(p.sync) + p.delta[g]);
            asm volatile("red.relaxed.sys.global.add.u64 [%0], 1;" :: "l"(dst) : "memory");
        }
    }
}

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

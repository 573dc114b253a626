// api.cu -- device / memory / stream / event / graph plumbing of the C ABI (include/b200_mm.h).
// These are what the ggml plugin's device, buffer-type, buffer and backend vtables call
// (ggml/src/ggml-backend-impl.h:17-182 in the reference).
#include "common.cuh"

#include <atomic>
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void b200_set_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void b200_count_launch(void) { g_launches.fetch_add(1, std::memory_order_relaxed); }

extern "C" {

const char * b200_last_error(void) { return g_err; }
const char * b200_version(void) { return "b200mm 0.1 (sm_100a)"; }
uint64_t b200_launch_count(void) { return g_launches.load(); }

int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int b200_device_info(int device, char * name, size_t name_len, size_t * free_bytes, size_t * total_bytes,
                     int * sm_count, int * cc_major, int * cc_minor) {
    cudaDeviceProp p;
    B200_CHECK(cudaGetDeviceProperties(&p, device));
    if (name && name_len) { strncpy(name, p.name, name_len - 1); name[name_len - 1] = 0; }
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    if (free_bytes || total_bytes) {
        int cur = 0;
        B200_CHECK(cudaGetDevice(&cur));
        B200_CHECK(cudaSetDevice(device));
        size_t f = 0, t = 0;
        B200_CHECK(cudaMemGetInfo(&f, &t));
        B200_CHECK(cudaSetDevice(cur));
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
    }
    return 0;
}

int b200_set_device(int device) { B200_CHECK(cudaSetDevice(device)); return 0; }

void * b200_malloc(int device, size_t bytes) {
    void * p = nullptr;
    if (cudaSetDevice(device) != cudaSuccess) { b200_set_error("cudaSetDevice(%d) failed", device); cudaGetLastError(); return nullptr; }
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { b200_set_error("cudaMalloc(%zu) on device %d: %s", bytes, device, cudaGetErrorString(e)); cudaGetLastError(); return nullptr; }
    return p;
}
int b200_free(int device, void * ptr) {
    if (!ptr) return 0;
    B200_CHECK(cudaSetDevice(device));
    B200_CHECK(cudaFree(ptr));
    return 0;
}
void * b200_malloc_host(size_t bytes) {
    void * p = nullptr;
    cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { b200_set_error("cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e)); cudaGetLastError(); return nullptr; }
    return p;
}
int b200_free_host(void * ptr) { if (ptr) B200_CHECK(cudaFreeHost(ptr)); return 0; }

b200_stream_t b200_stream_create(int device) {
    cudaStream_t s = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
        b200_set_error("stream create on device %d failed: %s", device, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return (b200_stream_t) s;
}
int b200_stream_destroy(b200_stream_t stream) { if (stream) B200_CHECK(cudaStreamDestroy((cudaStream_t) stream)); return 0; }
int b200_stream_synchronize(b200_stream_t stream) { B200_CHECK(cudaStreamSynchronize((cudaStream_t) stream)); return 0; }
int b200_device_synchronize(int device) { B200_CHECK(cudaSetDevice(device)); B200_CHECK(cudaDeviceSynchronize()); return 0; }

int b200_memcpy_h2d(void * dst, const void * src, size_t n, b200_stream_t s) {
    if (n) B200_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, (cudaStream_t) s));
    return 0;
}
int b200_memcpy_d2h(void * dst, const void * src, size_t n, b200_stream_t s) {
    if (n) B200_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, (cudaStream_t) s));
    return 0;
}
int b200_memcpy_d2d(void * dst, const void * src, size_t n, b200_stream_t s) {
    if (n) B200_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToDevice, (cudaStream_t) s));
    return 0;
}
int b200_memcpy_peer(void * dst, int dst_device, const void * src, int src_device, size_t n, b200_stream_t s) {
    if (n) B200_CHECK(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, n, (cudaStream_t) s));
    return 0;
}
int b200_memcpy2d(void * dst, size_t dpitch, const void * src, size_t spitch, size_t width, size_t height, b200_stream_t s) {
    if (width && height) B200_CHECK(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyDefault, (cudaStream_t) s));
    return 0;
}
int b200_enable_peer_access(int device, int peer) {
    if (device == peer) return 0;
    int can = 0;
    B200_CHECK(cudaDeviceCanAccessPeer(&can, device, peer));
    if (!can) { b200_set_error("device %d cannot access the memory of device %d", device, peer); return -1; }
    int cur = 0;
    B200_CHECK(cudaGetDevice(&cur));
    B200_CHECK(cudaSetDevice(device));
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) (void) cudaGetLastError();
    else if (e != cudaSuccess) { cudaSetDevice(cur); b200_set_error("cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e)); return -1; }
    B200_CHECK(cudaSetDevice(cur));
    return 0;
}
int b200_memset(void * dst, int value, size_t n, b200_stream_t s) {
    if (n) B200_CHECK(cudaMemsetAsync(dst, value, n, (cudaStream_t) s));
    return 0;
}

b200_event_t b200_event_create(int device) {
    cudaEvent_t e = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaEventCreate(&e) != cudaSuccess) {
        b200_set_error("event create failed: %s", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return (b200_event_t) e;
}
int b200_event_destroy(b200_event_t ev) { if (ev) B200_CHECK(cudaEventDestroy((cudaEvent_t) ev)); return 0; }
int b200_event_record(b200_event_t ev, b200_stream_t s) { B200_CHECK(cudaEventRecord((cudaEvent_t) ev, (cudaStream_t) s)); return 0; }
int b200_event_synchronize(b200_event_t ev) { B200_CHECK(cudaEventSynchronize((cudaEvent_t) ev)); return 0; }
int b200_stream_wait_event(b200_stream_t s, b200_event_t ev) { B200_CHECK(cudaStreamWaitEvent((cudaStream_t) s, (cudaEvent_t) ev, 0)); return 0; }
int b200_event_elapsed_ms(b200_event_t a, b200_event_t b, float * ms) { B200_CHECK(cudaEventElapsedTime(ms, (cudaEvent_t) a, (cudaEvent_t) b)); return 0; }

// ---- CUDA graphs --------------------------------------------------------------------------------
int b200_graph_begin(b200_stream_t stream) {
    B200_REQUIRE(stream != nullptr, "b200_graph_begin: capture needs a non-default stream");
    B200_CHECK(cudaStreamBeginCapture((cudaStream_t) stream, cudaStreamCaptureModeThreadLocal));
    return 0;
}
b200_graph_t b200_graph_end(b200_stream_t stream) {
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture((cudaStream_t) stream, &g);
    if (e != cudaSuccess || !g) { b200_set_error("cudaStreamEndCapture: %s", cudaGetErrorString(e)); cudaGetLastError(); return nullptr; }
    cudaGraphExec_t x = nullptr;
    e = cudaGraphInstantiate(&x, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) { b200_set_error("cudaGraphInstantiate: %s", cudaGetErrorString(e)); cudaGetLastError(); return nullptr; }
    return (b200_graph_t) x;
}
int b200_graph_launch(b200_graph_t graph, b200_stream_t stream) {
    B200_CHECK(cudaGraphLaunch((cudaGraphExec_t) graph, (cudaStream_t) stream));
    return 0;
}
int b200_graph_destroy(b200_graph_t graph) { if (graph) B200_CHECK(cudaGraphExecDestroy((cudaGraphExec_t) graph)); return 0; }

// ---- formats (ggml/src/ggml-common.h:170-431) ---------------------------------------------------
int b200_type_is_weight(int t) {
    return t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K || t == B200_TYPE_IQ4_XS;
}
int b200_block_elems(int t) {
    switch (t) {
        case B200_TYPE_F32: case B200_TYPE_I32: return 1;
        case B200_TYPE_Q4_0: case B200_TYPE_Q8_0: case B200_TYPE_Q8_1: return 32;
        case B200_TYPE_Q4_K: case B200_TYPE_Q5_K: case B200_TYPE_Q6_K: case B200_TYPE_Q8_K: case B200_TYPE_IQ4_XS: return 256;
    }
    return 0;
}
size_t b200_block_bytes(int t) {
    switch (t) {
        case B200_TYPE_F32: case B200_TYPE_I32: return 4;
        case B200_TYPE_Q4_0: return 18;  case B200_TYPE_Q8_0: return 34;  case B200_TYPE_Q8_1: return 36;
        case B200_TYPE_Q4_K: return 144; case B200_TYPE_Q5_K: return 176; case B200_TYPE_Q6_K: return 210; case B200_TYPE_Q8_K: return 292;
        case B200_TYPE_IQ4_XS: return 136;
    }
    return 0;
}
size_t b200_row_size(int t, int64_t k) {
    const int be = b200_block_elems(t);
    return be ? (size_t)(k / be) * b200_block_bytes(t) : 0;
}
int b200_act_type(int t) {
    switch (t) {
        case B200_TYPE_Q4_0: case B200_TYPE_Q8_0: return B200_TYPE_Q8_0;
        case B200_TYPE_Q4_K: case B200_TYPE_Q5_K: case B200_TYPE_Q6_K: case B200_TYPE_IQ4_XS: return B200_TYPE_Q8_K;
    }
    return -1;
}

// rows of an m-row matrix owned by `rank`: equal shares rounded to 128-row GEMM tiles, the last
// rank takes the remainder (cf. get_row_split, ggml-cuda/ggml-cuda.cu:790-803)
int b200_row_split(int64_t m, int world, int rank, int64_t * lo, int64_t * hi) {
    B200_REQUIRE(world > 0 && rank >= 0 && rank < world && lo && hi, "b200_row_split: bad arguments");
    const int64_t tile = 128;
    auto bound = [&](int r) -> int64_t {
        if (r >= world) return m;
        int64_t b = (m * r / world) / tile * tile;
        return b > m ? m : b;
    };
    *lo = bound(rank);
    *hi = bound(rank + 1);
    return 0;
}

} // extern "C"

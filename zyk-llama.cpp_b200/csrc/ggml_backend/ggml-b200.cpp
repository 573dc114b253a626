// ggml-b200.cpp -- the ggml backend plugin (libggml-b200.so): reg / device / buffer_type / buffer /
// backend vtables of /root/reference/ggml/src/ggml-backend-impl.h:11-251 implemented on top of the
// C ABI in include/b200_mm.h. Plain C++ (no CUDA headers): every device action is a b200_* call.
//
// What the device claims (supports_op): the layout no-ops (NONE / VIEW / RESHAPE / PERMUTE /
// TRANSPOSE) and GGML_OP_MUL_MAT / GGML_OP_MUL_MAT_ID with Q4_0 / Q8_0 / Q4_K / Q6_K / IQ4_XS
// weights (also Q5_K), f32 activations and f32 results, plus the glue ops of a llama-arch graph (csrc/glue_ops.cu). Everything else
// stays with the caller's other backends (ggml_backend_sched splits on supports_op, ggml-backend.cpp:787-842).
//
// Several GPUs in one process: every B200 is a device of the registry; proc "ggml_backend_split_buffer_type" hands out the row-split
// weight buffer type behind `-sm row` (see the split_* functions).
//
// Load-time repack: a whole-tensor set_tensor of a contiguous, non-view weight tensor with
// ne0 % 256 == 0 is stored in the B200 plane layout (b200_repack); get_tensor hands canonical
// ggml bytes back (b200_unrepack). The reference hook this replaces is
// ggml_backend_cpu_repack_buffer_set_tensor (ggml-cpu/repack.cpp:3141-3151). A tensor is demoted
// back to canonical bytes the moment something needs them in place (partial write, memset, a
// differently-shaped view used as MUL_MAT src0).
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"

#include "b200_mm.h"
#include "ggml_b200.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#define B200_MAX_DEVICES 16
#define B200_ALIGNMENT   256

namespace {

// ---------------------------------------------------------------------------------------------------
// contexts
// ---------------------------------------------------------------------------------------------------
struct device_ctx {
    int         index = 0;
    std::string name;         // "B200_0"
    std::string description;  // "NVIDIA B200"
    ggml_backend_buffer_type buft{};
    ggml_backend_buffer_type host_buft{};
    std::string buft_name, host_buft_name;
};

struct plane_entry {   // a tensor currently stored in the plane layout
    int     type;
    int64_t m, k, nmat;
    size_t  nbytes;
};

struct buffer_ctx {
    int    device = 0;
    void * base = nullptr;
    size_t size = 0;
    std::mutex mu;
    std::unordered_map<const void *, plane_entry> planes;  // key: tensor->data
    void * staging = nullptr;
    size_t staging_size = 0;

    void * stage(size_t n) {   // grow-only scratch for repack / unrepack (load time, get_tensor)
        if (n > staging_size) {
            if (staging) { b200_device_synchronize(device); b200_free(device, staging); }
            staging = b200_malloc(device, n);
            staging_size = staging ? n : 0;
        }
        return staging;
    }
    // the scratch of a big tensor (431 MB for an 8B output matrix, ~1 GB at 70B) is not kept for the life of the buffer
    void trim() {
        if (staging && staging_size > (size_t(32) << 20)) { b200_device_synchronize(device); b200_free(device, staging); staging = nullptr; staging_size = 0; }
    }
};

struct backend_ctx {
    int           device = 0;
    std::string   name;
    b200_stream_t stream = nullptr;
    void *        workspace = nullptr;
    size_t        workspace_size = 0;
    // decode chains built from runs of MUL_MAT groups (b200_chain_*), keyed by the exact argument list: pointers, shapes and
    // strides are baked into a chain like into a graph plan (ggml_backend_graph_plan_create, ggml-backend-impl.h:103-110)
    struct cached_chain { std::vector<b200_mul_mat_args> args; std::vector<int32_t> sizes; b200_chain_t chain; };
    std::vector<cached_chain> chains;
    // whole-graph CUDA graphs (counterpart of GGML_CUDA_USE_GRAPHS, ggml-cuda.cu:2930-3100): a cgraph whose nodes (ops, shapes, strides,
    // pointers, op_params) equal those of an earlier call is replayed as ONE cudaGraphLaunch; captured the second time a key is seen
    struct cached_graph { std::vector<uint64_t> key; b200_graph_t exec = nullptr; int seen = 0; bool bad = false; uint64_t stamp = 0; };
    std::vector<cached_graph> graphs;
    uint64_t graph_stamp = 0;
    bool capturing = false;        // a stream capture is open: nothing may allocate or synchronise
    // row-split weights (ggml_backend_split_buffer_type): one helper stream, workspace and copy of X per OTHER device that holds a slice
    struct aux_device {
        int           device = -1;
        b200_stream_t stream = nullptr;
        b200_event_t  done = nullptr;
        void *        workspace = nullptr; size_t workspace_size = 0;
        void *        x = nullptr;         size_t x_size = 0;
    };
    std::vector<aux_device> aux;
    b200_event_t fork = nullptr;   // recorded on `stream` when X is ready, awaited by the helper streams
};

// ---- row-split weight tensors (one slice of rows per device of this process) -----------------------------------------------------------
struct split_slice { void * data = nullptr; int64_t lo = 0, hi = 0; bool planes = false; };
struct split_tensor {
    int type = 0; int64_t m = 0, k = 0;
    int n_dev = 0;
    split_slice s[B200_MAX_DEVICES];     // by position in the registry's device list
};
struct split_buft_ctx {
    int main = 0, n_dev = 0;
    int cuda_index[B200_MAX_DEVICES];
    float cum[B200_MAX_DEVICES + 1];     // cumulative row fractions: device i owns [cum[i], cum[i + 1]) of the rows
    std::string name;
    ggml_backend_buffer_type buft{};
};
struct split_buffer_ctx {
    split_buft_ctx * bt = nullptr;
    std::vector<split_tensor *> tensors;
    void * staging[B200_MAX_DEVICES] = {};
    size_t staging_size[B200_MAX_DEVICES] = {};
};
const char * split_buft_get_name(ggml_backend_buffer_type_t buft);
// the slices of a MUL_MAT weight that lives in a split buffer, else nullptr
const split_tensor * split_of(const ggml_tensor * w) {
    if (!w || !w->buffer || w->buffer->buft->iface.get_name != split_buft_get_name) return nullptr;
    return (const split_tensor *) w->extra;
}
bool in_split_buffer(const ggml_tensor * t) {
    const ggml_tensor * b = t && t->view_src ? t->view_src : t;
    return b && b->buffer && b->buffer->buft->iface.get_name == split_buft_get_name;
}

// GGML_B200_PDL=0: launch the matmul kernels without programmatic dependent launch (their weight prefetch then starts only when the
// previous kernel of the stream has drained)
int32_t mm_flags() { static const int32_t f = [] { const char * e = getenv("GGML_B200_PDL"); return (e && e[0] == '0') ? 0 : (int32_t) B200_MM_PDL; }(); return f; }

bool repack_enabled() {
    static const bool on = [] { const char * e = getenv("GGML_B200_REPACK"); return !(e && e[0] == '0'); }();
    return on;
}

device_ctx * dev_of(ggml_backend_buffer_type_t buft) { return (device_ctx *) buft->device->context; }

bool weight_type(enum ggml_type t) { return b200_type_is_weight((int) t) != 0; }

// ---------------------------------------------------------------------------------------------------
// plane-layout bookkeeping
// ---------------------------------------------------------------------------------------------------
bool repack_candidate(const ggml_tensor * t) {
    return repack_enabled() && weight_type(t->type) && t->view_src == nullptr && ggml_is_contiguous(t) &&
           t->ne[0] % 256 == 0 && t->ne[1] > 0 && (((uintptr_t) t->data) & 15) == 0 &&
           b200_repack_supported((int) t->type, t->ne[1], t->ne[0]);
}

// canonical -> planes, in place through the staging buffer. `canon_dev` holds the canonical bytes.
bool planes_store(buffer_ctx * bc, const ggml_tensor * t, const void * canon_dev) {
    const int64_t m = t->ne[1], k = t->ne[0], nmat = t->ne[2] * t->ne[3];
    const size_t mat = (size_t) m * b200_row_size((int) t->type, k);
    for (int64_t i = 0; i < nmat; ++i) {
        if (b200_repack((int) t->type, m, k, (const char *) canon_dev + i * mat, (char *) t->data + i * mat, nullptr) != 0) {
            GGML_LOG_ERROR("ggml-b200: repack of %s failed: %s\n", t->name, b200_last_error());
            return false;
        }
    }
    bc->planes[t->data] = plane_entry{(int) t->type, m, k, nmat, mat * (size_t) nmat};
    return true;
}

// planes -> canonical bytes in `dst_dev` (device memory, may NOT alias the tensor)
bool planes_load(const plane_entry & e, const void * data, void * dst_dev) {
    const size_t mat = e.nbytes / (size_t) e.nmat;
    for (int64_t i = 0; i < e.nmat; ++i) {
        if (b200_unrepack(e.type, e.m, e.k, (const char *) data + i * mat, (char *) dst_dev + i * mat, nullptr) != 0) {
            GGML_LOG_ERROR("ggml-b200: unrepack failed: %s\n", b200_last_error());
            return false;
        }
    }
    return true;
}

// make the bytes at `data` canonical again (caller holds bc->mu)
void planes_demote_locked(buffer_ctx * bc, const void * data) {
    auto it = bc->planes.find(data);
    if (it == bc->planes.end()) return;
    const plane_entry e = it->second;
    b200_device_synchronize(bc->device);   // kernels on backend streams may still be reading the planes
    void * tmp = bc->stage(e.nbytes);
    GGML_ASSERT(tmp && "ggml-b200: out of device memory while demoting a repacked tensor");
    planes_load(e, data, tmp);
    b200_memcpy_d2d((void *) data, tmp, e.nbytes, nullptr);
    b200_stream_synchronize(nullptr);
    bc->planes.erase(it);
    bc->trim();
    static bool told = false;
    if (!told) { told = true; GGML_LOG_INFO("ggml-b200: a repacked weight tensor (%lld x %lld, type %d) went back to the GGUF byte layout (partial write, memset or a differently shaped view); its matmuls take the canonical-layout kernels from now on\n", (long long) e.m, (long long) e.k, e.type); }
}

// the buffer context that owns a tensor allocated in one of OUR device buffers, else nullptr
const char * buft_get_name(ggml_backend_buffer_type_t buft);
buffer_ctx * owner_ctx(const ggml_tensor * t) {
    ggml_backend_buffer_t buf = t->view_src ? t->view_src->buffer : t->buffer;
    if (!buf || buf->buft->iface.get_name != buft_get_name) return nullptr;
    return (buffer_ctx *) buf->context;
}

// Is MUL_MAT src0 `w` stored as planes right now? Demotes when a view needs canonical bytes.
int weight_layout(const ggml_tensor * w) {
    buffer_ctx * bc = owner_ctx(w);
    if (!bc) return B200_LAYOUT_CANONICAL;
    const ggml_tensor * base = w->view_src ? w->view_src : w;
    std::lock_guard<std::mutex> lock(bc->mu);
    auto it = bc->planes.find(base->data);
    if (it == bc->planes.end()) return B200_LAYOUT_CANONICAL;
    const plane_entry & e = it->second;
    const bool same = w->data == base->data && w->ne[0] == e.k && w->ne[1] == e.m && w->ne[2] * w->ne[3] == e.nmat &&
                      (int) w->type == e.type && ggml_is_contiguous(w);
    if (same) return B200_LAYOUT_PLANES;
    planes_demote_locked(bc, base->data);
    return B200_LAYOUT_CANONICAL;
}

// ---------------------------------------------------------------------------------------------------
// buffer
// ---------------------------------------------------------------------------------------------------
void buffer_free(ggml_backend_buffer_t buffer) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    if (bc->staging) b200_free(bc->device, bc->staging);
    b200_free(bc->device, bc->base);
    delete bc;
}

void * buffer_get_base(ggml_backend_buffer_t buffer) { return ((buffer_ctx *) buffer->context)->base; }

enum ggml_status buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    // (re)initialising a non-view tensor at an address forgets any plane state recorded there
    if (tensor->view_src == nullptr && tensor->data) {
        buffer_ctx * bc = (buffer_ctx *) buffer->context;
        std::lock_guard<std::mutex> lock(bc->mu);
        bc->planes.erase(tensor->data);
    }
    return GGML_STATUS_SUCCESS;
}

// The buffer-level set / get / memset are SYNCHRONOUS calls of the interface and run on the device's default stream, while graph_compute is
// asynchronous on the backend's own (non-blocking) stream -- and ggml_backend_tensor_get_async falls back to the synchronous get WITHOUT
// synchronising the backend when a backend has no get_tensor_async (ggml-backend.cpp:267-278; llama_context::decode reads the logits that
// way). Every one of them therefore first waits for all work queued on the device; the async variants below stay in stream order instead.
void buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    if (size == 0) return;
    b200_device_synchronize(bc->device);
    const ggml_tensor * base = tensor->view_src ? tensor->view_src : tensor;
    std::lock_guard<std::mutex> lock(bc->mu);
    const bool whole = offset == 0 && size == ggml_nbytes(tensor);
    if (whole && repack_candidate(tensor)) {
        void * tmp = bc->stage(size);
        if (tmp) {
            int rc = b200_memcpy_h2d(tmp, data, size, nullptr);
            if (rc == 0 && planes_store(bc, tensor, tmp)) {
                b200_stream_synchronize(nullptr);
                bc->trim();
                return;
            }
        }
        bc->planes.erase(tensor->data);   // fall through to the canonical write
    } else {
        planes_demote_locked(bc, base->data);
    }
    if (b200_memcpy_h2d((char *) tensor->data + offset, data, size, nullptr) != 0 || b200_stream_synchronize(nullptr) != 0)
        GGML_ABORT("ggml-b200: set_tensor(%s) failed: %s", tensor->name, b200_last_error());
}

void buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    if (size == 0) return;
    b200_device_synchronize(bc->device);
    const ggml_tensor * base = tensor->view_src ? tensor->view_src : tensor;
    std::lock_guard<std::mutex> lock(bc->mu);
    auto it = bc->planes.find(base->data);
    const char * src = (const char *) tensor->data + offset;
    if (it != bc->planes.end()) {
        // hand canonical bytes back: inverse repack of the owning tensor into scratch, copy the range out
        void * tmp = bc->stage(it->second.nbytes);
        GGML_ASSERT(tmp && "ggml-b200: out of device memory in get_tensor");
        planes_load(it->second, base->data, tmp);
        src = (const char *) tmp + ((const char *) tensor->data - (const char *) base->data) + offset;
    }
    if (b200_memcpy_d2h(data, src, size, nullptr) != 0 || b200_stream_synchronize(nullptr) != 0)
        GGML_ABORT("ggml-b200: get_tensor(%s) failed: %s", tensor->name, b200_last_error());
    bc->trim();
}

void buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    if (size == 0) return;
    b200_device_synchronize(bc->device);
    const ggml_tensor * base = tensor->view_src ? tensor->view_src : tensor;
    std::lock_guard<std::mutex> lock(bc->mu);
    planes_demote_locked(bc, base->data);
    b200_memset((char *) tensor->data + offset, value, size, nullptr);
    b200_stream_synchronize(nullptr);
}

bool buffer_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    b200_device_synchronize(((buffer_ctx *) buffer->context)->device);
    // device-to-device for plain (never repacked) data of our own buffers; everything else takes the
    // generic get_tensor + set_tensor route (ggml-backend.cpp:389-410), which keeps layouts consistent
    buffer_ctx * dc = (buffer_ctx *) buffer->context;
    buffer_ctx * sc = owner_ctx(src);
    if (!sc || ggml_is_quantized(src->type) || !ggml_is_contiguous(src) || !ggml_is_contiguous(dst) || ggml_nbytes(src) != ggml_nbytes(dst))
        return false;
    int rc = sc->device == dc->device
                 ? b200_memcpy_d2d(dst->data, src->data, ggml_nbytes(src), nullptr)
                 : b200_memcpy_peer(dst->data, dc->device, src->data, sc->device, ggml_nbytes(src), nullptr);
    if (rc != 0) return false;
    b200_stream_synchronize(nullptr);
    return true;
}

void buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    b200_device_synchronize(bc->device);
    std::lock_guard<std::mutex> lock(bc->mu);
    bc->planes.clear();
    b200_memset(bc->base, value, bc->size, nullptr);
    b200_stream_synchronize(nullptr);
}

void buffer_reset(ggml_backend_buffer_t buffer) {
    buffer_ctx * bc = (buffer_ctx *) buffer->context;
    std::lock_guard<std::mutex> lock(bc->mu);
    bc->planes.clear();
}

const ggml_backend_buffer_i buffer_iface = {
    /* .free_buffer   = */ buffer_free,
    /* .get_base      = */ buffer_get_base,
    /* .init_tensor   = */ buffer_init_tensor,
    /* .memset_tensor = */ buffer_memset_tensor,
    /* .set_tensor    = */ buffer_set_tensor,
    /* .get_tensor    = */ buffer_get_tensor,
    /* .cpy_tensor    = */ buffer_cpy_tensor,
    /* .clear         = */ buffer_clear,
    /* .reset         = */ buffer_reset,
};

// ---------------------------------------------------------------------------------------------------
// buffer types: device memory (default) and pinned host memory
// ---------------------------------------------------------------------------------------------------
const char * buft_get_name(ggml_backend_buffer_type_t buft) { return dev_of(buft)->buft_name.c_str(); }

ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    device_ctx * dc = dev_of(buft);
    void * p = b200_malloc(dc->index, size);
    if (!p) {
        GGML_LOG_ERROR("ggml-b200: allocating %.2f MiB on device %d failed: %s\n", size / 1048576.0, dc->index, b200_last_error());
        return nullptr;
    }
    buffer_ctx * bc = new buffer_ctx;
    bc->device = dc->index;
    bc->base = p;
    bc->size = size;
    return ggml_backend_buffer_init(buft, buffer_iface, bc, size);
}

size_t buft_get_alignment(ggml_backend_buffer_type_t) { return B200_ALIGNMENT; }

size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    // quantised rows are read with 128-bit loads: keep 512 B of slack behind every quantised tensor
    size_t n = ggml_nbytes(tensor);
    if (ggml_is_quantized(tensor->type)) n += 512;
    return n;
}

bool buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};

// pinned host memory: a CPU buffer whose storage came from b200_malloc_host (faster H2D/D2H for the
// scheduler's input/output copies; ggml_backend_dev_host_buffer_type, ggml-backend.h:181)
const char * host_buft_get_name(ggml_backend_buffer_type_t buft) { return dev_of(buft)->host_buft_name.c_str(); }

void host_buffer_free(ggml_backend_buffer_t buffer) { b200_free_host(buffer->context); }

ggml_backend_buffer_t host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * p = b200_malloc_host(size);
    if (!p) {   // pinned memory exhausted: let the caller fall back to a pageable CPU buffer
        GGML_LOG_WARN("ggml-b200: pinned host allocation of %.2f MiB failed, using pageable memory\n", size / 1048576.0);
        return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);
    }
    ggml_backend_buffer_t buf = ggml_backend_cpu_buffer_from_ptr(p, size);
    buf->buft = buft;
    buf->iface.free_buffer = host_buffer_free;
    return buf;
}

size_t host_buft_get_alignment(ggml_backend_buffer_type_t) { return 64; }
bool   host_buft_is_host(ggml_backend_buffer_type_t) { return true; }

const ggml_backend_buffer_type_i host_buft_iface = {
    /* .get_name       = */ host_buft_get_name,
    /* .alloc_buffer   = */ host_buft_alloc_buffer,
    /* .get_alignment  = */ host_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ nullptr,
    /* .is_host        = */ host_buft_is_host,
};

// ---------------------------------------------------------------------------------------------------
// row-split buffer type: proc "ggml_backend_split_buffer_type" (ggml-backend.h:199-200; caller src/llama-model.cpp:388-410, `-sm row`).
// The reference's counterpart is ggml_backend_cuda_split_buffer_type (ggml-cuda.cu:811-1107). A weight matrix placed here is cut into
// one run of rows per device of this process (tensor_split fractions, boundaries on multiples of 128 rows); every slice is an ordinary
// plane-layout matrix on its device. The buffer itself owns no storage (its base is a dummy address, like the reference's): slices are
// allocated per tensor in init_tensor. Only whole-tensor set / get are legal, as in the reference.
// ---------------------------------------------------------------------------------------------------
const char * split_buft_get_name(ggml_backend_buffer_type_t buft) { return ((split_buft_ctx *) buft->context)->name.c_str(); }

void split_rows(const split_buft_ctx * bt, int64_t m, int64_t * lo, int64_t * hi) {
    for (int i = 0; i < bt->n_dev; ++i) {
        int64_t a = i == 0 ? 0 : (int64_t) ((double) m * bt->cum[i]) / 128 * 128;
        int64_t b = i == bt->n_dev - 1 ? m : (int64_t) ((double) m * bt->cum[i + 1]) / 128 * 128;
        if (a > m) a = m;
        if (b > m) b = m;
        if (b < a) b = a;
        lo[i] = a; hi[i] = b;
    }
}

void split_buffer_free(ggml_backend_buffer_t buffer) {
    split_buffer_ctx * sc = (split_buffer_ctx *) buffer->context;
    for (split_tensor * t : sc->tensors) {
        for (int i = 0; i < t->n_dev; ++i) if (t->s[i].data) { b200_device_synchronize(sc->bt->cuda_index[i]); b200_free(sc->bt->cuda_index[i], t->s[i].data); }
        delete t;
    }
    for (int i = 0; i < sc->bt->n_dev; ++i) if (sc->staging[i]) b200_free(sc->bt->cuda_index[i], sc->staging[i]);
    delete sc;
}
void * split_buffer_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }   // never dereferenced: the data live in the slices

enum ggml_status split_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    split_buffer_ctx * sc = (split_buffer_ctx *) buffer->context;
    GGML_ASSERT(tensor->view_src == nullptr && "ggml-b200: views of row-split tensors are not supported");
    GGML_ASSERT(weight_type(tensor->type) && ggml_is_contiguous(tensor) && tensor->ne[2] == 1 && tensor->ne[3] == 1 &&
                "ggml-b200: the split buffer type holds 2-D quantised MUL_MAT weights only");
    split_tensor * t = new split_tensor;
    t->type = (int) tensor->type; t->m = tensor->ne[1]; t->k = tensor->ne[0]; t->n_dev = sc->bt->n_dev;
    int64_t lo[B200_MAX_DEVICES], hi[B200_MAX_DEVICES];
    split_rows(sc->bt, t->m, lo, hi);
    const size_t row = b200_row_size(t->type, t->k);
    for (int i = 0; i < t->n_dev; ++i) {
        t->s[i].lo = lo[i]; t->s[i].hi = hi[i];
        if (hi[i] == lo[i]) continue;
        t->s[i].data = b200_malloc(sc->bt->cuda_index[i], (size_t) (hi[i] - lo[i]) * row + 512);
        if (!t->s[i].data) {
            GGML_LOG_ERROR("ggml-b200: allocating a %.1f MiB slice of %s on device %d failed: %s\n", (hi[i] - lo[i]) * row / 1048576.0, tensor->name,
                           sc->bt->cuda_index[i], b200_last_error());
            for (int j = 0; j < i; ++j) if (t->s[j].data) b200_free(sc->bt->cuda_index[j], t->s[j].data);
            delete t;
            return GGML_STATUS_ALLOC_FAILED;
        }
    }
    sc->tensors.push_back(t);
    tensor->extra = t;
    return GGML_STATUS_SUCCESS;
}

void * split_stage(split_buffer_ctx * sc, int i, size_t n) {
    if (n > sc->staging_size[i]) {
        const int dev = sc->bt->cuda_index[i];
        if (sc->staging[i]) { b200_device_synchronize(dev); b200_free(dev, sc->staging[i]); }
        sc->staging[i] = b200_malloc(dev, n);
        sc->staging_size[i] = sc->staging[i] ? n : 0;
    }
    return sc->staging[i];
}

void split_buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    split_buffer_ctx * sc = (split_buffer_ctx *) buffer->context;
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "ggml-b200: row-split tensors are written whole (ggml-cuda.cu:1019-1021 has the same rule)");
    split_tensor * t = (split_tensor *) tensor->extra;
    const size_t row = b200_row_size(t->type, t->k);
    for (int i = 0; i < t->n_dev; ++i) {
        split_slice & s = t->s[i];
        const int64_t rows = s.hi - s.lo;
        if (rows == 0) continue;
        const int dev = sc->bt->cuda_index[i];
        b200_set_device(dev);
        b200_device_synchronize(dev);                      // kernels of an earlier graph may still read the slice
        const size_t bytes = (size_t) rows * row;
        const char * src = (const char *) data + (size_t) s.lo * row;
        const bool repack = repack_enabled() && t->k % 256 == 0 && b200_repack_supported(t->type, rows, t->k);
        void * tmp = repack ? split_stage(sc, i, bytes) : nullptr;
        int rc;
        if (tmp) {
            rc = b200_memcpy_h2d(tmp, src, bytes, nullptr);
            if (rc == 0) rc = b200_repack(t->type, rows, t->k, tmp, s.data, nullptr);
            s.planes = rc == 0;
        } else {
            rc = b200_memcpy_h2d(s.data, src, bytes, nullptr);
            s.planes = false;
        }
        if (rc != 0 || b200_stream_synchronize(nullptr) != 0) GGML_ABORT("ggml-b200: set_tensor(%s) on the split buffer failed: %s", tensor->name, b200_last_error());
    }
}

void split_buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    split_buffer_ctx * sc = (split_buffer_ctx *) buffer->context;
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "ggml-b200: row-split tensors are read whole");
    const split_tensor * t = (const split_tensor *) tensor->extra;
    const size_t row = b200_row_size(t->type, t->k);
    for (int i = 0; i < t->n_dev; ++i) {
        const split_slice & s = t->s[i];
        const int64_t rows = s.hi - s.lo;
        if (rows == 0) continue;
        const int dev = sc->bt->cuda_index[i];
        b200_set_device(dev);
        b200_device_synchronize(dev);
        const size_t bytes = (size_t) rows * row;
        const void * src = s.data;
        if (s.planes) {                                   // canonical ggml bytes go back out
            void * tmp = split_stage(sc, i, bytes);
            GGML_ASSERT(tmp && "ggml-b200: out of device memory in get_tensor");
            if (b200_unrepack(t->type, rows, t->k, s.data, tmp, nullptr) != 0) GGML_ABORT("ggml-b200: unrepack failed: %s", b200_last_error());
            src = tmp;
        }
        if (b200_memcpy_d2h((char *) data + (size_t) s.lo * row, src, bytes, nullptr) != 0 || b200_stream_synchronize(nullptr) != 0)
            GGML_ABORT("ggml-b200: get_tensor(%s) on the split buffer failed: %s", tensor->name, b200_last_error());
    }
}

void split_buffer_clear(ggml_backend_buffer_t, uint8_t) {}      // nothing is allocated before init_tensor

const ggml_backend_buffer_i split_buffer_iface = {
    /* .free_buffer   = */ split_buffer_free,
    /* .get_base      = */ split_buffer_get_base,
    /* .init_tensor   = */ split_buffer_init_tensor,
    /* .memset_tensor = */ nullptr,
    /* .set_tensor    = */ split_buffer_set_tensor,
    /* .get_tensor    = */ split_buffer_get_tensor,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ split_buffer_clear,
    /* .reset         = */ nullptr,
};

ggml_backend_buffer_t split_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    split_buffer_ctx * sc = new split_buffer_ctx;
    sc->bt = (split_buft_ctx *) buft->context;
    return ggml_backend_buffer_init(buft, split_buffer_iface, sc, size);
}
size_t split_buft_get_alignment(ggml_backend_buffer_type_t) { return B200_ALIGNMENT; }
size_t split_buft_get_alloc_size(ggml_backend_buffer_type_t buft, const ggml_tensor * tensor) {
    const split_buft_ctx * bt = (const split_buft_ctx *) buft->context;
    return ggml_nbytes(tensor) + (size_t) bt->n_dev * 512;          // what init_tensor allocates over all devices
}
bool split_buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i split_buft_iface = {
    /* .get_name       = */ split_buft_get_name,
    /* .alloc_buffer   = */ split_buft_alloc_buffer,
    /* .get_alignment  = */ split_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ split_buft_get_alloc_size,
    /* .is_host        = */ split_buft_is_host,
};

// ---------------------------------------------------------------------------------------------------
// op support
// ---------------------------------------------------------------------------------------------------
bool mul_mat_operands_ok(const ggml_tensor * w, const ggml_tensor * x, const ggml_tensor * y) {
    if (!weight_type(w->type) || x->type != GGML_TYPE_F32 || y->type != GGML_TYPE_F32) return false;
    if (w->nb[0] != ggml_type_size(w->type)) return false;            // quantised rows must be contiguous
    if (x->nb[0] != sizeof(float) || y->nb[0] != sizeof(float)) return false;
    if (x->nb[1] % sizeof(float) || x->nb[2] % sizeof(float) || x->nb[3] % sizeof(float)) return false;
    if (w->ne[0] != x->ne[0] || w->ne[0] % ggml_blck_size(w->type) != 0) return false;
    return true;
}

// ---- the other ops of a llama-arch graph (csrc/glue_ops.cu): exactly the variants implemented there ----------------------
bool glue_on() { static const bool on = [] { const char * e = getenv("GGML_B200_GLUE"); return !(e && e[0] == '0'); }(); return on; }
bool f32_rows(const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float); }
bool f_any(const ggml_tensor * t) { return t && (t->type == GGML_TYPE_F32 || t->type == GGML_TYPE_F16) && t->nb[0] == ggml_type_size(t->type); }
bool glue_supports(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    switch (op->op) {
        case GGML_OP_GET_ROWS:
            return a && b && b->type == GGML_TYPE_I32 && op->type == GGML_TYPE_F32 && (a->type == GGML_TYPE_F32 || a->type == GGML_TYPE_F16 || weight_type(a->type)) &&
                   a->nb[0] == ggml_type_size(a->type);
        case GGML_OP_RMS_NORM:
            return f32_rows(a) && f32_rows(op);
        case GGML_OP_ADD:
        case GGML_OP_MUL:
        case GGML_OP_DIV:
            return a && b && a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ggml_can_repeat(b, a);
        case GGML_OP_SCALE:
        case GGML_OP_CLAMP:
            return a && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ggml_are_same_shape(a, op);
        case GGML_OP_SUM_ROWS:
            return a && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32;
        case GGML_OP_ARGSORT:
            return a && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_I32 && a->ne[0] <= 1024 && op->nb[0] == sizeof(int32_t);
        case GGML_OP_ROPE: {
            if (!f32_rows(a) || !f32_rows(op) || !b || b->type != GGML_TYPE_I32) return false;
            const int32_t * p = (const int32_t *) op->op_params;
            float ext_factor; memcpy(&ext_factor, p + 7, sizeof(float));
            const int n_dims = p[1], mode = p[2];
            if ((mode != 0 && mode != GGML_ROPE_TYPE_NEOX) || ext_factor != 0.0f || n_dims <= 0 || n_dims % 2 || n_dims > a->ne[0] || a->ne[0] % 2) return false;
            const ggml_tensor * ff = op->src[2];
            return !ff || (ff->type == GGML_TYPE_F32 && ggml_is_contiguous(ff) && ff->ne[0] >= n_dims / 2);
        }
        case GGML_OP_SET_ROWS:
            return a && b && a->type == GGML_TYPE_F32 && (b->type == GGML_TYPE_I64 || b->type == GGML_TYPE_I32) && (op->type == GGML_TYPE_F32 || op->type == GGML_TYPE_F16) &&
                   op->nb[0] == ggml_type_size(op->type);
        case GGML_OP_CPY:
        case GGML_OP_CONT:
        case GGML_OP_DUP:
            return a && (a->type == GGML_TYPE_F32 || a->type == GGML_TYPE_F16) && (op->type == GGML_TYPE_F32 || op->type == GGML_TYPE_F16) && ggml_nelements(a) == ggml_nelements(op);
        case GGML_OP_SOFT_MAX: {
            if (!f32_rows(a) || !f32_rows(op) || !ggml_is_contiguous(op) || op->src[2]) return false;
            float max_bias; memcpy(&max_bias, (const float *) op->op_params + 1, sizeof(float));
            if (max_bias != 0.0f) return false;
            return !b || (f_any(b) && b->ne[0] >= a->ne[0] && b->ne[1] >= a->ne[1]);
        }
        case GGML_OP_GLU: {
            const int glu = ggml_get_op_params_i32(op, 0);
            if (glu != GGML_GLU_OP_REGLU && glu != GGML_GLU_OP_GEGLU && glu != GGML_GLU_OP_SWIGLU) return false;
            return f32_rows(a) && f32_rows(op) && (!b || (f32_rows(b) && ggml_are_same_shape(a, b)));
        }
        default:
            return false;
    }
}

bool device_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    // a row-split tensor can only be src0 of a plain 2-D MUL_MAT (the reference has the same restriction: ggml-cuda.cu:4530-4545, :2274)
    for (int j = 0; j < GGML_MAX_SRC; ++j) {
        if (!op->src[j] || !in_split_buffer(op->src[j])) continue;
        const ggml_tensor * w = op->src[0], * x = op->src[1];
        if (op->op != GGML_OP_MUL_MAT || j != 0 || w->view_src || !x || x->ne[2] != 1 || x->ne[3] != 1 || w->ne[2] != 1 || w->ne[3] != 1) return false;
    }
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_VIEW:
        case GGML_OP_RESHAPE:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT: {
            const ggml_tensor * w = op->src[0], * x = op->src[1];
            if (!w || !x) return false;
            if (glue_on() && (w->type == GGML_TYPE_F16 || w->type == GGML_TYPE_F32)) {
                // attention over the KV cache: f16 / f32 rows x f32 rows (b200_op_mul_mat_f)
                return x->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && w->nb[0] == ggml_type_size(w->type) && x->nb[0] == sizeof(float) &&
                       op->nb[0] == sizeof(float) && w->ne[0] == x->ne[0] && x->ne[2] % w->ne[2] == 0 && x->ne[3] % w->ne[3] == 0 && op->ne[2] * op->ne[3] <= 65535;
            }
            if (!mul_mat_operands_ok(w, x, op)) return false;
            return x->ne[2] % w->ne[2] == 0 && x->ne[3] % w->ne[3] == 0;   // broadcast (ggml.c:3168-3174)
        }
        case GGML_OP_MUL_MAT_ID: {
            const ggml_tensor * as = op->src[0], * b = op->src[1], * ids = op->src[2];
            if (!as || !b || !ids || !mul_mat_operands_ok(as, b, op)) return false;
            if (ids->type != GGML_TYPE_I32 || ids->nb[0] != sizeof(int32_t) || ids->nb[1] % sizeof(int32_t)) return false;
            if (as->ne[3] != 1 || b->ne[3] != 1) return false;
            if (as->nb[1] != ggml_row_size(as->type, as->ne[0])) return false;   // rows of one expert packed
            return b->ne[1] == 1 || b->ne[1] == ids->ne[0];
        }
        default:
            return glue_on() && glue_supports(op);
    }
}

// ---------------------------------------------------------------------------------------------------
// backend (one stream)
// ---------------------------------------------------------------------------------------------------
const char * backend_get_name(ggml_backend_t backend) { return ((backend_ctx *) backend->context)->name.c_str(); }

void backend_free(ggml_backend_t backend) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    b200_set_device(bc->device);
    if (bc->stream) { b200_stream_synchronize(bc->stream); b200_stream_destroy(bc->stream); }
    if (bc->workspace) b200_free(bc->device, bc->workspace);
    for (auto & a : bc->aux) {
        b200_set_device(a.device);
        b200_stream_synchronize(a.stream);
        if (a.workspace) b200_free(a.device, a.workspace);
        if (a.x) b200_free(a.device, a.x);
        b200_event_destroy(a.done);
        b200_stream_destroy(a.stream);
    }
    if (bc->fork) b200_event_destroy(bc->fork);
    b200_set_device(bc->device);
    for (auto & c : bc->chains) b200_chain_destroy(c.chain);
    for (auto & g : bc->graphs) if (g.exec) b200_graph_destroy(g.exec);
    delete bc;
    delete backend;
}

void backend_synchronize(ggml_backend_t backend) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    if (b200_stream_synchronize(bc->stream) != 0) GGML_ABORT("ggml-b200: synchronize failed: %s", b200_last_error());
}

bool backend_reserve(backend_ctx * bc, size_t need) {
    if (need <= bc->workspace_size) return true;
    if (bc->capturing) return false;       // (the eager pass before the capture sized the workspace; a larger request fails the capture, not the graph)
    b200_stream_synchronize(bc->stream);   // kernels in flight may still read the old workspace
    if (bc->workspace) b200_free(bc->device, bc->workspace);
    for (auto & c : bc->chains) b200_chain_destroy(c.chain);     // (chains do not use the workspace; dropped with it to keep one rule)
    bc->chains.clear();
    for (auto & g : bc->graphs) if (g.exec) b200_graph_destroy(g.exec);   // captured launches carry the old workspace pointer
    bc->graphs.clear();
    need = (need + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
    bc->workspace = b200_malloc(bc->device, need);
    bc->workspace_size = bc->workspace ? need : 0;
    return bc->workspace != nullptr;
}

// ---- MUL_MAT whose weight rows are spread over the devices of this process --------------------------------------------------------------
// Every device multiplies its rows by the same X and writes its part of Y straight into dst on the backend's device (peer stores over
// NVLink; the reference gathers with ggml_cuda_Memcpy2DPeerAsync instead, ggml-cuda.cu:1725-1744). X reaches the other devices with one
// strided peer copy per node. Fork / join are events, so the whole thing is capturable into the backend's CUDA graph.
backend_ctx::aux_device * aux_for(backend_ctx * bc, int device) {
    for (auto & a : bc->aux) if (a.device == device) return &a;
    if (bc->capturing) return nullptr;
    if (b200_enable_peer_access(device, bc->device) != 0 || b200_enable_peer_access(bc->device, device) != 0) return nullptr;
    backend_ctx::aux_device a;
    a.device = device;
    a.stream = b200_stream_create(device);
    a.done = b200_event_create(device);
    if (!a.stream || !a.done) return nullptr;
    bc->aux.push_back(a);
    return &bc->aux.back();
}
bool aux_reserve(backend_ctx * bc, backend_ctx::aux_device * a, void ** buf, size_t * have, size_t need) {
    if (need <= *have) return true;
    if (bc->capturing) return false;
    b200_stream_synchronize(a->stream);
    if (*buf) b200_free(a->device, *buf);
    for (auto & g : bc->graphs) if (g.exec) b200_graph_destroy(g.exec);   // captured launches carry the old pointers
    bc->graphs.clear();
    need = (need + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
    *buf = b200_malloc(a->device, need);
    *have = *buf ? need : 0;
    return *buf != nullptr;
}

enum ggml_status compute_mul_mat_split(backend_ctx * bc, const ggml_tensor * dst, const split_tensor * st) {
    const ggml_tensor * w = dst->src[0], * x = dst->src[1];
    const split_buft_ctx * bt = (const split_buft_ctx *) w->buffer->buft->context;
    const int64_t k = st->k, n = x->ne[1];
    const size_t need = b200_mul_mat_workspace(st->type, n, k);
    if (!backend_reserve(bc, need)) return GGML_STATUS_ALLOC_FAILED;
    if (!bc->fork) {
        if (bc->capturing) return GGML_STATUS_FAILED;
        bc->fork = b200_event_create(bc->device);
        if (!bc->fork) return GGML_STATUS_FAILED;
    }
    enum ggml_status result = GGML_STATUS_SUCCESS;
    std::vector<backend_ctx::aux_device *> used;
    bool forked = false;
    auto fill = [&](b200_mul_mat_args & a, const split_slice & s) {
        a.type = st->type; a.layout = s.planes ? B200_LAYOUT_PLANES : B200_LAYOUT_CANONICAL;
        a.m = s.hi - s.lo; a.n = n; a.k = k;
        a.w = s.data; a.w_row_stride = 0;
        a.y = (float *) dst->data + s.lo; a.y_row_stride = (int64_t) (dst->nb[1] / sizeof(float));
        a.flavour = B200_Q8_X86; a.flags = mm_flags();
    };
    for (int i = 0; i < st->n_dev && result == GGML_STATUS_SUCCESS; ++i) {
        const split_slice & s = st->s[i];
        const int dev = bt->cuda_index[i];
        if (s.hi == s.lo || dev == bc->device) continue;
        backend_ctx::aux_device * ax = aux_for(bc, dev);
        if (!ax) { GGML_LOG_ERROR("ggml-b200: no peer path between devices %d and %d: %s\n", bc->device, dev, b200_last_error()); result = GGML_STATUS_FAILED; break; }
        if (!aux_reserve(bc, ax, &ax->workspace, &ax->workspace_size, need) || !aux_reserve(bc, ax, &ax->x, &ax->x_size, (size_t) n * k * sizeof(float))) {
            result = GGML_STATUS_ALLOC_FAILED; break;
        }
        if (!forked) { b200_event_record(bc->fork, bc->stream); forked = true; }
        b200_set_device(dev);
        b200_stream_wait_event(ax->stream, bc->fork);
        b200_mul_mat_args a = {};
        fill(a, s);
        a.x = (const float *) ax->x; a.x_row_stride = k;
        a.workspace = ax->workspace; a.workspace_size = ax->workspace_size;
        if (b200_memcpy2d(ax->x, (size_t) k * sizeof(float), x->data, x->nb[1], (size_t) k * sizeof(float), (size_t) n, ax->stream) != 0 ||
            b200_mul_mat(&a, ax->stream) != 0) {
            GGML_LOG_ERROR("ggml-b200: MUL_MAT %s, rows [%lld, %lld) on device %d failed: %s\n", dst->name, (long long) s.lo, (long long) s.hi, dev, b200_last_error());
            result = GGML_STATUS_FAILED;
        }
        b200_event_record(ax->done, ax->stream);
        used.push_back(ax);
    }
    b200_set_device(bc->device);
    for (int i = 0; i < st->n_dev && result == GGML_STATUS_SUCCESS; ++i) {
        const split_slice & s = st->s[i];
        if (s.hi == s.lo || bt->cuda_index[i] != bc->device) continue;
        b200_mul_mat_args a = {};
        fill(a, s);
        a.x = (const float *) x->data; a.x_row_stride = (int64_t) (x->nb[1] / sizeof(float));
        a.workspace = bc->workspace; a.workspace_size = bc->workspace_size;
        if (b200_mul_mat(&a, bc->stream) != 0) {
            GGML_LOG_ERROR("ggml-b200: MUL_MAT %s, local rows failed: %s\n", dst->name, b200_last_error());
            result = GGML_STATUS_FAILED;
        }
    }
    for (auto * ax : used) b200_stream_wait_event(bc->stream, ax->done);      // join (also on failure: an open capture must see every fork rejoin)
    return result;
}

enum ggml_status compute_mul_mat(backend_ctx * bc, const ggml_tensor * dst) {
    if (const split_tensor * st = split_of(dst->src[0])) return compute_mul_mat_split(bc, dst, st);
    const ggml_tensor * w = dst->src[0], * x = dst->src[1];
    const int64_t k = w->ne[0], m = w->ne[1], n = x->ne[1];
    const int layout = weight_layout(w);
    const size_t need = b200_mul_mat_workspace((int) w->type, n, k);
    if (!backend_reserve(bc, need)) return GGML_STATUS_ALLOC_FAILED;
    const int64_t r2 = x->ne[2] / w->ne[2], r3 = x->ne[3] / w->ne[3];
    for (int64_t i13 = 0; i13 < x->ne[3]; ++i13) {
        for (int64_t i12 = 0; i12 < x->ne[2]; ++i12) {
            const int64_t i03 = i13 / r3, i02 = i12 / r2;
            b200_mul_mat_args a = {};
            a.type = (int32_t) w->type;
            a.layout = layout;
            a.m = m; a.n = n; a.k = k;
            a.w = (const char *) w->data + i02 * w->nb[2] + i03 * w->nb[3];
            a.w_row_stride = layout == B200_LAYOUT_PLANES ? 0 : (int64_t) w->nb[1];
            a.x = (const float *) ((const char *) x->data + i12 * x->nb[2] + i13 * x->nb[3]);
            a.x_row_stride = (int64_t) (x->nb[1] / sizeof(float));
            a.y = (float *) ((char *) dst->data + i12 * dst->nb[2] + i13 * dst->nb[3]);
            a.y_row_stride = (int64_t) (dst->nb[1] / sizeof(float));
            a.workspace = bc->workspace;
            a.workspace_size = bc->workspace_size;
            a.flavour = B200_Q8_X86; a.flags = mm_flags();
            if (b200_mul_mat(&a, bc->stream) != 0) {
                GGML_LOG_ERROR("ggml-b200: MUL_MAT %s failed: %s\n", dst->name, b200_last_error());
                return GGML_STATUS_FAILED;
            }
        }
    }
    return GGML_STATUS_SUCCESS;
}

enum ggml_status compute_mul_mat_id(backend_ctx * bc, const ggml_tensor * dst, bool x_prequantised = false) {
    const ggml_tensor * as = dst->src[0], * b = dst->src[1], * ids = dst->src[2];
    const int64_t k = as->ne[0], m = as->ne[1], n_as = as->ne[2];
    const int64_t n_used = ids->ne[0], n_tokens = ids->ne[1], nb1 = b->ne[1];
    const int layout = weight_layout(as);
    const int64_t chunk = n_used > 0 ? (65535 / n_used > 0 ? 65535 / n_used : 1) : n_tokens;   // (token, slot) pairs per launch
    const int64_t tmax = n_tokens < chunk ? n_tokens : chunk;
    // room for whichever path the call takes: per-pair GEMV (few pairs per expert) or expert gather + grouped tensor-core GEMM
    b200_mul_mat_id_args probe = {};
    probe.type = (int32_t) as->type; probe.layout = layout; probe.m = m; probe.k = k; probe.n_as = n_as; probe.n_used = n_used; probe.nb1 = nb1; probe.n_tokens = tmax;
    const size_t need = b200_mul_mat_id_workspace_ex(&probe);
    if (!backend_reserve(bc, need)) return GGML_STATUS_ALLOC_FAILED;
    for (int64_t t0 = 0; t0 < n_tokens; t0 += chunk) {
        const int64_t nt = n_tokens - t0 < chunk ? n_tokens - t0 : chunk;
        b200_mul_mat_id_args a = {};
        a.type = (int32_t) as->type; a.layout = layout;
        a.m = m; a.k = k; a.n_as = n_as; a.n_used = n_used; a.nb1 = nb1; a.n_tokens = nt;
        a.as = as->data; a.as_expert_stride = (int64_t) as->nb[2];
        a.b = (const float *) ((const char *) b->data + t0 * b->nb[2]);
        a.b_row_stride = (int64_t) (b->nb[1] / sizeof(float)); a.b_token_stride = (int64_t) (b->nb[2] / sizeof(float));
        a.ids = (const int32_t *) ((const char *) ids->data + t0 * ids->nb[1]);
        a.ids_row_stride = (int64_t) (ids->nb[1] / sizeof(int32_t));
        a.y = (float *) ((char *) dst->data + t0 * dst->nb[2]);
        a.y_row_stride = (int64_t) (dst->nb[1] / sizeof(float)); a.y_token_stride = (int64_t) (dst->nb[2] / sizeof(float));
        a.workspace = bc->workspace; a.workspace_size = bc->workspace_size;
        a.flavour = B200_Q8_X86; a.flags = mm_flags() | (x_prequantised && n_tokens <= chunk ? (int32_t) B200_MM_X_PREQUANTISED : 0);
        if (b200_mul_mat_id(&a, bc->stream) != 0) {
            GGML_LOG_ERROR("ggml-b200: MUL_MAT_ID %s failed: %s\n", dst->name, b200_last_error());
            return GGML_STATUS_FAILED;
        }
    }
    return GGML_STATUS_SUCCESS;
}

// ---- decode fast paths: runs of plain 2-D MUL_MAT nodes ---------------------------------------------------------------
// A run = consecutive MUL_MAT nodes (no-op nodes in between skipped) with n <= 4 and no batch dimensions. Nodes that read the
// same X form a group (wq / wk / wv, ffn_gate / ffn_up): ONE launch (b200_mul_mat_group, counterpart of the reference's fused
// mul_mat_vec_q, ggml-cuda.cu:2144-2180). Two or more consecutive groups of K-quant matrices: ONE persistent launch for all of
// them (b200_chain_*), cached per argument list. GGML_B200_GROUP=0 / GGML_B200_CHAIN=0 switch the two off.
bool is_noop(const ggml_tensor * t) {
    return ggml_is_empty(t) || t->op == GGML_OP_NONE || t->op == GGML_OP_VIEW || t->op == GGML_OP_RESHAPE || t->op == GGML_OP_PERMUTE ||
           t->op == GGML_OP_TRANSPOSE;
}
// a node the graph does not ask for (GGML_TENSOR_FLAG_COMPUTE unset: ggml-cpu.c:2946, ggml-cuda.cu:3386 skip those as well)
bool skipped(const ggml_tensor * t) { return (t->flags & GGML_TENSOR_FLAG_COMPUTE) == 0; }
// a 2-D MUL_MAT without batch dimensions: decode (n <= 4) and prefill alike
bool plain_mul_mat(const ggml_tensor * t) {
    if (t->op != GGML_OP_MUL_MAT || skipped(t)) return false;
    const ggml_tensor * w = t->src[0], * x = t->src[1];
    if (!weight_type(w->type) || in_split_buffer(w)) return false;
    return x->ne[2] == 1 && x->ne[3] == 1 && w->ne[2] == 1 && w->ne[3] == 1 && x->nb[0] == sizeof(float) && t->nb[0] == sizeof(float);
}
b200_mul_mat_args plain_args(const ggml_tensor * dst) {
    const ggml_tensor * w = dst->src[0], * x = dst->src[1];
    b200_mul_mat_args a = {};
    a.type = (int32_t) w->type; a.layout = weight_layout(w);
    a.m = w->ne[1]; a.n = x->ne[1]; a.k = w->ne[0];
    a.w = w->data; a.w_row_stride = a.layout == B200_LAYOUT_PLANES ? 0 : (int64_t) w->nb[1];
    a.x = (const float *) x->data; a.x_row_stride = (int64_t) (x->nb[1] / sizeof(float));
    a.y = (float *) dst->data; a.y_row_stride = (int64_t) (dst->nb[1] / sizeof(float));
    a.flavour = B200_Q8_X86; a.flags = mm_flags();
    return a;
}
// two argument lists describe the same chain (field by field: the structs carry padding)
bool same_args(const std::vector<b200_mul_mat_args> & a, const std::vector<b200_mul_mat_args> & b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); ++i) {
        const b200_mul_mat_args & p = a[i], & q = b[i];
        if (p.type != q.type || p.layout != q.layout || p.m != q.m || p.n != q.n || p.k != q.k || p.w != q.w || p.w_row_stride != q.w_row_stride ||
            p.x != q.x || p.x_row_stride != q.x_row_stride || p.y != q.y || p.y_row_stride != q.y_row_stride || p.flavour != q.flavour) return false;
    }
    return true;
}
bool env_on(const char * name) { const char * e = getenv(name); return !(e && e[0] == '0'); }

// tries to run nodes [i, ...) as groups / a chain; returns the index of the first node NOT consumed (== i: nothing done).
// Decode (n <= 4): groups become single launches, two or more consecutive K-quant groups ONE persistent chain launch. Prefill (n > 4):
// the nodes that read the same X share the fp16 copy of X and, per weight type, ONE tensor-core GEMM launch (b200_mul_mat_group).
// byte ranges of two tensors overlap (views included: the range of the bytes the tensor itself spans)
bool ranges_overlap(const ggml_tensor * a, const ggml_tensor * b) {
    if (!a || !b || !a->data || !b->data) return false;
    const char * a0 = (const char *) a->data, * b0 = (const char *) b->data;
    return a0 < b0 + ggml_nbytes(b) && b0 < a0 + ggml_nbytes(a);
}
// may node `h` run BEFORE the nodes [from, h_idx) that precede it in the graph? (it does not read what they write, they neither read nor write
// what it writes: ggml-alloc reuses buffers, so a later node's output may sit on an earlier node's input)
bool can_hoist(const ggml_cgraph * g, int from, int h_idx) {
    const ggml_tensor * h = g->nodes[h_idx];
    for (int s = from; s < h_idx; ++s) {
        const ggml_tensor * t = g->nodes[s];
        if (ggml_is_empty(t)) continue;
        if (ranges_overlap(h, t)) return false;
        for (int k = 0; k < GGML_MAX_SRC; ++k) {
            if (t->src[k] && ranges_overlap(h, t->src[k])) return false;
            if (h->src[k] && (h->src[k] == t || ranges_overlap(h->src[k], t))) return false;
        }
    }
    return true;
}

int compute_decode_run(backend_ctx * bc, ggml_cgraph * cgraph, int i, enum ggml_status & st, std::vector<char> & done) {
    static const bool group_on = env_on("GGML_B200_GROUP"), chain_on = env_on("GGML_B200_CHAIN"), hoist_on = env_on("GGML_B200_HOIST");
    st = GGML_STATUS_SUCCESS;
    if (!group_on) return i;
    std::vector<b200_mul_mat_args> args;
    std::vector<int32_t> sizes;
    int j = i, last = i;
    for (; j < cgraph->n_nodes; ++j) {
        const ggml_tensor * t = cgraph->nodes[j];
        if (done[j] || is_noop(t) || skipped(t) || ggml_is_empty(t)) continue;
        if (!plain_mul_mat(t)) break;
        b200_mul_mat_args a = plain_args(t);
        if (!args.empty() && a.n != args[0].n) break;
        const b200_mul_mat_args * head = sizes.empty() ? nullptr : &args[args.size() - (size_t) sizes.back()];
        const bool same_x = head && head->x == a.x && head->k == a.k && head->x_row_stride == a.x_row_stride && sizes.back() < 4;
        if (same_x) ++sizes.back(); else sizes.push_back(1);
        args.push_back(a);
        last = j + 1;
    }
    // wq ... rope ... wv ... wk: the other matmuls of the same X sit a few nodes further down the graph. Pull them into this group when nothing
    // in between depends on them or touches their output (one launch, one quantisation / fp16 copy of X)
    if (hoist_on && sizes.size() == 1) {
        const b200_mul_mat_args head = args[0];
        for (int h = j; h < cgraph->n_nodes && h < j + 12 && sizes.back() < 4; ++h) {
            const ggml_tensor * t = cgraph->nodes[h];
            if (done[h] || !plain_mul_mat(t)) continue;
            const b200_mul_mat_args a = plain_args(t);
            if (a.x != head.x || a.n != head.n || a.k != head.k || a.x_row_stride != head.x_row_stride) continue;
            if (!can_hoist(cgraph, last, h)) continue;
            args.push_back(a); ++sizes.back(); done[h] = 1;
        }
    }
    // a single group also takes the chain kernel (one persistent launch for one group: its block-per-lane items out of a shared-memory ring beat
    // the per-node GEMV in a real graph: llama-bench 8B tg128 353 -> 386); GGML_B200_CHAIN_SINGLE=0: only runs of two or more groups
    static const bool chain_single = env_on("GGML_B200_CHAIN_SINGLE");
    static const bool chain_q40 = [] { const char * e = getenv("GGML_B200_CHAIN_Q4_0"); return e && e[0] == '1'; }();
    static const bool chain_q80 = [] { const char * e = getenv("GGML_B200_CHAIN_Q8_0"); return e && e[0] == '1'; }();
    static const bool chain_iq4 = env_on("GGML_B200_CHAIN_IQ4_XS");      // IQ4_XS groups too (llama-bench 8B IQ4_XS tg128 298 -> 310); Q4_0 / Q8_0 stay per node (1B Q4_0: 1059 vs 988)
    if (args.size() < 2 && !(chain_single && args[0].n <= 4)) return i;                       // a single node: the ordinary path
    // workspace for the group launches (the chain does not need one)
    size_t need = 0;
    for (const auto & a : args) { const size_t w = b200_mul_mat_workspace(a.type, a.n, a.k); need = w > need ? w : need; }
    if (!backend_reserve(bc, need)) { st = GGML_STATUS_ALLOC_FAILED; return last; }
    for (auto & a : args) { a.workspace = bc->workspace; a.workspace_size = bc->workspace_size; }

    bool kquant = true;
    for (const auto & a : args) kquant = kquant && (a.type == B200_TYPE_Q4_K || a.type == B200_TYPE_Q5_K || a.type == B200_TYPE_Q6_K || (chain_q40 && a.type == B200_TYPE_Q4_0) || (chain_q80 && a.type == B200_TYPE_Q8_0) || (chain_iq4 && a.type == B200_TYPE_IQ4_XS));
    if (chain_on && kquant && args[0].n <= 4 && (sizes.size() >= 2 || chain_single) && b200_chain_supported(args.data(), sizes.data(), (int) sizes.size())) {
        b200_chain_t chain = nullptr;
        for (auto & c : bc->chains)
            if (c.sizes == sizes && same_args(c.args, args)) { chain = c.chain; break; }
        if (!chain) {
            chain = b200_chain_create(bc->device, args.data(), sizes.data(), (int) sizes.size());
            if (chain) {
                if (bc->chains.size() >= (chain_single ? 1024u : 16u)) { b200_stream_synchronize(bc->stream); b200_chain_destroy(bc->chains.front().chain); bc->chains.erase(bc->chains.begin()); }
                bc->chains.push_back({args, sizes, chain});
            }
        }
        if (chain) {
            if (b200_chain_launch(chain, bc->stream) != 0) { GGML_LOG_ERROR("ggml-b200: decode chain failed: %s\n", b200_last_error()); st = GGML_STATUS_FAILED; }
            return last;
        }
    }
    size_t pos = 0;
    for (int32_t cnt : sizes) {
        if (b200_mul_mat_group(&args[pos], cnt, bc->stream) != 0) {
            GGML_LOG_ERROR("ggml-b200: MUL_MAT group failed: %s\n", b200_last_error());
            st = GGML_STATUS_FAILED;
            return last;
        }
        pos += (size_t) cnt;
    }
    return last;
}

b200_tensor bt(const ggml_tensor * t) {
    b200_tensor r = {};
    r.data = t->data; r.type = (int32_t) t->type;
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = (int64_t) t->nb[i]; }
    return r;
}
// a tensor another op reads byte for byte must hold the canonical ggml bytes (GET_ROWS on a repacked token_embd)
void ensure_canonical(const ggml_tensor * t) {
    buffer_ctx * bc = owner_ctx(t);
    if (!bc) return;
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    std::lock_guard<std::mutex> lock(bc->mu);
    planes_demote_locked(bc, base->data);
}
b200_rope_params rope_params_of(const ggml_tensor * node) {
    const int32_t * p = (const int32_t *) node->op_params;
    b200_rope_params rp = {};
    rp.n_dims = p[1]; rp.mode = p[2];
    memcpy(&rp.freq_base, p + 5, sizeof(float)); memcpy(&rp.freq_scale, p + 6, sizeof(float));
    memcpy(&rp.ext_factor, p + 7, sizeof(float)); memcpy(&rp.attn_factor, p + 8, sizeof(float));
    return rp;
}
enum ggml_status compute_glue(backend_ctx * bc, const ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * b = node->src[1];
    const b200_tensor d = bt(node);
    int rc = 0;
    switch (node->op) {
        case GGML_OP_GET_ROWS: { ensure_canonical(a); const b200_tensor ta = bt(a), tb = bt(b); rc = b200_op_get_rows(&ta, &tb, &d, bc->stream); } break;
        case GGML_OP_RMS_NORM: { float eps; memcpy(&eps, node->op_params, sizeof(float)); const b200_tensor ta = bt(a); rc = b200_op_rms_norm(&ta, &d, eps, bc->stream); } break;
        case GGML_OP_ADD:
        case GGML_OP_MUL:
        case GGML_OP_DIV: { const b200_tensor ta = bt(a), tb = bt(b); rc = b200_op_binary(node->op == GGML_OP_ADD ? 0 : node->op == GGML_OP_MUL ? 1 : 2, &ta, &tb, &d, bc->stream); } break;
        case GGML_OP_SCALE:
        case GGML_OP_CLAMP: {
            float p0, p1; memcpy(&p0, node->op_params, sizeof(float)); memcpy(&p1, (const float *) node->op_params + 1, sizeof(float));
            const b200_tensor ta = bt(a); rc = b200_op_unary(node->op == GGML_OP_SCALE ? 0 : 1, &ta, &d, p0, p1, bc->stream);
        } break;
        case GGML_OP_SUM_ROWS: { const b200_tensor ta = bt(a); rc = b200_op_sum_rows(&ta, &d, bc->stream); } break;
        case GGML_OP_ARGSORT: { const b200_tensor ta = bt(a); rc = b200_op_argsort(&ta, &d, ggml_get_op_params_i32(node, 0), bc->stream); } break;
        case GGML_OP_ROPE: {
            const b200_rope_params rp = rope_params_of(node);
            const b200_tensor ta = bt(a), tb = bt(b);
            b200_tensor tf = {}; if (node->src[2]) tf = bt(node->src[2]);
            rc = b200_op_rope(&ta, &tb, node->src[2] ? &tf : nullptr, &d, &rp, bc->stream);
        } break;
        case GGML_OP_SET_ROWS: { const b200_tensor ta = bt(a), tb = bt(b); rc = b200_op_set_rows(&ta, &tb, &d, bc->stream); } break;
        case GGML_OP_CPY:
        case GGML_OP_CONT:
        case GGML_OP_DUP: { const b200_tensor ta = bt(a); rc = b200_op_cpy(&ta, &d, bc->stream); } break;
        case GGML_OP_SOFT_MAX: {
            float scale; memcpy(&scale, node->op_params, sizeof(float));
            const b200_tensor ta = bt(a); b200_tensor tm = {}; if (b) tm = bt(b);
            rc = b200_op_soft_max(&ta, b ? &tm : nullptr, &d, scale, bc->stream);
        } break;
        case GGML_OP_GLU: {
            const b200_tensor ta = bt(a); b200_tensor tb = {}; if (b) tb = bt(b);
            rc = b200_op_glu(&ta, b ? &tb : nullptr, &d, ggml_get_op_params_i32(node, 0), ggml_get_op_params_i32(node, 1), bc->stream);
        } break;
        case GGML_OP_MUL_MAT: { const b200_tensor ta = bt(a), tb = bt(b); rc = b200_op_mul_mat_f(&ta, &tb, &d, bc->stream); } break;
        default:
            GGML_LOG_ERROR("ggml-b200: op %s (%s) is not supported by this backend\n", ggml_op_name(node->op), node->name);
            return GGML_STATUS_FAILED;
    }
    if (rc != 0) {
        GGML_LOG_ERROR("ggml-b200: %s (%s) failed: %s\n", ggml_op_name(node->op), node->name, b200_last_error());
        return GGML_STATUS_FAILED;
    }
    return GGML_STATUS_SUCCESS;
}

enum ggml_status graph_issue(backend_ctx * bc, ggml_cgraph * cgraph);

// one 64-bit word per node: everything a captured launch sequence depends on
uint64_t mix64(uint64_t h, uint64_t v) { h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2); return h * 0xff51afd7ed558ccdull; }
uint64_t tensor_sig(uint64_t h, const ggml_tensor * t) {
    if (!t) return mix64(h, 0x1234);
    h = mix64(h, (uint64_t) (uintptr_t) t->data); h = mix64(h, (uint64_t) t->type);
    for (int i = 0; i < 4; ++i) { h = mix64(h, (uint64_t) t->ne[i]); h = mix64(h, (uint64_t) t->nb[i]); }
    return h;
}
void graph_key(const ggml_cgraph * g, std::vector<uint64_t> & key) {
    key.clear();
    key.reserve((size_t) g->n_nodes);
    for (int i = 0; i < g->n_nodes; ++i) {
        const ggml_tensor * n = g->nodes[i];
        uint64_t h = mix64((uint64_t) n->op, (uint64_t) n->flags);
        h = tensor_sig(h, n);
        for (int j = 0; j < GGML_MAX_SRC; ++j) if (n->src[j]) h = tensor_sig(mix64(h, (uint64_t) j), n->src[j]);
        const uint64_t * p = (const uint64_t *) n->op_params;
        for (size_t j = 0; j < sizeof(n->op_params) / 8; ++j) h = mix64(h, p[j]);
        key.push_back(h);
    }
}

enum ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    b200_set_device(bc->device);
    static const bool graphs_on = env_on("GGML_B200_GRAPHS");
    if (!graphs_on || cgraph->n_nodes < 8) return graph_issue(bc, cgraph);
    std::vector<uint64_t> key;
    graph_key(cgraph, key);
    backend_ctx::cached_graph * hit = nullptr;
    for (auto & c : bc->graphs) if (c.key == key) { hit = &c; break; }
    if (!hit) {
        if (bc->graphs.size() >= 8) {                              // drop the least recently used entry
            size_t lru = 0;
            for (size_t i = 1; i < bc->graphs.size(); ++i) if (bc->graphs[i].stamp < bc->graphs[lru].stamp) lru = i;
            if (bc->graphs[lru].exec) { b200_stream_synchronize(bc->stream); b200_graph_destroy(bc->graphs[lru].exec); }
            bc->graphs.erase(bc->graphs.begin() + (long) lru);
        }
        bc->graphs.emplace_back();
        hit = &bc->graphs.back();
        hit->key = std::move(key);
    }
    hit->stamp = ++bc->graph_stamp;
    if (hit->exec) {
        if (b200_graph_launch(hit->exec, bc->stream) == 0) return GGML_STATUS_SUCCESS;
        GGML_LOG_WARN("ggml-b200: graph replay failed (%s), running the nodes one by one\n", b200_last_error());
        b200_graph_destroy(hit->exec); hit->exec = nullptr; hit->bad = true;
    }
    if (hit->bad) return graph_issue(bc, cgraph);
    if (++hit->seen < 2) {
        // first sight: eager (allocations, chains, layout changes happen here) -- and, right behind it, the same launch sequence once more under stream capture
        // (nothing executes), so that the SECOND call of this graph is already a replay (GGML_B200_GRAPH_EARLY=0: capture at the second call instead)
        static const bool early = env_on("GGML_B200_GRAPH_EARLY");
        const std::vector<uint64_t> key_copy = hit->key;
        const enum ggml_status st = graph_issue(bc, cgraph);
        if (st != GGML_STATUS_SUCCESS || !early) return st;
        hit = nullptr;                                                  // (the eager pass may have dropped the cache: a grown workspace invalidates every capture)
        for (auto & c : bc->graphs) if (c.key == key_copy) { hit = &c; break; }
        if (!hit) { bc->graphs.emplace_back(); hit = &bc->graphs.back(); hit->key = key_copy; hit->seen = 1; hit->stamp = ++bc->graph_stamp; }
        if (b200_graph_begin(bc->stream) != 0) return st;
        bc->capturing = true;
        const enum ggml_status st2 = graph_issue(bc, cgraph);
        bc->capturing = false;
        b200_graph_t exec = b200_graph_end(bc->stream);
        if (st2 == GGML_STATUS_SUCCESS && exec) hit->exec = exec;
        else { if (exec) b200_graph_destroy(exec); }                    // (no harm done: the second call tries again the old way)
        return st;
    }
    // second sight: capture the very same launch sequence
    if (b200_graph_begin(bc->stream) != 0) { hit->bad = true; return graph_issue(bc, cgraph); }
    bc->capturing = true;
    const enum ggml_status st = graph_issue(bc, cgraph);
    bc->capturing = false;
    b200_graph_t exec = b200_graph_end(bc->stream);
    if (st != GGML_STATUS_SUCCESS || !exec) {
        if (exec) b200_graph_destroy(exec);
        hit->bad = true;
        GGML_LOG_WARN("ggml-b200: graph capture failed (%s), this graph stays on per-node launches\n", b200_last_error());
        return graph_issue(bc, cgraph);
    }
    hit->exec = exec;
    if (b200_graph_launch(exec, bc->stream) != 0) { hit->bad = true; b200_graph_destroy(exec); hit->exec = nullptr; return graph_issue(bc, cgraph); }
    return GGML_STATUS_SUCCESS;
}

// ---- fusions over the node list (GGML_B200_FUSE=0: off) ----------------------------------------------------------------------------------
bool norm_weight_ok(const ggml_tensor * w, const ggml_tensor * x) {
    return w && w->type == GGML_TYPE_F32 && w->nb[0] == sizeof(float) && w->ne[0] == x->ne[0] && ggml_can_repeat(w, x);
}
// [ADD ->] RMS_NORM -> MUL(weight) as one kernel; returns the number of nodes consumed (0: no match)
int try_fuse_norm(backend_ctx * bc, ggml_cgraph * g, int i, enum ggml_status & st) {
    const ggml_tensor * n0 = g->nodes[i];
    int r = i;                                            // index of the RMS_NORM node
    const ggml_tensor * add = nullptr;
    if (n0->op == GGML_OP_ADD) {
        if (i + 2 >= g->n_nodes || !f32_rows(n0) || !f32_rows(n0->src[0]) || !f32_rows(n0->src[1]) || !ggml_can_repeat(n0->src[1], n0->src[0]) ||
            n0->src[1]->ne[0] != n0->src[0]->ne[0]) return 0;
        add = n0; r = i + 1;
        if (g->nodes[r]->op != GGML_OP_RMS_NORM || g->nodes[r]->src[0] != add || skipped(g->nodes[r])) return 0;
    }
    static const enum ggml_op ops[2] = {GGML_OP_RMS_NORM, GGML_OP_MUL};
    if (!ggml_can_fuse(g, r, ops, 2)) return 0;
    const ggml_tensor * rms = g->nodes[r], * mul = g->nodes[r + 1];
    const ggml_tensor * w = mul->src[0] == rms ? mul->src[1] : mul->src[0];
    if (!f32_rows(rms) || !f32_rows(rms->src[0]) || !f32_rows(mul) || !norm_weight_ok(w, rms)) return 0;
    float eps; memcpy(&eps, rms->op_params, sizeof(float));
    const b200_tensor d = bt(mul), tw = bt(w);
    int rc;
    if (add) { const b200_tensor ta = bt(add->src[0]), tb = bt(add->src[1]), ts = bt(add); rc = b200_op_add_rms_norm_mul(&ta, &tb, &ts, &tw, &d, eps, bc->stream); }
    else     { const b200_tensor ta = bt(rms->src[0]); rc = b200_op_add_rms_norm_mul(&ta, nullptr, nullptr, &tw, &d, eps, bc->stream); }
    if (rc != 0) { GGML_LOG_ERROR("ggml-b200: fused norm (%s) failed: %s\n", mul->name, b200_last_error()); st = GGML_STATUS_FAILED; }
    return add ? 3 : 2;
}
int next_real(const ggml_cgraph * g, int i) {
    for (++i; i < g->n_nodes; ++i) { const ggml_tensor * t = g->nodes[i]; if (!is_noop(t) && !skipped(t) && !ggml_is_empty(t)) return i; }
    return -1;
}
// MUL_MAT(K, q) -> SOFT_MAX -> MUL_MAT(V^T, P) -> PERMUTE -> CONT for a few query tokens as one kernel; returns the index of the CONT node (-1: no match)
int try_fuse_attn(backend_ctx * bc, ggml_cgraph * g, int i, enum ggml_status & st) {
    const ggml_tensor * kq = g->nodes[i];
    if (kq->op != GGML_OP_MUL_MAT || kq->src[0]->type != GGML_TYPE_F16 || kq->src[1]->type != GGML_TYPE_F32) return -1;
    const int is = next_real(g, i); if (is < 0) return -1;
    const int iv = next_real(g, is); if (iv < 0) return -1;
    const int ic = next_real(g, iv); if (ic < 0) return -1;
    const ggml_tensor * sm = g->nodes[is], * kqv = g->nodes[iv], * cont = g->nodes[ic];
    if (sm->op != GGML_OP_SOFT_MAX || sm->src[0] != kq || sm->src[2] || kqv->op != GGML_OP_MUL_MAT || kqv->src[1] != sm || kqv->src[0]->type != GGML_TYPE_F16 ||
        cont->op != GGML_OP_CONT) return -1;
    const ggml_tensor * perm = cont->src[0];
    if (!perm || perm->op != GGML_OP_PERMUTE || perm->src[0] != kqv) return -1;
    if (!ggml_node_has_n_uses(g, i, 1) || !ggml_node_has_n_uses(g, is, 1) || !ggml_node_has_n_uses(g, iv, 1)) return -1;
    const ggml_tensor * K = kq->src[0], * q = kq->src[1], * V = kqv->src[0], * mask = sm->src[1];
    const int64_t hd = K->ne[0], n_kv = K->ne[1], n_tok = q->ne[1], n_head = q->ne[2];
    float scale, max_bias; memcpy(&scale, sm->op_params, sizeof(float)); memcpy(&max_bias, (const float *) sm->op_params + 1, sizeof(float));
    static const bool prefill_on = env_on("GGML_B200_FUSE_PREFILL_ATTN");
    const bool decode = n_tok <= 16 && n_kv <= 11000;
    auto al16 = [](const ggml_tensor * t) { return ((uintptr_t) t->data % 16) == 0 && t->nb[1] % 16 == 0 && t->nb[2] % 16 == 0; };
    // more query tokens: the tensor-core kernel (b200_op_attn_prefill) with its own shape rules
    const bool prefill = !decode && prefill_on && (hd == 64 || hd == 128) && n_kv >= 64 && n_kv % 64 == 0 && n_head <= 65535 && V->ne[3] == 1 && al16(K) && al16(V) && al16(q) &&
                         (!mask || (mask->ne[2] == 1 && mask->ne[3] == 1 && ((uintptr_t) mask->data % 8) == 0 && mask->nb[1] % 8 == 0));
    if (max_bias != 0.0f || (!decode && !prefill) || n_kv < 1 || K->ne[3] != 1 || q->ne[3] != 1 || K->nb[0] != 2 || V->nb[0] != 2 || q->nb[0] != 4 ||
        V->ne[0] != n_kv || V->ne[1] != hd || V->ne[2] != K->ne[2] || n_head % K->ne[2] != 0 || hd % 32 != 0 || K->nb[1] % 16 || K->nb[2] % 16 ||
        ((uintptr_t) K->data % 16)) return -1;
    if (mask && !(f_any(mask) && mask->ne[0] >= n_kv && mask->ne[1] >= n_tok)) return -1;
    // the permuted view must be [head_dim, n_head, n_tokens] of kqv [head_dim, n_tokens, n_head], and CONT its contiguous copy
    if (perm->ne[0] != hd || perm->ne[1] != n_head || perm->ne[2] != n_tok || perm->nb[1] != kqv->nb[2] || perm->nb[2] != kqv->nb[1] ||
        cont->type != GGML_TYPE_F32 || !ggml_is_contiguous(cont) || ggml_nelements(cont) != hd * n_head * n_tok) return -1;
    const b200_tensor tk = bt(K), tq = bt(q), tv = bt(V), td = bt(cont);
    b200_tensor tm = {}; if (mask) tm = bt(mask);
    const int rc = decode ? b200_op_attn_decode(&tk, &tq, mask ? &tm : nullptr, &tv, &td, scale, bc->stream)
                          : b200_op_attn_prefill(&tk, &tq, mask ? &tm : nullptr, &tv, &td, scale, bc->stream);
    if (rc != 0) {
        GGML_LOG_ERROR("ggml-b200: fused attention (%s) failed: %s\n", cont->name, b200_last_error()); st = GGML_STATUS_FAILED;
    }
    return ic;
}

// ROPE(k) -> SET_ROWS(k cache) [-> SET_ROWS(v cache)] as one launch (b200_op_rope_set_rows); returns the index of the last node consumed (-1: no match)
int try_fuse_kv(backend_ctx * bc, ggml_cgraph * g, int i, enum ggml_status & st) {
    const ggml_tensor * R = g->nodes[i];
    const int is1 = next_real(g, i);
    if (is1 < 0) return -1;
    const ggml_tensor * S1 = g->nodes[is1];
    if (S1->op != GGML_OP_SET_ROWS || !glue_supports(R) || !glue_supports(S1)) return -1;
    const ggml_tensor * v = S1->src[0];
    if (!v || v->data != R->data || v->type != GGML_TYPE_F32 || !ggml_is_contiguous(v) || !ggml_is_contiguous(R) || ggml_nelements(v) != ggml_nelements(R) || R->view_src) return -1;
    const ggml_tensor * root = v;
    while (root->view_src) root = root->view_src;
    if (root != R || ranges_overlap(S1, R) || ranges_overlap(S1, R->src[0])) return -1;
    int last = is1;
    const ggml_tensor * S2 = nullptr;
    const int is2 = next_real(g, is1);
    if (is2 >= 0 && g->nodes[is2]->op == GGML_OP_SET_ROWS && glue_supports(g->nodes[is2])) {
        const ggml_tensor * c = g->nodes[is2];
        // independent of the pair: it neither reads what they write nor writes what they read or write
        if (!ranges_overlap(c->src[0], R) && !ranges_overlap(c->src[0], S1) && !ranges_overlap(c, S1) && !ranges_overlap(c, R) && !ranges_overlap(c, R->src[0]) &&
            !ranges_overlap(c->src[1], R) && !ranges_overlap(c, S1->src[1])) { S2 = c; last = is2; }
    }
    const b200_rope_params rp = rope_params_of(R);
    const b200_tensor ta = bt(R->src[0]), tp = bt(R->src[1]), td = bt(R), tv = bt(v), ti = bt(S1->src[1]), tc = bt(S1);
    b200_tensor tf = {}; if (R->src[2]) tf = bt(R->src[2]);
    b200_tensor v2 = {}, i2 = {}, c2 = {};
    if (S2) { v2 = bt(S2->src[0]); i2 = bt(S2->src[1]); c2 = bt(S2); }
    if (b200_op_rope_set_rows(&ta, &tp, R->src[2] ? &tf : nullptr, &td, &rp, &tv, &ti, &tc, S2 ? &v2 : nullptr, S2 ? &i2 : nullptr, S2 ? &c2 : nullptr, bc->stream) != 0) {
        GGML_LOG_ERROR("ggml-b200: fused rope + set_rows (%s) failed: %s\n", R->name, b200_last_error()); st = GGML_STATUS_FAILED;
    }
    return last;
}

// GLU(gate, up) -> MUL_MAT(w, .) at decode: the gate goes into the activation quantiser of the matmul (b200_mul_mat_glu); returns the index of the
// MUL_MAT node (-1: no match)
int try_fuse_glu(backend_ctx * bc, ggml_cgraph * g, int i, enum ggml_status & st) {
    const ggml_tensor * G = g->nodes[i];
    static const bool glu_on = env_on("GGML_B200_FUSE_GLU");
    if (!glu_on || !G->src[1] || !glue_supports(G) || !ggml_node_has_n_uses(g, i, 1)) return -1;
    const int im = next_real(g, i);
    if (im < 0) return -1;
    const ggml_tensor * M = g->nodes[im];
    if (M->op == GGML_OP_MUL_MAT_ID && M->src[1] == G && !skipped(M)) {
        // ffn_down_exps: per-pair GEMV with the gate in its quantiser (decode: at most 8 (token, slot) pairs)
        const ggml_tensor * as = M->src[0], * ids = M->src[2], * a = G->src[0], * b = G->src[1];
        const int64_t n_used = ids->ne[0], n_tokens = ids->ne[1];
        if (n_used * n_tokens > 8 || G->ne[1] != n_used || in_split_buffer(as) || !ggml_is_contiguous(G) || !ggml_are_same_shape(a, b) || a->ne[0] != G->ne[0] ||
            a->nb[0] != sizeof(float) || b->nb[0] != sizeof(float) || a->nb[2] != a->nb[1] * a->ne[1] || b->nb[2] != b->nb[1] * b->ne[1] || a->ne[3] != 1) return -1;
        b200_mul_mat_id_args q = {};
        q.type = (int32_t) as->type; q.layout = weight_layout(as);
        q.m = as->ne[1]; q.k = as->ne[0]; q.n_as = as->ne[2]; q.n_used = n_used; q.nb1 = G->ne[1]; q.n_tokens = n_tokens;
        q.as = as->data; q.as_expert_stride = (int64_t) as->nb[2];
        q.b = (const float *) a->data; q.b_row_stride = (int64_t) (a->nb[1] / sizeof(float)); q.b_token_stride = (int64_t) (a->nb[2] / sizeof(float));
        q.ids = (const int32_t *) ids->data; q.ids_row_stride = (int64_t) (ids->nb[1] / sizeof(int32_t));
        q.y = (float *) M->data; q.y_row_stride = (int64_t) (M->nb[1] / sizeof(float)); q.y_token_stride = (int64_t) (M->nb[2] / sizeof(float));
        const size_t need = b200_mul_mat_id_workspace_ex(&q);
        if (!backend_reserve(bc, need)) { st = GGML_STATUS_ALLOC_FAILED; return im; }
        q.workspace = bc->workspace; q.workspace_size = bc->workspace_size; q.flavour = B200_Q8_X86; q.flags = mm_flags(); q.algo = 1;
        if (b200_mul_mat_id_glu(&q, (const float *) b->data, (int64_t) (b->nb[1] / sizeof(float)), ggml_get_op_params_i32(G, 0), ggml_get_op_params_i32(G, 1), bc->stream) != 0) {
            GGML_LOG_ERROR("ggml-b200: fused glu + MUL_MAT_ID (%s) failed: %s\n", M->name, b200_last_error()); st = GGML_STATUS_FAILED;
        }
        return im;
    }
    if (!plain_mul_mat(M) || M->src[1] != G || M->src[1]->ne[1] > 8) return -1;
    const ggml_tensor * a = G->src[0], * b = G->src[1];
    if (a->ne[2] != 1 || a->ne[3] != 1 || !ggml_are_same_shape(a, b) || a->ne[0] != G->ne[0] || a->nb[0] != sizeof(float) || b->nb[0] != sizeof(float) ||
        a->nb[1] % sizeof(float) || b->nb[1] % sizeof(float)) return -1;
    b200_mul_mat_args args = plain_args(M);
    args.x = (const float *) a->data; args.x_row_stride = (int64_t) (a->nb[1] / sizeof(float));
    const size_t need = b200_mul_mat_workspace(args.type, args.n, args.k);
    if (!backend_reserve(bc, need)) { st = GGML_STATUS_ALLOC_FAILED; return im; }
    args.workspace = bc->workspace; args.workspace_size = bc->workspace_size;
    if (b200_mul_mat_glu(&args, (const float *) b->data, (int64_t) (b->nb[1] / sizeof(float)), ggml_get_op_params_i32(G, 0), ggml_get_op_params_i32(G, 1), bc->stream) != 0) {
        GGML_LOG_ERROR("ggml-b200: fused glu + MUL_MAT (%s) failed: %s\n", M->name, b200_last_error()); st = GGML_STATUS_FAILED;
    }
    return im;
}

// SOFT_MAX -> ARGSORT(desc) -> GET_ROWS(probs, first n_used of the order) -> SUM_ROWS -> CLAMP -> DIV: the router of build_moe_ffn (src/llama-graph.cpp:1040-1140,
// softmax gating with renormalised weights) as one launch; returns the index of the DIV node (-1: no match)
const ggml_tensor * view_root(const ggml_tensor * t) { while (t && t->view_src) t = t->view_src; return t; }
#define MR_FAIL(n) do { static const bool dbg = getenv("GGML_B200_DEBUG_FUSE") != nullptr; if (dbg) fprintf(stderr, "ggml-b200: moe router not fused at node %d: check %d\n", i, n); } while (0)
int try_fuse_moe_route(backend_ctx * bc, ggml_cgraph * g, int i, enum ggml_status & st) {
    const ggml_tensor * sm = g->nodes[i];
    if (sm->src[1] || sm->src[2] || !sm->src[0] || sm->src[0]->type != GGML_TYPE_F32 || sm->ne[0] > 32 || sm->ne[2] != 1 || sm->ne[3] != 1) { MR_FAIL(1); return -1; }
    float scale, max_bias; memcpy(&scale, sm->op_params, sizeof(float)); memcpy(&max_bias, (const float *) sm->op_params + 1, sizeof(float));
    if (scale != 1.0f || max_bias != 0.0f) { MR_FAIL(2); return -1; }
    const int ia = next_real(g, i);  if (ia < 0) { MR_FAIL(3); return -1; }
    const int ig = next_real(g, ia); if (ig < 0) { MR_FAIL(4); return -1; }
    const int is = next_real(g, ig); if (is < 0) { MR_FAIL(5); return -1; }
    const int ic = next_real(g, is); if (ic < 0) { MR_FAIL(6); return -1; }
    const int id = next_real(g, ic); if (id < 0) { MR_FAIL(7); return -1; }
    const ggml_tensor * as = g->nodes[ia], * gr = g->nodes[ig], * su = g->nodes[is], * cl = g->nodes[ic], * dv = g->nodes[id];
    if (as->op != GGML_OP_ARGSORT || as->src[0] != sm || ggml_get_op_params_i32(as, 0) != (int32_t) GGML_SORT_ORDER_DESC) { MR_FAIL(8); return -1; }
    if (gr->op != GGML_OP_GET_ROWS || view_root(gr->src[0]) != sm || gr->src[0]->data != sm->data || view_root(gr->src[1]) != as || gr->src[1]->data != as->data) { MR_FAIL(9); return -1; }
    const int64_t n_expert = sm->ne[0], n_tok = sm->ne[1], n_used = gr->src[1]->ne[0];
    if (gr->src[0]->ne[0] != 1 || gr->src[0]->ne[1] != n_expert || gr->src[0]->ne[2] != n_tok || gr->src[1]->ne[1] != n_tok || gr->src[1]->nb[1] != as->nb[1] || n_used > n_expert ||
        gr->ne[0] != 1 || gr->ne[1] != n_used || gr->ne[2] != n_tok) { MR_FAIL(10); return -1; }
    if (su->op != GGML_OP_SUM_ROWS || view_root(su->src[0]) != gr || su->src[0]->data != gr->data || su->src[0]->ne[0] != n_used || su->src[0]->ne[1] != n_tok) { MR_FAIL(11); return -1; }
    if (cl->op != GGML_OP_CLAMP || view_root(cl->src[0]) != su || cl->src[0]->data != su->data || cl->data != su->data) { MR_FAIL(12); return -1; }       // (ggml_clamp works in place)
    if (dv->op != GGML_OP_DIV || view_root(dv->src[0]) != gr || dv->src[0]->data != gr->data || view_root(dv->src[1]) != su || dv->src[1]->data != su->data ||
        dv->ne[0] != n_used || dv->ne[1] != n_tok) { MR_FAIL(13); return -1; }
    { const ggml_tensor * cs[6] = {sm, sm->src[0], as, gr, su, dv}; for (const ggml_tensor * t : cs) if (!ggml_is_contiguous(t)) { MR_FAIL(14); return -1; } }
    if (as->type != GGML_TYPE_I32 || gr->type != GGML_TYPE_F32 || su->type != GGML_TYPE_F32 || dv->type != GGML_TYPE_F32 || sm->type != GGML_TYPE_F32) { MR_FAIL(15); return -1; }
    { const int qs[6] = {i, ia, ig, is, ic, id}; for (int q : qs) if (skipped(g->nodes[q])) { MR_FAIL(16); return -1; } }
    // probabilities, picked values and their sum are dead after these six nodes unless something further down reads them: ggml-alloc then hands their memory
    // to each other and to later tensors, so they are only written when they have readers (the order and the weights always are)
    bool need_p = false, need_g = false, need_s = false;
    for (int q = id + 1; q < g->n_nodes; ++q)
        for (int k = 0; k < GGML_MAX_SRC; ++k) {
            const ggml_tensor * r = view_root(g->nodes[q]->src[k]);
            need_p = need_p || (r && r == sm); need_g = need_g || (r && r == gr); need_s = need_s || (r && r == su);
        }
    const ggml_tensor * outs[5] = {need_p ? sm : nullptr, as, need_g ? gr : nullptr, need_s ? su : nullptr, dv};
    for (int a = 0; a < 5; ++a) for (int b = a + 1; b < 5; ++b)
        if (outs[a] && outs[b] && !(a == 2 && b == 4) && ranges_overlap(outs[a], outs[b])) { MR_FAIL(17); return -1; }
    // a warp reads its token's logits before it writes anything, but with several tokens another warp may already be writing: then no output may sit on the
    // logits (ggml-alloc reuses the memory of the dead logits / probabilities for the later tensors), except the probabilities of an in-place softmax
    if (n_tok > 1)
        for (int a = 0; a < 5; ++a)
            if (outs[a] && !(a == 0 && sm->data == sm->src[0]->data) && ranges_overlap(sm->src[0], outs[a])) { MR_FAIL(18); return -1; }
    float lo, hi; memcpy(&lo, cl->op_params, sizeof(float)); memcpy(&hi, (const float *) cl->op_params + 1, sizeof(float));
    const b200_tensor tl = bt(sm->src[0]), tp = bt(sm), to = bt(as), tg = bt(gr), ts = bt(su), tw = bt(dv);
    if (b200_op_moe_route(&tl, need_p ? &tp : nullptr, &to, need_g ? &tg : nullptr, need_s ? &ts : nullptr, &tw, (int) n_used, lo, hi, bc->stream) != 0) {
        GGML_LOG_ERROR("ggml-b200: fused MoE router (%s) failed: %s\n", dv->name, b200_last_error()); st = GGML_STATUS_FAILED;
    }
    return id;
}

enum ggml_status graph_issue(backend_ctx * bc, ggml_cgraph * cgraph) {
    static const bool fuse_on = env_on("GGML_B200_FUSE") && glue_on();
    std::vector<char> done((size_t) cgraph->n_nodes, 0);          // matmuls hoisted into an earlier group
    const ggml_tensor * last_id = nullptr, * keep_id = nullptr;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        ggml_tensor * node = cgraph->nodes[i];
        if (done[i] || ggml_is_empty(node) || skipped(node)) continue;
        enum ggml_status st = GGML_STATUS_SUCCESS;
        if (!is_noop(node)) { last_id = keep_id; keep_id = nullptr; }       // last_id: the MUL_MAT_ID that ran immediately before this node, if any
        { static const char * pre = getenv("GGML_B200_PRESYNC"); if (pre && !bc->capturing && !strcmp(pre, ggml_op_name(node->op))) b200_stream_synchronize(bc->stream); }
        if (plain_mul_mat(node)) {
            const int next = compute_decode_run(bc, cgraph, i, st, done);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (next > i) { i = next - 1; continue; }
        }
        if (fuse_on && (node->op == GGML_OP_ADD || node->op == GGML_OP_RMS_NORM)) {
            const int used = try_fuse_norm(bc, cgraph, i, st);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (used > 0) { i += used - 1; continue; }
        }
        if (fuse_on && node->op == GGML_OP_SOFT_MAX && !node->src[1]) {
            const int id = try_fuse_moe_route(bc, cgraph, i, st);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (id > i) { i = id; continue; }
        }
        if (fuse_on && node->op == GGML_OP_GLU) {
            const int im = try_fuse_glu(bc, cgraph, i, st);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (im > i) { i = im; continue; }
        }
        if (fuse_on && node->op == GGML_OP_ROPE) {
            const int last = try_fuse_kv(bc, cgraph, i, st);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (last > i) { i = last; continue; }
        }
        if (fuse_on && node->op == GGML_OP_MUL_MAT && node->src[0]->type == GGML_TYPE_F16) {
            const int ic = try_fuse_attn(bc, cgraph, i, st);
            if (st != GGML_STATUS_SUCCESS) return st;
            if (ic > i) { i = ic; continue; }
        }
        switch (node->op) {
            case GGML_OP_NONE:
            case GGML_OP_VIEW:
            case GGML_OP_RESHAPE:
            case GGML_OP_PERMUTE:
            case GGML_OP_TRANSPOSE:
                break;
            case GGML_OP_MUL_MAT:    st = weight_type(node->src[0]->type) ? compute_mul_mat(bc, node) : compute_glue(bc, node); break;
            case GGML_OP_MUL_MAT_ID: {
                // ffn_gate_exps after ffn_up_exps (or the other way round): same activations, same shapes, nothing in between -> quantised once
                const ggml_tensor * p = last_id;
                const bool same = p && p->src[1] == node->src[1] && p->src[2] == node->src[2] && p->src[0]->ne[0] == node->src[0]->ne[0] && p->src[0]->ne[1] == node->src[0]->ne[1] &&
                                  p->src[0]->ne[2] == node->src[0]->ne[2] && b200_act_type((int) p->src[0]->type) == b200_act_type((int) node->src[0]->type) &&
                                  node->src[2]->ne[0] * node->src[2]->ne[1] <= 8;       // (decode: the per-pair GEMV path on both calls)
                st = compute_mul_mat_id(bc, node, same);
                keep_id = node;
            } break;
            default:                 st = compute_glue(bc, node); break;
        }
        if (st != GGML_STATUS_SUCCESS) return st;
        // debugging aid: GGML_B200_SYNC=<op name|all> synchronises the stream after every node of that op (never during a capture)
        static const char * sync_env = getenv("GGML_B200_SYNC");
        if (sync_env && !bc->capturing && (!strcmp(sync_env, "all") || !strcmp(sync_env, ggml_op_name(node->op)) ||
                                           (node->op == GGML_OP_MUL_MAT && !strcmp(sync_env, weight_type(node->src[0]->type) ? "MUL_MAT_Q" : "MUL_MAT_F"))))
            b200_stream_synchronize(bc->stream);
    }
    return GGML_STATUS_SUCCESS;
}

// set / get in stream order (ggml_backend_i::set_tensor_async / get_tensor_async, ggml-backend-impl.h:92-95): plain tensors of this device's
// buffers are copied on the backend's stream; a tensor stored in the plane layout takes the synchronous, repack-aware buffer path
bool plane_tracked(const ggml_tensor * t) {
    buffer_ctx * bc = owner_ctx(t);
    if (!bc) return true;                      // not ours: let the buffer decide
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    std::lock_guard<std::mutex> lock(bc->mu);
    return bc->planes.count(base->data) != 0;
}
void backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    if (size == 0) return;
    const bool whole = offset == 0 && size == ggml_nbytes(tensor);
    if (plane_tracked(tensor) || (whole && repack_candidate(tensor))) { ggml_backend_tensor_set(tensor, data, offset, size); return; }
    b200_set_device(bc->device);
    if (b200_memcpy_h2d((char *) tensor->data + offset, data, size, bc->stream) != 0) GGML_ABORT("ggml-b200: set_tensor_async(%s) failed: %s", tensor->name, b200_last_error());
}
void backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    if (size == 0) return;
    if (plane_tracked(tensor)) { ggml_backend_tensor_get(tensor, data, offset, size); return; }
    b200_set_device(bc->device);
    if (b200_memcpy_d2h(data, (const char *) tensor->data + offset, size, bc->stream) != 0) GGML_ABORT("ggml-b200: get_tensor_async(%s) failed: %s", tensor->name, b200_last_error());
}

void backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    b200_event_record((b200_event_t) event->context, bc->stream);
}

void backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    backend_ctx * bc = (backend_ctx *) backend->context;
    b200_stream_wait_event(bc->stream, (b200_event_t) event->context);
}

const ggml_backend_i backend_iface = {
    /* .get_name           = */ backend_get_name,
    /* .free               = */ backend_free,
    /* .set_tensor_async   = */ backend_set_tensor_async,
    /* .get_tensor_async   = */ backend_get_tensor_async,
    /* .cpy_tensor_async   = */ nullptr,
    /* .synchronize        = */ backend_synchronize,
    /* .graph_plan_create  = */ nullptr,
    /* .graph_plan_free    = */ nullptr,
    /* .graph_plan_update  = */ nullptr,
    /* .graph_plan_compute = */ nullptr,
    /* .graph_compute      = */ backend_graph_compute,
    /* .event_record       = */ backend_event_record,
    /* .event_wait         = */ backend_event_wait,
    /* .graph_optimize     = */ nullptr,
};

ggml_guid_t backend_guid() {
    static ggml_guid guid = {0xb2, 0x00, 0x51, 0x00, 0xa0, 0x6d, 0x4d, 0x41, 0x54, 0x5f, 0x51, 0x34, 0x4b, 0x36, 0x4b, 0x38};
    return &guid;
}

// ---------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------
const char * device_get_name(ggml_backend_dev_t dev) { return ((device_ctx *) dev->context)->name.c_str(); }
const char * device_get_description(ggml_backend_dev_t dev) { return ((device_ctx *) dev->context)->description.c_str(); }

void device_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    device_ctx * dc = (device_ctx *) dev->context;
    size_t f = 0, t = 0;
    b200_device_info(dc->index, nullptr, 0, &f, &t, nullptr, nullptr, nullptr);
    *free = f; *total = t;
}

enum ggml_backend_dev_type device_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

void device_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name = device_get_name(dev);
    props->description = device_get_description(dev);
    props->type = GGML_BACKEND_DEVICE_TYPE_GPU;
    props->device_id = nullptr;
    device_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = {
        /* .async                = */ false,
        /* .host_buffer          = */ true,
        /* .buffer_from_host_ptr = */ false,
        /* .events               = */ true,
    };
}

ggml_backend_t device_init_backend(ggml_backend_dev_t dev, const char *) {
    device_ctx * dc = (device_ctx *) dev->context;
    backend_ctx * bc = new backend_ctx;
    bc->device = dc->index;
    bc->name = dc->name;
    bc->stream = b200_stream_create(dc->index);
    if (!bc->stream) {
        GGML_LOG_ERROR("ggml-b200: cannot create a stream on device %d: %s\n", dc->index, b200_last_error());
        delete bc;
        return nullptr;
    }
    return new ggml_backend{
        /* .guid    = */ backend_guid(),
        /* .iface   = */ backend_iface,
        /* .device  = */ dev,
        /* .context = */ bc,
    };
}

ggml_backend_buffer_type_t device_get_buffer_type(ggml_backend_dev_t dev) { return &((device_ctx *) dev->context)->buft; }
ggml_backend_buffer_type_t device_get_host_buffer_type(ggml_backend_dev_t dev) { return &((device_ctx *) dev->context)->host_buft; }

bool device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    // our own device memory on this device, or any host-visible buffer type is NOT enough: the kernels
    // dereference device pointers, so only buffers of this very device qualify
    if (buft->iface.get_name == split_buft_get_name) return buft->device == dev;       // the split type made for this device as main device
    return buft->iface.get_name == buft_get_name && buft->device == dev;
}

bool device_offload_op(ggml_backend_dev_t, const ggml_tensor *) { return false; }

ggml_backend_event_t device_event_new(ggml_backend_dev_t dev) {
    device_ctx * dc = (device_ctx *) dev->context;
    b200_event_t ev = b200_event_create(dc->index);
    if (!ev) return nullptr;
    return new ggml_backend_event{ /* .device = */ dev, /* .context = */ ev };
}
void device_event_free(ggml_backend_dev_t, ggml_backend_event_t event) {
    b200_event_destroy((b200_event_t) event->context);
    delete event;
}
void device_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) { b200_event_synchronize((b200_event_t) event->context); }

const ggml_backend_device_i device_iface = {
    /* .get_name             = */ device_get_name,
    /* .get_description      = */ device_get_description,
    /* .get_memory           = */ device_get_memory,
    /* .get_type             = */ device_get_type,
    /* .get_props            = */ device_get_props,
    /* .init_backend         = */ device_init_backend,
    /* .get_buffer_type      = */ device_get_buffer_type,
    /* .get_host_buffer_type = */ device_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ device_supports_op,
    /* .supports_buft        = */ device_supports_buft,
    /* .offload_op           = */ device_offload_op,
    /* .event_new            = */ device_event_new,
    /* .event_free           = */ device_event_free,
    /* .event_synchronize    = */ device_event_synchronize,
};

// ---------------------------------------------------------------------------------------------------
// registry
// ---------------------------------------------------------------------------------------------------
struct reg_ctx {
    std::vector<ggml_backend_device> devices;
    std::vector<device_ctx *>        ctxs;
};

const char * reg_get_name(ggml_backend_reg_t) { return "B200"; }
size_t reg_get_device_count(ggml_backend_reg_t reg) { return ((reg_ctx *) reg->context)->devices.size(); }
ggml_backend_dev_t reg_get_device(ggml_backend_reg_t reg, size_t index) {
    reg_ctx * rc = (reg_ctx *) reg->context;
    GGML_ASSERT(index < rc->devices.size());
    return &rc->devices[index];
}

ggml_backend_feature * reg_get_features(ggml_backend_reg_t) {
    static ggml_backend_feature features[] = {
        {"ARCH", "sm_100a"}, {"TCGEN05", "1"}, {"TMA", "1"}, {"REPACK", "planes"}, {nullptr, nullptr},
    };
    return features;
}

// proc "ggml_backend_split_buffer_type" (ggml-backend.h:199-200): main_device indexes this registry's device list, tensor_split holds one
// weight per device (all zero: equal shares). One buffer type object per distinct (main_device, split), kept for the life of the process.
ggml_backend_buffer_type_t split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mu;
    static std::vector<split_buft_ctx *> made;
    ggml_backend_reg_t reg = ggml_backend_init();
    reg_ctx * rc = (reg_ctx *) reg->context;
    const int n = (int) rc->devices.size();
    if (n < 1 || main_device < 0 || main_device >= n) return nullptr;
    float share[B200_MAX_DEVICES];
    float sum = 0.0f;
    for (int i = 0; i < n; ++i) { share[i] = tensor_split ? tensor_split[i] : 0.0f; if (share[i] < 0.0f) share[i] = 0.0f; sum += share[i]; }
    if (sum <= 0.0f) { for (int i = 0; i < n; ++i) share[i] = 1.0f; sum = (float) n; }
    float cum[B200_MAX_DEVICES + 1];
    cum[0] = 0.0f;
    for (int i = 0; i < n; ++i) cum[i + 1] = cum[i] + share[i] / sum;
    cum[n] = 1.0f;
    std::lock_guard<std::mutex> lock(mu);
    for (split_buft_ctx * c : made)
        if (c->main == main_device && c->n_dev == n && memcmp(c->cum, cum, sizeof(float) * (size_t) (n + 1)) == 0) return &c->buft;
    split_buft_ctx * c = new split_buft_ctx;
    c->main = main_device; c->n_dev = n;
    for (int i = 0; i < n; ++i) c->cuda_index[i] = rc->ctxs[i]->index;
    memcpy(c->cum, cum, sizeof(float) * (size_t) (n + 1));
    c->name = rc->ctxs[main_device]->name + "_Split";
    c->buft = ggml_backend_buffer_type{ /* .iface = */ split_buft_iface, /* .device = */ &rc->devices[main_device], /* .context = */ c };
    made.push_back(c);
    return &c->buft;
}

void * reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) reg_get_features;
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) split_buffer_type;
    return nullptr;
}

const ggml_backend_reg_i reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

int usable_devices(int * idx, int cap) {
    int n = b200_device_count(), out = 0;
    for (int i = 0; i < n && out < cap; ++i) {
        int major = 0, minor = 0;
        if (b200_device_info(i, nullptr, 0, nullptr, nullptr, nullptr, &major, &minor) == 0 && major == 10) idx[out++] = i;
    }
    return out;
}

} // namespace

// ---------------------------------------------------------------------------------------------------
// dlsym entry points (ggml-backend-reg.cpp:262-306)
// ---------------------------------------------------------------------------------------------------
extern "C" {

GGML_B200_API int ggml_backend_score(void) {
    int idx[B200_MAX_DEVICES];
    return usable_devices(idx, B200_MAX_DEVICES) > 0 ? 100 : 0;
}

GGML_B200_API ggml_backend_reg_t ggml_backend_init(void) {
    static ggml_backend_reg reg;
    static std::once_flag once;
    std::call_once(once, [] {
        reg_ctx * rc = new reg_ctx;
        int idx[B200_MAX_DEVICES];
        const int n = usable_devices(idx, B200_MAX_DEVICES);
        rc->devices.reserve(n);
        for (int i = 0; i < n; ++i) {
            device_ctx * dc = new device_ctx;
            char name[128] = "";
            b200_device_info(idx[i], name, sizeof(name), nullptr, nullptr, nullptr, nullptr, nullptr);
            dc->index = idx[i];
            dc->name = "B200_" + std::to_string(i);
            dc->description = name;
            dc->buft_name = dc->name;
            dc->host_buft_name = dc->name + "_Host";
            rc->ctxs.push_back(dc);
            rc->devices.push_back(ggml_backend_device{ /* .iface = */ device_iface, /* .reg = */ &reg, /* .context = */ dc });
        }
        for (int i = 0; i < n; ++i) {
            device_ctx * dc = rc->ctxs[i];
            dc->buft = ggml_backend_buffer_type{ /* .iface = */ buft_iface, /* .device = */ &rc->devices[i], /* .context = */ nullptr };
            dc->host_buft = ggml_backend_buffer_type{ /* .iface = */ host_buft_iface, /* .device = */ &rc->devices[i], /* .context = */ nullptr };
        }
        reg = ggml_backend_reg{ /* .api_version = */ GGML_BACKEND_API_VERSION, /* .iface = */ reg_iface, /* .context = */ rc };
    });
    return &reg;
}

} // extern "C"

/* b200_mm.h -- C ABI of the B200-native quantised matrix-multiply path.
 *
 * This is the drop-in boundary below the reference's backend plugin interface
 * (/root/reference/ggml/src/ggml-backend-impl.h:17-210). Every vtable entry of the ggml plugin
 * (include/ggml_b200.h, zyk-llama.cpp_b200/csrc/ggml_backend/ggml-b200.cpp) is a thin call into
 * one of the functions below, and the same functions are what tests/ and bench.py drive through
 * ctypes. Plain pointers and sizes only: no torch, no ggml and no CUDA types in any signature.
 *
 * Conventions
 *   - type ids are the reference's `enum ggml_type` values (ggml/include/ggml.h:389-430).
 *   - W is a quantised weight matrix of M rows x K elements (ggml ne00 = K, ne01 = M),
 *     X is f32 activations, N rows of K elements, Y is f32, N rows of M elements:
 *         Y[n][m] = sum_k W[m][k] * X[n][k]          (ggml_mul_mat, ggml/src/ggml.c:3176-3191)
 *   - all data pointers are DEVICE pointers unless the name says `host`.
 *   - `stream` is a b200_stream_t (0 = the device's default stream). Calls are asynchronous on
 *     that stream unless stated otherwise.
 *   - return value 0 = success, negative = error; b200_last_error() has the message. There is
 *     no CPU fallback anywhere: without a usable sm_100 device every compute entry point fails.
 */
#ifndef B200_MM_H
#define B200_MM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define B200_API
#else
#define B200_API __attribute__((visibility("default")))
#endif

/* ggml_type ids this path understands (ggml.h:389-430) */
enum b200_type {
    B200_TYPE_F32 = 0,
    B200_TYPE_F16 = 1,
    B200_TYPE_Q4_0 = 2,
    B200_TYPE_Q8_0 = 8,
    B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q4_K = 12,
    B200_TYPE_Q5_K = 13,
    B200_TYPE_Q6_K = 14,
    B200_TYPE_Q8_K = 15,
    B200_TYPE_IQ4_XS = 23,
    B200_TYPE_I32 = 26,
    B200_TYPE_I64 = 27,
};

/* how a weight matrix is stored in HBM */
enum b200_layout {
    B200_LAYOUT_CANONICAL = 0, /* ggml block structs, exactly the GGUF bytes (ggml-common.h:170-431) */
    B200_LAYOUT_PLANES = 1,    /* load-time repack: one plane per block field, see DESIGN.md          */
};

/* which of the reference's two non-identical Q8_0/Q8_1 activation quantisers to reproduce */
enum b200_q8_flavour {
    B200_Q8_REF = 0, /* quantize_row_q8_0_ref / _q8_1_ref   (ggml/src/ggml-quants.c:199-258)            */
    B200_Q8_X86 = 1, /* quantize_row_q8_0 / _q8_1 AVX2 path (ggml/src/ggml-cpu/arch/x86/quants.c:290-480)
                        -- what the reference CPU backend runs on x86, hence the default everywhere   */
};

typedef void * b200_stream_t;
typedef void * b200_event_t;
typedef void * b200_graph_t;

/* ---- errors ------------------------------------------------------------------------------- */
B200_API const char * b200_last_error(void);
B200_API const char * b200_version(void);

/* ---- devices, memory, streams (replaces ggml_backend_device_i::get_memory/get_props and
 *      ggml_backend_buffer_type_i::alloc_buffer, ggml-backend-impl.h:17-35,140-182) ------------ */
B200_API int  b200_device_count(void);
B200_API int  b200_device_info(int device, char * name, size_t name_len, size_t * free_bytes, size_t * total_bytes,
                               int * sm_count, int * cc_major, int * cc_minor);
B200_API int  b200_set_device(int device);
B200_API void * b200_malloc(int device, size_t bytes);        /* NULL on OOM */
B200_API int  b200_free(int device, void * ptr);
B200_API void * b200_malloc_host(size_t bytes);               /* pinned host memory */
B200_API int  b200_free_host(void * ptr);
B200_API b200_stream_t b200_stream_create(int device);
B200_API int  b200_stream_destroy(b200_stream_t stream);
B200_API int  b200_stream_synchronize(b200_stream_t stream);
B200_API int  b200_device_synchronize(int device);
/* buffer_i::set_tensor / get_tensor / memset_tensor / cpy_tensor / clear (ggml-backend-impl.h:41-58) */
B200_API int  b200_memcpy_h2d(void * dst, const void * host_src, size_t bytes, b200_stream_t stream);
B200_API int  b200_memcpy_d2h(void * host_dst, const void * src, size_t bytes, b200_stream_t stream);
B200_API int  b200_memcpy_d2d(void * dst, const void * src, size_t bytes, b200_stream_t stream);
B200_API int  b200_memcpy_peer(void * dst, int dst_device, const void * src, int src_device, size_t bytes, b200_stream_t stream);
B200_API int  b200_memset(void * dst, int value, size_t bytes, b200_stream_t stream);
/* strided copy between any two devices of this process / the host, unified addressing (ggml_cuda_Memcpy2DPeerAsync, ggml-cuda.cu:1403-1436) */
B200_API int  b200_memcpy2d(void * dst, size_t dpitch, const void * src, size_t spitch, size_t width, size_t height, b200_stream_t stream);
/* kernels of `device` may load / store the memory of `peer` afterwards (ggml_cuda_set_peer_access, ggml-cuda.cu:1438-1480); -1 when the
   two devices have no peer path */
B200_API int  b200_enable_peer_access(int device, int peer);
/* device_i::event_new/free/synchronize, backend_i::event_record/event_wait (ggml-backend-impl.h:112-116,178-181) */
B200_API b200_event_t b200_event_create(int device);
B200_API int  b200_event_destroy(b200_event_t ev);
B200_API int  b200_event_record(b200_event_t ev, b200_stream_t stream);
B200_API int  b200_event_synchronize(b200_event_t ev);
B200_API int  b200_stream_wait_event(b200_stream_t stream, b200_event_t ev);
B200_API int  b200_event_elapsed_ms(b200_event_t start, b200_event_t stop, float * ms);

/* ---- CUDA-graph capture of a sequence of calls on one stream (tg is launch bound) ----------- */
B200_API int  b200_graph_begin(b200_stream_t stream);
B200_API b200_graph_t b200_graph_end(b200_stream_t stream);   /* NULL on failure */
B200_API int  b200_graph_launch(b200_graph_t graph, b200_stream_t stream);
B200_API int  b200_graph_destroy(b200_graph_t graph);

/* ---- formats -------------------------------------------------------------------------------- */
B200_API int    b200_type_is_weight(int type);               /* Q4_0 Q8_0 Q4_K Q5_K Q6_K IQ4_XS     */
B200_API int    b200_block_elems(int type);                  /* ggml_blck_size                      */
B200_API size_t b200_block_bytes(int type);                  /* ggml_type_size                      */
B200_API size_t b200_row_size(int type, int64_t k);          /* ggml_row_size                       */
B200_API int    b200_act_type(int weight_type);              /* type_traits_cpu[].vec_dot_type      */

/* ---- (3) activation quantiser: f32 rows -> CANONICAL ggml q8 blocks, bit-exact ---------------
 * replaces quantize_row_q8_0 / _q8_1 / _q8_K (arch/x86/quants.c:290-495, ggml-quants.c:199-258,
 * 2555-2591) as called from ggml_compute_forward_mul_mat (ggml-cpu.c:1287-1322).
 *   x        : f32, `nrows` rows of `k` elements, row stride `x_row_stride` ELEMENTS
 *   y        : nrows * b200_row_size(q8_type, k) bytes, rows packed
 *   q8_type  : B200_TYPE_Q8_0 | Q8_1 | Q8_K;  flavour: enum b200_q8_flavour (ignored for Q8_K) */
B200_API int b200_quantize_rows(int q8_type, int flavour, const float * x, int64_t x_row_stride,
                                void * y, int64_t nrows, int64_t k, b200_stream_t stream);

/* ---- (5) load-time repack into the plane layout and its exact inverse ------------------------
 * replaces the buffer set_tensor repack hook (ggml_backend_cpu_repack_buffer_set_tensor,
 * ggml-cpu/repack.cpp:3141-3151); unrepack is what get_tensor needs to hand canonical bytes back.
 * src and dst must not overlap. One call handles one [m][k] matrix. */
B200_API int b200_repack_supported(int type, int64_t m, int64_t k);
B200_API int b200_repack(int type, int64_t m, int64_t k, const void * canonical, void * planes, b200_stream_t stream);
B200_API int b200_unrepack(int type, int64_t m, int64_t k, const void * planes, void * canonical, b200_stream_t stream);

/* ---- (1)(2) MUL_MAT -------------------------------------------------------------------------- */
typedef struct b200_mul_mat_args {
    int32_t type;          /* weight type                                                        */
    int32_t layout;        /* enum b200_layout of `w`                                            */
    int64_t m, n, k;
    const void *  w;       /* m rows; canonical: row stride = w_row_stride bytes (0 = packed)    */
    int64_t w_row_stride;  /* bytes; must be 0/packed for B200_LAYOUT_PLANES                     */
    const float * x;       /* n rows of k f32                                                    */
    int64_t x_row_stride;  /* elements                                                           */
    float *       y;       /* n rows of m f32                                                    */
    int64_t y_row_stride;  /* elements                                                           */
    void *  workspace;     /* >= b200_mul_mat_workspace(type, n, k) bytes, 256 B aligned         */
    size_t  workspace_size;
    int32_t flavour;       /* enum b200_q8_flavour for the Q8_0 activation quantiser             */
    int32_t algo;          /* 0 = auto, 1 = register GEMV, 2 = tensor-core GEMM, 3 = streaming GEMV */
    int32_t flags;         /* B200_MM_* bits                                                     */
    int32_t reserved;
} b200_mul_mat_args;

enum {
    B200_MM_X_PREQUANTISED = 1, /* workspace already holds the quantised X of a previous call with the
                                   same (x, n, k, activation type): skip the quantise kernel (QKV, gate/up) */
    B200_MM_PDL = 2,            /* launch with programmatic dependent launch: the kernel prefetches its
                                   weights while the previous kernel on the stream drains               */
};

B200_API size_t b200_mul_mat_workspace(int type, int64_t n, int64_t k);
/* replaces ggml_compute_forward_mul_mat (ggml-cpu.c:1227-1427) for a 2-D slice */
B200_API int    b200_mul_mat(const b200_mul_mat_args * args, b200_stream_t stream);
/* MUL_MAT whose activations are still to be gated: Y = W * (act(x) * x2), act = reglu / geglu / swiglu (enum ggml_glu_op 0 / 1 / 2;
   swapped: act(x2) * x). The GLU node in front of ffn_down (ggml_compute_forward_glu, ggml-cpu/ops.cpp) disappears into the activation
   quantiser; the reference CUDA backend fuses the same pair from the other side (mul_mat_vec_q<..., has_fusion>, mmvq.cu:173-218).
   args->x is x, n <= 8. */
B200_API int    b200_mul_mat_glu(const b200_mul_mat_args * args, const float * x2, int64_t x2_row_stride, int32_t glu_op, int32_t swapped,
                                 b200_stream_t stream);
/* `count` MUL_MAT nodes that read the same X (wq/wk/wv, ffn_gate/ffn_up: same x, n, k, strides): issued as ONE
 * streaming launch when n <= 4 and every matrix is in the plane layout, else as `count` b200_mul_mat calls.
 * Counterpart of the reference's fused gate/up GEMV (ggml_cuda_should_fuse_mul_mat_vec_q, ggml-cuda.cu:2144-2180). */
B200_API int    b200_mul_mat_group(const b200_mul_mat_args * args, int count, b200_stream_t stream);

/* ---- (1) decode CHAIN: a sequence of DEPENDENT MUL_MAT groups in one persistent launch ----------
 * replaces the per-node loop of graph_compute (ggml_backend_sched_compute_splits, ggml-backend.cpp:1443-1620, over
 * ggml_compute_forward_mul_mat, ggml-cpu.c:1227-1427) for the decode case (n <= 4): `nodes` lists the MUL_MAT nodes in
 * graph order, `group_sizes[g]` consecutive nodes form group g (they read the same X: wq/wk/wv, ffn_gate/ffn_up). Group g + 1
 * starts only when every row of group g has been written, exactly as if the groups had been issued one after the other
 * with b200_mul_mat_group on one stream, so a group may read what an earlier group wrote. One CTA per SM stays resident, the
 * weight stream (bulk copies into a shared-memory ring) runs ahead across the group boundaries. Pointers, shapes and
 * strides are baked in at create time (like a captured CUDA graph); workspace/algo/PDL fields are ignored.
 * Requirements (b200_chain_supported): plane layout (or canonical Q4_K), k % 256 == 0, the same n <= 4 everywhere. */
typedef struct b200_chain * b200_chain_t;
/* Row split of a chain over `world` GPUs, one process per GPU: nodes[i].m is THIS rank's row count (may be 0),
 * nodes[i].y points at column row_lo[i] of the full Y inside the symmetric arena (as for b200_mul_mat_gather). Every finished
 * row, every tagged word and every barrier arrival is stored into every rank's arena over NVLink by the kernel itself; the
 * launch completes on a rank only when all ranks' stores into its arena are done. `scratch`: zero-initialised region inside
 * the symmetric arena (same offset on every rank), 256 B aligned, >= b200_chain_scratch_bytes(max m_full, n) bytes. */
typedef struct b200_chain_split {
    int32_t world, rank;
    int64_t peer_delta[8];        /* as in b200_row_gather                                   */
    void *  scratch;
    size_t  scratch_bytes;
    const int64_t * row_lo;       /* per node: first row of this rank's slice                */
    const int64_t * m_full;       /* per node: rows of the whole matrix                      */
} b200_chain_split;
B200_API int          b200_chain_supported(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups);
B200_API b200_chain_t b200_chain_create(int device, const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups); /* NULL on failure */
B200_API size_t       b200_chain_scratch_bytes(int64_t max_m_full, int n);
B200_API b200_chain_t b200_chain_create_split(int device, const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups,
                                              const b200_chain_split * split);      /* split == NULL: b200_chain_create */
/* host-only view of the plan a chain would get on a device with `sms` SMs (no CUDA call, usable without a GPU; pointers are only
 * compared): modes[g] = how group g learns that its X is complete -- 0 first group (stream order), 1 arrival counter with release / acquire
 * fences, 2 relaxed arrival counter, 3 self-validating tagged words; *xreg = 1 when the activations-in-registers kernel takes the chain */
B200_API int          b200_chain_plan(const b200_mul_mat_args * nodes, const int32_t * group_sizes, int n_groups, int sms,
                                      const b200_chain_split * split, int32_t * modes, int32_t * xreg);
B200_API int          b200_chain_launch(b200_chain_t chain, b200_stream_t stream);   /* asynchronous, CUDA-graph capturable */
B200_API int          b200_chain_info(b200_chain_t chain, int * ctas, int * stage_bytes, int * smem_bytes, int * n_groups);
B200_API int          b200_chain_timeline(b200_chain_t chain, long long * host, size_t count);  /* debugging aid, see gemv_chain.cu */
B200_API int          b200_chain_destroy(b200_chain_t chain);

/* ---- (4) MUL_MAT_ID (ggml.c:3217-3243; CPU driver ggml-cpu.c:1500-1676) ----------------------
 *   as   : n_as expert matrices, each m rows x k, `as_expert_stride` bytes apart
 *   b    : f32 [n_tokens][nb1][k] (nb1 == n_used, or 1 = same row for every slot), strides in elements
 *   ids  : i32 [n_tokens][n_used] on the DEVICE, row stride `ids_row_stride` elements
 *   y    : f32 [n_tokens][n_used][m], strides in elements
 * No host synchronisation: expert routing is resolved on the device. */
typedef struct b200_mul_mat_id_args {
    int32_t type, layout;
    int64_t m, k, n_as, n_used, nb1, n_tokens;
    const void * as;   int64_t as_expert_stride;
    const float * b;   int64_t b_row_stride, b_token_stride;
    const int32_t * ids; int64_t ids_row_stride;
    float * y;         int64_t y_row_stride, y_token_stride;
    void * workspace;  size_t workspace_size;
    int32_t flavour, algo, flags, reserved;
} b200_mul_mat_id_args;

B200_API size_t b200_mul_mat_id_workspace(int type, int64_t k, int64_t n_used, int64_t nb1, int64_t n_tokens);
/* workspace for every path b200_mul_mat_id may take for these shapes (pointers in `args` are ignored). With n_tokens * n_used >= 8 pairs per
 * expert the call sorts the pairs by expert on the device (counterpart of mm_ids_helper, ggml-cuda/mmid.cu:28-117, and of the row grouping in
 * ggml_compute_forward_mul_mat_id, ggml-cpu/ggml-cpu.c:1589-1602), gathers their activation rows and runs ONE tensor-core GEMM launch over
 * (expert, tile) with the Y rows scattered back -- provided the workspace has this size; with the smaller b200_mul_mat_id_workspace() size it
 * stays on the per-pair GEMV. args->algo: 0 auto, 1 GEMV, 2 GEMM. */
B200_API size_t b200_mul_mat_id_workspace_ex(const b200_mul_mat_id_args * args);
B200_API int    b200_mul_mat_id(const b200_mul_mat_id_args * args, b200_stream_t stream);
/* MUL_MAT_ID whose activations are still to be gated (the GLU in front of ffn_down_exps), as b200_mul_mat_glu: b rows packed token by token, at most 8
   (token, slot) pairs */
B200_API int    b200_mul_mat_id_glu(const b200_mul_mat_id_args * args, const float * b2, int64_t b2_row_stride, int32_t glu_op, int32_t swapped,
                                    b200_stream_t stream);

/* ---- (e) multi-GPU row split ----------------------------------------------------------------
 * rows [*row_lo, *row_hi) of an m-row weight matrix owned by `rank` of `world` (tile rounded,
 * cf. get_row_split, ggml-cuda/ggml-cuda.cu:790-803). */
B200_API int b200_row_split(int64_t m, int world, int rank, int64_t * row_lo, int64_t * row_hi);

/* The exchange step, fused into the kernels (replaces the copy-back of ggml_cuda_op_mul_mat,
 * ggml-cuda/ggml-cuda.cu:1725-1744, and an ncclAllGather per node): every rank keeps its activations in a
 * SYMMETRIC ARENA (same size and layout on every rank, one process per GPU, mapped into the peers with
 * b200_ipc_import). The kernel that produces rows [lo, hi) of Y stores them into the arena of every rank over
 * NVLink (local address + peer_delta[g]); when its last CTA is done it adds one arrival to `sync[0]` of every rank.
 * The kernel that reads X of call (step, seq) first waits until (step * calls_per_step + seq) * world arrivals
 * are in, i.e. every rank has finished every earlier call. No host involvement, CUDA-graph replayable: the only
 * per-pass state is the device word `*step`, advanced by b200_row_gather_step once per pass. */
#define B200_MAX_PEERS 8
typedef struct b200_row_gather {
    int32_t  world, rank;
    int64_t  peer_delta[B200_MAX_PEERS]; /* bytes from a local arena address to the same location on rank g (0 for g == rank) */
    uint64_t * sync;                     /* two zero-initialised words INSIDE the local arena: [0] arrivals, [1] CTA counter  */
    const uint64_t * step;               /* local device word: passes started so far (>= 1 while a pass runs)                 */
    int64_t  seq;                        /* index of this call among the gather calls of one pass, 0 .. calls_per_step - 1    */
    int64_t  calls_per_step;
} b200_row_gather;

/* *step += 1, in stream order: issue once at the start of every pass (decode step / prefill ubatch) */
B200_API int b200_row_gather_step(uint64_t * step, b200_stream_t stream);
/* b200_mul_mat_group(args, count) on this rank's row slices (args[i].y points at column `lo` of the full Y inside the
 * arena, args[i].m = hi - lo) + the fused exchange described above: one wait, one arrival per call. */
B200_API int b200_mul_mat_gather(const b200_mul_mat_args * args, int count, const b200_row_gather * gather, b200_stream_t stream);
/* block the stream until every rank has finished calls 0 .. gather->seq - 1 of the current pass (seq == calls_per_step: the
 * whole pass): issue before the host or a foreign kernel reads a gathered Y */
B200_API int b200_row_gather_wait(const b200_row_gather * gather, b200_stream_t stream);
/* peer mapping of an arena allocated with b200_malloc: 64-byte handle out / in (cudaIpcMemHandle_t) */
B200_API int b200_ipc_export(const void * dev_ptr, void * handle64);
B200_API int b200_ipc_import(const void * handle64, void ** dev_ptr);
B200_API int b200_ipc_close(void * dev_ptr);

/* ---- the other ops of a llama-arch graph: what sits between the quantised MUL_MATs of a layer, so that the reference's
 * scheduler keeps the whole graph on this device. Semantics of the reference CPU backend (ggml/src/ggml-cpu/ops.cpp), restricted to the
 * variants the plugin claims in supports_op. A tensor is described the way ggml describes it (ggml.h struct ggml_tensor: data, type, ne[4]
 * elements per dimension, nb[4] strides in BYTES); all pointers are device pointers. */
typedef struct b200_tensor {
    void *  data;
    int32_t type;        /* enum b200_type (= enum ggml_type) */
    int32_t pad;
    int64_t ne[4];
    int64_t nb[4];
} b200_tensor;
/* ggml_compute_forward_get_rows: dst f32 [ne00, n...] = rows ids[i] (i32) of src0 (f32 / f16 / any of the five block formats, CANONICAL layout) */
B200_API int b200_op_get_rows(const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst, b200_stream_t stream);
/* ggml_compute_forward_rms_norm_f32 (ops.cpp:3640-3710): y = x / sqrt(mean(x^2) + eps) per row */
B200_API int b200_op_rms_norm(const b200_tensor * src, const b200_tensor * dst, float eps, b200_stream_t stream);
/* fused [ADD ->] RMS_NORM -> MUL(weight): sum = a (+ b); dst = sum / sqrt(mean(sum^2) + eps) * w; `b` and `sum` may be NULL (no add / sum not kept).
 * The residual pattern of every llama layer; the reference CUDA backend fuses the same nodes (ggml-cuda.cu ggml_cuda_can_fuse, rms_norm + mul [+ add]) */
B200_API int b200_op_add_rms_norm_mul(const b200_tensor * a, const b200_tensor * b, const b200_tensor * sum, const b200_tensor * w, const b200_tensor * dst,
                                      float eps, b200_stream_t stream);
/* fused decode attention for a few query tokens: dst[head_dim, n_head, n_tokens] = softmax(scale * K q + mask) V, replacing the
 * MUL_MAT -> SOFT_MAX -> MUL_MAT -> PERMUTE -> CONT sequence of build_attn_mha (src/llama-graph.cpp) when flash attention is off.
 * k: f16 [head_dim, n_kv, n_head_kv]; v: f16 [n_kv, head_dim, n_head_kv] (transposed V cache); q: f32 [head_dim, n_tokens, n_head]; n_kv <= 12000 */
B200_API int b200_op_attn_decode(const b200_tensor * k, const b200_tensor * q, const b200_tensor * mask, const b200_tensor * v, const b200_tensor * dst,
                                 float scale, b200_stream_t stream);
/* the same run of nodes for a prefill ubatch (any number of query tokens): one pass over the KV cache with an online softmax, both products on
   the tensor cores, the score matrix never written; head_dim 64 or 128, n_kv a multiple of 64, 16-byte aligned rows. The reference reaches
   this only through FLASH_ATTN_EXT (ggml-cuda/fattn*.cu); here the plain MUL_MAT / SOFT_MAX / MUL_MAT graph gets it */
B200_API int b200_op_attn_prefill(const b200_tensor * k, const b200_tensor * q, const b200_tensor * mask, const b200_tensor * v, const b200_tensor * dst,
                                  float scale, b200_stream_t stream);
/* ggml_compute_forward_add / _mul / _div, f32, src1 broadcast over src0 (ggml_can_repeat); op: 0 = add, 1 = mul, 2 = div */
B200_API int b200_op_binary(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, b200_stream_t stream);
/* the router of a mixture-of-experts layer (build_moe_ffn, src/llama-graph.cpp:1000-1150): ggml_compute_forward_scale (op 0: x * p0 + p1) and
   _clamp (op 1: to [p0, p1]); _sum_rows (f32 [k, ...] -> [1, ...]); _argsort_f32 (i32 indices that sort each row; order 0 ascending, 1
   descending as enum ggml_sort_order; rows of up to 1024 elements) -- ggml-cpu/ops.cpp */
B200_API int b200_op_unary(int op, const b200_tensor * src, const b200_tensor * dst, float p0, float p1, b200_stream_t stream);
B200_API int b200_op_sum_rows(const b200_tensor * src, const b200_tensor * dst, b200_stream_t stream);
B200_API int b200_op_argsort(const b200_tensor * src, const b200_tensor * dst, int order, b200_stream_t stream);
/* the six router nodes of build_moe_ffn in one launch (softmax gating, weights renormalised over the used experts: Mixtral): probs = softmax(logits),
   order = argsort(probs, descending), picked = probs at the first n_used indices, sum = clamp(sum(picked), lo, hi), weights = picked / sum.
   logits [n_expert <= 32, n_tokens]; every output is the contiguous tensor of its node; probs, picked and sum may be NULL (not needed by anyone else). */
B200_API int b200_op_moe_route(const b200_tensor * logits, const b200_tensor * probs, const b200_tensor * order, const b200_tensor * picked, const b200_tensor * sum,
                               const b200_tensor * weights, int n_used, float clamp_lo, float clamp_hi, b200_stream_t stream);
/* ggml_compute_forward_rope_flt (ops.cpp:5655-5790) for mode 0 (normal) and 2 (neox), ext_factor == 0; pos i32 [n_tokens], freq_factors f32 [n_dims / 2] or NULL */
typedef struct b200_rope_params { int32_t n_dims, mode; float freq_base, freq_scale, ext_factor, attn_factor; } b200_rope_params;
B200_API int b200_op_rope(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                          const b200_rope_params * params, b200_stream_t stream);
/* ggml_compute_forward_set_rows: dst row idx[i] (i64 / i32) = src row i; src f32, dst f32 / f16 (KV cache writes) */
B200_API int b200_op_set_rows(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst, b200_stream_t stream);
/* ROPE whose result also goes, converted, into the rows of a cache tensor: the ROPE(k) -> SET_ROWS(k cache) pair of build_attn
   (src/llama-graph.cpp; reference kernels ggml-cuda/rope.cu + set-rows.cu, two launches there), plus optionally an independent second
   SET_ROWS (v cache) in the same launch. sr_src is the SET_ROWS source: a contiguous reshape of dst (same data pointer). */
B200_API int b200_op_rope_set_rows(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                                   const b200_rope_params * params, const b200_tensor * sr_src, const b200_tensor * sr_idx, const b200_tensor * sr_dst,
                                   const b200_tensor * v_src, const b200_tensor * v_idx, const b200_tensor * v_dst, b200_stream_t stream);
/* ggml_compute_forward_dup (CPY / CONT / DUP) between f32 and f16, any strides */
B200_API int b200_op_cpy(const b200_tensor * src, const b200_tensor * dst, b200_stream_t stream);
/* ggml_compute_forward_soft_max_f32 (ops.cpp:5160-5280): softmax(src * scale + mask) per row; mask f32 / f16 or NULL; max_bias == 0, no sinks */
B200_API int b200_op_soft_max(const b200_tensor * src, const b200_tensor * mask, const b200_tensor * dst, float scale, b200_stream_t stream);
/* ggml_compute_forward_glu: dst = act(x) * g; glu_op 0 reglu, 1 geglu, 2 swiglu (enum ggml_glu_op); b == NULL: x, g are the halves of a's rows */
B200_API int b200_op_glu(const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, int glu_op, int swapped, b200_stream_t stream);
/* ggml_compute_forward_mul_mat for f16 / f32 src0 and f32 src1 (attention over the KV cache: K^T Q and V^T softmax), batched with broadcast */
B200_API int b200_op_mul_mat_f(const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, b200_stream_t stream);

/* ---- MUL_MAT without any allocation --------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it). Same result as b200_mul_mat with
 * algo = 1 (activation quantiser + register GEMV; meant for decode, any n is accepted while the scratch fits), but the
 * quantised activations go into a fixed file-scope device array of b200_noalloc_scratch_max() bytes (1 MiB, held by every
 * GPU that loads the library until exit) instead of args->workspace, which is ignored. The call allocates nothing (no
 * cudaMalloc / cudaFree / stream or event creation). Shapes whose scratch (b200_mul_mat_noalloc_scratch(type, n, k)) exceeds
 * the array are rejected with an error. All calls on a device share the array, which keeps the last call's contents (every
 * byte read is written first, nothing must be cleared): calls that overlap on the GPU race; calls on one stream do not. */
B200_API size_t b200_noalloc_scratch_max(void);
B200_API size_t b200_mul_mat_noalloc_scratch(int type, int64_t n, int64_t k);
B200_API int    b200_mul_mat_noalloc(const b200_mul_mat_args * args, b200_stream_t stream);

/* number of kernels this library has launched since load (bench.py's gpu_launches) */
B200_API uint64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_MM_H */

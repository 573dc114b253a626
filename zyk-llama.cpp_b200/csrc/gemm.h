// gemm.h -- internal interface of the prefill tensor-core GEMM (gemm.cu).
#pragma once
#include "common.cuh"

size_t b200_gemm_workspace(int type, int64_t n, int64_t k);
// does the GEMM take this shape? (force = caller set algo == 2)
bool   b200_gemm_wants(int type, int64_t m, int64_t n, int64_t k, bool force);
int    b200_gemm_launch(const b200_mul_mat_args * a, cudaStream_t s);
// up to 3 nodes of the same weight type that read the same X (and share the workspace): one launch, one tile list
bool   b200_gemm_groupable(const b200_mul_mat_args * args, int count);
int    b200_gemm_launch_group(const b200_mul_mat_args * args, int count, cudaStream_t s);
// same; `after_prologue` (may be null) is recorded on `s` once the fp16 copy of X is in the workspace (before the GEMM itself)
int    b200_gemm_launch_group_ev(const b200_mul_mat_args * args, int count, cudaStream_t s, cudaEvent_t after_prologue, const PeerGather * peer = nullptr);
// row split: one node whose epilogue stores Y into every rank's arena and whose last CTA announces the end of the call (peer->signal)
int    b200_gemm_launch_peer(const b200_mul_mat_args * a, cudaStream_t s, const PeerGather * peer);
// MUL_MAT_ID on the tensor cores: device-side sort of the (token, slot) pairs by expert, gathered fp16 activations, ONE grouped launch
// over (expert, weight tile, n-tile) with scattered Y rows. y_span = largest element offset of a Y row + m (must fit 31 bits).
size_t b200_gemm_moe_workspace(int type, int64_t m, int64_t k, int64_t n_as, int64_t pairs);
bool   b200_gemm_moe_wants(int type, int64_t m, int64_t k, int64_t n_as, int64_t n_used, int64_t n_tokens, int64_t y_span, bool force);
int    b200_gemm_moe_launch(const b200_mul_mat_id_args * a, cudaStream_t s);

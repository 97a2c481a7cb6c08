// Internal declarations for the cross-entropy and scoring/top-k paths.
#pragma once
#include "common.cuh"

namespace bsa {

// ---- FFMA (fp32) chunked paths: ce.cu / topk.cu
int ce_chunk_cols(int B, long long N);
size_t ce_simt_workspace_bytes(int B, long long N);
int ce_fwd_simt(int B, int d, long long N, const float* s, const float* E, const long long* answers, float* loss,
                float* lse, void* ws, size_t ws_bytes, cudaStream_t st);
int ce_bwd_simt(int B, int d, long long N, const float* s, const float* E, const long long* answers, const float* lse,
                const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, cudaStream_t st);

size_t topk_simt_workspace_bytes(int B, long long N);
int score_topk_simt(int B, int d, int k, long long rows, long long offset, const float* s, const float* E,
                    const long long* hist_indptr, const long long* hist_indices, const long long* hist_rows, float* top_val,
                    long long* top_idx, void* ws, size_t ws_bytes, cudaStream_t st);
int topk_merge(const float* cand_val, const long long* cand_idx, int G, int B, int k, float* top_val,
               long long* top_idx, cudaStream_t st);

}  // namespace bsa

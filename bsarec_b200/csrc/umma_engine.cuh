// tcgen05 (UMMA) 3xTF32 engine: user x item contractions with fused epilogues.
#pragma once
#include "common.cuh"

namespace bsa {

bool umma_ce_supported(int B, int d, long long N);
size_t umma_ce_workspace_bytes(int B, int d, long long N);
int umma_ce_fwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, float* loss,
                float* lse, void* ws, size_t ws_bytes, cudaStream_t st);
int umma_ce_bwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, const float* lse,
                const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, cudaStream_t st);

bool umma_topk_supported(int B, int d, int k, long long rows);
size_t umma_topk_workspace_bytes(int B, int d, int k, long long rows);
int umma_score_topk(int B, int d, int k, long long rows, long long offset, const float* s, const float* E,
                    const long long* hist_indptr, const long long* hist_indices, float* top_val, long long* top_idx,
                    void* ws, size_t ws_bytes, cudaStream_t st);

}  // namespace bsa

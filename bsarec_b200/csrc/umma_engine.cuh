// tcgen05 (UMMA) 3xTF32 engine: user x item contractions with fused epilogues.
#pragma once
#include "common.cuh"

namespace bsa {

bool umma_ce_supported(int B, int d, long long N);
size_t umma_ce_workspace_bytes(int B, int d, long long N);
int umma_ce_fwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, float* loss,
                float* lse, void* ws, size_t ws_bytes, cudaStream_t st, float* tgt_out = nullptr);
int umma_ce_bwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, const float* lse,
                const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, bool reuse_split, cudaStream_t st);

void set_ce_fused(int mode);
int get_ce_fused();

bool umma_topk_supported(int B, int d, int k, long long rows);
size_t umma_topk_workspace_bytes(int B, int d, int k, long long rows);
int umma_score_topk(int B, int d, int k, long long rows, long long offset, const float* s, const float* E,
                    const long long* hist_indptr, const long long* hist_indices, const long long* hist_rows, float* top_val,
                    long long* top_idx, void* ws, size_t ws_bytes, bool reuse_split, cudaStream_t st);

// generic C = alpha * A B^T (+bias)(act)(+C) on the tcgen05 engine; operands given unsplit (see umma_engine.cu)
size_t umma_gemm_workspace_bytes(long long a_elems, long long b_elems, int M, int N, int K);
int umma_gemm(int M, int N, int K, const float* A, long long a_rows, long long a_cols, long long lda, bool a_mn,
              const float* B, long long b_rows, long long b_cols, long long ldb, bool b_mn, float* C, long long ldc,
              const float* bias, int act, float* aux_out, const float* aux_in, float alpha, int accumulate, void* ws,
              size_t ws_bytes, cudaStream_t st);

// the same with operands supplied as split-TF32 (hi, lo) pairs (produced once by umma_split or by a producer
// epilogue and reused by several GEMMs); out_hi/out_lo (optional) also receive the result as a split pair.
int umma_split(const float* x, float* hi, float* lo, long long n, cudaStream_t st);
// lo = x - truncate_tf32(x): with it the fp32 tensor x itself is the hi half of its pair (the tensor core reads only the upper
// 19 bits of a tf32 operand word), so the split writes 4 B per element and no hi copy exists
int umma_split_lo(const float* x, float* lo, long long n, cudaStream_t st);
// the same for up to 8 tensors in one launch (16-byte aligned)
int umma_split_multi(int count, const float* const* x, float* const* hi, float* const* lo, const long long* n, cudaStream_t st);
size_t umma_gemm_partial_bytes(int M, int N);
// One pass: split x [M, N] into (hi, lo) and leave per-128-row column sums in part [umma_split_colsum_rows(M), N]
// (reduce_partials turns them into the bias gradient of the linear whose output gradient x is).
bool umma_split_colsum_ok(int N, const float* x, const float* hi, const float* lo);
int umma_split_colsum_rows(long long M);
// out_pitch / in_pitch (floats, 0 = N): the pair / x may be column windows of wider [M, pitch] buffers.  hi == nullptr: x is
// its own hi half, only lo = x - truncate_tf32(x) is written (see umma_split_lo)
int umma_split_colsum(const float* x, float* hi, float* lo, long long M, int N, float* part, cudaStream_t st, long long out_pitch = 0,
                      long long in_pitch = 0);
int umma_gemm_ps(int M, int N, int K, const float* a_hi, const float* a_lo, long long a_rows, long long a_cols,
                 long long lda, bool a_mn, const float* b_hi, const float* b_lo, long long b_rows, long long b_cols,
                 long long ldb, bool b_mn, float* C, long long ldc, const float* bias, int act, float* aux_out,
                 const float* aux_in, float alpha, int accumulate, float* out_hi, float* out_lo, float* partial,
                 size_t partial_bytes, cudaStream_t st, float* colpart = nullptr,
                 // segmented output: columns [s * seg_cols, (s + 1) * seg_cols) go to C + s * seg_stride (pitch ldc) with
                 // bias / bias1 / bias2 — q, k, v in one launch over one read of A (plain epilogue, seg_cols % 32 == 0)
                 int seg_cols = 0, long long seg_stride = 0, const float* bias1 = nullptr, const float* bias2 = nullptr,
                 // row-segmented output (split-K launches only, see umma_gemm_splits): rows [s * seg_rows, (s + 1) * seg_rows)
                 // go to C / C1 / C2 — the three weight gradients of q, k, v from one GEMM
                 int seg_rows = 0, float* C1 = nullptr, float* C2 = nullptr);
int umma_gemm_splits(int M, int N, int K);

}  // namespace bsa

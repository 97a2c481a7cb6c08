// Generic batched fp32 FFMA GEMM used wherever a contraction has not (yet) been moved to
// the tcgen05 engine, and as the on-GPU fp32 cross-check for it.
//   C[z](m,n) = alpha * sum_k A[z](m,k) * B[z](n,k)  (+ bias[n]) (act) (+ C[z](m,n) if accumulate)
// a_kmajor: A(m,k) = A[m*lda + k]   else A(m,k) = A[k*lda + m]
// b_kmajor: B(n,k) = B[n*ldb + k]   else B(n,k) = B[k*ldb + n]
// z = zo*inner + zi ; operand offsets zo*s?o + zi*s?i  (two-level batch: sequence x head)
#pragma once
#include "common.cuh"

namespace bsa {

enum GemmAct : int {
    ACT_NONE = 0,
    ACT_GELU_SAVE_PRE = 1,    // aux_out = pre-activation, C = gelu(pre)
    ACT_MUL_GELU_GRAD = 2,    // C = value * gelu'(aux_in)
    ACT_GELU_SAVE_GRAD = 3,   // aux_out = gelu'(pre-activation), C = gelu(pre): erf is evaluated once, in the forward
    ACT_MUL_AUX = 4,          // C = value * aux_in (the derivative saved by ACT_GELU_SAVE_GRAD)
};

struct GemmP {
    const float* A = nullptr;
    const float* B = nullptr;
    float* C = nullptr;
    int M = 0, N = 0, K = 0;
    int lda = 0, ldb = 0, ldc = 0;
    bool a_kmajor = true, b_kmajor = true;
    int batch = 1, inner = 1;
    long long sAo = 0, sAi = 0, sBo = 0, sBi = 0, sCo = 0, sCi = 0;
    float alpha = 1.f;
    const float* bias = nullptr;
    int accumulate = 0;
    int act = ACT_NONE;
    float* aux_out = nullptr;        // same indexing as C
    const float* aux_in = nullptr;   // same indexing as C
    // split-K (batch must be 1): partial sums go to `partial` [splitk][M][N], then a fixed-order
    // reduction applies alpha/accumulate -> deterministic, atomic-free.
    int splitk = 1;
    float* partial = nullptr;
};

int launch_gemm(const GemmP& p, cudaStream_t stream);

// number of K-splits that fills the machine for an [M,N] output contracted over K
int choose_splitk(int M, int N, int K);

}  // namespace bsa

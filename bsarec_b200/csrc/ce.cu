// Full-catalog cross-entropy (model/bsarec.py:32-35) without a [B, item_size] logits tensor.
// FFMA path: the catalog is swept in L2-sized column chunks; per chunk a GEMM produces logits that stay
// in L2, an online (max, sum-exp) update folds them into the running log-sum-exp, and the chunk is dropped.
#include "../../include/bsarec_b200.h"
#include "common.cuh"
#include "gemm_simt.cuh"
#include "ce_topk.cuh"

namespace bsa {

namespace {

constexpr int ROW_THREADS = 256;

__device__ __forceinline__ float block_max(float v, float* sh) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sh[0];
    for (int i = 1; i < ROW_THREADS / 32; ++i) r = fmaxf(r, sh[i]);
    __syncthreads();
    return r;
}
__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    for (int i = 0; i < ROW_THREADS / 32; ++i) r += sh[i];     // fixed order
    __syncthreads();
    return r;
}

__global__ void ce_init_kernel(float* run_m, float* run_s, float* tgt, int B) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B) { run_m[i] = -INFINITY; run_s[i] = 0.f; tgt[i] = 0.f; }
}

// one CTA per row: fold logits[row, 0:ncols) (global columns n0..) into (run_m, run_s); pick the target logit
__global__ void __launch_bounds__(ROW_THREADS)
ce_lse_chunk_kernel(const float* __restrict__ logits, int ld, long long n0, int ncols, const long long* __restrict__ answers,
                    float* __restrict__ run_m, float* __restrict__ run_s, float* __restrict__ tgt) {
    __shared__ float sh[ROW_THREADS / 32];
    const int row = blockIdx.x;
    const float* __restrict__ lr = logits + (long long)row * ld;
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < ncols; j += ROW_THREADS) mx = fmaxf(mx, lr[j]);
    mx = block_max(mx, sh);
    const float m_old = run_m[row];
    const float m_new = fmaxf(m_old, mx);
    float s = 0.f;
    for (int j = threadIdx.x; j < ncols; j += ROW_THREADS) s += expf(lr[j] - m_new);
    s = block_sum(s, sh);
    if (threadIdx.x == 0) {
        const float s_old = run_s[row];
        run_s[row] = (s_old > 0.f ? s_old * expf(m_old - m_new) : 0.f) + s;
        run_m[row] = m_new;
        const long long a = answers[row] - n0;
        if (a >= 0 && a < ncols) tgt[row] = lr[a];
    }
}

__global__ void ce_finalize_kernel(const float* __restrict__ run_m, const float* __restrict__ run_s,
                                   const float* __restrict__ tgt, float* __restrict__ lse, float* __restrict__ loss, int B) {
    __shared__ float sh[ROW_THREADS / 32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < B; i += ROW_THREADS) {
        const float l = run_m[i] + logf(run_s[i]);
        lse[i] = l;
        acc += l - tgt[i];
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) *loss = acc / (float)B;
}

// G = (softmax - onehot) * d_loss / B, in place on the logits chunk
__global__ void ce_grad_chunk_kernel(float* __restrict__ logits, int ld, long long n0, int ncols, int B,
                                     const long long* __restrict__ answers, const float* __restrict__ lse,
                                     const float* __restrict__ d_loss) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * ncols) return;
    const int row = (int)(i / ncols), j = (int)(i - (long long)row * ncols);
    const float scale = __ldg(d_loss) / (float)B;
    float p = expf(logits[(long long)row * ld + j] - lse[row]);
    if (answers[row] == n0 + j) p -= 1.0f;
    logits[(long long)row * ld + j] = p * scale;
}

}  // namespace

int ce_chunk_cols(int B, long long N) {
    long long nc = (32ll << 20) / (4ll * (B > 0 ? B : 1));     // <= 32 MiB of logits: stays in the 126 MB L2
    nc = nc / 128 * 128;
    if (nc < 1024) nc = 1024;
    if (nc > N) nc = N;
    return (int)nc;
}

size_t ce_simt_workspace_bytes(int B, long long N) {
    return ((size_t)B * ce_chunk_cols(B, N) + 3 * (size_t)B) * sizeof(float) + 1024;
}

int ce_fwd_simt(int B, int d, long long N, const float* s, const float* E, const long long* answers, float* loss,
                float* lse, void* ws, size_t ws_bytes, cudaStream_t st) {
    BSA_REQUIRE(ws_bytes >= ce_simt_workspace_bytes(B, N), "bsa_ce_fwd: workspace too small");
    const int nc = ce_chunk_cols(B, N);
    float* run_m = (float*)ws;
    float* run_s = run_m + B;
    float* tgt = run_s + B;
    float* logits = tgt + B + (64 - (3 * B) % 64) % 64;
    ce_init_kernel<<<cdiv(B, 256), 256, 0, st>>>(run_m, run_s, tgt, B);
    BSA_LAUNCH_CHECK();
    for (long long n0 = 0; n0 < N; n0 += nc) {
        const int cols = (int)((N - n0) < nc ? (N - n0) : nc);
        GemmP p;
        p.A = s; p.B = E + n0 * d; p.C = logits; p.M = B; p.N = cols; p.K = d; p.lda = d; p.ldb = d; p.ldc = nc;
        int rc = launch_gemm(p, st);
        if (rc) return rc;
        ce_lse_chunk_kernel<<<B, ROW_THREADS, 0, st>>>(logits, nc, n0, cols, answers, run_m, run_s, tgt);
        BSA_LAUNCH_CHECK();
    }
    ce_finalize_kernel<<<1, ROW_THREADS, 0, st>>>(run_m, run_s, tgt, lse, loss, B);
    BSA_LAUNCH_CHECK();
    return 0;
}

int ce_bwd_simt(int B, int d, long long N, const float* s, const float* E, const long long* answers, const float* lse,
                const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, cudaStream_t st) {
    BSA_REQUIRE(ws_bytes >= ce_simt_workspace_bytes(B, N), "bsa_ce_bwd: workspace too small");
    const int nc = ce_chunk_cols(B, N);
    float* logits = (float*)ws + 3 * B + (64 - (3 * B) % 64) % 64;
    BSA_CHECK_CUDA(cudaMemsetAsync(ds, 0, (size_t)B * d * sizeof(float), st));
    for (long long n0 = 0; n0 < N; n0 += nc) {
        const int cols = (int)((N - n0) < nc ? (N - n0) : nc);
        int rc;
        {
            GemmP p;
            p.A = s; p.B = E + n0 * d; p.C = logits; p.M = B; p.N = cols; p.K = d; p.lda = d; p.ldb = d; p.ldc = nc;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        ce_grad_chunk_kernel<<<cdiv((long long)B * cols, 256), 256, 0, st>>>(logits, nc, n0, cols, B, answers, lse, d_loss);
        BSA_LAUNCH_CHECK();
        {   // ds += G E_chunk
            GemmP p;
            p.A = logits; p.B = E + n0 * d; p.C = ds; p.M = B; p.N = d; p.K = cols; p.lda = nc; p.ldb = d; p.ldc = d;
            p.b_kmajor = false; p.accumulate = 1;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        {   // dE_chunk = G^T s
            GemmP p;
            p.A = logits; p.B = s; p.C = dE + n0 * d; p.M = cols; p.N = d; p.K = B; p.lda = nc; p.ldb = d; p.ldc = d;
            p.a_kmajor = false; p.b_kmajor = false;
            if ((rc = launch_gemm(p, st))) return rc;
        }
    }
    return 0;
}

}  // namespace bsa

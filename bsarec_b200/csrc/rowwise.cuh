// Row-wise (HBM-bound) kernels of the BSARec block: embedding gather + LayerNorm + dropout and
// its backward, residual+LayerNorm(+blend), the low-rank sequence filter, masked softmax,
// deterministic column reductions and the sorted scatter-add into the item table.
#pragma once
#include "common.cuh"

namespace bsa {

// ---- embedding: h0 = drop(LN_tf(E[ids] + P[t]))                      (_abstract_model.py:14-24)
int embed_ln_dropout_fwd(const long long* ids, const float* E, const float* Ppos, const float* gamma,
                         const float* beta, float* out, float* mean, float* rstd, int B, int L, int d,
                         long long item_size, Dropout dr, cudaStream_t st);
// dz = dLN-input ; pgamma/pbeta = per-block partials [nblk][d] ; returns nblk through *nblk
int embed_ln_dropout_bwd(const float* dout, const long long* ids, const float* E, const float* Ppos,
                         const float* gamma, const float* mean, const float* rstd, float* dz, float* pgamma,
                         float* pbeta, int* nblk, int B, int L, int d, Dropout dr, cudaStream_t st);
int rowwise_partial_blocks(long long rows);   // how many partial rows the *_bwd kernels write

// ---- z = drop(tmp) + resid ; ln = LN(z) ; out = blend_a*blend_with + blend_b*ln  (blend optional)
int residual_ln_fwd(const float* tmp, const float* resid, const float* gamma, const float* beta, float* out,
                    float* z_save, float* mean, float* rstd, const float* blend_with, float blend_a, float blend_b,
                    long long M, int d, Dropout dr, cudaStream_t st, float* out_hi = nullptr, float* out_lo = nullptr);
// out_hi / out_lo (optional): `out` also leaves as its split-TF32 pair (the A operand of the tcgen05 linear that follows)
// LN backward: upstream = dy * dy_scale.  dz (grad wrt LN input) always written; ddrop = dz * dropout
// multiplier (grad wrt the pre-dropout tensor) written when ddrop != nullptr.
int ln_bwd(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
           float* dz, float* ddrop, float* pgamma, float* pbeta, int* nblk, long long M, int d, Dropout dr,
           cudaStream_t st);
// Same, and t = (dropout on ? ddrop : dz) also leaves as the split-TF32 pair (t_hi, t_lo) with per-block column partials
// pcol [nblk][d] (sum = the bias gradient of the nn.Linear that produced the LN input).  Needs ln_bwd_split_ok(...).
bool ln_bwd_split_ok(int d, const float* dy, const float* z, const float* gamma, const float* dz, const float* ddrop,
                     const float* t_hi, const float* t_lo);
int ln_bwd_split(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
                 float* dz, float* ddrop, float* pgamma, float* pbeta, int* nblk, long long M, int d, Dropout dr, float* t_hi,
                 float* t_lo, float* pcol, cudaStream_t st);

// ---- filter branch (bsarec.py:90-104) in its exact low-rank real form low = C (C^T x)
int filter_ln_fwd(const float* x, const float* Cb, const float* sqrt_beta, const float* gamma, const float* beta,
                  float* dsp, float* z_save, float* mean, float* rstd, int B, int L, int d, int r, Dropout dr,
                  cudaStream_t st, float* x_lo = nullptr);
// x_lo (optional): the kernel also leaves x - truncate_tf32(x) — x is the A operand of the q / k / v projection that follows
// dx += beta2*du + C C^T((1-beta2)*du), du = dz_f * dropout ; also dx += dz_f (residual) ;
// pbeta2 [B][d] partial sums of du*(x-low)
int filter_bwd(const float* x, const float* Cb, const float* sqrt_beta, const float* dz_f, float* dx, float* pbeta2,
               int B, int L, int d, int r, Dropout dr, cudaStream_t st);
int build_filter_basis(float* Cb, int L, int kept_bins, int* r_out, cudaStream_t st);
int filter_rank(int L, int kept_bins);

// ---- attention probabilities over materialised scores (rows of length L)
int softmax_mask_dropout_fwd(const float* S, const long long* ids, float* P, float* Pd, int B, int H, int L,
                             Dropout dr, cudaStream_t st);
int softmax_bwd(const float* P, float* dPd_to_dS, int B, int H, int L, Dropout dr, cudaStream_t st);
int apply_dropout(const float* x, float* y, long long n, Dropout dr, cudaStream_t st);

// ---- deterministic reductions
// out[n] (+)= scale * sum_m X[m][n]   (two fixed-order passes; part needs colsum_partial_rows(M)*N floats)
int colsum(const float* X, long long M, int N, float* part, float* out, int accumulate, cudaStream_t st);
int colsum_partial_rows(long long M);
// out[n] (+)= mult[n]*scale * sum_{k<nblk} part[k][n]
int reduce_partials(const float* part, int nblk, int n, float* out, const float* mult, float scale, int accumulate,
                    cudaStream_t st);
// The same reductions collected and run as ONE launch (reduce_batch_flush).  Every job's `part` must stay untouched and
// every `out` unread until the flush; the flush is stream-ordered like any other launch.
struct ReduceBatch {
    static constexpr int MAXJ = 24;
    const float* part[MAXJ]; float* out[MAXJ]; const float* mult[MAXJ]; int nblk[MAXJ]; int n[MAXJ]; float scale[MAXJ]; int acc[MAXJ];
    int count = 0;
};
int reduce_batch_add(ReduceBatch& b, const float* part, int nblk, int n, float* out, const float* mult, float scale,
                     int accumulate, cudaStream_t st);
int reduce_batch_flush(ReduceBatch& b, cudaStream_t st);
// dP[t][j] (+)= sum_b dz[b][t][j]
int pos_grad(const float* dz, float* dP, int B, int L, int d, int accumulate, cudaStream_t st);

// ---- sorted, atomic-free scatter-add of dz rows into the item-table gradient (skips id 0)
size_t scatter_plan_bytes(long long M);
int scatter_plan(const long long* ids, long long M, long long item_size, void* plan, size_t plan_bytes, cudaStream_t st);
int scatter_add_rows(const void* plan, long long M, const float* dz, float* dE, int d, cudaStream_t st);

int fill_zero(float* p, long long n, cudaStream_t st);

// ---- rowwise_vec.cu: float4 / shared-memory-staged versions, used when rowwise_vec_ok(d) (d % 4 == 0)
bool rowwise_vec_ok(int d);
int embed_fwd_vec(const long long* ids, const float* E, const float* Ppos, const float* gamma, const float* beta, float* out,
                  float* mean, float* rstd, long long M, int L, int d, Dropout dr, int grid, cudaStream_t st);
int embed_bwd_vec(const float* dout, const long long* ids, const float* E, const float* Ppos, const float* gamma,
                  const float* mean, const float* rstd, float* dz, float* pgamma, float* pbeta, long long M, int L, int d,
                  Dropout dr, int grid, cudaStream_t st);
int residual_ln_fwd_vec(const float* tmp, const float* resid, const float* gamma, const float* beta, float* out, float* z_save,
                        float* mean, float* rstd, const float* blend_with, float blend_a, float blend_b, long long M, int d,
                        Dropout dr, int grid, cudaStream_t st, float* y_hi = nullptr, float* y_lo = nullptr);
int ln_bwd_vec(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
               float* dz, float* ddrop, float* pgamma, float* pbeta, long long M, int d, Dropout dr, int grid, float* t_hi,
               float* t_lo, float* pcol, cudaStream_t st);
int filter_ln_fwd_vec(const float* x, const float* Cb, const float* sqrt_beta, const float* gamma, const float* beta, float* dsp,
                      float* z_save, float* mean, float* rstd, int B, int L, int d, int r, Dropout dr, cudaStream_t st,
                      float* x_lo = nullptr);
int filter_bwd_vec(const float* x, const float* Cb, const float* sqrt_beta, const float* dz_f, float* dx, float* pbeta2, int B,
                   int L, int d, int r, Dropout dr, cudaStream_t st);
// dst[b*L + (L-1)][:] = src[b][:], everything else 0
int scatter_last_rows(const float* src, float* dst, int B, int L, int d, cudaStream_t st);
int gather_last_rows(const float* src, float* dst, int B, int L, int d, cudaStream_t st);
// pairwise BCE at the last position (SASRec's loss, model/sasrec.py:42-63): bce.cu
size_t bce_pair_workspace_bytes(int B, int d);
int bce_pair_fwd(int B, int d, long long N, const float* s, const float* E, const long long* pos, const long long* neg,
                 float* loss, void* ws, int variant, cudaStream_t st);
int bce_pair_bwd(int B, int d, long long N, const float* s, const float* E, const long long* pos, const long long* neg,
                 const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, int variant, cudaStream_t st);
int scatter_rows(const long long* idx, long long M, long long dst_rows, const float* src, int d, float* dst, cudaStream_t st);
// FMLPRec's learned circular filter: fmlp.cu
bool fmlp_supported(int L, int d);
int fmlp_trig(float* Tc, float* Ts, int L, cudaStream_t st);
int fmlp_kernel_fwd(const float* W, const float* Tc, const float* Ts, float* K, int L, int d, cudaStream_t st);
int fmlp_kernel_bwd(const float* dK, const float* Tc, const float* Ts, float* dW, int L, int d, cudaStream_t st);
int circ_apply(const float* in, const float* kern, bool kern_per_batch, const float* add, float* out, int B, int L, int d,
               bool adjoint, cudaStream_t st);
// in-batch InfoNCE of DuoRec (model/duorec.py:47-75, sim = 'dot'): bce.cu
size_t nce_workspace_bytes(int B);
int nce_fwd(int B, int d, float tau, const float* z, float* loss, void* ws, cudaStream_t st);
int nce_bwd(int B, int d, float tau, const float* z, const float* d_loss, float scale, float* dz, const void* ws, cudaStream_t st);
int gather_rows(const long long* idx, long long M, long long src_rows, const float* src, int d, float* dst, cudaStream_t st);

}  // namespace bsa

// Kernels of the last block's last-position path (see lastrow.cu).
#pragma once
#include "common.cuh"

namespace bsa {

bool lastrow_supported(int L, int d, int H);

// per sequence: p [B,H,L] (softmax of query row L-1), xs [B,H,d] = sum_j pd_j x_j, sp [B,H] = sum_j pd_j,
// low [B,d] = (F x)[L-1]
int last_attn_fwd(const float* x, const long long* ids, const float* gq, const float* ql, const float* bk, const float* Cb,
                  int r, float* p, float* xs, float* sp, float* low, int B, int H, int L, int d, Dropout dr, cudaStream_t st);
// adjoint: dx [B,L,d] (every position, overwritten), dgq [B,H,d], dcq [B,H]
int last_attn_bwd(const float* x, const float* p, const float* gq, const float* dxs, const float* dsp, const float* dlow,
                  const float* Cb, int r, float* dx, float* dgq, float* dcq, int B, int H, int L, int d, Dropout dr,
                  cudaStream_t st);

int last_filter_u(float* u, const float* low, const float* xl, const float* sqrt_beta, int B, int d, cudaStream_t st);
int last_ctx_bias(float* ctx, const float* bv, const float* sp, int B, int d, int H, cudaStream_t st);
int last_filter_bwd(const float* dzf, const float* xl, const float* low, const float* sqrt_beta, float* dlow, float* dxl_acc,
                    float* pb2, int B, int d, Dropout dr_rows, cudaStream_t st);
int last_head_scale(float* out, const float* a, const float* shead, int B, int d, int H, cudaStream_t st);
int last_add_head_bias(float* out, const float* shead, const float* vcol, int B, int d, int H, cudaStream_t st);
int last_dsp(const float* dctx, const float* bv, float* dsp, int B, int d, int H, cudaStream_t st);
int last_add_rows(float* dx, const float* add, int B, int L, int d, cudaStream_t st);

}  // namespace bsa

// Fused causal multi-head attention (model/_modules.py:108-133 + get_attention_mask, _abstract_model.py:53-69):
// scores, additive -10000 mask from ids, softmax, dropout and the value contraction in one kernel; the [B,H,L,L]
// probabilities are never written.  Backward recomputes them from the saved row log-sum-exp.
#pragma once
#include "common.cuh"

namespace bsa {

// head sizes the fused kernels are instantiated for (other shapes use the unfused GEMM + softmax path)
bool attn_fused_supported(int L, int d, int H);

// q, k, v, ctx: [B*L, d] row-major, head h in columns [h*dh, (h+1)*dh).  lse: [B, H, L, 2] = (row max, log row sum) of the
// masked, scaled scores of each query row, saved for the backward pass (kept apart: rows whose every key is masked
// sit near -10000, where max + log sum would lose the low bits the softmax needs).
// ctx_lo (optional): also leave ctx - truncate_tf32(ctx), the lo half of ctx as a tcgen05 A operand (ctx is its own hi half)
int attn_fwd(const float* q, const float* k, const float* v, const long long* ids, float* ctx, float* lse, int B, int H,
             int L, int d, Dropout dr, cudaStream_t st, float* ctx_lo = nullptr);

// The same forward on tcgen05 / TMEM (attn_umma.cu): head size 64 and L <= 208, i.e. the whole key axis in one accumulator.
extern int g_attn_fwd_umma;      // 1 (default): use it where the shape allows; 0: always the mma.sync kernel
bool attn_umma_supported(int L, int d, int H);
int attn_fwd_umma(const float* q, const float* k, const float* v, const long long* ids, float* ctx, float* lse, int B, int H,
                  int L, int d, Dropout dr, cudaStream_t st, float* ctx_lo = nullptr);

// Query-owner half of the backward on tcgen05 (dQ and delta; same shapes as attn_fwd_umma).  g_attn_bwd_umma = 1 makes
// attn_bwd use it and keep only the key-owner launch (dK, dV) on the mma.sync kernel.
extern int g_attn_bwd_umma;
int attn_bwd_q_umma(const float* q, const float* k, const float* v, const long long* ids, const float* ctx, const float* lse,
                    const float* dctx, float* dq, float* delta, float* pd_out, float* ds_out, int B, int H, int L, int d,
                    Dropout dr, cudaStream_t st, int ldo = 0);
// Key-owner half on tcgen05 (attn_kv_umma.cu): dK, dV from the Pd / dS tiles [B*H*L, NK] the query-owner kernel left in HBM.
// g_attn_bwd_umma = 2 (default) selects it; 1 keeps the key-owner launch on the mma.sync kernel (recomputation).
int attn_bwd_kv_umma(const float* q, const float* dctx, const float* pd, const float* ds, float* dk, float* dv, int B, int H,
                     int L, int d, cudaStream_t st, int ldo = 0);
// floats of scratch attn_bwd needs behind `delta` for (B, H, L, d) under the current settings
long long attn_bwd_scratch_floats(int B, int H, int L, int d);

// dq, dk, dv [B*L, d] are overwritten.  delta: [B, H, L] scratch (row sums of dctx * ctx).
int attn_bwd(const float* q, const float* k, const float* v, const long long* ids, const float* ctx, const float* lse,
             const float* dctx, float* dq, float* dk, float* dv, float* delta, int B, int H, int L, int d, Dropout dr,
             cudaStream_t st, int ldo = 0);
// ldo (0 = d): row pitch of dq / dk / dv — 3d when the three are column windows of one packed [B*L, 3d] tensor; only the
// two-kernel tcgen05 backward writes strided outputs: attn_bwd_strided_ok(L, d, H)
bool attn_bwd_strided_ok(int L, int d, int H);

}  // namespace bsa

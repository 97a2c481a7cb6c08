/* bsarec_b200 — C ABI of the B200 (sm_100a) kernels behind BSARec's training and full-ranking
 * evaluation hot path.  This is the drop-in boundary: plain pointers and sizes, no torch types.
 *
 * The reference has no FFI of its own (it is pure Python/PyTorch); each entry point below replaces a
 * group of eager ATen call sites, cited as file:line relative to the reference's src/ directory.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero (a cudaError_t or -1) on failure; the message is
 *    read with bsa_last_error_string().  Nothing throws, nothing synchronises the device.
 *  - all pointers are DEVICE pointers unless the name ends in _host; tensors are dense row-major fp32,
 *    ids are int64 (torch.long) exactly as the reference's DataLoader yields them.
 *  - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream).
 *  - the caller owns every buffer; the library keeps no reference after the call returns and allocates
 *    no device memory.  Scratch comes in through (workspace, workspace_bytes); the *_bytes() queries
 *    give the sizes.
 *  - dropout uses the stateless counter RNG restated in oracle/rng_ref.py; `seed` selects the stream of
 *    one forward/backward pair, `training == 0` disables it (nn.Module.eval()).
 */
#ifndef BSAREC_B200_H
#define BSAREC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

#define BSA_ABI_VERSION 4      /* bumped whenever a desc struct or an entry point's argument list changes */

/* ------------------------------------------------------------------------------------------------
 * Library
 * ---------------------------------------------------------------------------------------------- */
int bsa_abi_version(void);
const char* bsa_last_error_string(void);
/* kernels launched by this library since load (bench.py's "gpu_launches") */
unsigned long long bsa_launch_count(void);
/* CUDA-graph capture of a whole step (bsarec_b200/graphs.py; "capture launch-bound inner loops in CUDA graphs"):
 * every launch argument is frozen at capture time, so the two per-step host values of the path — the dropout seed
 * and Adam's step number (trainers.py:103-107 draws fresh masks and advances the optimizer every iteration) — can be
 * taken from device memory instead.  state_dev points to two uint64 words that the caller advances between replays:
 *   [0] is added to every `seed` (a launch made with seed s behaves exactly like a launch made with s + state[0]),
 *   [1] is added to every Adam `step`.
 * Launches made while the pointer is set read the words when they RUN.  Null (the default) restores host values.
 * Process-global (one process per GPU), not thread-safe against concurrent launches from other threads. */
void bsa_set_replay_state(const uint64_t* state_dev);
const uint64_t* bsa_get_replay_state(void);
/* Per-kernel device timing for the roofline report: when enabled, tagged launches are bracketed by CUDA
 * events on the launching stream.  bsa_profile_get synchronises them and returns the tag's name, total
 * device milliseconds, algorithmic work (FLOPs or bytes as the tag defines) and launch count. */
void bsa_profile_enable(int on);
void bsa_profile_reset(void);
int bsa_profile_num_tags(void);
int bsa_profile_get(int tag, const char** name, double* ms, double* work, long long* launches);

/* ------------------------------------------------------------------------------------------------
 * (1) Embedding: h0 = dropout(LayerNorm_tf(E[ids] + P[t]))
 *     replaces SequentialRecModel.add_position_embedding  (model/_abstract_model.py:14-24)
 *     and the custom LayerNorm                             (model/_modules.py:16-20)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int B, L, d;               /* batch, max_seq_length, hidden_size */
    int64_t item_size;         /* rows of the item table */
    float p_hidden;            /* hidden_dropout_prob */
    int training;
    uint64_t seed;
} bsa_embed_desc;

/* out [B,L,d]; mean/rstd [B,L] are saved for the backward pass */
int bsa_embed_ln_dropout_fwd(const bsa_embed_desc* desc, const int64_t* ids, const float* item_emb,
                             const float* pos_emb, const float* ln_w, const float* ln_b, float* out,
                             float* mean, float* rstd, void* stream);

size_t bsa_embed_bwd_workspace_bytes(const bsa_embed_desc* desc);
/* d_item_emb [item_size,d] is ACCUMULATED into (the dense cross-entropy gradient is already there; the
 * gather adds rows for ids != 0 only — padding_idx=0, _abstract_model.py:10) through a sorted,
 * atomic-free segmented sum: fixed summation order, bit-reproducible.
 * d_pos_emb [L,d], d_ln_w [d], d_ln_b [d] are overwritten. */
int bsa_embed_ln_dropout_bwd(const bsa_embed_desc* desc, const int64_t* ids, const float* item_emb,
                             const float* pos_emb, const float* ln_w, const float* mean, const float* rstd,
                             const float* d_out, float* d_item_emb, float* d_pos_emb, float* d_ln_w,
                             float* d_ln_b, void* workspace, size_t workspace_bytes, void* stream);

/* The same backward in two halves, for data-parallel training: the dense cross-entropy part of the
 * item-table gradient is all-reduced while the encoder backward runs, the sparse gather part is exchanged as rows.
 *   bsa_embed_ln_dropout_bwd_rows: everything above except the scatter; dz_rows [B*L,d] (out) is the gradient of the
 *     gathered rows E[ids] (the operand of ATen's embedding_dense_backward, _abstract_model.py:17).
 *   bsa_scatter_add_rows: d_item_emb[ids[i]] += rows[i] for ids[i] != 0, i < n_rows, through the sorted atomic-free
 *     segmented sum (ties summed in row order -> every rank that feeds the same (ids, rows) gets the same bits).
 * Workspace of the first: bsa_embed_bwd_workspace_bytes. */
int bsa_embed_ln_dropout_bwd_rows(const bsa_embed_desc* desc, const int64_t* ids, const float* item_emb,
                                  const float* pos_emb, const float* ln_w, const float* mean, const float* rstd,
                                  const float* d_out, float* dz_rows, float* d_pos_emb, float* d_ln_w, float* d_ln_b,
                                  void* workspace, size_t workspace_bytes, void* stream);
size_t bsa_scatter_add_workspace_bytes(int64_t n_rows);
int bsa_scatter_add_rows(const int64_t* ids, int64_t n_rows, int64_t item_size, const float* rows, int d,
                         float* d_item_emb, void* workspace, size_t workspace_bytes, void* stream);

/* helper: stable (id, row) sort plan used by the scatter-add; exposed for tests.
 * sorted_ids/sorted_rows are int32 [B*L]. */
size_t bsa_sort_scatter_plan_bytes(int64_t n);
int bsa_sort_scatter_plan(const int64_t* ids, int64_t n, int64_t item_size, void* plan, size_t plan_bytes,
                          int32_t* sorted_ids, int32_t* sorted_rows, void* stream);

/* ------------------------------------------------------------------------------------------------
 * (2) One BSARecBlock: filter branch + causal multi-head attention + alpha blend + GELU FFN
 *     replaces BSARecBlock/BSARecLayer/FrequencyLayer.forward   (model/bsarec.py:62-104),
 *              MultiHeadAttention.forward                        (model/_modules.py:108-140),
 *              FeedForward.forward                               (model/_modules.py:61-69),
 *              get_attention_mask (computed from ids in-kernel)  (model/_abstract_model.py:53-69)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int B, L, d, H;            /* batch, sequence length, hidden size, attention heads */
    int kept_bins;             /* c//2 + 1 rfft bins kept by the low-pass (bsarec.py:87) */
    float alpha;               /* blend: alpha*filter + (1-alpha)*attention (bsarec.py:78) */
    float p_hidden, p_attn;    /* hidden_dropout_prob, attention_probs_dropout_prob */
    int training;
    uint64_t seed;
    int layer;                 /* block index: selects the dropout sites 1+4*layer .. 4+4*layer */
    int inference;             /* != 0: forward only — what exists only for bsa_block_bwd (LayerNorm inputs and statistics,
                                  GELU pre-activation) is not written; `saved` is scratch and bsa_block_bwd must not follow */
    int kind;                  /* 0: BSARecBlock.  1: FMLPRecBlock (sibling model, model/fmlprec.py:84-118): no attention; the
                                  filter layer is a learned per-channel circular convolution whose taps K [L,d]
                                  (bsa_fmlp_kernel_fwd) arrive as `sqrt_beta`, f_ln_* is its LayerNorm, the attention
                                  members are ignored; bsa_block_bwd leaves dK [L,d] in grads.sqrt_beta */
} bsa_block_desc;

/* nn.Linear weights are [out,in] row-major, exactly the state_dict tensors */
typedef struct {
    const float *sqrt_beta, *f_ln_w, *f_ln_b;                 /* filter_layer.{sqrt_beta,LayerNorm.*} */
    const float *wq, *bq, *wk, *bk, *wv, *bv, *wo, *bo;       /* attention_layer.{query,key,value,dense} */
    const float *a_ln_w, *a_ln_b;                             /* attention_layer.LayerNorm (nn.LayerNorm) */
    const float *w1, *b1, *w2, *b2, *o_ln_w, *o_ln_b;         /* feed_forward.{dense_1,dense_2,LayerNorm} */
} bsa_block_weights;

typedef struct {
    float *sqrt_beta, *f_ln_w, *f_ln_b;
    float *wq, *bq, *wk, *bk, *wv, *bv, *wo, *bo;
    float *a_ln_w, *a_ln_b;
    float *w1, *b1, *w2, *b2, *o_ln_w, *o_ln_b;
} bsa_block_grads;

/* orthonormal basis C [L, r] of the exact real low-pass operator F = C C^T
 * (== irfft(zero_{k>=kept}(rfft(x, norm='ortho')), norm='ortho'), bsarec.py:93-97) */
int bsa_filter_rank(int L, int kept_bins);
int bsa_build_filter_basis(float* basis, int L, int kept_bins, void* stream);

/* how the tcgen05 cross-entropy backward gets its logits: 2 (default) = bsa_ce_fwd leaves the fp32 logits in the workspace
 * and bsa_ce_bwd (called with that workspace, reuse_split = 1) reads them back — ds and dE by one A-from-TMEM kernel each;
 * 1 = logits recomputed, gradient tile + ds fused; 0 = logits recomputed, separate GEMMs.  Process-wide; set it before
 * the forward call.  Modes 1 and 0 stay as the on-GPU cross-checks of mode 2. */
void bsa_set_ce_mode(int mode);
int bsa_get_ce_mode(void);
/* where the block's nn.Linear contractions run: 2 = tcgen05 3xTF32 engine (default), 1 = fp32 FFMA kernels
 * (the on-GPU cross-check).  Process-wide. */
void bsa_set_linear_impl(int impl);
int bsa_get_linear_impl(void);
/* how the attention branch runs: 2 = fused kernel (scores, mask, softmax, dropout and P V in one pass, backward by
 * recomputation from the saved row log-sum-exp; head sizes 32 and 64), 1 = unfused GEMM + softmax kernels that
 * materialise the [B,H,L,L] probabilities (every head size; the on-GPU cross-check).  With 2 the forward runs on
 * tcgen05/TMEM (attn_umma.cu: S = Q K^T into TMEM, softmax in place, P fed back as the TMEM A operand of P V) when the head
 * size is 64 and L <= 208, else on the mma.sync kernel; 3 = fused, mma.sync kernels only.  Process-wide; must not change
 * between a bsa_block_fwd and its bsa_block_bwd. */
void bsa_set_attn_impl(int impl);
int bsa_get_attn_impl(void);

size_t bsa_block_saved_bytes(const bsa_block_desc* desc);      /* activations kept for backward */
size_t bsa_block_workspace_bytes(const bsa_block_desc* desc);  /* scratch for fwd and bwd */

/* x_out [B,L,d] = block(x_in); `saved` receives the activations backward needs (always required:
 * inference may reuse one buffer for all layers).  On the tcgen05 path `saved` also receives the split-TF32 pairs of the
 * block's six weight matrices (one launch at the top of the forward pass) and the lo halves of x_in / ctx / y, whose hi
 * halves are those fp32 tensors themselves: bsa_block_bwd multiplies by the weights the forward pass saw and must be given
 * the same x_in. */
int bsa_block_fwd(const bsa_block_desc* desc, const bsa_block_weights* w, const float* basis,
                  const int64_t* ids, const float* x_in, float* x_out, void* saved, void* workspace,
                  size_t workspace_bytes, void* stream);

/* d_x_in [B,L,d] and every member of `g` are OVERWRITTEN with this block's gradients (the parameter gradients by one batched
 * reduction launch at the end of the call). */
int bsa_block_bwd(const bsa_block_desc* desc, const bsa_block_weights* w, const float* basis,
                  const int64_t* ids, const float* x_in, const void* saved, const float* d_x_out,
                  float* d_x_in, const bsa_block_grads* g, void* workspace, size_t workspace_bytes,
                  void* stream);

/* The LAST block evaluated at position L-1 only.  calculate_loss (model/bsarec.py:32) and the trainer's eval branch
 * (trainers.py:127) read nothing but out[:, -1, :] of the last block, so there only one query row per sequence is
 * live; with the key/value projections folded into the query side the block needs no [B*L, d] GEMM (csrc/lastrow.cu).
 * Same values as bsa_block_fwd(...)[:, L-1, :] up to fp32 rounding, same dropout masks, exact adjoint.
 *   out_last [B,d];  bsa_block_last_bwd: d_out_last [B,d] in, d_x_in [B,L,d] (every position) and `g` overwritten.
 * Supported for heads <= 8, d <= 1024 (bsa_block_last_supported); the host falls back to the full block otherwise. */
int bsa_block_last_supported(const bsa_block_desc* desc);
size_t bsa_block_last_saved_bytes(const bsa_block_desc* desc);
size_t bsa_block_last_workspace_bytes(const bsa_block_desc* desc);
int bsa_block_last_fwd(const bsa_block_desc* desc, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                       const float* x_in, float* out_last, void* saved, void* workspace, size_t workspace_bytes,
                       void* stream);
int bsa_block_last_bwd(const bsa_block_desc* desc, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                       const float* x_in, const void* saved, const float* d_out_last, float* d_x_in,
                       const bsa_block_grads* g, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * (3) Full-catalog cross-entropy without materialising the [B, item_size] logits
 *     replaces BSARecModel.calculate_loss's matmul + CrossEntropyLoss   (model/bsarec.py:32-35)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int B, d;
    int64_t item_size;
    int impl;                  /* 0 = auto, 1 = fp32 FFMA, 2 = tcgen05 3xTF32 */
    int reuse_split;           /* bsa_ce_bwd only: `workspace` is the buffer bsa_ce_fwd just used with the same
                                  seq_out / item_emb and still holds their split-TF32 copies (skips re-splitting the
                                  item table: 3 GB of traffic at 1 M items) */
} bsa_ce_desc;

size_t bsa_ce_workspace_bytes(const bsa_ce_desc* desc);
/* seq_out [B,d] (last position), answers int64 [B].  loss: one float = mean_b(lse_b - logit_b[ans_b]);
 * lse [B] is saved for the backward pass. */
int bsa_ce_fwd(const bsa_ce_desc* desc, const float* seq_out, const float* item_emb, const int64_t* answers,
               float* loss, float* lse, void* workspace, size_t workspace_bytes, void* stream);
/* One ROW SHARD of a vocabulary-parallel cross-entropy (data-parallel training with the item table's CE sharded over the
 * ranks): item_shard [desc.item_size = shard rows, d]; answers_local int64 [B] = answer - shard offset, or any
 * NEGATIVE value when the target lies in another shard.  lse_part [B] = log-sum-exp over the shard's rows, tgt_part [B] =
 * the target logit (0 when not in this shard): the caller combines them across shards, loss = mean_b(logsumexp_r lse_part -
 * sum_r tgt_part).  bsa_ce_bwd then takes the GLOBAL lse, the same answers_local and this workspace (reuse_split = 1) and
 * returns the shard's complete d_item_emb rows and this shard's partial d_seq_out (summed over shards by the caller).
 * tcgen05 path only. */
int bsa_ce_fwd_partial(const bsa_ce_desc* desc, const float* seq_out, const float* item_shard, const int64_t* answers_local,
                       float* lse_part, float* tgt_part, void* workspace, size_t workspace_bytes, void* stream);
/* d_loss: device scalar (upstream gradient).  d_seq_out [B,d] overwritten; d_item_emb [item_size,d]
 * OVERWRITTEN with the dense gradient (every row, including row 0). */
int bsa_ce_bwd(const bsa_ce_desc* desc, const float* seq_out, const float* item_emb, const int64_t* answers,
               const float* lse, const float* d_loss, float* d_seq_out, float* d_item_emb, void* workspace,
               size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * (4) Full-catalog scoring + history masking + per-user top-k (+ cross-shard merge)
 *     replaces Trainer.predict_full                (trainers.py:62-68) and the numpy masking /
 *              argpartition / argsort eval branch   (trainers.py:130-149)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int B, d, k;               /* users in the batch, hidden size, list length (<= 128) */
    int64_t shard_rows;        /* rows of this item-table shard */
    int64_t shard_offset;      /* global id of the shard's row 0 */
    int impl;                  /* 0 = auto, 1 = fp32 FFMA, 2 = tcgen05 3xTF32 */
    int reuse_split;           /* 1: `workspace` was last used by bsa_score_topk with the SAME item_shard contents and
                                  shard_rows/d/B/k: its split-TF32 copy of the shard is reused (evaluation sweeps score
                                  millions of users against a table that does not change, trainers.py:118-158) */
} bsa_topk_desc;

size_t bsa_score_topk_workspace_bytes(const bsa_topk_desc* desc);
/* History is CSR: hist_indices int64 (GLOBAL item ids, sorted ascending within a row); batch row b owns CSR row
 * hist_rows[b] (int64 [B]; the trainer's `user_ids` into the rating matrix, trainers.py:131-133) or, when hist_rows
 * is NULL, CSR row b (hist_indptr int64 [B+1]).  History items score exactly 0.0f (trainers.py:134), they are not removed.
 * Outputs: top_val fp32 [B,k], top_idx int64 [B,k] (global ids), descending score, ties by ascending id;
 * unused slots (fewer than k candidates in the shard) hold -inf / -1. */
int bsa_score_topk(const bsa_topk_desc* desc, const float* seq_out, const float* item_shard,
                   const int64_t* hist_indptr, const int64_t* hist_indices, const int64_t* hist_rows, float* top_val,
                   int64_t* top_idx, void* workspace, size_t workspace_bytes, void* stream);
/* merge G candidate lists [G,B,k] (one per shard) into [B,k] */
int bsa_topk_merge(const float* cand_val, const int64_t* cand_idx, int G, int B, int k, float* top_val,
                   int64_t* top_idx, void* stream);

/* ------------------------------------------------------------------------------------------------
 * The tcgen05 engine as a plain GEMM (what cuBLAS sgemm is to the reference's nn.Linear / torch.matmul
 * call sites, model/_modules.py:62-66,109-136): C[M,N] = A B^T with error-compensated TF32 (3 tcgen05.mma
 * per k-step, fp32 accumulate in TMEM), results within fp32 rounding of an FFMA GEMM.
 *   a_mn == 0: A is [M rows, K cols] row-major (lda = row pitch);  a_mn != 0: A is [K rows, M cols].
 *   b_mn == 0: B is [N rows, K cols] row-major;                    b_mn != 0: B is [K rows, N cols].
 * impl: 2 = tcgen05, 1 = fp32 FFMA reference kernel (same semantics; used as the on-GPU cross-check).
 * ---------------------------------------------------------------------------------------------- */
size_t bsa_gemm_workspace_bytes(int M, int N, int K, int64_t lda, int64_t ldb, int a_mn, int b_mn);
int bsa_gemm(int impl, int M, int N, int K, const float* A, int64_t lda, int a_mn, const float* B, int64_t ldb,
             int b_mn, float* C, int64_t ldc, const float* bias, float alpha, int accumulate, void* workspace,
             size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Optimizer step: torch.optim.Adam's update — Trainer.__init__ / optim.step(),
 * trainers.py:27-28 and :107 — as one pass over (param, grad, exp_avg, exp_avg_sq), all fp32 [n], updated in place.
 * No amsgrad / maximize; weight_decay is Adam's L2 term (grad += wd * param).  `step` counts from 1.
 * ---------------------------------------------------------------------------------------------- */
int bsa_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int64_t step, void* stream);
/* The same update for `count` tensors in one launch per 48 tensors (the model's 41 small parameters).  The five
 * arrays are HOST arrays of length count holding DEVICE pointers / element counts; all tensors share `step`. */
int bsa_adam_step_multi(int count, float* const* params_host, const float* const* grads_host, float* const* exp_avg_host,
                        float* const* exp_avg_sq_host, const int64_t* sizes_host, float lr, float beta1, float beta2, float eps,
                        float weight_decay, int64_t step, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Either side of the model call: integer kernels, bit-exact.
 * ---------------------------------------------------------------------------------------------- */
/* Batch construction from a device-resident CSR of the user sequences — replaces RecDataset.__init__'s prefix
 * expansion and __getitem__'s slicing / left padding (dataset.py:18-28, 61-72) plus the DataLoader collate.
 *   seq_indptr int64 [n_users+1], seq_items int64 [nnz]: the lines of <data_name>.txt (dataset.py:184-197).
 *   mode 0 = train: sample s is the (s - sample_start[u])-th prefix of user u's window seq[-(L+2):-2];
 *            sample_start int64 [n_users+1] = exclusive prefix sum of min(max(len_u - 2, 0), L).
 *   mode 1 = valid (items = seq[:-1]), mode 2 = test (items = seq): sample s is user s; sample_start unused.
 *   sample_idx int64 [B] (the sampler's indices).  Outputs: input_ids int64 [B,L] left-padded with 0,
 *   answers int64 [B] (= items[-1]), user_ids int64 [B] (train: the owning user; valid/test: s), as the reference's
 *   batch tuple (dataset.py:74-113).  Every sequence must have len >= 3 (the caller checks once on the host). */
int bsa_build_batch(int mode, const int64_t* seq_indptr, const int64_t* seq_items, const int64_t* sample_start, int64_t n_users,
                    const int64_t* sample_idx, int B, int L, int64_t* input_ids, int64_t* answers, int64_t* user_ids,
                    void* stream);
/* Hit-rank histogram — what recall_at_k / ndcg_k (metrics.py:3-23) are sums over: hist[r] += number of users whose
 * answer is at position r of pred[u, :k] (lists hold distinct ids), hist[k] += users without a hit.  pred int64 [U,k],
 * answers int64 [U], hist int64 [k+1] ACCUMULATED (zero it once per evaluation).  Integer atomics only. */
int bsa_hit_rank_hist(const int64_t* pred, const int64_t* answers, int64_t U, int k, int64_t* hist, void* stream);

/* FMLPRec's learned filter (model/fmlprec.py:95-118): irfft(rfft(x) * W) == circular convolution with taps
 * K[:, c] = irfft(W[:, c], n = L).  tables: Tc, Ts fp32 [L, L/2+1] (once per L).  kernel_fwd: complex_weight [L/2+1, d, 2]
 * -> K [L, d].  kernel_bwd: dK [L, d] (from bsa_block_bwd, kind 1) -> d complex_weight [L/2+1, d, 2] (overwritten). */
int bsa_fmlp_supported(int L, int d);
int bsa_fmlp_tables(float* Tc, float* Ts, int L, void* stream);
int bsa_fmlp_kernel_fwd(const float* complex_weight, const float* Tc, const float* Ts, float* K, int L, int d, void* stream);
int bsa_fmlp_kernel_bwd(const float* dK, const float* Tc, const float* Ts, float* d_complex_weight, int L, int d, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Sibling model on the same kernels: SASRec = the block with the filter branch blended out
 * (alpha = 0) + this pairwise BCE at the last position, replacing SASRecModel.calculate_loss (model/sasrec.py:42-63):
 *   xp_b = <E[pos_b], s_b>, xn_b = <E[neg_b], s_b>; loss = mean_{pos_b != 0} softplus(-xp_b) + mean_{pos_b != 0} softplus(xn_b).
 * fwd keeps the logits in `workspace` for bwd (same workspace); bwd OVERWRITES d_seq_out [B,d] and ACCUMULATES the 2B
 * gathered-row gradients into d_item_emb through the sorted atomic-free scatter (id 0 skipped: padding_idx).
 * variant 0: SASRec as above.  variant 1: FMLPRec (model/fmlprec.py:40-60): loss = mean over ALL users of
 *   -log(sigmoid(xp) + 1e-24) - log(1 - sigmoid(xn) + 1e-24).
 * ---------------------------------------------------------------------------------------------- */
size_t bsa_bce_pair_workspace_bytes(int B, int d);
int bsa_bce_pair_fwd(int B, int d, int64_t item_size, const float* seq_out, const float* item_emb, const int64_t* pos_ids,
                     const int64_t* neg_ids, float* loss, void* workspace, size_t workspace_bytes, int variant, void* stream);
int bsa_bce_pair_bwd(int B, int d, int64_t item_size, const float* seq_out, const float* item_emb, const int64_t* pos_ids,
                     const int64_t* neg_ids, const float* d_loss, float* d_seq_out, float* d_item_emb, void* workspace,
                     size_t workspace_bytes, int variant, void* stream);

/* In-batch InfoNCE of the sibling DuoRec (model/duorec.py:47-75, sim = 'dot'; its cross-entropy term is bsa_ce_fwd/bwd):
 * z [2B, d] = the two views' last-position vectors stacked; row r's positive is row (r + B) mod 2B, its negatives every
 * other row but itself; loss = mean_r( -sim[r, pair] + logsumexp_{c != r} sim[r, c] ), sim = z z^T / tau.
 * fwd keeps the row log-sum-exps in `workspace`; bwd writes dz [2B, d] = scale * d_loss * d loss / d z. */
size_t bsa_info_nce_workspace_bytes(int B);
int bsa_info_nce_fwd(int B, int d, float tau, const float* z, float* loss, void* workspace, size_t workspace_bytes, void* stream);
int bsa_info_nce_bwd(int B, int d, float tau, const float* z, const float* d_loss, float scale, float* dz, const void* workspace,
                     size_t workspace_bytes, void* stream);
/* No-allocation variant of bsa_info_nce_fwd.  NO counterpart in the reference: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same inputs and `loss` output as bsa_info_nce_fwd, no workspace argument: the scratch (row
 * log-sum-exps and per-row losses, 2 x [2B] floats) is a file-scope device array of the library, sized for
 * B <= BSA_INFO_NCE_STATIC_MAX_B (64 KB; every GPU that loads the library holds it until exit).  Larger B, or
 * 2B + d > 12288 floats, is rejected with an error (-1, bsa_last_error_string).  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine).  All calls share the array and it keeps the last call's contents; every element
 * read is written first by the same call, so nothing needs clearing, but calls that OVERLAP on the GPU (two streams) race on
 * it — serialise them.  The log-sum-exps are not handed back: use bsa_info_nce_fwd when bsa_info_nce_bwd follows. */
#define BSA_INFO_NCE_STATIC_MAX_B 4096
int bsa_info_nce_static_max_b(void);
int bsa_info_nce_fwd_static(int B, int d, float tau, const float* z, float* loss, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Small utilities shared by the host wrappers
 * ---------------------------------------------------------------------------------------------- */
/* dst [B,d] = src[:, L-1, :] */
int bsa_gather_last(const float* src, float* dst, int B, int L, int d, void* stream);
/* dst [n_rows,d] = src[idx[i], :] — the row gather of nn.Embedding (_abstract_model.py:17-18) on its own: bit-exact
 * copies.  Used by the data-parallel exchange to pack the rows that carry an item-table gradient; idx int64 [n_rows]
 * in [0, src_rows), d % 4 == 0. */
int bsa_gather_rows(const int64_t* idx, int64_t n_rows, int64_t src_rows, const float* src, int d, float* dst, void* stream);
/* dst[idx[i], :] = src[i, :] — rows fetched from their owning rank written into the local replica of the item table
 * (vocabulary-parallel training); idx int64 [n_rows] in [0, dst_rows), duplicates must carry identical rows. */
int bsa_scatter_rows(const int64_t* idx, int64_t n_rows, int64_t dst_rows, const float* src, int d, float* dst, void* stream);
/* dst [B,L,d] = 0 except dst[:, L-1, :] = src [B,d] */
int bsa_scatter_last(const float* src, float* dst, int B, int L, int d, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* BSAREC_B200_H */

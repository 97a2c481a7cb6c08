// extern "C" surface declared in include/bsarec_b200.h.  Argument checking and workspace carving only;
// the kernels live in the other translation units.
#include "../../include/bsarec_b200.h"
#include "common.cuh"
#include "rowwise.cuh"
#include "ce_topk.cuh"
#include "umma_engine.cuh"
#include "gemm_simt.cuh"

#include <cstdarg>
#include <utility>
#include <vector>

namespace bsa {

static thread_local char g_err[1024] = "";
unsigned long long g_launch_count = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int block_fwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
              const float* x, float* x_out, void* saved_p, void* ws, size_t ws_bytes, cudaStream_t st);
int block_bwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
              const float* x, const void* saved_p, const float* g_out, float* dx, const bsa_block_grads& g, void* ws,
              size_t ws_bytes, cudaStream_t st);
size_t block_saved_bytes(const bsa_block_desc& c);
size_t block_workspace_bytes(const bsa_block_desc& c);
bool block_last_supported(const bsa_block_desc& c);
int block_last_fwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
                   const float* x, float* out_last, void* saved_p, void* ws, size_t ws_bytes, cudaStream_t st);
int block_last_bwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
                   const float* x, const void* saved_p, const float* d_out_last, float* dx, const bsa_block_grads& g,
                   void* ws, size_t ws_bytes, cudaStream_t st);
size_t block_last_saved_bytes(const bsa_block_desc& c);
size_t block_last_workspace_bytes(const bsa_block_desc& c);
void set_linear_impl(int impl);
int linear_impl();
void set_attn_impl(int impl);
int attn_impl();

static inline size_t al(size_t x) { return (x + 255) / 256 * 256; }

// ---- kernel timing registry
int g_profile_on = 0;
const unsigned long long* g_replay_state = nullptr;
namespace {
struct ProfRec {
    const char* name;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
    cudaEvent_t open = nullptr;
    double ms = 0, work = 0;
    long long launches = 0;
};
ProfRec g_prof[PROF_NUM_TAGS] = {{"gemm_simt"}, {"ce_fwd_umma"}, {"ce_bwd_umma"}, {"score_topk_umma"}, {"linear_umma"},
                                 {"embed_fwd"}, {"embed_bwd"}, {"scatter_add"}, {"attn_fwd"}, {"attn_bwd"}, {"adam"}, {"last_row_fwd"}, {"last_row_bwd"},
                                 {"filter_fwd"}, {"filter_bwd"}, {"residual_ln_fwd"}, {"ln_bwd"}};
std::vector<cudaEvent_t> g_event_pool;
cudaEvent_t get_event() {
    if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
void drain(ProfRec& r) {
    for (auto& pr : r.pending) {
        cudaEventSynchronize(pr.second);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) r.ms += ms;
        g_event_pool.push_back(pr.first);
        g_event_pool.push_back(pr.second);
    }
    r.pending.clear();
}
}  // namespace
void prof_begin(int tag, cudaStream_t st) {
    ProfRec& r = g_prof[tag];
    r.open = get_event();
    cudaEventRecord(r.open, st);
}
void prof_end(int tag, cudaStream_t st, double work) {
    ProfRec& r = g_prof[tag];
    if (!r.open) return;
    cudaEvent_t e = get_event();
    cudaEventRecord(e, st);
    r.pending.emplace_back(r.open, e);
    r.open = nullptr;
    r.work += work;
    r.launches += 1;
    if (r.pending.size() > 4096) drain(r);
}

}  // namespace bsa

using namespace bsa;

extern "C" {

int bsa_abi_version(void) { return BSA_ABI_VERSION; }
const char* bsa_last_error_string(void) { return g_err; }
unsigned long long bsa_launch_count(void) { return g_launch_count; }

void bsa_set_replay_state(const uint64_t* state_dev) { g_replay_state = (const unsigned long long*)state_dev; }
const uint64_t* bsa_get_replay_state(void) { return (const uint64_t*)g_replay_state; }

void bsa_profile_enable(int on) { g_profile_on = on ? 1 : 0; }
void bsa_profile_reset(void) {
    for (auto& r : g_prof) { drain(r); r.ms = 0; r.work = 0; r.launches = 0; }
}
int bsa_profile_num_tags(void) { return PROF_NUM_TAGS; }
int bsa_profile_get(int tag, const char** name, double* ms, double* work, long long* launches) {
    BSA_REQUIRE(tag >= 0 && tag < PROF_NUM_TAGS, "bsa_profile_get: bad tag %d", tag);
    ProfRec& r = g_prof[tag];
    drain(r);
    if (name) *name = r.name;
    if (ms) *ms = r.ms;
    if (work) *work = r.work;
    if (launches) *launches = r.launches;
    return 0;
}

// ------------------------------------------------------------------------------------------ embedding
int bsa_embed_ln_dropout_fwd(const bsa_embed_desc* c, const int64_t* ids, const float* item_emb, const float* pos_emb,
                             const float* ln_w, const float* ln_b, float* out, float* mean, float* rstd, void* stream) {
    BSA_REQUIRE(c && ids && item_emb && pos_emb && ln_w && ln_b && out && mean && rstd, "bsa_embed_ln_dropout_fwd: null argument");
    BSA_REQUIRE(c->d % 4 == 0, "hidden size must be a multiple of 4 (got %d)", c->d);
    return embed_ln_dropout_fwd((const long long*)ids, item_emb, pos_emb, ln_w, ln_b, out, mean, rstd, c->B, c->L, c->d,
                                c->item_size, make_dropout(c->seed, 0, c->p_hidden, c->training), (cudaStream_t)stream);
}

size_t bsa_embed_bwd_workspace_bytes(const bsa_embed_desc* c) {
    const long long M = (long long)c->B * c->L;
    return al((size_t)M * c->d * sizeof(float)) + 2 * al((size_t)rowwise_partial_blocks(M) * c->d * sizeof(float)) +
           al(scatter_plan_bytes(M)) + 1024;
}

// Everything of the embedding backward except the scatter into the item table; dz_rows [B*L,d] is the gradient of the
// gathered rows (what the scatter-add consumes).
static int embed_bwd_rows(const bsa_embed_desc* c, const int64_t* ids, const float* item_emb, const float* pos_emb,
                          const float* ln_w, const float* mean, const float* rstd, const float* d_out, float* dz,
                          float* d_pos_emb, float* d_ln_w, float* d_ln_b, char* p, cudaStream_t st) {
    const long long M = (long long)c->B * c->L;
    const size_t pb = al((size_t)rowwise_partial_blocks(M) * c->d * sizeof(float));
    float* pg = (float*)p; p += pb;
    float* pbt = (float*)p; p += pb;
    int nblk = 0, rc;
    if ((rc = embed_ln_dropout_bwd(d_out, (const long long*)ids, item_emb, pos_emb, ln_w, mean, rstd, dz, pg, pbt, &nblk,
                                   c->B, c->L, c->d, make_dropout(c->seed, 0, c->p_hidden, c->training), st))) return rc;
    // both LayerNorm parameter gradients and the position-table gradient (column sums of dz seen as [B, L*d]) in one launch
    ReduceBatch rb;
    if ((rc = reduce_batch_add(rb, pg, nblk, c->d, d_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbt, nblk, c->d, d_ln_b, nullptr, 1.f, 0, st))) return rc;
    if ((long long)c->L * c->d < (1ll << 30)) {
        if ((rc = reduce_batch_add(rb, dz, c->B, c->L * c->d, d_pos_emb, nullptr, 1.f, 0, st))) return rc;
        return reduce_batch_flush(rb, st);
    }
    if ((rc = reduce_batch_flush(rb, st))) return rc;
    return pos_grad(dz, d_pos_emb, c->B, c->L, c->d, 0, st);
}

int bsa_embed_ln_dropout_bwd(const bsa_embed_desc* c, const int64_t* ids, const float* item_emb, const float* pos_emb,
                             const float* ln_w, const float* mean, const float* rstd, const float* d_out,
                             float* d_item_emb, float* d_pos_emb, float* d_ln_w, float* d_ln_b, void* workspace,
                             size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && ids && item_emb && pos_emb && ln_w && mean && rstd && d_out && d_item_emb && d_pos_emb && d_ln_w && d_ln_b && workspace,
                "bsa_embed_ln_dropout_bwd: null argument");
    BSA_REQUIRE(workspace_bytes >= bsa_embed_bwd_workspace_bytes(c), "bsa_embed_ln_dropout_bwd: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const long long M = (long long)c->B * c->L;
    if (M == 0) return 0;
    char* p = (char*)workspace;
    float* dz = (float*)p; p += al((size_t)M * c->d * sizeof(float));
    int rc;
    // the scatter plan (key extraction + radix sort) depends on the ids only: it runs on the side stream beside the row kernel
    // and its reductions, and the main stream joins before the scatter-add reads it
    void* plan = p + 2 * al((size_t)rowwise_partial_blocks(M) * c->d * sizeof(float));
    const size_t plan_bytes = al(scatter_plan_bytes(M));
    Fork fk(st);
    if ((rc = scatter_plan((const long long*)ids, M, c->item_size, plan, plan_bytes, fk.leaf()))) return rc;
    if ((rc = embed_bwd_rows(c, ids, item_emb, pos_emb, ln_w, mean, rstd, d_out, dz, d_pos_emb, d_ln_w, d_ln_b, p, st))) return rc;
    if ((rc = fk.join())) return rc;
    return scatter_add_rows(plan, M, dz, d_item_emb, c->d, st);
}

int bsa_embed_ln_dropout_bwd_rows(const bsa_embed_desc* c, const int64_t* ids, const float* item_emb, const float* pos_emb,
                                  const float* ln_w, const float* mean, const float* rstd, const float* d_out,
                                  float* dz_rows, float* d_pos_emb, float* d_ln_w, float* d_ln_b, void* workspace,
                                  size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && ids && item_emb && pos_emb && ln_w && mean && rstd && d_out && dz_rows && d_pos_emb && d_ln_w && d_ln_b && workspace,
                "bsa_embed_ln_dropout_bwd_rows: null argument");
    BSA_REQUIRE(workspace_bytes >= bsa_embed_bwd_workspace_bytes(c), "bsa_embed_ln_dropout_bwd_rows: workspace too small");
    if ((long long)c->B * c->L == 0) return 0;
    return embed_bwd_rows(c, ids, item_emb, pos_emb, ln_w, mean, rstd, d_out, dz_rows, d_pos_emb, d_ln_w, d_ln_b,
                          (char*)workspace, (cudaStream_t)stream);
}

size_t bsa_scatter_add_workspace_bytes(int64_t n_rows) { return al(scatter_plan_bytes(n_rows)) + 256; }

int bsa_scatter_add_rows(const int64_t* ids, int64_t n_rows, int64_t item_size, const float* rows, int d, float* d_item_emb,
                         void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(ids && rows && d_item_emb && workspace, "bsa_scatter_add_rows: null argument");
    BSA_REQUIRE(workspace_bytes >= bsa_scatter_add_workspace_bytes(n_rows), "bsa_scatter_add_rows: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows == 0) return 0;
    int rc;
    if ((rc = scatter_plan((const long long*)ids, n_rows, item_size, workspace, al(scatter_plan_bytes(n_rows)), st))) return rc;
    return scatter_add_rows(workspace, n_rows, rows, d_item_emb, d, st);
}

size_t bsa_sort_scatter_plan_bytes(int64_t n) { return al(scatter_plan_bytes(n)); }

int bsa_sort_scatter_plan(const int64_t* ids, int64_t n, int64_t item_size, void* plan, size_t plan_bytes,
                          int32_t* sorted_ids, int32_t* sorted_rows, void* stream) {
    BSA_REQUIRE(ids && plan, "bsa_sort_scatter_plan: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    int rc = scatter_plan((const long long*)ids, n, item_size, plan, plan_bytes, st);
    if (rc || n == 0) return rc;
    const size_t arr = al((size_t)n * sizeof(int));
    if (sorted_ids) BSA_CHECK_CUDA(cudaMemcpyAsync(sorted_ids, (char*)plan + 2 * arr, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
    if (sorted_rows) BSA_CHECK_CUDA(cudaMemcpyAsync(sorted_rows, (char*)plan + 3 * arr, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return 0;
}

// ------------------------------------------------------------------------------------------ block
int bsa_filter_rank(int L, int kept_bins) { return filter_rank(L, kept_bins); }

int bsa_build_filter_basis(float* basis, int L, int kept_bins, void* stream) {
    BSA_REQUIRE(basis && L >= 1 && kept_bins >= 1, "bsa_build_filter_basis: bad argument");
    return build_filter_basis(basis, L, kept_bins, nullptr, (cudaStream_t)stream);
}

void bsa_set_ce_mode(int mode) { set_ce_fused(mode); }
int bsa_get_ce_mode(void) { return get_ce_fused(); }
void bsa_set_linear_impl(int impl) { set_linear_impl(impl); }
int bsa_get_linear_impl(void) { return linear_impl(); }
void bsa_set_attn_impl(int impl) { set_attn_impl(impl); }
int bsa_get_attn_impl(void) { return attn_impl(); }

size_t bsa_block_saved_bytes(const bsa_block_desc* c) { return block_saved_bytes(*c); }
size_t bsa_block_workspace_bytes(const bsa_block_desc* c) { return block_workspace_bytes(*c); }

int bsa_block_fwd(const bsa_block_desc* c, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                  const float* x_in, float* x_out, void* saved, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && w && basis && ids && x_in && x_out && saved && workspace, "bsa_block_fwd: null argument");
    return block_fwd(*c, *w, basis, (const long long*)ids, x_in, x_out, saved, workspace, workspace_bytes, (cudaStream_t)stream);
}

int bsa_block_bwd(const bsa_block_desc* c, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                  const float* x_in, const void* saved, const float* d_x_out, float* d_x_in, const bsa_block_grads* g,
                  void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && w && basis && ids && x_in && saved && d_x_out && d_x_in && g && workspace, "bsa_block_bwd: null argument");
    return block_bwd(*c, *w, basis, (const long long*)ids, x_in, saved, d_x_out, d_x_in, *g, workspace, workspace_bytes, (cudaStream_t)stream);
}

int bsa_block_last_supported(const bsa_block_desc* c) { return c && block_last_supported(*c) ? 1 : 0; }
size_t bsa_block_last_saved_bytes(const bsa_block_desc* c) { return block_last_saved_bytes(*c); }
size_t bsa_block_last_workspace_bytes(const bsa_block_desc* c) { return block_last_workspace_bytes(*c); }

int bsa_block_last_fwd(const bsa_block_desc* c, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                       const float* x_in, float* out_last, void* saved, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && w && basis && ids && x_in && out_last && saved && workspace, "bsa_block_last_fwd: null argument");
    return block_last_fwd(*c, *w, basis, (const long long*)ids, x_in, out_last, saved, workspace, workspace_bytes, (cudaStream_t)stream);
}

int bsa_block_last_bwd(const bsa_block_desc* c, const bsa_block_weights* w, const float* basis, const int64_t* ids,
                       const float* x_in, const void* saved, const float* d_out_last, float* d_x_in, const bsa_block_grads* g,
                       void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && w && basis && ids && x_in && saved && d_out_last && d_x_in && g && workspace, "bsa_block_last_bwd: null argument");
    return block_last_bwd(*c, *w, basis, (const long long*)ids, x_in, saved, d_out_last, d_x_in, *g, workspace, workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ cross-entropy
static int ce_impl(const bsa_ce_desc* c) {
    if (c->impl == 1 || c->impl == 2) return c->impl;
    return umma_ce_supported(c->B, c->d, c->item_size) ? 2 : 1;
}

size_t bsa_ce_workspace_bytes(const bsa_ce_desc* c) {
    size_t a = ce_simt_workspace_bytes(c->B, c->item_size);
    size_t b = umma_ce_workspace_bytes(c->B, c->d, c->item_size);
    return a > b ? a : b;
}

int bsa_ce_fwd(const bsa_ce_desc* c, const float* seq_out, const float* item_emb, const int64_t* answers, float* loss,
               float* lse, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && seq_out && item_emb && answers && loss && lse && workspace, "bsa_ce_fwd: null argument");
    BSA_REQUIRE(c->B > 0 && c->item_size > 0, "bsa_ce_fwd: empty batch or catalog");
    if (ce_impl(c) == 2) {
        BSA_REQUIRE(umma_ce_supported(c->B, c->d, c->item_size), "bsa_ce_fwd: tcgen05 path does not support B=%d d=%d", c->B, c->d);
        return umma_ce_fwd(c->B, c->d, c->item_size, seq_out, item_emb, (const long long*)answers, loss, lse, workspace, workspace_bytes, (cudaStream_t)stream);
    }
    return ce_fwd_simt(c->B, c->d, c->item_size, seq_out, item_emb, (const long long*)answers, loss, lse, workspace, workspace_bytes, (cudaStream_t)stream);
}

int bsa_ce_fwd_partial(const bsa_ce_desc* c, const float* seq_out, const float* item_shard, const int64_t* answers_local,
                       float* lse_part, float* tgt_part, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && seq_out && item_shard && answers_local && lse_part && tgt_part && workspace, "bsa_ce_fwd_partial: null argument");
    BSA_REQUIRE(c->B > 0 && c->item_size > 0, "bsa_ce_fwd_partial: empty batch or shard");
    BSA_REQUIRE(umma_ce_supported(c->B, c->d, c->item_size), "bsa_ce_fwd_partial: tcgen05 path does not support B=%d d=%d", c->B, c->d);
    return umma_ce_fwd(c->B, c->d, c->item_size, seq_out, item_shard, (const long long*)answers_local, nullptr, lse_part, workspace,
                       workspace_bytes, (cudaStream_t)stream, tgt_part);
}

int bsa_ce_bwd(const bsa_ce_desc* c, const float* seq_out, const float* item_emb, const int64_t* answers, const float* lse,
               const float* d_loss, float* d_seq_out, float* d_item_emb, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && seq_out && item_emb && answers && lse && d_loss && d_seq_out && d_item_emb && workspace, "bsa_ce_bwd: null argument");
    if (ce_impl(c) == 2) {
        BSA_REQUIRE(umma_ce_supported(c->B, c->d, c->item_size), "bsa_ce_bwd: tcgen05 path does not support B=%d d=%d", c->B, c->d);
        return umma_ce_bwd(c->B, c->d, c->item_size, seq_out, item_emb, (const long long*)answers, lse, d_loss, d_seq_out, d_item_emb, workspace, workspace_bytes, c->reuse_split != 0, (cudaStream_t)stream);
    }
    return ce_bwd_simt(c->B, c->d, c->item_size, seq_out, item_emb, (const long long*)answers, lse, d_loss, d_seq_out, d_item_emb, workspace, workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ scoring / top-k
static int topk_impl(const bsa_topk_desc* c) {
    if (c->impl == 1 || c->impl == 2) return c->impl;
    return umma_topk_supported(c->B, c->d, c->k, c->shard_rows) ? 2 : 1;
}

size_t bsa_score_topk_workspace_bytes(const bsa_topk_desc* c) {
    size_t a = topk_simt_workspace_bytes(c->B, c->shard_rows);
    size_t b = umma_topk_workspace_bytes(c->B, c->d, c->k, c->shard_rows);
    return a > b ? a : b;
}

int bsa_score_topk(const bsa_topk_desc* c, const float* seq_out, const float* item_shard, const int64_t* hist_indptr,
                   const int64_t* hist_indices, const int64_t* hist_rows, float* top_val, int64_t* top_idx, void* workspace,
                   size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(c && seq_out && item_shard && top_val && top_idx && workspace, "bsa_score_topk: null argument");
    BSA_REQUIRE((hist_indptr == nullptr) == (hist_indices == nullptr) || hist_indptr, "bsa_score_topk: hist_indices without hist_indptr");
    if (c->B == 0) return 0;
    if (topk_impl(c) == 2) {
        BSA_REQUIRE(umma_topk_supported(c->B, c->d, c->k, c->shard_rows), "bsa_score_topk: tcgen05 path does not support this shape");
        return umma_score_topk(c->B, c->d, c->k, c->shard_rows, c->shard_offset, seq_out, item_shard, (const long long*)hist_indptr,
                               (const long long*)hist_indices, (const long long*)hist_rows, top_val, (long long*)top_idx, workspace, workspace_bytes, c->reuse_split != 0, (cudaStream_t)stream);
    }
    return score_topk_simt(c->B, c->d, c->k, c->shard_rows, c->shard_offset, seq_out, item_shard, (const long long*)hist_indptr,
                           (const long long*)hist_indices, (const long long*)hist_rows, top_val, (long long*)top_idx, workspace, workspace_bytes, (cudaStream_t)stream);
}

int bsa_topk_merge(const float* cand_val, const int64_t* cand_idx, int G, int B, int k, float* top_val, int64_t* top_idx, void* stream) {
    BSA_REQUIRE(cand_val && cand_idx && top_val && top_idx, "bsa_topk_merge: null argument");
    return topk_merge(cand_val, (const long long*)cand_idx, G, B, k, top_val, (long long*)top_idx, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ generic GEMM
size_t bsa_gemm_workspace_bytes(int M, int N, int K, int64_t lda, int64_t ldb, int a_mn, int b_mn) {
    const long long a_rows = a_mn ? K : M, b_rows = b_mn ? K : N;
    return umma_gemm_workspace_bytes(a_rows * lda, b_rows * ldb, M, N, K);
}

int bsa_gemm(int impl, int M, int N, int K, const float* A, int64_t lda, int a_mn, const float* B, int64_t ldb, int b_mn,
             float* C, int64_t ldc, const float* bias, float alpha, int accumulate, void* workspace, size_t workspace_bytes,
             void* stream) {
    BSA_REQUIRE(A && B && C, "bsa_gemm: null argument");
    if (impl == 1) {
        GemmP p;
        p.A = A; p.B = B; p.C = C; p.M = M; p.N = N; p.K = K; p.lda = (int)lda; p.ldb = (int)ldb; p.ldc = (int)ldc;
        p.a_kmajor = !a_mn; p.b_kmajor = !b_mn; p.bias = bias; p.alpha = alpha; p.accumulate = accumulate;
        return launch_gemm(p, (cudaStream_t)stream);
    }
    BSA_REQUIRE(workspace && workspace_bytes >= bsa_gemm_workspace_bytes(M, N, K, lda, ldb, a_mn, b_mn), "bsa_gemm: workspace too small");
    const long long a_rows = a_mn ? K : M, a_cols = a_mn ? M : K, b_rows = b_mn ? K : N, b_cols = b_mn ? N : K;
    return umma_gemm(M, N, K, A, a_rows, a_cols, lda, a_mn != 0, B, b_rows, b_cols, ldb, b_mn != 0, C, ldc, bias, 0, nullptr,
                     nullptr, alpha, accumulate, workspace, workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ sibling model loss
size_t bsa_bce_pair_workspace_bytes(int B, int d) { return al(bce_pair_workspace_bytes(B, d)); }

int bsa_bce_pair_fwd(int B, int d, int64_t item_size, const float* seq_out, const float* item_emb, const int64_t* pos_ids,
                     const int64_t* neg_ids, float* loss, void* workspace, size_t workspace_bytes, int variant, void* stream) {
    BSA_REQUIRE(seq_out && item_emb && pos_ids && neg_ids && loss && workspace, "bsa_bce_pair_fwd: null argument");
    BSA_REQUIRE(B >= 1 && workspace_bytes >= bce_pair_workspace_bytes(B, d), "bsa_bce_pair_fwd: workspace too small");
    return bce_pair_fwd(B, d, item_size, seq_out, item_emb, (const long long*)pos_ids, (const long long*)neg_ids, loss, workspace,
                        variant, (cudaStream_t)stream);
}

int bsa_bce_pair_bwd(int B, int d, int64_t item_size, const float* seq_out, const float* item_emb, const int64_t* pos_ids,
                     const int64_t* neg_ids, const float* d_loss, float* d_seq_out, float* d_item_emb, void* workspace,
                     size_t workspace_bytes, int variant, void* stream) {
    BSA_REQUIRE(seq_out && item_emb && pos_ids && neg_ids && d_loss && d_seq_out && d_item_emb && workspace,
                "bsa_bce_pair_bwd: null argument");
    return bce_pair_bwd(B, d, item_size, seq_out, item_emb, (const long long*)pos_ids, (const long long*)neg_ids, d_loss,
                        d_seq_out, d_item_emb, workspace, workspace_bytes, variant, (cudaStream_t)stream);
}

int bsa_fmlp_supported(int L, int d) { return fmlp_supported(L, d) ? 1 : 0; }
int bsa_fmlp_tables(float* Tc, float* Ts, int L, void* stream) {
    BSA_REQUIRE(Tc && Ts && L >= 1, "bsa_fmlp_tables: bad argument");
    return fmlp_trig(Tc, Ts, L, (cudaStream_t)stream);
}
int bsa_fmlp_kernel_fwd(const float* W, const float* Tc, const float* Ts, float* K, int L, int d, void* stream) {
    BSA_REQUIRE(W && Tc && Ts && K, "bsa_fmlp_kernel_fwd: null argument");
    return fmlp_kernel_fwd(W, Tc, Ts, K, L, d, (cudaStream_t)stream);
}
int bsa_fmlp_kernel_bwd(const float* dK, const float* Tc, const float* Ts, float* dW, int L, int d, void* stream) {
    BSA_REQUIRE(dK && Tc && Ts && dW, "bsa_fmlp_kernel_bwd: null argument");
    return fmlp_kernel_bwd(dK, Tc, Ts, dW, L, d, (cudaStream_t)stream);
}

size_t bsa_info_nce_workspace_bytes(int B) { return al(nce_workspace_bytes(B)); }

int bsa_info_nce_fwd(int B, int d, float tau, const float* z, float* loss, void* workspace, size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(z && loss && workspace, "bsa_info_nce_fwd: null argument");
    BSA_REQUIRE(workspace_bytes >= nce_workspace_bytes(B), "bsa_info_nce_fwd: workspace too small");
    return nce_fwd(B, d, tau, z, loss, workspace, (cudaStream_t)stream);
}

int bsa_info_nce_bwd(int B, int d, float tau, const float* z, const float* d_loss, float scale, float* dz, const void* workspace,
                     size_t workspace_bytes, void* stream) {
    BSA_REQUIRE(z && d_loss && dz && workspace, "bsa_info_nce_bwd: null argument");
    BSA_REQUIRE(workspace_bytes >= nce_workspace_bytes(B), "bsa_info_nce_bwd: workspace too small");
    return nce_bwd(B, d, tau, z, d_loss, scale, dz, workspace, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ utilities
int bsa_gather_last(const float* src, float* dst, int B, int L, int d, void* stream) {
    BSA_REQUIRE(src && dst, "bsa_gather_last: null argument");
    if (B == 0) return 0;
    return gather_last_rows(src, dst, B, L, d, (cudaStream_t)stream);
}

int bsa_gather_rows(const int64_t* idx, int64_t n_rows, int64_t src_rows, const float* src, int d, float* dst, void* stream) {
    BSA_REQUIRE(n_rows >= 0 && src_rows >= 1 && d >= 4, "bsa_gather_rows: bad sizes");
    if (n_rows == 0) return 0;
    BSA_REQUIRE(idx && src && dst, "bsa_gather_rows: null argument");
    return gather_rows((const long long*)idx, n_rows, src_rows, src, d, dst, (cudaStream_t)stream);
}

int bsa_scatter_rows(const int64_t* idx, int64_t n_rows, int64_t dst_rows, const float* src, int d, float* dst, void* stream) {
    BSA_REQUIRE(n_rows >= 0 && dst_rows >= 1 && d >= 4, "bsa_scatter_rows: bad sizes");
    if (n_rows == 0) return 0;
    BSA_REQUIRE(idx && src && dst, "bsa_scatter_rows: null argument");
    return scatter_rows((const long long*)idx, n_rows, dst_rows, src, d, dst, (cudaStream_t)stream);
}

int bsa_scatter_last(const float* src, float* dst, int B, int L, int d, void* stream) {
    BSA_REQUIRE(src && dst, "bsa_scatter_last: null argument");
    if (B == 0) return 0;
    return scatter_last_rows(src, dst, B, L, d, (cudaStream_t)stream);
}

}  // extern "C"

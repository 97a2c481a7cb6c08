// Full-catalog scoring + history masking + per-user top-k (trainers.py:62-68,130-149).
// History items score exactly 0.0 (they are NOT removed); ordering is (score desc, id asc).
#include "../../include/bsarec_b200.h"
#include "common.cuh"
#include "gemm_simt.cuh"
#include "ce_topk.cuh"

namespace bsa {

namespace {

constexpr int TK_THREADS = 256;
constexpr int TK_TILE = TK_THREADS * 4;     // scores examined between two overflow checks
constexpr int TK_CAP = 4096;                // candidate buffer (entries)
constexpr int TK_FLUSH = 2048;              // compact when more than this many candidates are buffered

template <class IdxT>
__device__ __forceinline__ bool before(float va, IdxT ia, float vb, IdxT ib) {
    return (va > vb) || (va == vb && ia < ib);
}

// in-place bitonic sort of n (power of two) entries into (score desc, id asc) order
template <class IdxT>
__device__ void bitonic_sort_desc(float* val, IdxT* idx, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const bool up = ((i & k) == 0);        // "up" block: smaller rank first
                    const float va = val[i], vb = val[ixj];
                    const IdxT ia = idx[i], ib = idx[ixj];
                    const bool a_first = before<IdxT>(va, ia, vb, ib);
                    if (up ? !a_first : a_first) { val[i] = vb; val[ixj] = va; idx[i] = ib; idx[ixj] = ia; }
                }
            }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ int next_pow2(int n) { int p = 1; while (p < n) p <<= 1; return p; }

// is global id `gid` in the sorted history slice [lo, hi) ?
__device__ __forceinline__ bool in_history(const long long* __restrict__ h, long long lo, long long hi, long long gid) {
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        const long long v = h[mid];
        if (v == gid) return true;
        if (v < gid) lo = mid + 1; else hi = mid;
    }
    return false;
}

// state := the user's history items that fall in this shard, each with score 0.0 (first k by id)
__global__ void topk_init_kernel(const long long* __restrict__ indptr, const long long* __restrict__ indices,
                                 const long long* __restrict__ hrows, long long offset, long long rows, int k, float* __restrict__ top_val,
                                 long long* __restrict__ top_idx) {
    const int row = blockIdx.x;
    if (threadIdx.x != 0) return;
    int n = 0;
    const long long hr = hrows ? hrows[row] : row;
    const long long lo = indptr ? indptr[hr] : 0, hi = indptr ? indptr[hr + 1] : 0;
    for (long long p = lo; p < hi && n < k; ++p) {
        const long long gid = indices[p];
        if (gid >= offset && gid < offset + rows) { top_val[(long long)row * k + n] = 0.0f; top_idx[(long long)row * k + n] = gid; ++n; }
    }
    for (; n < k; ++n) { top_val[(long long)row * k + n] = -INFINITY; top_idx[(long long)row * k + n] = -1; }
}

// one CTA per user: fold scores[row, 0:ncols) (shard-local ids n0..) into the running top-k state.
// Candidates are kept in shared memory as (score, shard-local id int32).
__global__ void __launch_bounds__(TK_THREADS)
topk_chunk_kernel(const float* __restrict__ scores, int ld, long long offset, int n0, int ncols,
                  const long long* __restrict__ indptr, const long long* __restrict__ indices,
                  const long long* __restrict__ hrows, int k, float* __restrict__ top_val, long long* __restrict__ top_idx) {
    constexpr int SENT = 0x7fffffff;
    __shared__ float cval[TK_CAP];
    __shared__ int cidx[TK_CAP];
    __shared__ int count;
    __shared__ float thr_v;
    __shared__ int thr_i;
    const int row = blockIdx.x;
    const float* __restrict__ sr = scores + (long long)row * ld;
    const long long hr = hrows ? hrows[row] : row;
    const long long hlo = indptr ? indptr[hr] : 0, hhi = indptr ? indptr[hr + 1] : 0;
    for (int i = threadIdx.x; i < k; i += TK_THREADS) {
        const long long gi = top_idx[(long long)row * k + i];
        cval[i] = top_val[(long long)row * k + i];
        cidx[i] = gi < 0 ? SENT : (int)(gi - offset);
    }
    __syncthreads();
    if (threadIdx.x == 0) { count = k; thr_v = cval[k - 1]; thr_i = cidx[k - 1]; }
    __syncthreads();
    bool dirty = false;
    for (int base = 0; base < ncols; base += TK_TILE) {
        const float tv = thr_v;
        const int ti = thr_i;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = base + u * TK_THREADS + threadIdx.x;
            if (j < ncols) {
                const float v = sr[j];
                const int lid = n0 + j;
                if (before<int>(v, lid, tv, ti) && !in_history(indices, hlo, hhi, offset + lid)) {
                    const int slot = atomicAdd(&count, 1);
                    cval[slot] = v; cidx[slot] = lid;
                }
            }
        }
        __syncthreads();
        const int c = count;
        if (c > k) dirty = true;
        if (c > TK_FLUSH) {
            const int n = next_pow2(c);
            for (int i = c + threadIdx.x; i < n; i += TK_THREADS) { cval[i] = -INFINITY; cidx[i] = SENT; }
            __syncthreads();
            bitonic_sort_desc<int>(cval, cidx, n);
            if (threadIdx.x == 0) { count = k; thr_v = cval[k - 1]; thr_i = cidx[k - 1]; }
        }
        __syncthreads();
    }
    if (!dirty) return;
    const int c = count;
    if (c > k) {
        const int n = next_pow2(c);
        for (int i = c + threadIdx.x; i < n; i += TK_THREADS) { cval[i] = -INFINITY; cidx[i] = SENT; }
        __syncthreads();
        bitonic_sort_desc<int>(cval, cidx, n);
    }
    for (int i = threadIdx.x; i < k; i += TK_THREADS) {
        top_val[(long long)row * k + i] = cval[i];
        top_idx[(long long)row * k + i] = cidx[i] == SENT ? -1 : offset + cidx[i];
    }
}

__global__ void __launch_bounds__(TK_THREADS)
topk_merge_kernel(const float* __restrict__ cand_val, const long long* __restrict__ cand_idx, int G, int B, int k,
                  float* __restrict__ top_val, long long* __restrict__ top_idx) {
    __shared__ float cval[1024];
    __shared__ long long cidx[1024];
    const int row = blockIdx.x;
    const int c = G * k;
    const int n = next_pow2(c);
    for (int i = threadIdx.x; i < n; i += TK_THREADS) {
        if (i < c) {
            const int g = i / k, q = i - g * k;
            const long long o = ((long long)g * B + row) * k + q;
            const long long id = cand_idx[o];
            cval[i] = id < 0 ? -INFINITY : cand_val[o];
            cidx[i] = id < 0 ? 0x7fffffffffffffffll : id;
        } else { cval[i] = -INFINITY; cidx[i] = 0x7fffffffffffffffll; }
    }
    __syncthreads();
    bitonic_sort_desc<long long>(cval, cidx, n);
    for (int i = threadIdx.x; i < k; i += TK_THREADS) {
        top_val[(long long)row * k + i] = cval[i];
        top_idx[(long long)row * k + i] = cidx[i] == 0x7fffffffffffffffll ? -1 : cidx[i];
    }
}

}  // namespace

size_t topk_simt_workspace_bytes(int B, long long N) { return (size_t)B * ce_chunk_cols(B, N) * sizeof(float) + 1024; }

int score_topk_simt(int B, int d, int k, long long rows, long long offset, const float* s, const float* E,
                    const long long* hist_indptr, const long long* hist_indices, const long long* hist_rows, float* top_val,
                    long long* top_idx, void* ws, size_t ws_bytes, cudaStream_t st) {
    BSA_REQUIRE(ws_bytes >= topk_simt_workspace_bytes(B, rows), "bsa_score_topk: workspace too small");
    BSA_REQUIRE(k >= 1 && k <= 128 && rows < (1ll << 31), "bsa_score_topk: need 1 <= k <= 128 and shard rows < 2^31");
    const int nc = ce_chunk_cols(B, rows);
    float* scores = (float*)ws;
    topk_init_kernel<<<B, 32, 0, st>>>(hist_indptr, hist_indices, hist_rows, offset, rows, k, top_val, top_idx);
    BSA_LAUNCH_CHECK();
    for (long long n0 = 0; n0 < rows; n0 += nc) {
        const int cols = (int)((rows - n0) < nc ? (rows - n0) : nc);
        GemmP p;
        p.A = s; p.B = E + n0 * d; p.C = scores; p.M = B; p.N = cols; p.K = d; p.lda = d; p.ldb = d; p.ldc = nc;
        int rc = launch_gemm(p, st);
        if (rc) return rc;
        topk_chunk_kernel<<<B, TK_THREADS, 0, st>>>(scores, nc, offset, (int)n0, cols, hist_indptr, hist_indices, hist_rows, k, top_val, top_idx);
        BSA_LAUNCH_CHECK();
    }
    return 0;
}

int topk_merge(const float* cand_val, const long long* cand_idx, int G, int B, int k, float* top_val,
               long long* top_idx, cudaStream_t st) {
    BSA_REQUIRE(G >= 1 && k >= 1 && (long long)G * k <= 1024, "bsa_topk_merge: G*k = %d exceeds 1024", G * k);
    if (B == 0) return 0;
    topk_merge_kernel<<<B, TK_THREADS, 0, st>>>(cand_val, cand_idx, G, B, k, top_val, top_idx);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

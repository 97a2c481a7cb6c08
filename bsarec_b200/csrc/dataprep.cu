// Integer kernels either side of the model call: batch construction from a
// device-resident CSR of user sequences (what RecDataset.__getitem__ + the DataLoader collate do on the host,
// dataset.py:18-27,61-117) and the hit-rank histogram that HR@k / NDCG@k are sums over (metrics.py:3-23).
// Both are bit-exact by construction (no floating point on the device).
#include "../../include/bsarec_b200.h"
#include "common.cuh"

namespace bsa {
namespace {

// One warp per batch row.  Row b draws sample `sample_idx[b]`:
//   mode 0 (train):  sample s belongs to user u with sample_start[u] <= s < sample_start[u+1]; with n = len_u - 2 and
//                    w0 = max(n - L, 0) (the window seq[-(L+2):-2], dataset.py:20), i = s - sample_start[u]:
//                    input = seq[w0 : w0+i], answer = seq[w0+i], user_id = u            (dataset.py:21-23,62-64)
//   mode 1 (valid):  u = s, items = seq[:-1]:  input = seq[:len-2][-L:], answer = seq[len-2], user_id = s (dataset.py:24-26)
//   mode 2 (test):   u = s, items = seq:       input = seq[:len-1][-L:], answer = seq[len-1], user_id = s (dataset.py:28)
// input_ids is left-padded with 0 to L (dataset.py:69-72).
__global__ void build_batch_kernel(int mode, const long long* __restrict__ seq_indptr, const long long* __restrict__ seq_items,
                                   const long long* __restrict__ sample_start, long long n_users,
                                   const long long* __restrict__ sample_idx, int B, int L, long long* __restrict__ input_ids,
                                   long long* __restrict__ answers, long long* __restrict__ user_ids) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B) return;
    const long long s = sample_idx[warp];
    long long u, begin, end;     // input = seq[begin:end], answer = seq[end]  (offsets inside the user's sequence)
    if (mode == 0) {
        long long lo = 0, hi = n_users;            // last u with sample_start[u] <= s
        while (hi - lo > 1) {
            const long long mid = (lo + hi) >> 1;
            if (sample_start[mid] <= s) lo = mid; else hi = mid;
        }
        u = lo;
        const long long len = seq_indptr[u + 1] - seq_indptr[u];
        const long long w0 = (len - 2 > L) ? (len - 2 - L) : 0;
        begin = w0;
        end = w0 + (s - sample_start[u]);
    } else {
        u = s;
        const long long len = seq_indptr[u + 1] - seq_indptr[u];
        end = len - (mode == 1 ? 2 : 1);
        begin = end > L ? end - L : 0;
    }
    const long long base = seq_indptr[u];
    const int n = (int)(end - begin), pad = L - n;
    long long* out = input_ids + (long long)warp * L;
    for (int t = lane; t < L; t += 32) out[t] = (t < pad) ? 0ll : seq_items[base + begin + (t - pad)];
    if (lane == 0) {
        answers[warp] = seq_items[base + end];
        user_ids[warp] = u;
    }
}

// hist[r] += #users whose answer sits at position r of their list (r < k), hist[k] += #users with no hit.
// Thread per user, block-level shared histogram, one integer atomic per bin per block: the result does not depend on
// scheduling.  Positions past the first match are ignored: the lists hold distinct ids, so metrics.py:19-20's sum over j
// has at most one non-zero term.
__global__ void hit_rank_hist_kernel(const long long* __restrict__ pred, const long long* __restrict__ answers, long long U,
                                     int k, unsigned long long* __restrict__ hist) {
    extern __shared__ unsigned int sh[];
    for (int i = threadIdx.x; i <= k; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < U; u += (long long)gridDim.x * blockDim.x) {
        const long long a = answers[u];
        const long long* p = pred + u * k;
        int r = k;
        for (int j = 0; j < k; ++j)
            if (p[j] == a) { r = j; break; }
        atomicAdd(&sh[r], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i <= k; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

}  // namespace
}  // namespace bsa

extern "C" {

int bsa_build_batch(int mode, const int64_t* seq_indptr, const int64_t* seq_items, const int64_t* sample_start, int64_t n_users,
                    const int64_t* sample_idx, int B, int L, int64_t* input_ids, int64_t* answers, int64_t* user_ids,
                    void* stream) {
    BSA_REQUIRE(mode >= 0 && mode <= 2, "bsa_build_batch: mode must be 0 (train), 1 (valid) or 2 (test)");
    BSA_REQUIRE(B >= 0 && L > 0 && n_users > 0, "bsa_build_batch: bad sizes");
    if (B == 0) return 0;
    BSA_REQUIRE(seq_indptr && seq_items && sample_idx && input_ids && answers && user_ids, "bsa_build_batch: null argument");
    BSA_REQUIRE(mode != 0 || sample_start, "bsa_build_batch: train mode needs sample_start");
    const int threads = 256, rows_per_block = threads / 32;
    bsa::build_batch_kernel<<<bsa::cdiv(B, rows_per_block), threads, 0, (cudaStream_t)stream>>>(
        mode, (const long long*)seq_indptr, (const long long*)seq_items, (const long long*)sample_start, n_users,
        (const long long*)sample_idx, B, L, (long long*)input_ids, (long long*)answers, (long long*)user_ids);
    BSA_LAUNCH_CHECK();
    return 0;
}

int bsa_hit_rank_hist(const int64_t* pred, const int64_t* answers, int64_t U, int k, int64_t* hist, void* stream) {
    BSA_REQUIRE(k > 0 && k <= 1024 && U >= 0, "bsa_hit_rank_hist: bad sizes");
    if (U == 0) return 0;
    BSA_REQUIRE(pred && answers && hist, "bsa_hit_rank_hist: null argument");
    const int threads = 256;
    long long blocks = (U + threads - 1) / threads;
    if (blocks > 148 * 8) blocks = 148 * 8;
    bsa::hit_rank_hist_kernel<<<(int)blocks, threads, (k + 1) * sizeof(unsigned int), (cudaStream_t)stream>>>(
        (const long long*)pred, (const long long*)answers, U, k, (unsigned long long*)hist);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"

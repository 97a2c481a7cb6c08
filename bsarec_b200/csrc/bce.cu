// Pairwise binary cross-entropy at the last position — the loss of the sibling model SASRec
// (model/sasrec.py:42-63): pos/neg logits = <E[pos_b], s_b>, <E[neg_b], s_b>, BCEWithLogits against ones / zeros, each
// averaged over the users whose positive id is not padding.  SURVEY.md §8(f) rank 4: the encoder underneath is the
// BSARec block with the filter branch blended out (alpha = 0), this file is the only new arithmetic.
// Warp per user, float4 lanes, fixed-order reductions (no atomics) -> bit-reproducible.
#include "../../include/bsarec_b200.h"
#include "common.cuh"
#include "rowwise.cuh"
#include <mutex>

namespace bsa {
namespace {

inline size_t al(size_t x) { return (x + 255) / 256 * 256; }

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }
__device__ __forceinline__ float sigmoid_f(float x) { return x >= 0.f ? 1.f / (1.f + expf(-x)) : expf(x) / (1.f + expf(x)); }

// logits[b] = (pos logit, neg logit); one block, warps stride over users; loss = mean_valid softplus(-xp) + mean_valid softplus(xn)
__global__ void __launch_bounds__(1024) bce_pair_fwd_kernel(const float* __restrict__ s, const float* __restrict__ E,
                                                             const long long* __restrict__ pos, const long long* __restrict__ neg,
                                                             int B, int d, long long N, float* __restrict__ logits,
                                                             float* __restrict__ loss, int* __restrict__ n_valid, int variant) {
    __shared__ float sh_p[32], sh_n[32];
    __shared__ int sh_c[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    float lp = 0.f, ln = 0.f;
    int cnt = 0;
    for (int b = warp; b < B; b += nw) {
        long long ip = pos[b], in = neg[b];
        if (ip < 0 || ip >= N) ip = 0;
        if (in < 0 || in >= N) in = 0;
        const float4* sv = reinterpret_cast<const float4*>(s + (long long)b * d);
        const float4* ep = reinterpret_cast<const float4*>(E + ip * d);
        const float4* en = reinterpret_cast<const float4*>(E + in * d);
        float xp = 0.f, xn = 0.f;
        for (int c = lane; c < d / 4; c += 32) {
            const float4 a = sv[c], p4 = __ldg(ep + c), n4 = __ldg(en + c);
            xp += a.x * p4.x + a.y * p4.y + a.z * p4.z + a.w * p4.w;
            xn += a.x * n4.x + a.y * n4.y + a.z * n4.z + a.w * n4.w;
        }
        xp = warp_sum_f(xp); xn = warp_sum_f(xn);
        if (lane == 0) { logits[2 * b] = xp; logits[2 * b + 1] = xn; }
        if (variant == 1) {          // fmlprec.py:55-58: -log(sigmoid(xp) + 1e-24) - log(1 - sigmoid(xn) + 1e-24), mean over ALL users
            lp += -logf(sigmoid_f(xp) + 1e-24f); ln += -logf(1.0f - sigmoid_f(xn) + 1e-24f); ++cnt;
        } else if (pos[b] != 0) { lp += softplus_f(-xp); ln += softplus_f(xn); ++cnt; }     // sasrec.py:58: indices = pos_ids != 0
    }
    if (lane == 0) { sh_p[warp] = lp; sh_n[warp] = ln; sh_c[warp] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float tp = 0.f, tn = 0.f;
        int c = 0;
        for (int w = 0; w < nw; ++w) { tp += sh_p[w]; tn += sh_n[w]; c += sh_c[w]; }
        *n_valid = c;
        *loss = c > 0 ? tp / (float)c + tn / (float)c : nanf("");        // torch: mean over an empty selection is nan
    }
}

// ds[b,:] = gp E[pos_b] + gn E[neg_b]; rows[b,:] = gp s_b (-> dE[pos_b]), rows[B+b,:] = gn s_b (-> dE[neg_b]); ids2 = (pos, neg)
// with users whose positive is padding zeroed out (id 0 rows are skipped by the scatter-add: padding_idx = 0).
__global__ void __launch_bounds__(256) bce_pair_bwd_kernel(const float* __restrict__ s, const float* __restrict__ E,
                                                            const long long* __restrict__ pos, const long long* __restrict__ neg,
                                                            const float* __restrict__ logits, const int* __restrict__ n_valid,
                                                            const float* __restrict__ d_loss, int B, int d, long long N,
                                                            float* __restrict__ ds, float* __restrict__ rows, long long* __restrict__ ids2,
                                                            int variant) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (b >= B) return;
    const bool ok = (variant == 1 || pos[b] != 0) && *n_valid > 0;
    const float g = ok ? __ldg(d_loss) / (float)(*n_valid) : 0.f;
    float gp, gn;
    if (variant == 1) {              // d/dx of -log(sigmoid(x) + eps) and of -log(1 - sigmoid(x) + eps)
        const float sp = sigmoid_f(logits[2 * b]), sn = sigmoid_f(logits[2 * b + 1]);
        gp = -sp * (1.0f - sp) / (sp + 1e-24f) * g;
        gn = sn * (1.0f - sn) / (1.0f - sn + 1e-24f) * g;
    } else {
        gp = (sigmoid_f(logits[2 * b]) - 1.f) * g;
        gn = sigmoid_f(logits[2 * b + 1]) * g;
    }
    long long ip = pos[b], in = neg[b];
    if (ip < 0 || ip >= N) ip = 0;
    if (in < 0 || in >= N) in = 0;
    if (lane == 0) { ids2[b] = ok ? ip : 0; ids2[B + b] = ok ? in : 0; }
    const float4* sv = reinterpret_cast<const float4*>(s + (long long)b * d);
    const float4* ep = reinterpret_cast<const float4*>(E + ip * d);
    const float4* en = reinterpret_cast<const float4*>(E + in * d);
    float4* dso = reinterpret_cast<float4*>(ds + (long long)b * d);
    float4* rp = reinterpret_cast<float4*>(rows + (long long)b * d);
    float4* rn = reinterpret_cast<float4*>(rows + (long long)(B + b) * d);
    for (int c = lane; c < d / 4; c += 32) {
        const float4 a = sv[c], p4 = __ldg(ep + c), n4 = __ldg(en + c);
        dso[c] = make_float4(gp * p4.x + gn * n4.x, gp * p4.y + gn * n4.y, gp * p4.z + gn * n4.z, gp * p4.w + gn * n4.w);
        rp[c] = make_float4(gp * a.x, gp * a.y, gp * a.z, gp * a.w);
        rn[c] = make_float4(gn * a.x, gn * a.y, gn * a.z, gn * a.w);
    }
}

// ------------------------------------------------------------------------------------------ in-batch InfoNCE (DuoRec)
// model/duorec.py:47-75 with sim = 'dot': z [2B, d] = (view i rows, view j rows); row r's positive is its pair
// (r + B) mod 2B, its negatives every other row but itself; CrossEntropy(label 0) over (positive, negatives) is
//   loss_r = -sim[r, pair] + logsumexp_{c != r} sim[r, c],   sim = z z^T / tau;   loss = mean_r loss_r.
// One block per row; sims are recomputed in the backward (2B x 2B x d FMAs: nothing).  Fixed-order reductions.
__global__ void __launch_bounds__(256) nce_fwd_kernel(const float* __restrict__ z, int n2, int B, int d, float inv_tau,
                                                       float* __restrict__ lse, float* __restrict__ loss_rows) {
    extern __shared__ float nsm[];                 // [d] this row | [n2] sims
    float* zi = nsm;
    float* sim = nsm + d;
    __shared__ float red[8];
    const int r = blockIdx.x;
    for (int c = threadIdx.x; c < d; c += blockDim.x) zi[c] = z[(long long)r * d + c];
    __syncthreads();
    for (int j = threadIdx.x; j < n2; j += blockDim.x) {
        const float4* zj = reinterpret_cast<const float4*>(z + (long long)j * d);
        float acc = 0.f;
        for (int c = 0; c < d / 4; ++c) {
            const float4 a = zj[c];
            acc += a.x * zi[4 * c] + a.y * zi[4 * c + 1] + a.z * zi[4 * c + 2] + a.w * zi[4 * c + 3];
        }
        sim[j] = acc * inv_tau;
    }
    __syncthreads();
    float m = -INFINITY;
    for (int j = threadIdx.x; j < n2; j += blockDim.x) if (j != r) m = fmaxf(m, sim[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = red[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = fmaxf(m, red[w]);
    __syncthreads();
    float su = 0.f;
    for (int j = threadIdx.x; j < n2; j += blockDim.x) if (j != r) su += expf(sim[j] - m);
    su = warp_sum_f(su);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = su;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        const float l = m + logf(t);
        lse[r] = l;
        loss_rows[r] = l - sim[(r + B) % n2];
    }
}

__global__ void nce_mean_kernel(const float* __restrict__ loss_rows, int n2, float* __restrict__ loss) {
    __shared__ float red[32];
    float s = 0.f;
    for (int j = threadIdx.x; j < n2; j += blockDim.x) s += loss_rows[j];
    s = warp_sum_f(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        *loss = t / (float)n2;
    }
}

// dz[r,:] = inv_tau * sum_c (G[r,c] + G[c,r]) z[c,:],  G[r,c] = (exp(sim[r,c] - lse_r) - [c == pair(r)]) * g / 2B for c != r
__global__ void __launch_bounds__(256) nce_bwd_kernel(const float* __restrict__ z, const float* __restrict__ lse, int n2, int B,
                                                       int d, float inv_tau, const float* __restrict__ d_loss, float scale,
                                                       float* __restrict__ dz) {
    extern __shared__ float nsm[];                 // [d] this row | [n2] coefficients
    float* zi = nsm;
    float* coef = nsm + d;
    const int r = blockIdx.x;
    const float g = __ldg(d_loss) * scale / (float)n2;
    for (int c = threadIdx.x; c < d; c += blockDim.x) zi[c] = z[(long long)r * d + c];
    __syncthreads();
    const float lr = lse[r];
    for (int j = threadIdx.x; j < n2; j += blockDim.x) {
        float cf = 0.f;
        if (j != r) {
            const float4* zj = reinterpret_cast<const float4*>(z + (long long)j * d);
            float acc = 0.f;
            for (int c = 0; c < d / 4; ++c) {
                const float4 a = zj[c];
                acc += a.x * zi[4 * c] + a.y * zi[4 * c + 1] + a.z * zi[4 * c + 2] + a.w * zi[4 * c + 3];
            }
            const float sm = acc * inv_tau;
            const float pair = (j == (r + B) % n2) ? 1.f : 0.f;        // pairing is symmetric: j == pair(r) <=> r == pair(j)
            cf = (expf(sm - lr) - pair) + (expf(sm - lse[j]) - pair);
        }
        coef[j] = cf * g * inv_tau;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
        float acc = 0.f;
        for (int j = 0; j < n2; ++j) acc += coef[j] * z[(long long)j * d + c];
        dz[(long long)r * d + c] = acc;
    }
}

}  // namespace

size_t nce_workspace_bytes(int B) { return al((size_t)2 * B * 4) * 2 + 256; }

// z [2B, d]; loss = mean InfoNCE; ws keeps lse [2B] for the backward
int nce_fwd(int B, int d, float tau, const float* z, float* loss, void* ws, cudaStream_t st) {
    BSA_REQUIRE(d % 4 == 0 && B >= 1 && tau > 0.f, "info_nce: need d %% 4 == 0, B >= 1, tau > 0");
    const int n2 = 2 * B;
    const size_t smem = (size_t)(d + n2) * 4;
    BSA_REQUIRE(smem <= 48 * 1024, "info_nce: 2B + d = %d floats exceed the shared-memory plan", d + n2);
    float* lse = (float*)ws;
    float* rows = (float*)((char*)ws + al((size_t)n2 * 4));
    nce_fwd_kernel<<<n2, 256, smem, st>>>(z, n2, B, d, 1.0f / tau, lse, rows);
    BSA_LAUNCH_CHECK();
    nce_mean_kernel<<<1, 1024, 0, st>>>(rows, n2, loss);
    BSA_LAUNCH_CHECK();
    return 0;
}

int nce_bwd(int B, int d, float tau, const float* z, const float* d_loss, float scale, float* dz, const void* ws, cudaStream_t st) {
    const int n2 = 2 * B;
    const size_t smem = (size_t)(d + n2) * 4;
    nce_bwd_kernel<<<n2, 256, smem, st>>>(z, (const float*)ws, n2, B, d, 1.0f / tau, d_loss, scale, dz);
    BSA_LAUNCH_CHECK();
    return 0;
}

size_t bce_pair_workspace_bytes(int B, int d) {
    return al((size_t)2 * B * 4) + al(16) + al((size_t)2 * B * d * 4) + al((size_t)2 * B * 8) + scatter_plan_bytes(2ll * B) + 1024;
}

int bce_pair_fwd(int B, int d, long long N, const float* s, const float* E, const long long* pos, const long long* neg,
                 float* loss, void* ws, int variant, cudaStream_t st) {
    BSA_REQUIRE(d % 4 == 0, "bce_pair: hidden size must be a multiple of 4");
    float* logits = (float*)ws;
    int* n_valid = (int*)((char*)ws + al((size_t)2 * B * 4));
    bce_pair_fwd_kernel<<<1, 1024, 0, st>>>(s, E, pos, neg, B, d, N, logits, loss, n_valid, variant);
    BSA_LAUNCH_CHECK();
    return 0;
}

int bce_pair_bwd(int B, int d, long long N, const float* s, const float* E, const long long* pos, const long long* neg,
                 const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, int variant, cudaStream_t st) {
    BSA_REQUIRE(ws_bytes >= bce_pair_workspace_bytes(B, d), "bce_pair_bwd: workspace too small");
    char* p = (char*)ws;
    float* logits = (float*)p; p += al((size_t)2 * B * 4);
    int* n_valid = (int*)p; p += al(16);
    float* rows = (float*)p; p += al((size_t)2 * B * d * 4);
    long long* ids2 = (long long*)p; p += al((size_t)2 * B * 8);
    bce_pair_bwd_kernel<<<cdiv(B, 8), 256, 0, st>>>(s, E, pos, neg, logits, n_valid, d_loss, B, d, N, ds, rows, ids2, variant);
    BSA_LAUNCH_CHECK();
    int rc = scatter_plan(ids2, 2ll * B, N, p, scatter_plan_bytes(2ll * B), st);
    if (rc) return rc;
    return scatter_add_rows(p, 2ll * B, rows, dE, d, st);       // dE accumulated: the caller zeroes it once per step
}

}  // namespace bsa

// ================================================================================================ no-allocation variant
// bsa_info_nce_fwd_static: the InfoNCE forward with its scratch (row log-sum-exps and per-row losses, 2 x [2B] floats) in
// a file-scope device array instead of a caller-provided workspace — for callers that cannot allocate (code under CUDA
// graph capture, tight serving loops).  Every GPU that loads the library holds the array until exit.  All calls share it
// and it keeps the last call's contents; both kernels overwrite every element they later read (nothing has to start at
// zero), but two calls that overlap on the GPU (different streams) race on it.
namespace bsa {
namespace {
constexpr int NCE_STATIC_MAX_B = BSA_INFO_NCE_STATIC_MAX_B;
constexpr size_t NCE_STATIC_HALF = ((size_t)2 * NCE_STATIC_MAX_B * 4 + 255) / 256 * 256;     // bytes of one [2B] array
static __device__ __align__(256) unsigned char g_nce_static_scratch[2 * NCE_STATIC_HALF];

constexpr int NCE_MAX_DEVICES = 64;
void* g_nce_static_ptr[NCE_MAX_DEVICES] = {};
std::mutex g_nce_static_mu;

// the array's NAME is a host-side stand-in in host code: its device address is looked up once per device, with that
// device current, and that pointer is what the kernels get
int nce_static_scratch(void** out) {
    int dev = -1;
    BSA_CHECK_CUDA(cudaGetDevice(&dev));
    BSA_REQUIRE(dev >= 0 && dev < NCE_MAX_DEVICES, "bsa_info_nce_fwd_static: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_nce_static_mu);
    if (!g_nce_static_ptr[dev]) {
        void* p = nullptr;
        BSA_CHECK_CUDA(cudaGetSymbolAddress(&p, g_nce_static_scratch));
        BSA_REQUIRE(p != nullptr, "bsa_info_nce_fwd_static: cudaGetSymbolAddress returned a null address");
        g_nce_static_ptr[dev] = p;
    }
    *out = g_nce_static_ptr[dev];
    return 0;
}
}  // namespace
}  // namespace bsa

extern "C" int bsa_info_nce_static_max_b(void) { return bsa::NCE_STATIC_MAX_B; }

extern "C" int bsa_info_nce_fwd_static(int B, int d, float tau, const float* z, float* loss, void* stream) {
    using namespace bsa;
    BSA_REQUIRE(z && loss, "bsa_info_nce_fwd_static: null argument");
    BSA_REQUIRE(B >= 1 && B <= NCE_STATIC_MAX_B,
                "bsa_info_nce_fwd_static: B = %d outside [1, %d] (the static scratch holds 2 x [2 * %d] floats; use "
                "bsa_info_nce_fwd with a workspace for larger batches)", B, NCE_STATIC_MAX_B, NCE_STATIC_MAX_B);
    BSA_REQUIRE(d >= 4 && d % 4 == 0 && tau > 0.f, "bsa_info_nce_fwd_static: need d %% 4 == 0, tau > 0");
    const int n2 = 2 * B;
    const size_t smem = (size_t)(d + n2) * 4;
    BSA_REQUIRE(smem <= 48 * 1024, "bsa_info_nce_fwd_static: 2B + d = %d floats exceed the shared-memory plan", d + n2);
    void* base = nullptr;
    int rc = nce_static_scratch(&base);
    if (rc) return rc;
    float* lse = (float*)base;
    float* rows = (float*)((char*)base + NCE_STATIC_HALF);
    cudaStream_t st = (cudaStream_t)stream;
    nce_fwd_kernel<<<n2, 256, smem, st>>>(z, n2, B, d, 1.0f / tau, lse, rows);
    BSA_LAUNCH_CHECK();
    nce_mean_kernel<<<1, 1024, 0, st>>>(rows, n2, loss);
    BSA_LAUNCH_CHECK();
    return 0;
}


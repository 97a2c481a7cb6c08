// Row-wise kernels (one warp per row of d or L values, values held in registers).
#include "rowwise.cuh"
#include "umma_engine.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cmath>
#include <vector>

namespace bsa {

namespace {

constexpr int WARPS = 8;                 // warps per block for the row-per-warp kernels
constexpr int MAX_ROW_BLOCKS = 148 * 4;  // persistent-style grid: <= 4 blocks per SM
constexpr float LN_EPS = 1e-12f;

#define BSA_DISPATCH_VPL(n, ...)                                                            \
    do {                                                                                    \
        if ((n) <= 64) { constexpr int VPL = 2; __VA_ARGS__; }                              \
        else if ((n) <= 256) { constexpr int VPL = 8; __VA_ARGS__; }                        \
        else if ((n) <= 1024) { constexpr int VPL = 32; __VA_ARGS__; }                      \
        else { set_error("row length %d > 1024 is not supported", (int)(n)); return -1; }   \
    } while (0)

// the float4 kernels of rowwise_vec.cu need d % 4 == 0 and 16-byte aligned tensors (null pointers = absent operands)
template <class... P>
inline bool vec_ok(int d, P... ptrs) {
    const uintptr_t bits = (... | (uintptr_t)ptrs);
    return rowwise_vec_ok(d) && (bits & 15u) == 0;
}

inline int row_grid(long long rows) {
    long long g = (rows + WARPS - 1) / WARPS;
    if (g > MAX_ROW_BLOCKS) g = MAX_ROW_BLOCKS;
    if (g < 1) g = 1;
    return (int)g;
}

// mean / rstd of one row held as v[VPL] per lane (two-pass, as _modules.py:16-20)
template <int VPL>
__device__ __forceinline__ void row_stats(const float (&v)[VPL], int d, int lane, float& mean, float& rstd) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) if (lane + 32 * i < d) s += v[i];
    mean = warp_sum(s) / (float)d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) if (lane + 32 * i < d) { const float c = v[i] - mean; q += c * c; }
    const float var = warp_sum(q) / (float)d;
    rstd = 1.0f / sqrtf(var + LN_EPS);
}

// block-level fixed-order reduction of per-warp/per-lane partials into part[blockIdx.x][0..d)
template <int VPL>
__device__ __forceinline__ void block_store_partials(const float (&a)[VPL], float* __restrict__ part, int d) {
    __shared__ float stage[WARPS][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        __syncthreads();
        stage[w][lane] = a[i];
        __syncthreads();
        if (w == 0) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < WARPS; ++k) s += stage[k][lane];
            const int j = lane + 32 * i;
            if (j < d) part[(long long)blockIdx.x * d + j] = s;
        }
    }
}

// ============================================================================ embedding
template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
embed_fwd_kernel(const long long* __restrict__ ids, const float* __restrict__ E, const float* __restrict__ Ppos,
                 const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ out,
                 float* __restrict__ mean_o, float* __restrict__ rstd_o, long long M, int L, int d, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        const long long id = ids[row];
        const int t = (int)(row % L);
        const float* __restrict__ e = E + id * d;
        const float* __restrict__ pp = Ppos + (long long)t * d;
        float v[VPL];
#pragma unroll
        for (int i = 0; i < VPL; ++i) { const int j = lane + 32 * i; v[i] = j < d ? __ldg(e + j) + __ldg(pp + j) : 0.f; }
        float mean, rstd;
        row_stats<VPL>(v, d, lane, mean, rstd);
        if (lane == 0) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) {
                const float y = __ldg(gamma + j) * ((v[i] - mean) * rstd) + __ldg(beta + j);
                out[row * d + j] = y * dropout_mult(dr, (unsigned long long)(row * d + j));
            }
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
embed_bwd_kernel(const float* __restrict__ dout, const long long* __restrict__ ids, const float* __restrict__ E,
                 const float* __restrict__ Ppos, const float* __restrict__ gamma, const float* __restrict__ mean_i,
                 const float* __restrict__ rstd_i, float* __restrict__ dz, float* __restrict__ pgamma,
                 float* __restrict__ pbeta, long long M, int L, int d, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    float ag[VPL], ab[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) ag[i] = ab[i] = 0.f;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        const long long id = ids[row];
        const int t = (int)(row % L);
        const float mean = mean_i[row], rstd = rstd_i[row];
        float xh[VPL], g[VPL];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            xh[i] = g[i] = 0.f;
            if (j < d) {
                const float zv = __ldg(E + id * d + j) + __ldg(Ppos + (long long)t * d + j);
                xh[i] = (zv - mean) * rstd;
                const float dy = dout[row * d + j] * dropout_mult(dr, (unsigned long long)(row * d + j));
                ag[i] += dy * xh[i];
                ab[i] += dy;
                g[i] = dy * __ldg(gamma + j);
                c1 += g[i];
                c2 += g[i] * xh[i];
            }
        }
        c1 = warp_sum(c1) / (float)d;
        c2 = warp_sum(c2) / (float)d;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) dz[row * d + j] = rstd * (g[i] - c1 - xh[i] * c2);
        }
    }
    block_store_partials<VPL>(ag, pgamma, d);
    block_store_partials<VPL>(ab, pbeta, d);
}

// ============================================================================ residual + LN (+ blend)
template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
residual_ln_fwd_kernel(const float* __restrict__ tmp, const float* __restrict__ resid, const float* __restrict__ gamma,
                       const float* __restrict__ beta, float* __restrict__ out, float* __restrict__ z_save,
                       float* __restrict__ mean_o, float* __restrict__ rstd_o, const float* __restrict__ blend_with,
                       float blend_a, float blend_b, long long M, int d, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        float v[VPL];
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            v[i] = 0.f;
            if (j < d) {
                const long long o = row * d + j;
                v[i] = tmp[o] * dropout_mult_row(dr, row, d, j) + resid[o];
                if (z_save) z_save[o] = v[i];
            }
        }
        float mean, rstd;
        row_stats<VPL>(v, d, lane, mean, rstd);
        if (lane == 0 && mean_o) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) {
                const long long o = row * d + j;
                float y = __ldg(gamma + j) * ((v[i] - mean) * rstd) + __ldg(beta + j);
                if (blend_with) y = blend_a * blend_with[o] + blend_b * y;
                out[o] = y;
            }
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
ln_bwd_kernel(const float* __restrict__ dy, float dy_scale, const float* __restrict__ z, const float* __restrict__ mean_i,
              const float* __restrict__ rstd_i, const float* __restrict__ gamma, float* __restrict__ dz,
              float* __restrict__ ddrop, float* __restrict__ pgamma, float* __restrict__ pbeta, long long M, int d,
              Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    float ag[VPL], ab[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) ag[i] = ab[i] = 0.f;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        const float mean = mean_i[row], rstd = rstd_i[row];
        float xh[VPL], g[VPL];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            xh[i] = g[i] = 0.f;
            if (j < d) {
                const long long o = row * d + j;
                xh[i] = (z[o] - mean) * rstd;
                const float up = dy[o] * dy_scale;
                ag[i] += up * xh[i];
                ab[i] += up;
                g[i] = up * __ldg(gamma + j);
                c1 += g[i];
                c2 += g[i] * xh[i];
            }
        }
        c1 = warp_sum(c1) / (float)d;
        c2 = warp_sum(c2) / (float)d;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) {
                const long long o = row * d + j;
                const float r = rstd * (g[i] - c1 - xh[i] * c2);
                dz[o] = r;
                if (ddrop) ddrop[o] = r * dropout_mult_row(dr, row, d, j);
            }
        }
    }
    block_store_partials<VPL>(ag, pgamma, d);
    block_store_partials<VPL>(ab, pbeta, d);
}

// ============================================================================ filter branch
// One CTA per sequence.  coef[r][d] = C^T x kept in shared memory, then one warp per position.
constexpr int RCH = 8;   // basis vectors accumulated per pass over the sequence

// coef[rr][j] = sum_t Cb[t][rr] * src(t,j)   (src supplied by a functor so fwd/bwd share it)
template <class Src>
__device__ __forceinline__ void project_onto_basis(float* __restrict__ coef, const float* __restrict__ Cb, int L, int d,
                                                   int r, Src src) {
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        for (int r0 = 0; r0 < r; r0 += RCH) {
            float acc[RCH];
#pragma unroll
            for (int q = 0; q < RCH; ++q) acc[q] = 0.f;
            for (int t = 0; t < L; ++t) {
                const float xv = src(t, j);
#pragma unroll
                for (int q = 0; q < RCH; ++q)
                    if (r0 + q < r) acc[q] = fmaf(__ldg(Cb + (long long)t * r + r0 + q), xv, acc[q]);
            }
#pragma unroll
            for (int q = 0; q < RCH; ++q)
                if (r0 + q < r) coef[(r0 + q) * d + j] = acc[q];
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
filter_ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ Cb, const float* __restrict__ sqrt_beta,
                     const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ dsp,
                     float* __restrict__ z_save, float* __restrict__ mean_o, float* __restrict__ rstd_o, int L, int d,
                     int r, Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ float coef[];     // [r][d]
    const int b = blockIdx.x;
    const float* __restrict__ xb = x + (long long)b * L * d;
    project_onto_basis(coef, Cb, L, d, r, [&](int t, int j) { return xb[(long long)t * d + j]; });
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int t = threadIdx.x >> 5; t < L; t += WARPS) {
        const long long row = (long long)b * L + t;
        float v[VPL];
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            v[i] = 0.f;
            if (j < d) {
                float low = 0.f;
                for (int q = 0; q < r; ++q) low = fmaf(__ldg(Cb + (long long)t * r + q), coef[q * d + j], low);
                const float xv = xb[(long long)t * d + j];
                const float sb = __ldg(sqrt_beta + j);
                const float u = low + (sb * sb) * (xv - low);
                const long long o = row * d + j;
                v[i] = u * dropout_mult(dr, (unsigned long long)o) + xv;
                if (z_save) z_save[o] = v[i];
            }
        }
        float mean, rstd;
        row_stats<VPL>(v, d, lane, mean, rstd);
        if (lane == 0 && mean_o) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) dsp[row * d + j] = __ldg(gamma + j) * ((v[i] - mean) * rstd) + __ldg(beta + j);
        }
    }
}

// u = beta2*x + (1-beta2)*F x  =>  dx += dz_f + beta2*du + F((1-beta2)*du)   (F symmetric)
// dbeta2[j] = sum_{t} du*(x - low)
__global__ void __launch_bounds__(WARPS * 32)
filter_bwd_kernel(const float* __restrict__ x, const float* __restrict__ Cb, const float* __restrict__ sqrt_beta,
                  const float* __restrict__ dz_f, float* __restrict__ dx, float* __restrict__ pbeta2, int L, int d, int r,
                  Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ float sm[];       // coef_x [r][d], coef_g [r][d]
    float* coef_x = sm;
    float* coef_g = sm + (long long)r * d;
    const int b = blockIdx.x;
    const long long base = (long long)b * L * d;
    project_onto_basis(coef_x, Cb, L, d, r, [&](int t, int j) { return x[base + (long long)t * d + j]; });
    project_onto_basis(coef_g, Cb, L, d, r, [&](int t, int j) {
        const long long o = base + (long long)t * d + j;
        const float sb = __ldg(sqrt_beta + j);
        return (1.0f - sb * sb) * dz_f[o] * dropout_mult(dr, (unsigned long long)o);
    });
    __syncthreads();
    // thread per channel walks the sequence: coalesced across threads, fixed order over t
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        const float sb = __ldg(sqrt_beta + j);
        const float b2 = sb * sb;
        float acc = 0.f;
        for (int t = 0; t < L; ++t) {
            const long long o = base + (long long)t * d + j;
            float low = 0.f, fg = 0.f;
            for (int q = 0; q < r; ++q) {
                const float c = __ldg(Cb + (long long)t * r + q);
                low = fmaf(c, coef_x[q * d + j], low);
                fg = fmaf(c, coef_g[q * d + j], fg);
            }
            const float dzf = dz_f[o];
            const float du = dzf * dropout_mult(dr, (unsigned long long)o);
            acc = fmaf(du, x[o] - low, acc);
            dx[o] += dzf + b2 * du + fg;
        }
        pbeta2[(long long)b * d + j] = acc;
    }
}

__global__ void filter_basis_kernel(float* __restrict__ Cb, int L, int kept, int r) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L * r) return;
    const int t = idx / r, q = idx - t * r;
    double v;
    if (q == 0) v = 1.0 / sqrt((double)L);
    else {
        const int k = (q + 1) / 2;
        const double w = 2.0 * 3.14159265358979323846 * (double)k * (double)t / (double)L;
        if ((L % 2 == 0) && k == L / 2) v = cos(w) / sqrt((double)L);      // Nyquist: cos only
        else v = ((q & 1) ? cos(w) : sin(w)) * sqrt(2.0 / (double)L);
    }
    Cb[idx] = (float)v;
}

// ============================================================================ attention probabilities
template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
softmax_fwd_kernel(const float* __restrict__ S, const long long* __restrict__ ids, float* __restrict__ P,
                   float* __restrict__ Pd, long long rows, int H, int L, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < rows; row += (long long)gridDim.x * WARPS) {
        const int i = (int)(row % L);
        const long long b = row / ((long long)H * L);
        const long long* __restrict__ idb = ids + b * L;
        float v[VPL];
        float mx = -INFINITY;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int j = lane + 32 * q;
            v[q] = -INFINITY;
            if (j < L) {
                // additive mask exactly as _abstract_model.py:53-69: 0 if (key is an item and j<=i) else -10000
                const float add = (j <= i && idb[j] > 0) ? 0.0f : -10000.0f;
                v[q] = S[row * L + j] + add;
                mx = fmaxf(mx, v[q]);
            }
        }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int j = lane + 32 * q;
            if (j < L) { v[q] = expf(v[q] - mx); sum += v[q]; }
        }
        sum = warp_sum(sum);
        const float inv = 1.0f / sum;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int j = lane + 32 * q;
            if (j < L) {
                const long long o = row * L + j;
                const float pr = v[q] * inv;
                P[o] = pr;
                if (Pd != P) Pd[o] = pr * dropout_mult(dr, (unsigned long long)o);
            }
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(WARPS * 32)
softmax_bwd_kernel(const float* __restrict__ P, float* __restrict__ G, long long rows, int L, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < rows; row += (long long)gridDim.x * WARPS) {
        float p[VPL], dp[VPL];
        float dot = 0.f;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int j = lane + 32 * q;
            p[q] = dp[q] = 0.f;
            if (j < L) {
                const long long o = row * L + j;
                p[q] = P[o];
                dp[q] = G[o] * dropout_mult(dr, (unsigned long long)o);
                dot += p[q] * dp[q];
            }
        }
        dot = warp_sum(dot);
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int j = lane + 32 * q;
            if (j < L) G[row * L + j] = p[q] * (dp[q] - dot);
        }
    }
}

__global__ void apply_dropout_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, Dropout dr) {
    dr = dropout_resolve(dr);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = x[i] * dropout_mult(dr, (unsigned long long)i);
}

// ============================================================================ reductions
constexpr int COLSUM_ROWS = 128;

__global__ void colsum_partial_kernel(const float* __restrict__ X, long long M, int N, int rows_per_blk,
                                      float* __restrict__ part) {
    const long long r0 = (long long)blockIdx.x * rows_per_blk;
    const long long r1 = min(M, r0 + rows_per_blk);
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        float s = 0.f;
        for (long long m = r0; m < r1; ++m) s += X[m * N + n];
        part[(long long)blockIdx.x * N + n] = s;
    }
}

// out[j] (+)= mult[j]*scale * sum_k part[k][j].  Block = 32 columns x 8 row groups: group g sums rows k = g, g+8, ...
// (independent loads, unrolled), then the 8 group sums are added in fixed order -> deterministic.
__global__ void __launch_bounds__(256)
reduce_partials_kernel(const float* __restrict__ part, int nblk, int n, float* __restrict__ out,
                       const float* __restrict__ mult, float scale, int accumulate) {
    __shared__ float sh[8][32];
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (j < n) {
        int k = g;
        for (; k + 24 < nblk; k += 32) {
            s0 += part[(long long)k * n + j];
            s1 += part[(long long)(k + 8) * n + j];
            s2 += part[(long long)(k + 16) * n + j];
            s3 += part[(long long)(k + 24) * n + j];
        }
        for (; k < nblk; k += 8) s0 += part[(long long)k * n + j];
    }
    sh[g][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (g == 0 && j < n) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += sh[k][lane];
        s *= scale;
        if (mult) s *= mult[j];
        out[j] = accumulate ? out[j] + s : s;
    }
}

// Up to ReduceBatch::MAXJ of those reductions in one launch (blockIdx.y = job): the parameter-gradient partials of a block's
// backward pass are reduced together at its end — ~25 dependent 5 - 20 us launches of 8 CTAs become one launch that fills the
// GPU.  Block = 32 columns x 32 row groups; fixed summation order -> deterministic.
struct ReduceJobsP {
    const float* part[ReduceBatch::MAXJ]; float* out[ReduceBatch::MAXJ]; const float* mult[ReduceBatch::MAXJ];
    int nblk[ReduceBatch::MAXJ]; int n[ReduceBatch::MAXJ]; float scale[ReduceBatch::MAXJ]; int acc[ReduceBatch::MAXJ];
};
__global__ void __launch_bounds__(1024)
reduce_partials_multi_kernel(const __grid_constant__ ReduceJobsP t) {
    __shared__ float sh[32][33];
    const int job = blockIdx.y;
    const int n = t.n[job], nblk = t.nblk[job];
    if ((int)blockIdx.x * 32 >= n) return;                       // block-uniform
    const float* __restrict__ part = t.part[job];
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (j < n) {
        int k = g;
        for (; k + 96 < nblk; k += 128) {
            s0 += part[(long long)k * n + j];
            s1 += part[(long long)(k + 32) * n + j];
            s2 += part[(long long)(k + 64) * n + j];
            s3 += part[(long long)(k + 96) * n + j];
        }
        for (; k < nblk; k += 32) s0 += part[(long long)k * n + j];
    }
    sh[g][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (g == 0 && j < n) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) s += sh[k][lane];
        s *= t.scale[job];
        if (t.mult[job]) s *= t.mult[job][j];
        float* out = t.out[job];
        out[j] = t.acc[job] ? out[j] + s : s;
    }
}

__global__ void pos_grad_kernel(const float* __restrict__ dz, float* __restrict__ dP, int B, long long LD, int accumulate) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= LD) return;
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dz[(long long)b * LD + i];
    dP[i] = accumulate ? dP[i] + s : s;
}

// ============================================================================ sorted scatter-add
__global__ void scatter_keys_kernel(const long long* __restrict__ ids, int* __restrict__ keys, int* __restrict__ vals,
                                    long long M) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M) { keys[i] = (int)ids[i]; vals[i] = (int)i; }
}

// one warp per segment start; the warp sums the segment's rows in ascending row order (stable sort)
__global__ void __launch_bounds__(256)
scatter_add_kernel(const int* __restrict__ keys, const int* __restrict__ rows, long long M, const float* __restrict__ dz,
                   float* __restrict__ dE, int d) {
    const long long i = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= M) return;
    const int key = keys[i];
    if (key == 0) return;                              // padding_idx=0: no gradient through the gather
    if (i > 0 && keys[i - 1] == key) return;           // not a segment start
    const int lane = threadIdx.x & 31;
    long long end = i + 1;
    while (end < M && keys[end] == key) ++end;
    float* __restrict__ out = dE + (long long)key * d;
    for (int j0 = 0; j0 < d; j0 += 128) {
        const int j = j0 + lane * 4;
        if (j >= d) continue;
        if (j + 3 < d) {
            float4 acc = *reinterpret_cast<const float4*>(out + j);
            for (long long s = i; s < end; ++s) {
                const float4 v = *reinterpret_cast<const float4*>(dz + (long long)rows[s] * d + j);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            *reinterpret_cast<float4*>(out + j) = acc;
        } else {
            for (int jj = j; jj < d; ++jj) {
                float acc = out[jj];
                for (long long s = i; s < end; ++s) acc += dz[(long long)rows[s] * d + jj];
                out[jj] = acc;
            }
        }
    }
}

__global__ void last_rows_kernel(const float* __restrict__ src, float* __restrict__ dst, int B, int L, int d, int scatter) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * d) return;
    const long long b = i / d, j = i - b * d;
    const long long full = (b * L + (L - 1)) * d + j;
    if (scatter) dst[full] = src[i]; else dst[i] = src[full];
}

struct ScatterPlan {        // laid out in the caller's workspace
    int* keys_in; int* vals_in; int* keys_out; int* vals_out; void* tmp; size_t tmp_bytes;
};

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

inline ScatterPlan carve_plan(void* plan, long long M, size_t cub_bytes) {
    char* p = (char*)plan;
    ScatterPlan s;
    const size_t arr = align_up((size_t)M * sizeof(int));
    s.keys_in = (int*)p; p += arr;
    s.vals_in = (int*)p; p += arr;
    s.keys_out = (int*)p; p += arr;
    s.vals_out = (int*)p; p += arr;
    s.tmp = p; s.tmp_bytes = cub_bytes;
    return s;
}

inline size_t cub_sort_bytes(long long M) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, (int)M, 0, 32, (cudaStream_t)0);
    return bytes;
}

}  // namespace

// ============================================================================ host launchers
int rowwise_partial_blocks(long long rows) { return row_grid(rows); }

int embed_ln_dropout_fwd(const long long* ids, const float* E, const float* Ppos, const float* gamma,
                         const float* beta, float* out, float* mean, float* rstd, int B, int L, int d,
                         long long item_size, Dropout dr, cudaStream_t st) {
    (void)item_size;
    const long long M = (long long)B * L;
    if (M == 0) return 0;
    ProfScope prof(PROF_EMBED_FWD, st, (double)M * (8.0 + 2.0 * 4.0 * d));          // ids + gathered row read + output written
    if (vec_ok(d, E, Ppos, gamma, beta, out)) return embed_fwd_vec(ids, E, Ppos, gamma, beta, out, mean, rstd, M, L, d, dr, row_grid(M), st);
    BSA_DISPATCH_VPL(d, (embed_fwd_kernel<VPL><<<row_grid(M), WARPS * 32, 0, st>>>(ids, E, Ppos, gamma, beta, out, mean,
                                                                                  rstd, M, L, d, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int embed_ln_dropout_bwd(const float* dout, const long long* ids, const float* E, const float* Ppos,
                         const float* gamma, const float* mean, const float* rstd, float* dz, float* pgamma,
                         float* pbeta, int* nblk, int B, int L, int d, Dropout dr, cudaStream_t st) {
    const long long M = (long long)B * L;
    const int g = row_grid(M);
    *nblk = g;
    ProfScope prof(PROF_EMBED_BWD, st, (double)M * (8.0 + 3.0 * 4.0 * d));          // ids + dout, row read + dz written
    if (vec_ok(d, dout, E, Ppos, gamma, dz)) return embed_bwd_vec(dout, ids, E, Ppos, gamma, mean, rstd, dz, pgamma, pbeta, M, L, d, dr, g, st);
    BSA_DISPATCH_VPL(d, (embed_bwd_kernel<VPL><<<g, WARPS * 32, 0, st>>>(dout, ids, E, Ppos, gamma, mean, rstd, dz,
                                                                       pgamma, pbeta, M, L, d, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int residual_ln_fwd(const float* tmp, const float* resid, const float* gamma, const float* beta, float* out,
                    float* z_save, float* mean, float* rstd, const float* blend_with, float blend_a, float blend_b,
                    long long M, int d, Dropout dr, cudaStream_t st, float* out_hi, float* out_lo) {
    if (M == 0) return 0;
    {
        ProfScope prof(PROF_RESID_LN_FWD, st, 4.0 * M * d * (3.0 + (z_save ? 1.0 : 0.0) + (blend_with ? 1.0 : 0.0)));      // the optional split pair is not credited
        // out_lo without out_hi: `out` is its own hi half, lo = out - truncate_tf32(out)
        if (vec_ok(d, tmp, resid, gamma, beta, out, z_save, blend_with) && vec_ok(d, out_hi, out_lo))
            return residual_ln_fwd_vec(tmp, resid, gamma, beta, out, z_save, mean, rstd, blend_with, blend_a, blend_b, M, d, dr,
                                       row_grid(M), st, out_hi, out_lo);
        BSA_DISPATCH_VPL(d, (residual_ln_fwd_kernel<VPL><<<row_grid(M), WARPS * 32, 0, st>>>(
                                tmp, resid, gamma, beta, out, z_save, mean, rstd, blend_with, blend_a, blend_b, M, d, dr)));
        BSA_LAUNCH_CHECK();
    }
    return out_hi ? umma_split(out, out_hi, out_lo, M * d, st) : (out_lo ? umma_split_lo(out, out_lo, M * d, st) : 0);
}

int ln_bwd(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
           float* dz, float* ddrop, float* pgamma, float* pbeta, int* nblk, long long M, int d, Dropout dr,
           cudaStream_t st) {
    const int g = row_grid(M);
    *nblk = g;
    ProfScope prof(PROF_LN_BWD, st, 4.0 * M * d * (3.0 + (ddrop ? 1.0 : 0.0)));
    if (vec_ok(d, dy, z, gamma, dz, ddrop))
        return ln_bwd_vec(dy, dy_scale, z, mean, rstd, gamma, dz, ddrop, pgamma, pbeta, M, d, dr, g, nullptr, nullptr, nullptr, st);
    BSA_DISPATCH_VPL(d, (ln_bwd_kernel<VPL><<<g, WARPS * 32, 0, st>>>(dy, dy_scale, z, mean, rstd, gamma, dz, ddrop,
                                                                    pgamma, pbeta, M, d, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

bool ln_bwd_split_ok(int d, const float* dy, const float* z, const float* gamma, const float* dz, const float* ddrop,
                     const float* t_hi, const float* t_lo) {
    return vec_ok(d, dy, z, gamma, dz, ddrop, t_hi, t_lo);
}

int ln_bwd_split(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
                 float* dz, float* ddrop, float* pgamma, float* pbeta, int* nblk, long long M, int d, Dropout dr, float* t_hi,
                 float* t_lo, float* pcol, cudaStream_t st) {
    BSA_REQUIRE(ln_bwd_split_ok(d, dy, z, gamma, dz, ddrop, t_hi, t_lo), "ln_bwd_split: needs d %% 4 == 0 and 16-byte aligned tensors");
    const int g = row_grid(M);
    *nblk = g;
    ProfScope prof(PROF_LN_BWD, st, 4.0 * M * d * (4.0 + (ddrop ? 1.0 : 0.0)));      // dy, z read; dz and t written (t counted once: its (hi, lo) pair is an implementation cost)
    return ln_bwd_vec(dy, dy_scale, z, mean, rstd, gamma, dz, ddrop, pgamma, pbeta, M, d, dr, g, t_hi, t_lo, pcol, st);
}

int filter_rank(int L, int kept_bins) {
    int kept = kept_bins < L / 2 + 1 ? kept_bins : L / 2 + 1;
    if (kept < 1) kept = 1;
    int r = 2 * kept - 1;
    if (L % 2 == 0 && kept == L / 2 + 1) r -= 1;       // Nyquist bin has no sin column
    return r;
}

int build_filter_basis(float* Cb, int L, int kept_bins, int* r_out, cudaStream_t st) {
    int kept = kept_bins < L / 2 + 1 ? kept_bins : L / 2 + 1;
    if (kept < 1) kept = 1;
    const int r = filter_rank(L, kept_bins);
    if (r_out) *r_out = r;
    filter_basis_kernel<<<cdiv((long long)L * r, 256), 256, 0, st>>>(Cb, L, kept, r);
    BSA_LAUNCH_CHECK();
    return 0;
}

int filter_ln_fwd(const float* x, const float* Cb, const float* sqrt_beta, const float* gamma, const float* beta,
                  float* dsp, float* z_save, float* mean, float* rstd, int B, int L, int d, int r, Dropout dr,
                  cudaStream_t st, float* x_lo) {
    if (B == 0) return 0;
    if (x_lo && !vec_ok(d, x, sqrt_beta, gamma, beta, dsp, z_save, x_lo)) {      // scalar kernel below: separate split pass
        const int rc = umma_split_lo(x, x_lo, (long long)B * L * d, st);
        if (rc) return rc;
        x_lo = nullptr;
    }
    ProfScope prof(PROF_FILTER_FWD, st, 4.0 * B * (double)L * d * (z_save ? 3.0 : 2.0));   // x read; z_save, dsp written (x_lo not credited)
    if (vec_ok(d, x, sqrt_beta, gamma, beta, dsp, z_save)) return filter_ln_fwd_vec(x, Cb, sqrt_beta, gamma, beta, dsp, z_save, mean, rstd, B, L, d, r, dr, st, x_lo);
    const size_t smem = (size_t)r * d * sizeof(float);
    BSA_REQUIRE(smem <= 200 * 1024, "filter: basis rank %d x hidden %d does not fit shared memory", r, d);
    BSA_DISPATCH_VPL(d, {
        if (smem > 48 * 1024)
            BSA_CHECK_CUDA(cudaFuncSetAttribute(filter_ln_fwd_kernel<VPL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)smem));
        filter_ln_fwd_kernel<VPL><<<B, WARPS * 32, smem, st>>>(x, Cb, sqrt_beta, gamma, beta, dsp, z_save, mean, rstd, L,
                                                             d, r, dr);
    });
    BSA_LAUNCH_CHECK();
    return 0;
}

int filter_bwd(const float* x, const float* Cb, const float* sqrt_beta, const float* dz_f, float* dx, float* pbeta2,
               int B, int L, int d, int r, Dropout dr, cudaStream_t st) {
    if (B == 0) return 0;
    ProfScope prof(PROF_FILTER_BWD, st, 4.0 * B * (double)L * d * 4.0);             // x, dz_f read; dx read + written
    if (vec_ok(d, x, sqrt_beta, dz_f, dx)) return filter_bwd_vec(x, Cb, sqrt_beta, dz_f, dx, pbeta2, B, L, d, r, dr, st);
    const size_t smem = 2 * (size_t)r * d * sizeof(float);
    BSA_REQUIRE(smem <= 200 * 1024, "filter bwd: basis rank %d x hidden %d does not fit shared memory", r, d);
    if (smem > 48 * 1024)
        BSA_CHECK_CUDA(cudaFuncSetAttribute(filter_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    filter_bwd_kernel<<<B, WARPS * 32, smem, st>>>(x, Cb, sqrt_beta, dz_f, dx, pbeta2, L, d, r, dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

int softmax_mask_dropout_fwd(const float* S, const long long* ids, float* P, float* Pd, int B, int H, int L,
                             Dropout dr, cudaStream_t st) {
    const long long rows = (long long)B * H * L;
    if (rows == 0) return 0;
    BSA_DISPATCH_VPL(L, (softmax_fwd_kernel<VPL><<<row_grid(rows), WARPS * 32, 0, st>>>(S, ids, P, Pd, rows, H, L, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int softmax_bwd(const float* P, float* G, int B, int H, int L, Dropout dr, cudaStream_t st) {
    const long long rows = (long long)B * H * L;
    if (rows == 0) return 0;
    BSA_DISPATCH_VPL(L, (softmax_bwd_kernel<VPL><<<row_grid(rows), WARPS * 32, 0, st>>>(P, G, rows, L, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int apply_dropout(const float* x, float* y, long long n, Dropout dr, cudaStream_t st) {
    if (n == 0) return 0;
    long long g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    apply_dropout_kernel<<<(int)g, 256, 0, st>>>(x, y, n, dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

int colsum_partial_rows(long long M) { return (int)((M + COLSUM_ROWS - 1) / COLSUM_ROWS); }

int colsum(const float* X, long long M, int N, float* part, float* out, int accumulate, cudaStream_t st) {
    const int nblk = colsum_partial_rows(M);
    if (nblk == 0) return 0;
    // few rows (the [B, d] tensors of the last-row path): the rows themselves are the partials — one launch instead of a
    // 4-CTA partial pass plus the reduction
    if (M <= 1024) return reduce_partials(X, (int)M, N, out, nullptr, 1.0f, accumulate, st);
    colsum_partial_kernel<<<nblk, 256, 0, st>>>(X, M, N, COLSUM_ROWS, part);
    BSA_LAUNCH_CHECK();
    return reduce_partials(part, nblk, N, out, nullptr, 1.0f, accumulate, st);
}

int reduce_partials(const float* part, int nblk, int n, float* out, const float* mult, float scale, int accumulate,
                    cudaStream_t st) {
    reduce_partials_kernel<<<cdiv(n, 32), 256, 0, st>>>(part, nblk, n, out, mult, scale, accumulate);
    BSA_LAUNCH_CHECK();
    return 0;
}

int reduce_batch_flush(ReduceBatch& b, cudaStream_t st) {
    if (b.count == 0) return 0;
    ReduceJobsP t;
    int nmax = 0;
    for (int i = 0; i < ReduceBatch::MAXJ; ++i) {
        const int j = i < b.count ? i : 0;
        t.part[i] = b.part[j]; t.out[i] = b.out[j]; t.mult[i] = b.mult[j]; t.nblk[i] = b.nblk[j]; t.n[i] = b.n[j];
        t.scale[i] = b.scale[j]; t.acc[i] = b.acc[j];
        if (i < b.count && b.n[i] > nmax) nmax = b.n[i];
    }
    const int count = b.count;
    b.count = 0;
    if (nmax == 0) return 0;
    reduce_partials_multi_kernel<<<dim3((unsigned)cdiv(nmax, 32), (unsigned)count), 1024, 0, st>>>(t);
    BSA_LAUNCH_CHECK();
    return 0;
}

int reduce_batch_add(ReduceBatch& b, const float* part, int nblk, int n, float* out, const float* mult, float scale,
                     int accumulate, cudaStream_t st) {
    if (n <= 0) return 0;
    if (b.count == ReduceBatch::MAXJ) { int rc = reduce_batch_flush(b, st); if (rc) return rc; }
    const int i = b.count++;
    b.part[i] = part; b.nblk[i] = nblk; b.n[i] = n; b.out[i] = out; b.mult[i] = mult; b.scale[i] = scale; b.acc[i] = accumulate;
    return 0;
}

int pos_grad(const float* dz, float* dP, int B, int L, int d, int accumulate, cudaStream_t st) {
    const long long LD = (long long)L * d;
    pos_grad_kernel<<<cdiv(LD, 256), 256, 0, st>>>(dz, dP, B, LD, accumulate);
    BSA_LAUNCH_CHECK();
    return 0;
}

size_t scatter_plan_bytes(long long M) {
    if (M <= 0) return 256;
    return 4 * align_up((size_t)M * sizeof(int)) + align_up(cub_sort_bytes(M)) + 256;
}

int scatter_plan(const long long* ids, long long M, long long item_size, void* plan, size_t plan_bytes, cudaStream_t st) {
    if (M == 0) return 0;
    BSA_REQUIRE(M < (1ll << 31) && item_size < (1ll << 31), "scatter_plan: sizes exceed int32");
    const size_t cb = cub_sort_bytes(M);
    BSA_REQUIRE(plan_bytes >= scatter_plan_bytes(M), "scatter_plan: workspace too small");
    ScatterPlan s = carve_plan(plan, M, cb);
    scatter_keys_kernel<<<cdiv(M, 256), 256, 0, st>>>(ids, s.keys_in, s.vals_in, M);
    BSA_LAUNCH_CHECK();
    int end_bit = 1;
    while ((1ll << end_bit) < item_size && end_bit < 31) ++end_bit;
    size_t tb = s.tmp_bytes;
    BSA_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(s.tmp, tb, s.keys_in, s.keys_out, s.vals_in, s.vals_out, (int)M, 0,
                                                   end_bit, st));
    bsa::g_launch_count += 3;   // radix sort passes (library kernels)
    return 0;
}

int scatter_add_rows(const void* plan, long long M, const float* dz, float* dE, int d, cudaStream_t st) {
    if (M == 0) return 0;
    BSA_REQUIRE(d % 4 == 0, "scatter_add_rows: hidden size must be a multiple of 4");
    ScatterPlan s = carve_plan(const_cast<void*>(plan), M, 0);
    scatter_add_kernel<<<cdiv(M, 8), 256, 0, st>>>(s.keys_out, s.vals_out, M, dz, dE, d);
    BSA_LAUNCH_CHECK();
    return 0;
}

int fill_zero(float* p, long long n, cudaStream_t st) {
    if (n == 0) return 0;
    BSA_CHECK_CUDA(cudaMemsetAsync(p, 0, (size_t)n * sizeof(float), st));
    return 0;
}

int scatter_last_rows(const float* src, float* dst, int B, int L, int d, cudaStream_t st) {
    int rc = fill_zero(dst, (long long)B * L * d, st);
    if (rc) return rc;
    last_rows_kernel<<<cdiv((long long)B * d, 256), 256, 0, st>>>(src, dst, B, L, d, 1);
    BSA_LAUNCH_CHECK();
    return 0;
}

// dst[i, :] = src[idx[i], :]: a warp per row, float4 lanes (d % 4 == 0) — pure copies, bit-exact by construction
__global__ void __launch_bounds__(256) gather_rows_kernel(const long long* __restrict__ idx, long long M, long long src_rows,
                                                          const float* __restrict__ src, int d, float* __restrict__ dst) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= M) return;
    long long r = idx[row];
    if (r < 0 || r >= src_rows) r = 0;          // out-of-range indices read row 0 (callers validate; never fault)
    const float4* s4 = reinterpret_cast<const float4*>(src + r * d);
    float4* d4 = reinterpret_cast<float4*>(dst + row * d);
    for (int c = threadIdx.x & 31; c < d / 4; c += 32) d4[c] = __ldg(s4 + c);
}

int gather_rows(const long long* idx, long long M, long long src_rows, const float* src, int d, float* dst, cudaStream_t st) {
    if (M == 0) return 0;
    BSA_REQUIRE(d % 4 == 0, "gather_rows: row length must be a multiple of 4");
    gather_rows_kernel<<<cdiv(M, 8), 256, 0, st>>>(idx, M, src_rows, src, d, dst);
    BSA_LAUNCH_CHECK();
    return 0;
}

// dst[idx[i], :] = src[i, :]: the "put" counterpart (rows fetched from their owner written into the local table replica);
// duplicate indices carry identical rows by construction, so the write order does not matter
__global__ void __launch_bounds__(256) scatter_rows_kernel(const long long* __restrict__ idx, long long M, long long dst_rows,
                                                           const float* __restrict__ src, int d, float* __restrict__ dst) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= M) return;
    const long long r = idx[row];
    if (r < 0 || r >= dst_rows) return;
    const float4* s4 = reinterpret_cast<const float4*>(src + row * d);
    float4* d4 = reinterpret_cast<float4*>(dst + r * d);
    for (int c = threadIdx.x & 31; c < d / 4; c += 32) d4[c] = s4[c];
}

int scatter_rows(const long long* idx, long long M, long long dst_rows, const float* src, int d, float* dst, cudaStream_t st) {
    if (M == 0) return 0;
    BSA_REQUIRE(d % 4 == 0, "scatter_rows: row length must be a multiple of 4");
    scatter_rows_kernel<<<cdiv(M, 8), 256, 0, st>>>(idx, M, dst_rows, src, d, dst);
    BSA_LAUNCH_CHECK();
    return 0;
}

int gather_last_rows(const float* src, float* dst, int B, int L, int d, cudaStream_t st) {
    last_rows_kernel<<<cdiv((long long)B * d, 256), 256, 0, st>>>(src, dst, B, L, d, 0);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

// 128-bit versions of the HBM-bound row kernels (hidden size a multiple of 4): every global access is a float4, the
// dropout hash is drawn once per four elements, and the filter kernels stage the sequence tile x[b] ([L, d] fp32) in
// shared memory with cp.async so HBM sees it once.  Same math and the same fixed summation orders per launch
// configuration as rowwise.cu (which remains the path for hidden sizes that are not a multiple of 4).
//
// Why (profiles/r1o_rowwise_ncu_summary.txt): the scalar warp-per-row kernels ran at 13-38 % of DRAM peak — 4-byte
// loads, one hash per element, and in the filter kernels one thread per channel walking the sequence serially.
#include <cstdlib>
#include "rowwise.cuh"

namespace bsa {
namespace {

constexpr int WARPS = 8;
constexpr float LN_EPS = 1e-12f;

#define BSA_DISPATCH_NV(n, ...)                                                             \
    do {                                                                                    \
        if ((n) <= 128) { constexpr int NV = 1; __VA_ARGS__; }                              \
        else if ((n) <= 256) { constexpr int NV = 2; __VA_ARGS__; }                         \
        else if ((n) <= 512) { constexpr int NV = 4; __VA_ARGS__; }                         \
        else if ((n) <= 1024) { constexpr int NV = 8; __VA_ARGS__; }                        \
        else { set_error("row length %d > 1024 is not supported", (int)(n)); return -1; }   \
    } while (0)

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4(float a) { return make_float4(a, a, a, a); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator-(float4 a, float4 b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 operator*(float a, float4 b) { return make_float4(a * b.x, a * b.y, a * b.z, a * b.w); }
__device__ __forceinline__ float hsum(float4 a) { return (a.x + a.y) + (a.z + a.w); }
__device__ __forceinline__ float4 fma4(float a, float4 b, float4 c) {
    return make_float4(fmaf(a, b.x, c.x), fmaf(a, b.y, c.y), fmaf(a, b.z, c.z), fmaf(a, b.w, c.w));
}
__device__ __forceinline__ float4 mult4(const Dropout& dr, unsigned long long i) {
    float m[4];
    dropout_mult4(dr, i, m);
    return make_float4(m[0], m[1], m[2], m[3]);
}
__device__ __forceinline__ void split1(float x, float& hi, float& lo) {      // as umma_engine.cu's split_tf32
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    lo = x - hi;
}
__device__ __forceinline__ float lo_trunc(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ unsigned long long row_index(const Dropout& dr, long long row, int d, int c) {
    return (unsigned long long)((row * dr.rmul + dr.radd) * d + c);
}

// mean / rstd of one row held as v[NV] float4 per lane (two-pass, as _modules.py:16-20)
template <int NV>
__device__ __forceinline__ void row_stats4(const float4 (&v)[NV], int d, int lane, float& mean, float& rstd) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) if (4 * (lane + 32 * i) < d) s += hsum(v[i]);
    mean = warp_sum(s) / (float)d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
        if (4 * (lane + 32 * i) < d) { const float4 c = v[i] - f4(mean); q += hsum(c * c); }
    rstd = 1.0f / sqrtf(warp_sum(q) / (float)d + LN_EPS);
}

// per-block column partials: part[blockIdx.x][0..d) = sum over the block's warps (fixed order) of a[NV]
template <int NV>
__device__ __forceinline__ void block_store_partials4(const float4 (&a)[NV], float* __restrict__ part, int d, float* stage) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int c = 4 * (lane + 32 * i);
        if (c < d) st4(stage + w * d + c, a[i]);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < d; j += WARPS * 32) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < WARPS; ++k) s += stage[k * d + j];
        part[(long long)blockIdx.x * d + j] = s;
    }
}

// ============================================================================ embedding
template <int NV>
__global__ void __launch_bounds__(WARPS * 32, NV <= 2 ? 4 : 2)
embed_fwd_vec_kernel(const long long* __restrict__ ids, const float* __restrict__ E, const float* __restrict__ Ppos,
                     const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ out,
                     float* __restrict__ mean_o, float* __restrict__ rstd_o, long long M, int L, int d, Dropout dr) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * WARPS;
    long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
    long long id = row < M ? ids[row] : 0;
    for (; row < M; row += stride) {
        const long long id_next = (row + stride < M) ? ids[row + stride] : 0;      // next gather's address early
        const int t = (int)(row % L);
        const float* __restrict__ e = E + id * d;
        const float* __restrict__ pp = Ppos + (long long)t * d;
        float4 v[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            v[i] = c < d ? ldg4(e + c) + ldg4(pp + c) : f4(0.f);
        }
        float mean, rstd;
        row_stats4<NV>(v, d, lane, mean, rstd);
        if (lane == 0) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            if (c < d) {
                const float4 y = ldg4(gamma + c) * (rstd * (v[i] - f4(mean))) + ldg4(beta + c);
                st4(out + row * d + c, y * mult4(dr, (unsigned long long)(row * d + c)));
            }
        }
        id = id_next;
    }
}

template <int NV>
__global__ void __launch_bounds__(WARPS * 32, NV <= 2 ? 4 : 2)
embed_bwd_vec_kernel(const float* __restrict__ dout, const long long* __restrict__ ids, const float* __restrict__ E,
                     const float* __restrict__ Ppos, const float* __restrict__ gamma, const float* __restrict__ mean_i,
                     const float* __restrict__ rstd_i, float* __restrict__ dz, float* __restrict__ pgamma,
                     float* __restrict__ pbeta, long long M, int L, int d, Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ float stage[];
    const int lane = threadIdx.x & 31;
    float4 ag[NV], ab[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) ag[i] = ab[i] = f4(0.f);
    const long long stride = (long long)gridDim.x * WARPS;
    long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
    long long id = row < M ? ids[row] : 0;
    for (; row < M; row += stride) {
        const long long id_next = (row + stride < M) ? ids[row + stride] : 0;
        const int t = (int)(row % L);
        const float mean = mean_i[row], rstd = rstd_i[row];
        float4 xh[NV], g[NV];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            xh[i] = g[i] = f4(0.f);
            if (c < d) {
                const float4 zv = ldg4(E + id * d + c) + ldg4(Ppos + (long long)t * d + c);
                xh[i] = rstd * (zv - f4(mean));
                const float4 dy = ld4(dout + row * d + c) * mult4(dr, (unsigned long long)(row * d + c));
                ag[i] = ag[i] + dy * xh[i];
                ab[i] = ab[i] + dy;
                g[i] = dy * ldg4(gamma + c);
                c1 += hsum(g[i]);
                c2 += hsum(g[i] * xh[i]);
            }
        }
        c1 = warp_sum(c1) / (float)d;
        c2 = warp_sum(c2) / (float)d;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            if (c < d) st4(dz + row * d + c, rstd * (g[i] - f4(c1) - c2 * xh[i]));
        }
        id = id_next;
    }
    block_store_partials4<NV>(ag, pgamma, d, stage);
    block_store_partials4<NV>(ab, pbeta, d, stage);
}

// ============================================================================ residual + LN (+ blend)
template <int NV>
__global__ void __launch_bounds__(WARPS * 32, NV <= 2 ? 4 : 2)
residual_ln_fwd_vec_kernel(const float* __restrict__ tmp, const float* __restrict__ resid, const float* __restrict__ gamma,
                           const float* __restrict__ beta, float* __restrict__ out, float* __restrict__ z_save,
                           float* __restrict__ mean_o, float* __restrict__ rstd_o, const float* __restrict__ blend_with,
                           float blend_a, float blend_b, long long M, int d, Dropout dr, float* __restrict__ y_hi,
                           float* __restrict__ y_lo) {
    dr = dropout_resolve(dr);
    const int lane = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        float4 v[NV], bw[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            v[i] = bw[i] = f4(0.f);
            if (c < d) {
                const long long o = row * d + c;
                v[i] = ld4(tmp + o) * mult4(dr, row_index(dr, row, d, c)) + ld4(resid + o);
                if (blend_with) bw[i] = ld4(blend_with + o);
                if (z_save) st4(z_save + o, v[i]);
            }
        }
        float mean, rstd;
        row_stats4<NV>(v, d, lane, mean, rstd);
        if (lane == 0 && mean_o) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            if (c < d) {
                float4 y = ldg4(gamma + c) * (rstd * (v[i] - f4(mean))) + ldg4(beta + c);
                if (blend_with) y = blend_a * bw[i] + blend_b * y;
                st4(out + row * d + c, y);
                if (y_lo) {          // the next tcgen05 linear's A operand: `out` itself is the hi half (the MMA reads its upper
                                     // 19 bits), lo = y - truncate(y) leaves here — no separate split pass over `out`
                    st4(y_lo + row * d + c, make_float4(lo_trunc(y.x), lo_trunc(y.y), lo_trunc(y.z), lo_trunc(y.w)));
                    if (y_hi) {      // a separate (rna hi, lo) pair was asked for
                        float4 h, l;
                        split1(y.x, h.x, l.x); split1(y.y, h.y, l.y); split1(y.z, h.z, l.z); split1(y.w, h.w, l.w);
                        st4(y_hi + row * d + c, h);
                        st4(y_lo + row * d + c, l);
                    }
                }
            }
        }
    }
}

template <int NV>
__global__ void __launch_bounds__(WARPS * 32, NV <= 2 ? 4 : 2)
ln_bwd_vec_kernel(const float* __restrict__ dy, float dy_scale, const float* __restrict__ z, const float* __restrict__ mean_i,
                  const float* __restrict__ rstd_i, const float* __restrict__ gamma, float* __restrict__ dz,
                  float* __restrict__ ddrop, float* __restrict__ pgamma, float* __restrict__ pbeta, long long M, int d,
                  Dropout dr, float* __restrict__ t_hi, float* __restrict__ t_lo, float* __restrict__ pcol) {
    dr = dropout_resolve(dr);
    extern __shared__ float stage[];
    const int lane = threadIdx.x & 31;
    float4 ag[NV], ab[NV], ac[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) ag[i] = ab[i] = ac[i] = f4(0.f);
    for (long long row = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5); row < M; row += (long long)gridDim.x * WARPS) {
        const float mean = mean_i[row], rstd = rstd_i[row];
        float4 xh[NV], g[NV];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            xh[i] = g[i] = f4(0.f);
            if (c < d) {
                const long long o = row * d + c;
                xh[i] = rstd * (ld4(z + o) - f4(mean));
                const float4 up = dy_scale * ld4(dy + o);
                ag[i] = ag[i] + up * xh[i];
                ab[i] = ab[i] + up;
                g[i] = up * ldg4(gamma + c);
                c1 += hsum(g[i]);
                c2 += hsum(g[i] * xh[i]);
            }
        }
        c1 = warp_sum(c1) / (float)d;
        c2 = warp_sum(c2) / (float)d;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int c = 4 * (lane + 32 * i);
            if (c < d) {
                const long long o = row * d + c;
                const float4 r = rstd * (g[i] - f4(c1) - c2 * xh[i]);
                st4(dz + o, r);
                if (ddrop || t_hi) {
                    // t = gradient w.r.t. the pre-dropout tensor (the nn.Linear output): written as fp32 (ddrop) and / or
                    // directly as the split-TF32 pair the tcgen05 dW / dX GEMMs consume; its column sums are the bias gradient
                    const float4 t = dr.thr ? r * mult4(dr, row_index(dr, row, d, c)) : r;
                    if (ddrop) st4(ddrop + o, t);
                    if (t_hi) {
                        float4 h, l;
                        split1(t.x, h.x, l.x); split1(t.y, h.y, l.y); split1(t.z, h.z, l.z); split1(t.w, h.w, l.w);
                        st4(t_hi + o, h);
                        st4(t_lo + o, l);
                        ac[i] = ac[i] + t;
                    }
                }
            }
        }
    }
    block_store_partials4<NV>(ag, pgamma, d, stage);
    block_store_partials4<NV>(ab, pbeta, d, stage);
    if (pcol) block_store_partials4<NV>(ac, pcol, d, stage);
}

// ============================================================================ filter branch
// One CTA of FT threads per sequence.  The d/4 float4 columns of a row are covered by LR = d/4 threads; the CTA's
// G = FT / LR thread groups take positions t = g, g+G, ...  Pass 1 projects onto the basis (coef = C^T x, partial sums per
// group, folded group after group -> fixed order); pass 2 applies C coef.  With STAGE the tile x[b] is copied to shared
// memory once (cp.async, all chunks in flight) and both passes read it there; otherwise pass 2 re-reads global memory.
constexpr int FT = 512;
constexpr int RCH = 8;   // basis vectors per pass over the tile

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// coef[q][c..c+3] = fold over groups of acc  (first group stores, the rest add, one group per barrier round)
__device__ __forceinline__ void fold_groups(float* __restrict__ coef, const float4 (&acc)[RCH], int r0, int r, int d, int c,
                                            int g, int G, bool active) {
    for (int gg = 0; gg < G; ++gg) {
        if (active && g == gg) {
#pragma unroll
            for (int q = 0; q < RCH; ++q)
                if (r0 + q < r) {
                    float* p = coef + (r0 + q) * d + c;
                    st4(p, gg == 0 ? acc[q] : ld4(p) + acc[q]);
                }
        }
        __syncthreads();
    }
}

// The same fold in ONE barrier round when shared memory has room for every group's partials (scratch [G][min(r, RCH)][d]):
// all groups store, then the thread of group q sums coefficient row q over the groups in the same fixed order (bit-identical
// to the sequential fold).  Small tiles (C1: L = 50, d = 64 -> G = 32 groups) spent most of their time in the G barrier
// rounds of the sequential version; at C3 the staged 205 KB tile leaves no room and the sequential fold stays.
__device__ __forceinline__ void fold_groups_scratch(float* __restrict__ scratch, float* __restrict__ coef, const float4 (&acc)[RCH],
                                                    int r0, int r, int d, int c, int g, int G, bool active) {
    const int rr = min(r - r0, RCH);
    if (active)
#pragma unroll
        for (int q = 0; q < RCH; ++q)
            if (q < rr) st4(scratch + ((long long)g * rr + q) * d + c, acc[q]);
    __syncthreads();
    if (active)
        for (int q = g; q < rr; q += G) {
            float4 sum = ld4(scratch + (long long)q * d + c);
            for (int gg = 1; gg < G; ++gg) sum = sum + ld4(scratch + ((long long)gg * rr + q) * d + c);
            st4(coef + (long long)(r0 + q) * d + c, sum);
        }
    __syncthreads();
}

template <int NV, bool STAGE>
__global__ void __launch_bounds__(FT)
filter_ln_fwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ Cb, const float* __restrict__ sqrt_beta,
                         const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ dsp,
                         float* __restrict__ z_save, float* __restrict__ mean_o, float* __restrict__ rstd_o, int L, int d,
                         int r, int cb_smem, Dropout dr, float* __restrict__ x_lo, int fold_off) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) float fsm[];
    float* coef = fsm;                       // [r][d]
    float* par = coef + (long long)r * d;    // [3][d]: beta^2, LayerNorm weight, LayerNorm bias
    float* cbs = par + 3 * d;                // [L][r] basis copy when cb_smem
    float* tile = cbs + (cb_smem ? ((L * r + 3) & ~3) : 0);    // [L][d] when STAGE
    const int b = blockIdx.x;
    const float* __restrict__ xb = x + (long long)b * L * d;
    if (STAGE)
        for (int i = threadIdx.x; i < L * d / 4; i += FT) cp_async16(tile + 4 * i, xb + 4 * i);
    for (int j = threadIdx.x; j < d; j += FT) {
        const float sb = __ldg(sqrt_beta + j);
        par[j] = sb * sb; par[d + j] = __ldg(gamma + j); par[2 * d + j] = __ldg(beta + j);
    }
    if (cb_smem)
        for (int i = threadIdx.x; i < L * r; i += FT) cbs[i] = __ldg(Cb + i);
    if (STAGE) cp_async_wait_all();
    __syncthreads();
    const float* src = STAGE ? tile : xb;
    const float* cb = cb_smem ? cbs : Cb;
    const int LR = d >> 2, G = FT / LR, g = threadIdx.x / LR, c = 4 * (threadIdx.x - g * LR);
    const bool active = g < G;
    for (int r0 = 0; r0 < r; r0 += RCH) {
        float4 acc[RCH];
#pragma unroll
        for (int q = 0; q < RCH; ++q) acc[q] = f4(0.f);
        if (active)
            for (int t = g; t < L; t += G) {
                const float4 xv = ld4(src + t * d + c);
#pragma unroll
                for (int q = 0; q < RCH; ++q)
                    if (r0 + q < r) acc[q] = fma4(cb[t * r + r0 + q], xv, acc[q]);
            }
        if (fold_off) fold_groups_scratch(fsm + fold_off, coef, acc, r0, r, d, c, g, G, active);
        else fold_groups(coef, acc, r0, r, d, c, g, G, active);
    }
    // pass 2: warp per position
    const int lane = threadIdx.x & 31;
    for (int t = threadIdx.x >> 5; t < L; t += FT / 32) {
        const long long row = (long long)b * L + t;
        float4 v[NV], low[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) low[i] = f4(0.f);
        for (int q = 0; q < r; ++q) {
            const float cq = cb[t * r + q];
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int cc = 4 * (lane + 32 * i);
                if (cc < d) low[i] = fma4(cq, ld4(coef + q * d + cc), low[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int cc = 4 * (lane + 32 * i);
            v[i] = f4(0.f);
            if (cc < d) {
                const float4 xv = ld4(src + t * d + cc);
                const float4 u = low[i] + ld4(par + cc) * (xv - low[i]);
                const long long o = row * d + cc;
                if (x_lo) st4(x_lo + o, make_float4(lo_trunc(xv.x), lo_trunc(xv.y), lo_trunc(xv.z), lo_trunc(xv.w)));
                v[i] = u * mult4(dr, (unsigned long long)o) + xv;
                if (z_save) st4(z_save + o, v[i]);
            }
        }
        float mean, rstd;
        row_stats4<NV>(v, d, lane, mean, rstd);
        if (lane == 0 && mean_o) { mean_o[row] = mean; rstd_o[row] = rstd; }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int cc = 4 * (lane + 32 * i);
            if (cc < d) st4(dsp + row * d + cc, ld4(par + d + cc) * (rstd * (v[i] - f4(mean))) + ld4(par + 2 * d + cc));
        }
    }
}

// u = beta2*x + (1-beta2)*F x  =>  dx += dz_f + beta2*du + F((1-beta2)*du)   (F = C C^T symmetric), du = dz_f * dropout
// dbeta2[j] = sum_t du*(x - low)  -> pbeta2 [B][d]
template <bool STAGE>
__global__ void __launch_bounds__(FT)
filter_bwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ Cb, const float* __restrict__ sqrt_beta,
                      const float* __restrict__ dz_f, float* __restrict__ dx, float* __restrict__ pbeta2, int L, int d, int r,
                      int cb_smem, Dropout dr, int fold_off) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) float fsm[];
    float* coef_x = fsm;                         // [r][d]
    float* coef_g = fsm + (long long)r * d;      // [r][d]
    float* pb = coef_g + (long long)r * d;       // [d]
    float* cbs = pb + d;                         // [L][r] basis copy when cb_smem
    float* tile = cbs + (cb_smem ? ((L * r + 3) & ~3) : 0);     // [L][d] when STAGE
    const int b = blockIdx.x;
    const long long base = (long long)b * L * d;
    const float* __restrict__ xb = x + base;
    const float* __restrict__ gb = dz_f + base;
    float* __restrict__ dxb = dx + base;
    if (STAGE)
        for (int i = threadIdx.x; i < L * d / 4; i += FT) cp_async16(tile + 4 * i, xb + 4 * i);
    if (cb_smem)
        for (int i = threadIdx.x; i < L * r; i += FT) cbs[i] = __ldg(Cb + i);
    if (STAGE) cp_async_wait_all();
    __syncthreads();
    const float* src = STAGE ? tile : xb;
    const float* cb = cb_smem ? cbs : Cb;
    const int LR = d >> 2, G = FT / LR, g = threadIdx.x / LR, c = 4 * (threadIdx.x - g * LR);
    const bool active = g < G;
    float4 b2 = f4(0.f);
    if (active) { const float4 sb = ldg4(sqrt_beta + c); b2 = sb * sb; }
    const float4 omb2 = f4(1.0f) - b2;
    for (int r0 = 0; r0 < r; r0 += RCH) {
        float4 ax[RCH], ag[RCH];
#pragma unroll
        for (int q = 0; q < RCH; ++q) ax[q] = ag[q] = f4(0.f);
        if (active)
            for (int t = g; t < L; t += G) {
                const float4 xv = ld4(src + t * d + c);
                const float4 gv = omb2 * (ld4(gb + t * d + c) * mult4(dr, (unsigned long long)(base + (long long)t * d + c)));
#pragma unroll
                for (int q = 0; q < RCH; ++q)
                    if (r0 + q < r) {
                        const float cq = cb[t * r + r0 + q];
                        ax[q] = fma4(cq, xv, ax[q]);
                        ag[q] = fma4(cq, gv, ag[q]);
                    }
            }
        if (fold_off) {
            fold_groups_scratch(fsm + fold_off, coef_x, ax, r0, r, d, c, g, G, active);
            fold_groups_scratch(fsm + fold_off, coef_g, ag, r0, r, d, c, g, G, active);
        } else {
            fold_groups(coef_x, ax, r0, r, d, c, g, G, active);
            fold_groups(coef_g, ag, r0, r, d, c, g, G, active);
        }
    }
    float4 acc = f4(0.f);
    if (active)
        for (int t = g; t < L; t += G) {
            float4 low = f4(0.f), fg = f4(0.f);
            for (int q = 0; q < r; ++q) {
                const float cq = cb[t * r + q];
                low = fma4(cq, ld4(coef_x + q * d + c), low);
                fg = fma4(cq, ld4(coef_g + q * d + c), fg);
            }
            const int o = t * d + c;
            const float4 dzf = ld4(gb + o);
            const float4 du = dzf * mult4(dr, (unsigned long long)(base + o));
            acc = acc + du * (ld4(src + o) - low);
            st4(dxb + o, ld4(dxb + o) + dzf + b2 * du + fg);
        }
    // fold the per-group dbeta2 partials, group after group (or through the scratch: same order, one barrier round)
    if (fold_off) {
        float* scratch = fsm + fold_off;
        if (active) st4(scratch + (long long)g * d + c, acc);
        __syncthreads();
        if (active && g == 0) {
            float4 sum = ld4(scratch + c);
            for (int gg = 1; gg < G; ++gg) sum = sum + ld4(scratch + (long long)gg * d + c);
            st4(pb + c, sum);
        }
        __syncthreads();
    } else
    for (int gg = 0; gg < G; ++gg) {
        if (active && g == gg) st4(pb + c, gg == 0 ? acc : ld4(pb + c) + acc);
        __syncthreads();
    }
    for (int j = threadIdx.x; j < d; j += FT) pbeta2[(long long)b * d + j] = pb[j];
}

constexpr size_t FOLD_SMEM_LIMIT = 110 * 1024;      // with the fold scratch two CTAs still share an SM
static const size_t STAGE_SMEM_LIMIT = [] { const char* e = getenv("BSA_FILTER_STAGE_KB"); return (size_t)(e ? atoi(e) : 100) * 1024; }();
// The tile is staged in shared memory only while two CTAs still fit an SM (<= 100 KB: C1, C2).  A 205 KB tile (C3) means one CTA
// per SM whose load, projection and store phases cannot overlap anything; unstaged, four CTAs share the SM, pass 2 re-reads the
// tile through L2, and the one-round fold's scratch fits (profiles/r2s_*, r4h_*).  BSA_FILTER_STAGE_KB=220 restores staging.

}  // namespace

// ============================================================================ host launchers
bool rowwise_vec_ok(int d) { return d % 4 == 0 && d >= 4 && d <= 1024; }

int embed_fwd_vec(const long long* ids, const float* E, const float* Ppos, const float* gamma, const float* beta, float* out,
                  float* mean, float* rstd, long long M, int L, int d, Dropout dr, int grid, cudaStream_t st) {
    BSA_DISPATCH_NV(d, (embed_fwd_vec_kernel<NV><<<grid, WARPS * 32, 0, st>>>(ids, E, Ppos, gamma, beta, out, mean, rstd, M, L, d, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int embed_bwd_vec(const float* dout, const long long* ids, const float* E, const float* Ppos, const float* gamma,
                  const float* mean, const float* rstd, float* dz, float* pgamma, float* pbeta, long long M, int L, int d,
                  Dropout dr, int grid, cudaStream_t st) {
    const size_t smem = (size_t)WARPS * d * sizeof(float);
    BSA_DISPATCH_NV(d, (embed_bwd_vec_kernel<NV><<<grid, WARPS * 32, smem, st>>>(dout, ids, E, Ppos, gamma, mean, rstd, dz, pgamma,
                                                                                pbeta, M, L, d, dr)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int residual_ln_fwd_vec(const float* tmp, const float* resid, const float* gamma, const float* beta, float* out, float* z_save,
                        float* mean, float* rstd, const float* blend_with, float blend_a, float blend_b, long long M, int d,
                        Dropout dr, int grid, cudaStream_t st, float* y_hi, float* y_lo) {
    BSA_DISPATCH_NV(d, (residual_ln_fwd_vec_kernel<NV><<<grid, WARPS * 32, 0, st>>>(tmp, resid, gamma, beta, out, z_save, mean, rstd,
                                                                                   blend_with, blend_a, blend_b, M, d, dr, y_hi, y_lo)));
    BSA_LAUNCH_CHECK();
    return 0;
}

int ln_bwd_vec(const float* dy, float dy_scale, const float* z, const float* mean, const float* rstd, const float* gamma,
               float* dz, float* ddrop, float* pgamma, float* pbeta, long long M, int d, Dropout dr, int grid, float* t_hi,
               float* t_lo, float* pcol, cudaStream_t st) {
    const size_t smem = (size_t)WARPS * d * sizeof(float);
    BSA_DISPATCH_NV(d, (ln_bwd_vec_kernel<NV><<<grid, WARPS * 32, smem, st>>>(dy, dy_scale, z, mean, rstd, gamma, dz, ddrop, pgamma,
                                                                             pbeta, M, d, dr, t_hi, t_lo, pcol)));
    BSA_LAUNCH_CHECK();
    return 0;
}

template <class K>
static int set_smem(K kernel, size_t smem) {
    if (smem > 48 * 1024) BSA_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return 0;
}

constexpr size_t CB_SMEM_LIMIT = 16 * 1024;     // the basis [L][r] rides in shared memory when it is this small (C1-C3: 4 KB)

int filter_ln_fwd_vec(const float* x, const float* Cb, const float* sqrt_beta, const float* gamma, const float* beta, float* dsp,
                      float* z_save, float* mean, float* rstd, int B, int L, int d, int r, Dropout dr, cudaStream_t st, float* x_lo) {
    const size_t cbb = (((size_t)L * r + 3) & ~(size_t)3) * sizeof(float);
    const int cb_smem = cbb <= CB_SMEM_LIMIT;
    const size_t coef = ((size_t)r * d + 3 * d) * sizeof(float) + (cb_smem ? cbb : 0), tile = (size_t)L * d * sizeof(float);
    BSA_REQUIRE(coef <= 200 * 1024, "filter: basis rank %d x hidden %d does not fit shared memory", r, d);
    const bool stage = coef + tile <= STAGE_SMEM_LIMIT;
    size_t smem = coef + (stage ? tile : 0);
    // one-round fold of the projection partials when its scratch [G][min(r, RCH)][d] still fits (small tiles: C1, C2)
    const size_t fold_bytes = (size_t)(FT / (d >> 2)) * (r < RCH ? r : RCH) * d * sizeof(float);
    int fold_off = 0;
    if (smem + fold_bytes <= FOLD_SMEM_LIMIT) { fold_off = (int)(smem / sizeof(float)); smem += fold_bytes; }
    int rc = 0;
    BSA_DISPATCH_NV(d, {
        if (stage) {
            if ((rc = set_smem(filter_ln_fwd_vec_kernel<NV, true>, smem))) return rc;
            filter_ln_fwd_vec_kernel<NV, true><<<B, FT, smem, st>>>(x, Cb, sqrt_beta, gamma, beta, dsp, z_save, mean, rstd, L, d, r, cb_smem, dr, x_lo, fold_off);
        } else {
            if ((rc = set_smem(filter_ln_fwd_vec_kernel<NV, false>, smem))) return rc;
            filter_ln_fwd_vec_kernel<NV, false><<<B, FT, smem, st>>>(x, Cb, sqrt_beta, gamma, beta, dsp, z_save, mean, rstd, L, d, r, cb_smem, dr, x_lo, fold_off);
        }
    });
    BSA_LAUNCH_CHECK();
    return 0;
}

int filter_bwd_vec(const float* x, const float* Cb, const float* sqrt_beta, const float* dz_f, float* dx, float* pbeta2, int B,
                   int L, int d, int r, Dropout dr, cudaStream_t st) {
    const size_t cbb = (((size_t)L * r + 3) & ~(size_t)3) * sizeof(float);
    const int cb_smem = cbb <= CB_SMEM_LIMIT;
    const size_t coef = (2 * (size_t)r * d + d) * sizeof(float) + (cb_smem ? cbb : 0), tile = (size_t)L * d * sizeof(float);
    BSA_REQUIRE(coef <= 200 * 1024, "filter bwd: basis rank %d x hidden %d does not fit shared memory", r, d);
    const bool stage = coef + tile <= STAGE_SMEM_LIMIT;
    size_t smem = coef + (stage ? tile : 0);
    const size_t fold_bytes = (size_t)(FT / (d >> 2)) * (r < RCH ? r : RCH) * d * sizeof(float);
    int fold_off = 0;
    if (smem + fold_bytes <= FOLD_SMEM_LIMIT) { fold_off = (int)(smem / sizeof(float)); smem += fold_bytes; }
    int rc = 0;
    if (stage) {
        if ((rc = set_smem(filter_bwd_vec_kernel<true>, smem))) return rc;
        filter_bwd_vec_kernel<true><<<B, FT, smem, st>>>(x, Cb, sqrt_beta, dz_f, dx, pbeta2, L, d, r, cb_smem, dr, fold_off);
    } else {
        if ((rc = set_smem(filter_bwd_vec_kernel<false>, smem))) return rc;
        filter_bwd_vec_kernel<false><<<B, FT, smem, st>>>(x, Cb, sqrt_beta, dz_f, dx, pbeta2, L, d, r, cb_smem, dr, fold_off);
    }
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

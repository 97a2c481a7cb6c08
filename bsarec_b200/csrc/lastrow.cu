// Last-position path of the LAST BSARecBlock.
//
// calculate_loss (model/bsarec.py:32) and the trainer's eval branch (trainers.py:127) consume only position L-1 of the
// last block's output, so in that block only ONE query row per sequence is live: its filter output is a fixed
// weighted sum over positions, its attention needs one score row per head, and everything after the attention is a
// [B, d] problem.  With the key/value projections folded into the query side
//     s_j   = scale * (q_h . (Wk_h x_j + bk_h)) = scale * ((Wk_h^T q_h) . x_j + q_h . bk_h)
//     ctx_h = sum_j pd_j (Wv_h x_j + bv_h)      = Wv_h (sum_j pd_j x_j) + bv_h sum_j pd_j
// the block needs no [B*L, d] GEMM at all: per sequence, one pass over x [L, d] for the H score rows and one for the
// H weighted sums (and the filter's).  Same result as the full block at row L-1 (up to fp32 rounding), same dropout
// masks (indices of the full tensors), and the backward is the exact adjoint: dx [B, L, d] for every position.
// SURVEY.md §7.4 ("last-layer pruning") lists this as a result-preserving shortcut of the reference's math.
#include <cstdlib>
#include "lastrow.cuh"

namespace bsa {
namespace {

constexpr int LR_THREADS = 512;
constexpr int LR_WARPS = LR_THREADS / 32;
constexpr int LR_MAXH = 8;          // heads
constexpr int LR_ROWS = 4;          // positions per warp iteration in the score pass (head vectors re-used across them)

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4(float a) { return make_float4(a, a, a, a); }
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 fma4(float a, float4 b, float4 c) {
    return make_float4(fmaf(a, b.x, c.x), fmaf(a, b.y, c.y), fmaf(a, b.z, c.z), fmaf(a, b.w, c.w));
}
__device__ __forceinline__ float dot4(float4 a, float4 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w))); }
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// filter row of the last position: f[s] = sum_r C[L-1][r] * C[s][r]
__device__ __forceinline__ void filter_last_weights(float* f, const float* __restrict__ Cb, int L, int r) {
    for (int s = threadIdx.x; s < L; s += LR_THREADS) {
        float a = 0.f;
        for (int q = 0; q < r; ++q) a += __ldg(Cb + (long long)(L - 1) * r + q) * __ldg(Cb + (long long)s * r + q);
        f[s] = a;
    }
}

// out[h][j] = bias[h] + vec[h] . x_j for every position j: one warp per LR_ROWS positions, lanes over float4 columns,
// the H head vectors loaded once per column chunk and used for all LR_ROWS positions
template <int HT>
__device__ __forceinline__ void rows_dot_heads(float* out, const float* src, const float* vec_s, const float* bias_s,
                                               int H, int L, int d) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j0 = warp; j0 < L; j0 += LR_WARPS * LR_ROWS) {
        float acc[LR_ROWS][HT];
#pragma unroll
        for (int rr = 0; rr < LR_ROWS; ++rr)
#pragma unroll
            for (int h = 0; h < HT; ++h) acc[rr][h] = 0.f;
        for (int c = 4 * lane; c < d; c += 128) {
            float4 v[HT];
#pragma unroll
            for (int h = 0; h < HT; ++h) v[h] = h < H ? ld4(vec_s + h * d + c) : f4(0.f);
#pragma unroll
            for (int rr = 0; rr < LR_ROWS; ++rr) {
                const int j = j0 + rr * LR_WARPS;
                if (j < L) {
                    const float4 xv = ld4(src + j * d + c);
#pragma unroll
                    for (int h = 0; h < HT; ++h) acc[rr][h] += dot4(xv, v[h]);
                }
            }
        }
#pragma unroll
        for (int rr = 0; rr < LR_ROWS; ++rr) {
            const int j = j0 + rr * LR_WARPS;
#pragma unroll
            for (int h = 0; h < HT; ++h) {
                const float t = warp_sum(acc[rr][h]);
                if (lane == 0 && j < L && h < H) out[h * L + j] = t + bias_s[h];
            }
        }
    }
}

// shared memory (floats): vec [H][d] | fold [H+1][d] | tile [L][d] (STAGE) | sc [H][L] | aux [H][L] (bwd) | f [L] | small [8]
struct LrSmem { float *vec, *fold, *tile, *sc, *aux, *f, *small; };
__device__ __forceinline__ LrSmem lr_carve(float* base, int H, int L, int d, bool stage, bool bwd) {
    LrSmem s;
    s.vec = base;
    s.fold = s.vec + H * d;
    s.tile = s.fold + (H + 1) * d;
    s.sc = s.tile + (stage ? L * d : 0);
    s.aux = s.sc + H * L;
    s.f = s.aux + (bwd ? H * L : 0);
    s.small = s.f + L;
    return s;
}

// ---------------------------------------------------------------------------- forward
// grid = B.  With STAGE the tile x[b] is copied to shared memory once (cp.async) and both passes read it there.
template <int HT, bool STAGE>
__global__ void __launch_bounds__(LR_THREADS)
last_attn_fwd_kernel(const float* __restrict__ x, const long long* __restrict__ ids, const float* __restrict__ gq,
                     const float* __restrict__ ql, const float* __restrict__ bk, const float* __restrict__ Cb, int r,
                     float* __restrict__ p_out, float* __restrict__ xs, float* __restrict__ sp, float* __restrict__ low,
                     int H, int L, int d, float scale, Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) float lr_smem[];
    const LrSmem m = lr_carve(lr_smem, H, L, d, STAGE, false);
    float *sc = m.sc, *f = m.f, *vec = m.vec, *cq = m.small;
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, dh = d / H;
    const float* xb = x + (long long)b * L * d;
    if (STAGE)
        for (int i = threadIdx.x; i < L * d / 4; i += LR_THREADS) cp_async16(m.tile + 4 * i, xb + 4 * i);
    filter_last_weights(f, Cb, L, r);
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) vec[i] = gq[(long long)b * H * d + i];
    for (int h = warp; h < H; h += LR_WARPS) {          // cq[h] = q_h . bk_h
        float a = 0.f;
        for (int c = lane; c < dh; c += 32) a += ql[(long long)b * d + h * dh + c] * __ldg(bk + h * dh + c);
        a = warp_sum(a);
        if (lane == 0) cq[h] = a;
    }
    if (STAGE) cp_async_wait_all();
    __syncthreads();
    const float* src = STAGE ? m.tile : xb;
    rows_dot_heads<HT>(sc, src, vec, cq, H, L, d);
    __syncthreads();
    // softmax of the masked, scaled scores of query row L-1 (every j <= L-1: only the padding mask applies)
    for (int h = warp; h < H; h += LR_WARPS) {
        float mx = -INFINITY;
        for (int j = lane; j < L; j += 32) {
            const float v = sc[h * L + j] * scale + ((ids[(long long)b * L + j] > 0) ? 0.f : -10000.f);
            sc[h * L + j] = v;
            mx = fmaxf(mx, v);
        }
        mx = warp_max(mx);
        float sum = 0.f;
        for (int j = lane; j < L; j += 32) { const float e = expf(sc[h * L + j] - mx); sc[h * L + j] = e; sum += e; }
        sum = warp_sum(sum);
        const float inv = 1.0f / sum;
        const unsigned long long pbase = (((unsigned long long)b * H + h) * L + (L - 1)) * L;
        float spd = 0.f;
        for (int j = lane; j < L; j += 32) {
            const float pr = sc[h * L + j] * inv;
            p_out[((long long)b * H + h) * L + j] = pr;
            const float pd = pr * dropout_mult(dr, pbase + j);
            sc[h * L + j] = pd;
            spd += pd;
        }
        spd = warp_sum(spd);
        if (lane == 0) sp[(long long)b * H + h] = spd;
    }
    __syncthreads();
    // weighted sums over positions: xs[h][c] = sum_j pd[h][j] x[j][c],  low[c] = sum_j f[j] x[j][c].
    // thread = (position group g, float4 column c); groups fold one after the other (fixed order)
    const int LRc = d >> 2, G = LR_THREADS / LRc, g = threadIdx.x / LRc, c = 4 * (threadIdx.x - g * LRc);
    const bool active = g < G;
    float4 acc[HT + 1];
#pragma unroll
    for (int h = 0; h <= HT; ++h) acc[h] = f4(0.f);
    if (active)
        for (int j = g; j < L; j += G) {
            const float4 xv = ld4(src + j * d + c);
            acc[HT] = fma4(f[j], xv, acc[HT]);
#pragma unroll
            for (int h = 0; h < HT; ++h)
                if (h < H) acc[h] = fma4(sc[h * L + j], xv, acc[h]);
        }
    for (int gg = 0; gg < G; ++gg) {
        if (active && g == gg) {
#pragma unroll
            for (int h = 0; h < HT; ++h)
                if (h < H) { float* p = m.fold + h * d + c; st4(p, gg == 0 ? acc[h] : add4(ld4(p), acc[h])); }
            float* p = m.fold + H * d + c;
            st4(p, gg == 0 ? acc[HT] : add4(ld4(p), acc[HT]));
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) xs[(long long)b * H * d + i] = m.fold[i];
    for (int i = threadIdx.x; i < d; i += LR_THREADS) low[(long long)b * d + i] = m.fold[H * d + i];
}

// ---------------------------------------------------------------------------- backward
// grid = B.
//   dpd_j = dxs_h . x_j + dsp_h ;  dp = dpd o mult ;  ds = p o (dp - sum(p o dp))
//   dx_j  = f[j] dlow + sum_h (pd_hj dxs_h + scale ds_hj gq_h) ;  dgq_h = scale sum_j ds_hj x_j ;  dcq_h = scale sum_j ds_hj
template <int HT, bool STAGE>
__global__ void __launch_bounds__(LR_THREADS)
last_attn_bwd_kernel(const float* __restrict__ x, const float* __restrict__ p_in, const float* __restrict__ gq,
                     const float* __restrict__ dxs, const float* __restrict__ dsp, const float* __restrict__ dlow,
                     const float* __restrict__ Cb, int r, float* __restrict__ dx, float* __restrict__ dgq,
                     float* __restrict__ dcq, int H, int L, int d, float scale, Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) float lr_smem[];
    const LrSmem m = lr_carve(lr_smem, H, L, d, STAGE, true);
    float *a = m.sc, *pdv = m.aux, *f = m.f, *vec = m.vec, *bias = m.small;
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* xb = x + (long long)b * L * d;
    if (STAGE)
        for (int i = threadIdx.x; i < L * d / 4; i += LR_THREADS) cp_async16(m.tile + 4 * i, xb + 4 * i);
    filter_last_weights(f, Cb, L, r);
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) vec[i] = dxs[(long long)b * H * d + i];
    if (threadIdx.x < H) bias[threadIdx.x] = dsp[(long long)b * H + threadIdx.x];
    if (STAGE) cp_async_wait_all();
    __syncthreads();
    const float* src = STAGE ? m.tile : xb;
    rows_dot_heads<HT>(a, src, vec, bias, H, L, d);
    __syncthreads();
    for (int h = warp; h < H; h += LR_WARPS) {
        const unsigned long long pbase = (((unsigned long long)b * H + h) * L + (L - 1)) * L;
        const float* pr = p_in + ((long long)b * H + h) * L;
        float dot = 0.f;
        for (int j = lane; j < L; j += 32) {
            const float mlt = dropout_mult(dr, pbase + j);
            const float pj = pr[j];
            const float dp = a[h * L + j] * mlt;
            pdv[h * L + j] = pj * mlt;
            a[h * L + j] = dp;
            dot += pj * dp;
        }
        dot = warp_sum(dot);
        float sds = 0.f;
        for (int j = lane; j < L; j += 32) {
            const float ds = pr[j] * (a[h * L + j] - dot);
            a[h * L + j] = ds;
            sds += ds;
        }
        sds = warp_sum(sds);
        if (lane == 0) dcq[(long long)b * H + h] = scale * sds;
    }
    __syncthreads();
    const int LRc = d >> 2, G = LR_THREADS / LRc, g = threadIdx.x / LRc, c = 4 * (threadIdx.x - g * LRc);
    const bool active = g < G;
    float4 rdxs[HT], rgq[HT], acc[HT], rdlow = f4(0.f);
#pragma unroll
    for (int h = 0; h < HT; ++h) {
        const bool ok = active && h < H;
        rdxs[h] = ok ? ld4(vec + h * d + c) : f4(0.f);
        rgq[h] = ok ? ld4(gq + ((long long)b * H + h) * d + c) : f4(0.f);
        rgq[h] = make_float4(scale * rgq[h].x, scale * rgq[h].y, scale * rgq[h].z, scale * rgq[h].w);
        acc[h] = f4(0.f);
    }
    if (active) {
        rdlow = ld4(dlow + (long long)b * d + c);
        for (int j = g; j < L; j += G) {
            const float4 xv = ld4(src + j * d + c);
            const float fj = f[j];
            float4 o = make_float4(fj * rdlow.x, fj * rdlow.y, fj * rdlow.z, fj * rdlow.w);
#pragma unroll
            for (int h = 0; h < HT; ++h) {
                if (h < H) {
                    const float ds = a[h * L + j];
                    o = fma4(pdv[h * L + j], rdxs[h], o);
                    o = fma4(ds, rgq[h], o);
                    acc[h] = fma4(ds, xv, acc[h]);
                }
            }
            st4(dx + ((long long)b * L + j) * d + c, o);
        }
    }
    for (int gg = 0; gg < G; ++gg) {
        if (active && g == gg) {
#pragma unroll
            for (int h = 0; h < HT; ++h)
                if (h < H) { float* p = m.fold + h * d + c; st4(p, gg == 0 ? acc[h] : add4(ld4(p), acc[h])); }
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) dgq[(long long)b * H * d + i] = scale * m.fold[i];
}

// ---------------------------------------------------------------------------- small element-wise pieces ([B, d])
enum LastEw : int {
    EW_FILTER_U = 0,      // out = low + beta2 (xl - low)
    EW_CTX_BIAS,          // out += bv[c] * sp[b][h(c)]
    EW_FILTER_BWD,        // du = dzf * mult ; dlow = (1-beta2) du ; dxl = dxl_in + dzf + beta2 du ; pb2 = du (xl - low)
    EW_HEAD_SCALE,        // out = a[b][c] * s[b][h(c)]          (rows of dbv / dbk before the column sum)
    EW_ADD_HEAD_BIAS,     // out += s[b][h(c)] * v[c]            (dql += dcq * bk)
};

__global__ void last_ew_kernel(int op, float* __restrict__ out, float* __restrict__ out2, float* __restrict__ out3,
                               const float* __restrict__ a, const float* __restrict__ b2, const float* __restrict__ c2,
                               const float* __restrict__ vcol, const float* __restrict__ shead, int B, int d, int H,
                               Dropout dr) {
    dr = dropout_resolve(dr);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * d) return;
    const int c = (int)(i % d);
    const long long row = i / d;
    const int h = c / (d / H);
    switch (op) {
        case EW_FILTER_U: {           // a = low, b2 = xl, vcol = sqrt_beta
            const float sb = vcol[c], be = sb * sb;
            out[i] = a[i] + be * (b2[i] - a[i]);
        } break;
        case EW_CTX_BIAS: out[i] += vcol[c] * shead[row * H + h]; break;
        case EW_FILTER_BWD: {         // a = dzf, b2 = xl, c2 = low, vcol = sqrt_beta ; out = dlow, out2 = dxl (+=), out3 = pb2
            const float sb = vcol[c], be = sb * sb;
            const float du = a[i] * dropout_mult_row(dr, row, d, c);
            out[i] = (1.0f - be) * du;
            out2[i] += a[i] + be * du;
            out3[i] = du * (b2[i] - c2[i]);
        } break;
        case EW_HEAD_SCALE: out[i] = a[i] * shead[row * H + h]; break;
        case EW_ADD_HEAD_BIAS: out[i] += shead[row * H + h] * vcol[c]; break;
    }
}

// dsp[b][h] = sum_{c in h} bv[c] * dctx[b][c]
__global__ void last_dsp_kernel(const float* __restrict__ dctx, const float* __restrict__ bv, float* __restrict__ dsp, int B, int d, int H) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * H) return;
    const int b = i / H, h = i - b * H, dh = d / H;
    float s = 0.f;
    for (int c = 0; c < dh; ++c) s += bv[h * dh + c] * dctx[(long long)b * d + h * dh + c];
    dsp[i] = s;
}

// dx[b][L-1][:] += add[b][:]
__global__ void last_add_rows_kernel(float* __restrict__ dx, const float* __restrict__ add, int B, int L, int d) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * d) return;
    const long long b = i / d, c = i - b * d;
    dx[(b * L + (L - 1)) * d + c] += add[i];
}

int ew(int op, float* out, float* out2, float* out3, const float* a, const float* b2, const float* c2, const float* vcol,
       const float* shead, int B, int d, int H, Dropout dr, cudaStream_t st) {
    const long long n = (long long)B * d;
    last_ew_kernel<<<cdiv(n, 256), 256, 0, st>>>(op, out, out2, out3, a, b2, c2, vcol, shead, B, d, H, dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

static const size_t LR_STAGE_LIMIT = [] { const char* e = getenv("BSA_LASTROW_STAGE_KB"); return (size_t)(e ? atoi(e) : 220) * 1024; }();   // 0: never stage the tile
size_t lr_smem_base(int H, int L, int d, bool bwd) { return ((size_t)H * d + (size_t)(H + 1) * d + (size_t)(bwd ? 2 : 1) * H * L + L + 8) * 4; }
size_t lr_smem_bytes(int H, int L, int d, bool bwd, bool* stage) {
    const size_t base = lr_smem_base(H, L, d, bwd), tile = (size_t)L * d * 4;
    *stage = base + tile <= LR_STAGE_LIMIT;
    return base + (*stage ? tile : 0);
}

template <class K>
int lr_set_smem(K kernel, size_t smem) {
    if (smem > 48 * 1024) BSA_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return 0;
}

#define LR_DISPATCH(H, stage, ...)                                                                      \
    do {                                                                                                \
        if ((H) <= 1) { constexpr int HT = 1; if (stage) { constexpr bool ST = true; __VA_ARGS__; } else { constexpr bool ST = false; __VA_ARGS__; } } \
        else if ((H) <= 2) { constexpr int HT = 2; if (stage) { constexpr bool ST = true; __VA_ARGS__; } else { constexpr bool ST = false; __VA_ARGS__; } } \
        else if ((H) <= 4) { constexpr int HT = 4; if (stage) { constexpr bool ST = true; __VA_ARGS__; } else { constexpr bool ST = false; __VA_ARGS__; } } \
        else { constexpr int HT = 8; if (stage) { constexpr bool ST = true; __VA_ARGS__; } else { constexpr bool ST = false; __VA_ARGS__; } } \
    } while (0)

}  // namespace

bool lastrow_supported(int L, int d, int H) {
    if (H < 1 || H > LR_MAXH || d % H != 0 || d % 4 != 0 || d / 4 > LR_THREADS || L < 1) return false;
    return lr_smem_base(H, L, d, true) <= 200 * 1024;
}

int last_attn_fwd(const float* x, const long long* ids, const float* gq, const float* ql, const float* bk, const float* Cb,
                  int r, float* p, float* xs, float* sp, float* low, int B, int H, int L, int d, Dropout dr, cudaStream_t st) {
    if (B == 0) return 0;
    bool stage;
    const size_t smem = lr_smem_bytes(H, L, d, false, &stage);
    ProfScope prof(PROF_LAST_FWD, st, 4.0 * B * (double)L * d + 8.0 * B * L);      // x read once + ids
    int rc = 0;
    LR_DISPATCH(H, stage, {
        if ((rc = lr_set_smem(last_attn_fwd_kernel<HT, ST>, smem))) return rc;
        last_attn_fwd_kernel<HT, ST><<<B, LR_THREADS, smem, st>>>(x, ids, gq, ql, bk, Cb, r, p, xs, sp, low, H, L, d,
                                                                 1.0f / sqrtf((float)(d / H)), dr);
    });
    BSA_LAUNCH_CHECK();
    return 0;
}

int last_attn_bwd(const float* x, const float* p, const float* gq, const float* dxs, const float* dsp, const float* dlow,
                  const float* Cb, int r, float* dx, float* dgq, float* dcq, int B, int H, int L, int d, Dropout dr,
                  cudaStream_t st) {
    if (B == 0) return 0;
    bool stage;
    const size_t smem = lr_smem_bytes(H, L, d, true, &stage);
    ProfScope prof(PROF_LAST_BWD, st, 8.0 * B * (double)L * d);                    // x read + dx written
    int rc = 0;
    LR_DISPATCH(H, stage, {
        if ((rc = lr_set_smem(last_attn_bwd_kernel<HT, ST>, smem))) return rc;
        last_attn_bwd_kernel<HT, ST><<<B, LR_THREADS, smem, st>>>(x, p, gq, dxs, dsp, dlow, Cb, r, dx, dgq, dcq, H, L, d,
                                                                 1.0f / sqrtf((float)(d / H)), dr);
    });
    BSA_LAUNCH_CHECK();
    return 0;
}

int last_filter_u(float* u, const float* low, const float* xl, const float* sqrt_beta, int B, int d, cudaStream_t st) {
    return ew(EW_FILTER_U, u, nullptr, nullptr, low, xl, nullptr, sqrt_beta, nullptr, B, d, 1, Dropout{}, st);
}
int last_ctx_bias(float* ctx, const float* bv, const float* sp, int B, int d, int H, cudaStream_t st) {
    return ew(EW_CTX_BIAS, ctx, nullptr, nullptr, nullptr, nullptr, nullptr, bv, sp, B, d, H, Dropout{}, st);
}
int last_filter_bwd(const float* dzf, const float* xl, const float* low, const float* sqrt_beta, float* dlow, float* dxl_acc,
                    float* pb2, int B, int d, Dropout dr_rows, cudaStream_t st) {
    return ew(EW_FILTER_BWD, dlow, dxl_acc, pb2, dzf, xl, low, sqrt_beta, nullptr, B, d, 1, dr_rows, st);
}
int last_head_scale(float* out, const float* a, const float* shead, int B, int d, int H, cudaStream_t st) {
    return ew(EW_HEAD_SCALE, out, nullptr, nullptr, a, nullptr, nullptr, nullptr, shead, B, d, H, Dropout{}, st);
}
int last_add_head_bias(float* out, const float* shead, const float* vcol, int B, int d, int H, cudaStream_t st) {
    return ew(EW_ADD_HEAD_BIAS, out, nullptr, nullptr, nullptr, nullptr, nullptr, vcol, shead, B, d, H, Dropout{}, st);
}
int last_dsp(const float* dctx, const float* bv, float* dsp, int B, int d, int H, cudaStream_t st) {
    if (B == 0) return 0;
    last_dsp_kernel<<<cdiv((long long)B * H, 128), 128, 0, st>>>(dctx, bv, dsp, B, d, H);
    BSA_LAUNCH_CHECK();
    return 0;
}
int last_add_rows(float* dx, const float* add, int B, int L, int d, cudaStream_t st) {
    if (B == 0) return 0;
    last_add_rows_kernel<<<cdiv((long long)B * d, 256), 256, 0, st>>>(dx, add, B, L, d);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

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
#include "lastrow.cuh"

namespace bsa {
namespace {

constexpr int LR_THREADS = 256;
constexpr int LR_WARPS = LR_THREADS / 32;
constexpr int LR_MAXH = 8;          // heads
constexpr int LR_CPT = 4;           // columns per thread (d <= 1024)

// filter row of the last position: f[s] = sum_r C[L-1][r] * C[s][r]
__device__ __forceinline__ void filter_last_weights(float* f, const float* __restrict__ Cb, int L, int r) {
    for (int s = threadIdx.x; s < L; s += LR_THREADS) {
        float a = 0.f;
        for (int q = 0; q < r; ++q) a += __ldg(Cb + (long long)(L - 1) * r + q) * __ldg(Cb + (long long)s * r + q);
        f[s] = a;
    }
}

// a[h][j] = bias[h] + vec[h] . x_j   for every position j (one warp per row, lanes over columns)
__device__ __forceinline__ void rows_dot_heads(float* out, const float* __restrict__ xb, const float* vec_s, const float* bias_s,
                                               int H, int L, int d) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j = warp; j < L; j += LR_WARPS) {
        float acc[LR_MAXH];
#pragma unroll
        for (int h = 0; h < LR_MAXH; ++h) acc[h] = 0.f;
        for (int c = lane; c < d; c += 32) {
            const float xv = xb[(long long)j * d + c];
#pragma unroll
            for (int h = 0; h < LR_MAXH; ++h)
                if (h < H) acc[h] += xv * vec_s[h * d + c];
        }
#pragma unroll
        for (int h = 0; h < LR_MAXH; ++h) {
            if (h < H) {
                const float t = warp_sum(acc[h]);
                if (lane == 0) out[h * L + j] = t + bias_s[h];
            }
        }
    }
}

// ---------------------------------------------------------------------------- forward
// grid = B.  smem: sc [H][L] | f [L] | vec [H][d] | cq [H]
__global__ void __launch_bounds__(LR_THREADS)
last_attn_fwd_kernel(const float* __restrict__ x, const long long* __restrict__ ids, const float* __restrict__ gq,
                     const float* __restrict__ ql, const float* __restrict__ bk, const float* __restrict__ Cb, int r,
                     float* __restrict__ p_out, float* __restrict__ xs, float* __restrict__ sp, float* __restrict__ low,
                     int H, int L, int d, float scale, Dropout dr) {
    extern __shared__ float lr_smem[];
    float* sc = lr_smem;
    float* f = sc + H * L;
    float* vec = f + L;
    float* cq = vec + H * d;
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, dh = d / H;
    const float* xb = x + (long long)b * L * d;

    filter_last_weights(f, Cb, L, r);
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) vec[i] = gq[(long long)b * H * d + i];
    for (int h = warp; h < H; h += LR_WARPS) {          // cq[h] = q_h . bk_h
        float a = 0.f;
        for (int c = lane; c < dh; c += 32) a += ql[(long long)b * d + h * dh + c] * __ldg(bk + h * dh + c);
        a = warp_sum(a);
        if (lane == 0) cq[h] = a;
    }
    __syncthreads();
    rows_dot_heads(sc, xb, vec, cq, H, L, d);
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
    // weighted sums over positions: xs[h][c] = sum_j pd[h][j] x[j][c],  low[c] = sum_j f[j] x[j][c]
    float acc[LR_CPT][LR_MAXH + 1];
#pragma unroll
    for (int q = 0; q < LR_CPT; ++q)
#pragma unroll
        for (int h = 0; h <= LR_MAXH; ++h) acc[q][h] = 0.f;
    for (int j = 0; j < L; ++j) {
        const float fj = f[j];
#pragma unroll
        for (int q = 0; q < LR_CPT; ++q) {
            const int c = threadIdx.x + q * LR_THREADS;
            if (c < d) {
                const float xv = xb[(long long)j * d + c];
                acc[q][LR_MAXH] += fj * xv;
#pragma unroll
                for (int h = 0; h < LR_MAXH; ++h)
                    if (h < H) acc[q][h] += sc[h * L + j] * xv;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < LR_CPT; ++q) {
        const int c = threadIdx.x + q * LR_THREADS;
        if (c < d) {
            low[(long long)b * d + c] = acc[q][LR_MAXH];
#pragma unroll
            for (int h = 0; h < LR_MAXH; ++h)
                if (h < H) xs[((long long)b * H + h) * d + c] = acc[q][h];
        }
    }
}

// ---------------------------------------------------------------------------- backward
// grid = B.  smem: a [H][L] (dpd -> ds) | pdv [H][L] | f [L] | vec [H][d] (dxs) | dsp [H]
//   dpd_j = dxs_h . x_j + dsp_h ;  dp = dpd o mult ;  ds = p o (dp - sum(p o dp))
//   dx_j  = f[j] dlow + sum_h (pd_hj dxs_h + scale ds_hj gq_h) ;  dgq_h = scale sum_j ds_hj x_j ;  dcq_h = scale sum_j ds_hj
__global__ void __launch_bounds__(LR_THREADS)
last_attn_bwd_kernel(const float* __restrict__ x, const float* __restrict__ p_in, const float* __restrict__ gq,
                     const float* __restrict__ dxs, const float* __restrict__ dsp, const float* __restrict__ dlow,
                     const float* __restrict__ Cb, int r, float* __restrict__ dx, float* __restrict__ dgq,
                     float* __restrict__ dcq, int H, int L, int d, float scale, Dropout dr) {
    extern __shared__ float lr_smem[];
    float* a = lr_smem;
    float* pdv = a + H * L;
    float* f = pdv + H * L;
    float* vec = f + L;
    float* bias = vec + H * d;
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* xb = x + (long long)b * L * d;

    filter_last_weights(f, Cb, L, r);
    for (int i = threadIdx.x; i < H * d; i += LR_THREADS) vec[i] = dxs[(long long)b * H * d + i];
    if (threadIdx.x < H) bias[threadIdx.x] = dsp[(long long)b * H + threadIdx.x];
    __syncthreads();
    rows_dot_heads(a, xb, vec, bias, H, L, d);
    __syncthreads();
    for (int h = warp; h < H; h += LR_WARPS) {
        const unsigned long long pbase = (((unsigned long long)b * H + h) * L + (L - 1)) * L;
        const float* pr = p_in + ((long long)b * H + h) * L;
        float dot = 0.f;
        for (int j = lane; j < L; j += 32) {
            const float m = dropout_mult(dr, pbase + j);
            const float pj = pr[j];
            const float dp = a[h * L + j] * m;
            pdv[h * L + j] = pj * m;
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
    float rdxs[LR_CPT][LR_MAXH], rgq[LR_CPT][LR_MAXH], acc[LR_CPT][LR_MAXH], rdlow[LR_CPT];
#pragma unroll
    for (int q = 0; q < LR_CPT; ++q) {
        const int c = threadIdx.x + q * LR_THREADS;
        rdlow[q] = c < d ? dlow[(long long)b * d + c] : 0.f;
#pragma unroll
        for (int h = 0; h < LR_MAXH; ++h) {
            const bool ok = c < d && h < H;
            rdxs[q][h] = ok ? vec[h * d + c] : 0.f;
            rgq[q][h] = ok ? scale * gq[((long long)b * H + h) * d + c] : 0.f;
            acc[q][h] = 0.f;
        }
    }
    for (int j = 0; j < L; ++j) {
        const float fj = f[j];
#pragma unroll
        for (int q = 0; q < LR_CPT; ++q) {
            const int c = threadIdx.x + q * LR_THREADS;
            if (c < d) {
                const float xv = xb[(long long)j * d + c];
                float o = fj * rdlow[q];
#pragma unroll
                for (int h = 0; h < LR_MAXH; ++h) {
                    if (h < H) {
                        const float ds = a[h * L + j];
                        o += pdv[h * L + j] * rdxs[q][h] + ds * rgq[q][h];
                        acc[q][h] += ds * xv;
                    }
                }
                dx[((long long)b * L + j) * d + c] = o;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < LR_CPT; ++q) {
        const int c = threadIdx.x + q * LR_THREADS;
        if (c < d) {
#pragma unroll
            for (int h = 0; h < LR_MAXH; ++h)
                if (h < H) dgq[((long long)b * H + h) * d + c] = scale * acc[q][h];
        }
    }
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

size_t lr_smem_fwd(int H, int L, int d) { return ((size_t)H * L + L + (size_t)H * d + H + 8) * 4; }
size_t lr_smem_bwd(int H, int L, int d) { return ((size_t)2 * H * L + L + (size_t)H * d + H + 8) * 4; }

}  // namespace

bool lastrow_supported(int L, int d, int H) {
    if (H < 1 || H > LR_MAXH || d % H != 0 || d > LR_THREADS * LR_CPT || L < 1) return false;
    return lr_smem_bwd(H, L, d) <= 200 * 1024;
}

int last_attn_fwd(const float* x, const long long* ids, const float* gq, const float* ql, const float* bk, const float* Cb,
                  int r, float* p, float* xs, float* sp, float* low, int B, int H, int L, int d, Dropout dr, cudaStream_t st) {
    if (B == 0) return 0;
    const size_t smem = lr_smem_fwd(H, L, d);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(last_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    ProfScope prof(PROF_LAST_FWD, st, 4.0 * B * (double)L * d + 8.0 * B * L);      // x read once + ids
    last_attn_fwd_kernel<<<B, LR_THREADS, smem, st>>>(x, ids, gq, ql, bk, Cb, r, p, xs, sp, low, H, L, d,
                                                      1.0f / sqrtf((float)(d / H)), dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

int last_attn_bwd(const float* x, const float* p, const float* gq, const float* dxs, const float* dsp, const float* dlow,
                  const float* Cb, int r, float* dx, float* dgq, float* dcq, int B, int H, int L, int d, Dropout dr,
                  cudaStream_t st) {
    if (B == 0) return 0;
    const size_t smem = lr_smem_bwd(H, L, d);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(last_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    ProfScope prof(PROF_LAST_BWD, st, 8.0 * B * (double)L * d);                    // x read + dx written
    last_attn_bwd_kernel<<<B, LR_THREADS, smem, st>>>(x, p, gq, dxs, dsp, dlow, Cb, r, dx, dgq, dcq, H, L, d,
                                                      1.0f / sqrtf((float)(d / H)), dr);
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

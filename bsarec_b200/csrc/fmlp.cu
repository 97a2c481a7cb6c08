// Sibling model FMLPRec (model/fmlprec.py:95-118): the filter layer  y = irfft(rfft(x, dim=1, 'ortho') * W, n=L, 'ortho')
// with a LEARNED complex weight W [L/2+1, d] per channel is, exactly, a per-channel circular convolution along the sequence:
//     y[t, c] = sum_s K[(t - s) mod L, c] x[s, c],     K[:, c] = irfft(W[:, c], n = L)   (the 1/L-normalised inverse),
// i.e. K = Tc W_re + Ts W_im with Tc[tau, f] = w_f cos(2 pi f tau / L) / L, Ts[tau, f] = -w_f sin(2 pi f tau / L) / L,
// w_0 = 1, w_f = 2, w_{L/2} = 1 (L even): "the same L x L operator kernel with d distinct matrices" (SURVEY.md §8f rank 4).
// Kernels: the trig tables, K from W and dW from dK (small dense sums), and ONE circular kernel used three ways —
// forward convolution, its adjoint (dx), and the per-sequence cross-correlation whose sum over the batch is dK.
// fp32 FFMA; a thread owns 4 consecutive outputs of one channel and slides the kernel taps through registers.
#include "common.cuh"
#include "rowwise.cuh"

namespace bsa {
namespace {

__global__ void fmlp_trig_kernel(float* __restrict__ Tc, float* __restrict__ Ts, int L, int F) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * F) return;
    const int tau = i / F, f = i - tau * F;
    const double w = (f == 0 || (L % 2 == 0 && f == L / 2)) ? 1.0 : 2.0;
    const long long k = ((long long)f * tau) % L;                 // exact angle reduction
    double sv, cv;
    sincospi(2.0 * (double)k / (double)L, &sv, &cv);
    Tc[i] = (float)(w * cv / L);
    Ts[i] = (float)(-w * sv / L);
}

// K[tau, c] = sum_f Tc[tau, f] W[f, c, 0] + Ts[tau, f] W[f, c, 1]        (W = complex_weight[0] as [F, d, 2])
__global__ void fmlp_kernel_fwd_kernel(const float* __restrict__ W, const float* __restrict__ Tc, const float* __restrict__ Ts,
                                       float* __restrict__ K, int L, int F, int d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * d) return;
    const int tau = i / d, c = i - tau * d;
    float acc = 0.f;
    for (int f = 0; f < F; ++f) {
        const float2 w = *reinterpret_cast<const float2*>(W + ((long long)f * d + c) * 2);
        acc += Tc[tau * F + f] * w.x + Ts[tau * F + f] * w.y;
    }
    K[i] = acc;
}

// dW[f, c, 0] = sum_tau Tc[tau, f] dK[tau, c];  dW[f, c, 1] = sum_tau Ts[tau, f] dK[tau, c]
__global__ void fmlp_kernel_bwd_kernel(const float* __restrict__ dK, const float* __restrict__ Tc, const float* __restrict__ Ts,
                                       float* __restrict__ dW, int L, int F, int d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F * d) return;
    const int f = i / d, c = i - f * d;
    float re = 0.f, im = 0.f;
    for (int tau = 0; tau < L; ++tau) {
        const float g = dK[(long long)tau * d + c];
        re += Tc[tau * F + f] * g;
        im += Ts[tau * F + f] * g;
    }
    *reinterpret_cast<float2*>(dW + ((long long)f * d + c) * 2) = make_float2(re, im);
}

// out[b, t, c] = (add ? add[b, t, c] : 0) + sum_s kern[(t - s) mod L, c] * in[b, s, c]          (adjoint == false)
//              = ...                      + sum_s kern[(s - t) mod L, c] * in[b, s, c]          (adjoint == true)
// kern is [L, d] shared by the batch (kern_bstride == 0) or per sequence (kern_bstride == L*d: cross-correlation).
// grid (B, d / 32); shared memory holds the sequence's 32-channel slab of `in` and of `kern` (kern stored twice in a
// row so that a window of taps never wraps); 256 threads = 8 position groups x 32 channels.
constexpr int FM_CH = 32, FM_TPB = 256, FM_OUT = 4;
__global__ void __launch_bounds__(FM_TPB)
circ_kernel(const float* __restrict__ in, const float* __restrict__ kern, long long kern_bstride, const float* __restrict__ add,
            float* __restrict__ out, int L, int d, int adjoint) {
    extern __shared__ float fsm[];
    float* xs = fsm;                       // [L][32]
    float* ks = fsm + (size_t)L * FM_CH;   // [2L][32]: taps, repeated
    const int b = blockIdx.x, c0 = blockIdx.y * FM_CH;
    const int c = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const float* inb = in + (long long)b * L * d + c0;
    const float* kb = kern + (long long)b * kern_bstride + c0;
    for (int i = threadIdx.x; i < L * FM_CH; i += FM_TPB) {
        const int s = i >> 5, cc = i & 31;
        xs[i] = (c0 + cc < d) ? inb[(long long)s * d + cc] : 0.f;
        const float kv = (c0 + cc < d) ? kb[(long long)s * d + cc] : 0.f;
        // forward reads taps at decreasing index as s grows; store the (possibly reversed) tap sequence so that BOTH
        // variants walk it the same way: q[j] = kern[j] (forward) / kern[(L - j) mod L] (adjoint)
        const int j = adjoint ? (s == 0 ? 0 : L - s) : s;
        ks[j * FM_CH + cc] = kv;
        ks[(j + L) * FM_CH + cc] = kv;
    }
    __syncthreads();
    if (c0 + c >= d) return;
    // out[t] = sum_s q[(t - s) mod L] x[s];  outputs t0 .. t0+3: taps q[t0 + o - s]; as s -> s + 1 every tap index drops by one
    for (int t0 = grp * FM_OUT; t0 < L; t0 += (FM_TPB / 32) * FM_OUT) {
        float acc[FM_OUT] = {0.f, 0.f, 0.f, 0.f};
        // window w[o] = q[(t0 + o - s) mod L]; start at s = 0: indices t0 .. t0+3 (+L to stay positive while sliding down)
        int base = t0 + L;                                   // index of w[0] in the doubled array
        float w0 = ks[(base) * FM_CH + c], w1 = ks[((base + 1) % (2 * L)) * FM_CH + c],
              w2 = ks[((base + 2) % (2 * L)) * FM_CH + c], w3 = ks[((base + 3) % (2 * L)) * FM_CH + c];
        for (int s = 0; s < L; ++s) {
            const float xv = xs[s * FM_CH + c];
            acc[0] += w0 * xv; acc[1] += w1 * xv; acc[2] += w2 * xv; acc[3] += w3 * xv;
            --base;
            if (base < 0) base += L;                          // stays a valid index of the doubled array, same value mod L
            w3 = w2; w2 = w1; w1 = w0; w0 = ks[base * FM_CH + c];
        }
#pragma unroll
        for (int o = 0; o < FM_OUT; ++o) {
            const int t = t0 + o;
            if (t < L) {
                const long long off = ((long long)b * L + t) * d + c0 + c;
                out[off] = acc[o] + (add ? add[off] : 0.f);
            }
        }
    }
}

}  // namespace

bool fmlp_supported(int L, int d) { return L >= 1 && d >= 1 && (size_t)3 * L * FM_CH * 4 <= 200 * 1024; }

int fmlp_trig(float* Tc, float* Ts, int L, cudaStream_t st) {
    const int F = L / 2 + 1;
    fmlp_trig_kernel<<<cdiv((long long)L * F, 256), 256, 0, st>>>(Tc, Ts, L, F);
    BSA_LAUNCH_CHECK();
    return 0;
}

int fmlp_kernel_fwd(const float* W, const float* Tc, const float* Ts, float* K, int L, int d, cudaStream_t st) {
    fmlp_kernel_fwd_kernel<<<cdiv((long long)L * d, 256), 256, 0, st>>>(W, Tc, Ts, K, L, L / 2 + 1, d);
    BSA_LAUNCH_CHECK();
    return 0;
}

int fmlp_kernel_bwd(const float* dK, const float* Tc, const float* Ts, float* dW, int L, int d, cudaStream_t st) {
    const int F = L / 2 + 1;
    fmlp_kernel_bwd_kernel<<<cdiv((long long)F * d, 256), 256, 0, st>>>(dK, Tc, Ts, dW, L, F, d);
    BSA_LAUNCH_CHECK();
    return 0;
}

int circ_apply(const float* in, const float* kern, bool kern_per_batch, const float* add, float* out, int B, int L, int d,
               bool adjoint, cudaStream_t st) {
    if (B == 0) return 0;
    BSA_REQUIRE(fmlp_supported(L, d), "fmlp filter: sequence length %d does not fit the shared-memory plan", L);
    const size_t smem = (size_t)3 * L * FM_CH * sizeof(float);
    static size_t configured = 0;
    if (smem > configured && smem > 48 * 1024) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(circ_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const dim3 grid(B, cdiv(d, FM_CH));
    circ_kernel<<<grid, FM_TPB, smem, st>>>(in, kern, kern_per_batch ? (long long)L * d : 0, add, out, L, d, adjoint ? 1 : 0);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

// bsa_adam_step: one fused pass of torch.optim.Adam's update (trainers.py:27-28, :107) over a parameter tensor.
// HBM-bound: reads p, g, m, v and writes p, m, v once (28 B per element) instead of the ~13 foreach passes of the
// stock optimizer.  Same arithmetic order as torch (lerp for exp_avg, mul+addcmul for exp_avg_sq, sqrt/bias-correction/
// eps, addcdiv), so results agree to fp32 rounding.
#include "../../include/bsarec_b200.h"
#include "common.cuh"

namespace bsa {
namespace {

struct AdamP {
    float lr_over_bc1, bc2_sqrt, beta1, beta2, eps, wd;
    // CUDA-graph replays: the step number is `step + replay[1]`, the bias corrections are formed on the device
    float lr = 0.f;
    long long step = 0;
    const unsigned long long* replay = nullptr;
};

__device__ __forceinline__ AdamP adam_resolve(AdamP a) {
    if (a.replay) {
        const double s = (double)(a.step + (long long)__ldg(a.replay + 1));
        const double bc1 = 1.0 - pow((double)a.beta1, s), bc2 = 1.0 - pow((double)a.beta2, s);
        a.lr_over_bc1 = (float)((double)a.lr / bc1);
        a.bc2_sqrt = (float)sqrt(bc2);
    }
    return a;
}

static AdamP make_adam(float lr, float beta1, float beta2, float eps, float wd, long long step) {
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    AdamP a{(float)((double)lr / bc1), (float)sqrt(bc2), beta1, beta2, eps, wd};
    a.lr = lr;
    a.step = step;
    a.replay = g_replay_state;
    return a;
}

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, const AdamP& a) {
    if (a.wd != 0.f) g = fmaf(a.wd, p, g);
    m = m + (g - m) * (1.0f - a.beta1);
    v = v * a.beta2 + (1.0f - a.beta2) * g * g;
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p - a.lr_over_bc1 * (m / denom);
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n, AdamP a_in) {
    const AdamP a = adam_resolve(a_in);
    const long long n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 P = reinterpret_cast<float4*>(p)[i];
        const float4 G = __ldcs(reinterpret_cast<const float4*>(g) + i);      // gradient is dead after this read
        float4 Mv = reinterpret_cast<float4*>(m)[i];
        float4 V = reinterpret_cast<float4*>(v)[i];
        adam1(P.x, G.x, Mv.x, V.x, a); adam1(P.y, G.y, Mv.y, V.y, a); adam1(P.z, G.z, Mv.z, V.z, a); adam1(P.w, G.w, Mv.w, V.w, a);
        reinterpret_cast<float4*>(p)[i] = P;
        reinterpret_cast<float4*>(m)[i] = Mv;
        reinterpret_cast<float4*>(v)[i] = V;
    }
    const long long t = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) adam1(p[t], g[t], m[t], v[t], a);
}

// Up to ADAM_MULTI tensors in one launch (blockIdx.y = tensor): the 41 small parameters of the model are one launch
// instead of 41.  Pointers travel by value in the kernel parameter block.
constexpr int ADAM_MULTI = 48;
struct AdamMulti {
    float* p[ADAM_MULTI]; const float* g[ADAM_MULTI]; float* m[ADAM_MULTI]; float* v[ADAM_MULTI];
    long long n[ADAM_MULTI];
};

__global__ void __launch_bounds__(256)
adam_multi_kernel(const __grid_constant__ AdamMulti t, AdamP a_in) {
    const AdamP a = adam_resolve(a_in);
    const int k = blockIdx.y;
    float* __restrict__ p = t.p[k];
    const float* __restrict__ g = t.g[k];
    float* __restrict__ m = t.m[k];
    float* __restrict__ v = t.v[k];
    const long long n = t.n[k], n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 P = reinterpret_cast<float4*>(p)[i];
        const float4 G = __ldcs(reinterpret_cast<const float4*>(g) + i);
        float4 Mv = reinterpret_cast<float4*>(m)[i];
        float4 V = reinterpret_cast<float4*>(v)[i];
        adam1(P.x, G.x, Mv.x, V.x, a); adam1(P.y, G.y, Mv.y, V.y, a); adam1(P.z, G.z, Mv.z, V.z, a); adam1(P.w, G.w, Mv.w, V.w, a);
        reinterpret_cast<float4*>(p)[i] = P;
        reinterpret_cast<float4*>(m)[i] = Mv;
        reinterpret_cast<float4*>(v)[i] = V;
    }
    const long long e = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) adam1(p[e], g[e], m[e], v[e], a);
}

}  // namespace

int adam_step_multi(int count, float* const* p, const float* const* g, float* const* m, float* const* v, const long long* n,
                    float lr, float beta1, float beta2, float eps, float wd, long long step, cudaStream_t st) {
    BSA_REQUIRE(step >= 1, "bsa_adam_step_multi: step counts from 1");
    const AdamP a = make_adam(lr, beta1, beta2, eps, wd, step);
    for (int base = 0; base < count; base += ADAM_MULTI) {
        AdamMulti t;
        const int c = count - base < ADAM_MULTI ? count - base : ADAM_MULTI;
        long long nmax = 0;
        double bytes = 0;
        for (int i = 0; i < ADAM_MULTI; ++i) {
            const int j = base + (i < c ? i : 0);
            t.p[i] = p[j]; t.g[i] = g[j]; t.m[i] = m[j]; t.v[i] = v[j]; t.n[i] = i < c ? n[j] : 0;
            if (i < c) {
                BSA_REQUIRE(((reinterpret_cast<uintptr_t>(p[j]) | reinterpret_cast<uintptr_t>(g[j]) | reinterpret_cast<uintptr_t>(m[j]) |
                              reinterpret_cast<uintptr_t>(v[j])) & 15) == 0, "bsa_adam_step_multi: tensors must be 16-byte aligned");
                if (n[j] > nmax) nmax = n[j];
                bytes += 28.0 * (double)n[j];
            }
        }
        if (nmax == 0) continue;
        long long bx = (nmax / 4 + 255) / 256;
        if (bx > 64) bx = 64;
        if (bx < 1) bx = 1;
        ProfScope prof(PROF_ADAM, st, bytes);
        adam_multi_kernel<<<dim3((unsigned)bx, (unsigned)c), 256, 0, st>>>(t, a);
        BSA_LAUNCH_CHECK();
    }
    return 0;
}

int adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps,
              float wd, long long step, cudaStream_t st) {
    if (n == 0) return 0;
    BSA_REQUIRE(step >= 1, "bsa_adam_step: step counts from 1");
    BSA_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                  reinterpret_cast<uintptr_t>(v)) & 15) == 0, "bsa_adam_step: tensors must be 16-byte aligned");
    const AdamP a = make_adam(lr, beta1, beta2, eps, wd, step);
    long long blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    ProfScope prof(PROF_ADAM, st, 28.0 * (double)n);
    adam_kernel<<<(int)blocks, 256, 0, st>>>(p, g, m, v, n, a);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

extern "C" int bsa_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int64_t step, void* stream) {
    using namespace bsa;
    BSA_REQUIRE(n == 0 || (param && grad && exp_avg && exp_avg_sq), "bsa_adam_step: null argument");
    return adam_step(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, (cudaStream_t)stream);
}

extern "C" int bsa_adam_step_multi(int count, float* const* params_host, const float* const* grads_host, float* const* exp_avg_host,
                                   float* const* exp_avg_sq_host, const int64_t* sizes_host, float lr, float beta1, float beta2,
                                   float eps, float weight_decay, int64_t step, void* stream) {
    using namespace bsa;
    if (count == 0) return 0;
    BSA_REQUIRE(count > 0 && params_host && grads_host && exp_avg_host && exp_avg_sq_host && sizes_host, "bsa_adam_step_multi: null argument");
    return adam_step_multi(count, params_host, grads_host, exp_avg_host, exp_avg_sq_host, (const long long*)sizes_host, lr, beta1, beta2,
                           eps, weight_decay, step, (cudaStream_t)stream);
}

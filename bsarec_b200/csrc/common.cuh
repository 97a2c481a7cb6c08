// Shared device/host helpers for the bsarec_b200 kernels (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

namespace bsa {

// ---------------------------------------------------------------------------- errors / counters
// Every C-ABI export returns int (0 = ok); the message is kept per thread and read back with
// bsa_last_error_string() (SURVEY.md §8b "Errors").
void set_error(const char* fmt, ...);
extern unsigned long long g_launch_count;      // kernels launched by this library (bench.py "gpu_launches")

#define BSA_CHECK_CUDA(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            bsa::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                                    \
        }                                                                                      \
    } while (0)

#define BSA_REQUIRE(cond, ...)                 \
    do {                                       \
        if (!(cond)) {                         \
            bsa::set_error(__VA_ARGS__);       \
            return -1;                         \
        }                                      \
    } while (0)

#define BSA_LAUNCH_CHECK()                     \
    do {                                       \
        ++bsa::g_launch_count;                 \
        BSA_CHECK_CUDA(cudaPeekAtLastError()); \
    } while (0)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------------------- kernel timing (bench.py roofline)
// When enabled through bsa_profile_enable(1), tagged launches are bracketed by CUDA events on the launching
// stream; bsa_profile_get() synchronises those events and reports total device time, algorithmic work
// (FLOPs or bytes, supplied by the launcher) and launch count per tag.  Off by default: zero overhead.
extern int g_profile_on;
void prof_begin(int tag, cudaStream_t st);
void prof_end(int tag, cudaStream_t st, double work);
enum ProfTag : int { PROF_GEMM_SIMT = 0, PROF_CE_FWD_UMMA, PROF_CE_BWD_UMMA, PROF_TOPK_UMMA, PROF_LINEAR_UMMA,
                     PROF_EMBED_FWD, PROF_EMBED_BWD, PROF_SCATTER_ADD, PROF_ATTN_FWD, PROF_ATTN_BWD, PROF_ADAM, PROF_LAST_FWD, PROF_LAST_BWD,
                     PROF_FILTER_FWD, PROF_FILTER_BWD, PROF_RESID_LN_FWD, PROF_LN_BWD, PROF_NUM_TAGS };

struct ProfScope {
    int tag; cudaStream_t st; double work;
    ProfScope(int t, cudaStream_t s, double w) : tag(t), st(s), work(w) { if (g_profile_on) prof_begin(tag, st); }
    ~ProfScope() { if (g_profile_on) prof_end(tag, st, work); }
};

// ---------------------------------------------------------------------------- dropout RNG
// Stateless counter RNG, restated bit-for-bit in oracle/rng_ref.py.
struct Dropout {
    unsigned long long key;   // site key
    unsigned int thr;         // keep iff r16 >= thr ; thr == 0 -> dropout off
    float scale;              // 1/(1-p)
    // row mapping for kernels that run on a row subset of a [rows, d] tensor (the last-position path): element (row, j)
    // draws the mask of element (row * rmul + radd, j) of the full tensor.  Identity by default.
    long long rmul = 1, radd = 0;
    // CUDA-graph replays (bsa_set_replay_state): when `epoch` is set the key is formed ON THE DEVICE from
    // pre + *epoch, i.e. the launch behaves as if it had been made with seed + *epoch — one captured launch, a new
    // dropout stream every replay.  Null (the default) leaves `key` as computed on the host.
    unsigned long long pre = 0;
    const unsigned long long* epoch = nullptr;
};

// device words of the replay state: [0] = seed offset added to every dropout seed, [1] = Adam step offset
extern const unsigned long long* g_replay_state;

__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline Dropout make_dropout(unsigned long long seed, int site, float p, int training) {
    Dropout d;
    d.pre = seed + 0x9E3779B97F4A7C15ull * (unsigned long long)(site + 1);
    d.key = splitmix64(d.pre);
    d.epoch = g_replay_state;
    d.thr = (training && p > 0.f) ? (unsigned int)(p * 65536.0f + 0.5f) : 0u;
    d.scale = (d.thr ? 1.0f / (1.0f - p) : 1.0f);
    return d;
}

static inline Dropout dropout_rows(Dropout d, long long rmul, long long radd) { d.rmul = rmul; d.radd = radd; return d; }

__device__ __forceinline__ unsigned long long dropout_key(const Dropout& dr) {
    return dr.epoch ? splitmix64(dr.pre + __ldg(dr.epoch)) : dr.key;
}

// Kernels call this once at entry: with a replay epoch the key is a second hash of a device word, and leaving that inside
// dropout_mult* repeats it for every group of four elements (ncu, attention forward: 20 IMAD per mask hash instead of 8).
__device__ __forceinline__ Dropout dropout_resolve(Dropout dr) {
    if (dr.epoch) { dr.key = splitmix64(dr.pre + __ldg(dr.epoch)); dr.epoch = nullptr; }
    return dr;
}

// multiplier (0 or scale) for linear element index i
__device__ __forceinline__ float dropout_mult(const Dropout& dr, unsigned long long i) {
    if (dr.thr == 0u) return 1.0f;
    unsigned long long w = splitmix64(dropout_key(dr) + (i >> 2));
    unsigned int r16 = (unsigned int)(w >> (16 * (unsigned int)(i & 3ull))) & 0xFFFFu;
    return r16 >= dr.thr ? dr.scale : 0.0f;
}

// element (row, j) of a [*, d] tensor under the row mapping
__device__ __forceinline__ float dropout_mult_row(const Dropout& dr, long long row, int d, int j) {
    return dropout_mult(dr, (unsigned long long)((row * dr.rmul + dr.radd) * d + j));
}

// four consecutive elements starting at i (i % 4 == 0): one hash for all four
__device__ __forceinline__ void dropout_mult4(const Dropout& dr, unsigned long long i, float m[4]) {
    if (dr.thr == 0u) { m[0] = m[1] = m[2] = m[3] = 1.0f; return; }
    unsigned long long w = splitmix64(dropout_key(dr) + (i >> 2));
#pragma unroll
    for (int j = 0; j < 4; ++j) m[j] = (((unsigned int)(w >> (16 * j)) & 0xFFFFu) >= dr.thr) ? dr.scale : 0.0f;
}

// ---------------------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// GELU, erf form (_modules.py:47-56): gelu(x) = x Phi(x), gelu'(x) = Phi(x) + x phi(x).
// Phi through erfc's rational form (Abramowitz & Stegun 7.1.26, |error| <= 1.5e-7 on erf, i.e. 7.5e-8 on Phi — below fp32
// rounding of the values it multiplies and four orders under the 1e-4 parity bar): one rcp.approx, one ex2.approx and five
// FMAs give the value AND the derivative, because erfc's exp(-x^2/2) is also phi's.  erff + expf cost ~2x the instructions,
// and the FFN epilogues of the tcgen05 GEMMs are bound by exactly that count.
__device__ __forceinline__ float gelu_cdf(float x, float& e) {       // returns Phi(x); e = exp(-x^2 / 2)
    float t;                                                         // 1 / (1 + p |x| / sqrt 2)
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.23164189f, fabsf(x), 1.0f)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752044448170368f));      // -0.5 * log2(e)
    float poly = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);  // the 0.5 of erfc / 2 rides in the coefficients
    poly = fmaf(t, poly, 0.5f * 1.421413741f);
    poly = fmaf(t, poly, 0.5f * -0.284496736f);
    poly = fmaf(t, poly, 0.5f * 0.254829592f);
    const float q = (poly * t) * e;                                  // 0.5 erfc(|x| / sqrt 2)
    return x < 0.f ? q : 1.0f - q;
}
__device__ __forceinline__ float gelu_erf(float x) { float e; return x * gelu_cdf(x, e); }
// gelu(x) and gelu'(x) from one evaluation: same expressions (and bits) as gelu_erf / gelu_erf_grad
__device__ __forceinline__ float gelu_erf_both(float x, float& grad) {
    float e;
    const float cdf = gelu_cdf(x, e);
    grad = fmaf(x * 0.39894228040143267794f, e, cdf);
    return x * cdf;
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
    float e;
    const float cdf = gelu_cdf(x, e);
    return fmaf(x * 0.39894228040143267794f, e, cdf);
}

// A second stream per device for leaf launches (Fork): the weight-gradient GEMMs of the small paths, the scatter plan.  They are leaves of the backward pass (their
// outputs are read by the optimizer only), so each one forks from the main stream where its inputs are complete and the call
// joins the side stream before it returns: callers see ordinary single-stream semantics, CUDA-graph capture records the fork
// and join as dependencies.  The [B, d]-sized launches of the last-row path leave most of the GPU idle; the weight gradients
// now run beside the input-gradient chain instead of inside it.  BSA_SIDE_STREAM=0 keeps everything on the caller's stream.
// One side stream and one set of events per device: like the rest of the library (one process per GPU, one host thread driving
// it — trainers.py's loop), this is not meant for several host threads issuing backward calls on the same device at once.
struct SideStream {
    cudaStream_t s = nullptr;
    cudaEvent_t fork[8] = {};
    cudaEvent_t join = nullptr;
    bool ok = false;
};
inline SideStream* side_stream_for_current_device() {
    static const bool enabled = [] { const char* e = getenv("BSA_SIDE_STREAM"); return !(e && atoi(e) == 0); }();
    if (!enabled) return nullptr;
    static SideStream pool[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    SideStream& ss = pool[dev];
    if (!ss.ok) {
        if (cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        for (auto& e : ss.fork)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        ss.ok = true;
    }
    return &ss;
}
struct Fork {
    cudaStream_t main; SideStream* ss; int used = 0;
    Fork(cudaStream_t m, bool enable = true) : main(m), ss(enable ? side_stream_for_current_device() : nullptr) {}
    ~Fork() {               // an error return between a fork and the join: still hand the side stream back to the main one
        if (ss && used) { if (cudaEventRecord(ss->join, ss->s) == cudaSuccess) cudaStreamWaitEvent(main, ss->join, 0); }
    }
    // the stream for a leaf launch whose inputs are complete on the main stream NOW
    cudaStream_t leaf() {
        if (!ss || used >= 8) return main;
        if (cudaEventRecord(ss->fork[used], main) != cudaSuccess || cudaStreamWaitEvent(ss->s, ss->fork[used], 0) != cudaSuccess) return main;
        ++used;
        return ss->s;
    }
    int join() {
        if (!ss || used == 0) return 0;
        BSA_CHECK_CUDA(cudaEventRecord(ss->join, ss->s));
        BSA_CHECK_CUDA(cudaStreamWaitEvent(main, ss->join, 0));
        used = 0;
        return 0;
    }
};

}  // namespace bsa

// tcgen05 3xTF32 engine: epilogues, launchers and the cross-entropy / scoring / linear entry points built on
// umma_gemm.cuh.
#include <cstdlib>
#include <cstring>
#include "umma_engine.cuh"
#include "umma_gemm.cuh"
#include "ce_topk.cuh"

namespace bsa {

using namespace umma;

namespace {

__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    lo = x - hi;                       // exact in fp32
}

__global__ void split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, long long n) {
    const long long n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = reinterpret_cast<const float4*>(x)[i];
        float4 h, l;
        split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
        reinterpret_cast<float4*>(hi)[i] = h;
        reinterpret_cast<float4*>(lo)[i] = l;
    }
    const long long t = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) split_tf32(x[t], hi[t], lo[t]);
}

// "hi = the tensor itself": a kind::tf32 MMA reads only the upper 19 bits of each fp32 operand word, so the fp32 tensor
// can serve as its own hi part when lo = x - truncate(x) (13 low mantissa bits cleared).  The split then writes 4 B per
// element instead of 8 and the hi copy does not exist: item table of the cross-entropy (0.51 -> 0.32 ms at C3, same-box
// A/B profiles/r3e_*), block inputs x / ctx / y.  Verified by the 1e-5 fp64 bars of test_fused_cross_entropy_fwd_bwd and
// test_ce_backward_modes_agree (a rounding MMA input stage would fail them at 2^-11).  BSA_E_TRUNC=0 restores the (rna hi, lo)
// pair for the item table.
__global__ void split_lo_trunc_kernel(const float* __restrict__ x, float* __restrict__ lo, long long n) {
    const long long n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    auto lo1 = [](float v) { return v - __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); };
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = reinterpret_cast<const float4*>(x)[i];
        reinterpret_cast<float4*>(lo)[i] = make_float4(lo1(v.x), lo1(v.y), lo1(v.z), lo1(v.w));
    }
    const long long t = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) lo[t] = lo1(x[t]);
}

// up to SplitMulti::MAXT tensors in one launch (blockIdx.y = tensor): the block's weight matrices are split together
struct SplitMultiP { const float* x[8]; float* hi[8]; float* lo[8]; long long n[8]; };
__global__ void split_tf32_multi_kernel(const __grid_constant__ SplitMultiP t) {
    const int k = blockIdx.y;
    const float* __restrict__ x = t.x[k];
    float* __restrict__ hi = t.hi[k];
    float* __restrict__ lo = t.lo[k];
    const long long n = t.n[k], n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = reinterpret_cast<const float4*>(x)[i];
        float4 h, l;
        split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
        reinterpret_cast<float4*>(hi)[i] = h;
        reinterpret_cast<float4*>(lo)[i] = l;
    }
    const long long e = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) split_tf32(x[e], hi[e], lo[e]);
}

// split + column sums of a [M, N] activation gradient in one pass (N / 4 threads per row, 128 rows per block):
// part[blockIdx.x][N] = sum over the block's rows, reduced afterwards by reduce_partials -> the bias gradient
constexpr int SPLIT_CS_ROWS = 128;
__global__ void __launch_bounds__(256)
split_colsum_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, long long M, int N,
                    float* __restrict__ part, long long out_pitch4, long long in_pitch4) {
    __shared__ float4 red[256];
    const int tpr = N >> 2, groups = 256 / tpr;               // threads per row, rows in flight
    const int c4 = threadIdx.x % tpr, rg = threadIdx.x / tpr;
    const long long r0 = (long long)blockIdx.x * SPLIT_CS_ROWS;
    float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int r = rg; r < SPLIT_CS_ROWS; r += groups) {
        const long long row = r0 + r;
        if (row >= M) break;
        const long long i = row * in_pitch4 + c4;             // x and the pair may be column windows of wider [M, pitch] buffers
        const long long o = row * out_pitch4 + c4;
        if (hi) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(x) + i);
            float4 h, l;
            split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
            reinterpret_cast<float4*>(hi)[o] = h;
            reinterpret_cast<float4*>(lo)[o] = l;
            cs.x += v.x; cs.y += v.y; cs.z += v.z; cs.w += v.w;
            continue;
        }
        // hi == nullptr: x is its own hi half (it is read again by the GEMMs), only lo = x - truncate_tf32(x) is written
        const float4 v = reinterpret_cast<const float4*>(x)[i];
        auto lo1 = [](float t) { return t - __uint_as_float(__float_as_uint(t) & 0xFFFFE000u); };
        reinterpret_cast<float4*>(lo)[o] = make_float4(lo1(v.x), lo1(v.y), lo1(v.z), lo1(v.w));
        cs.x += v.x; cs.y += v.y; cs.z += v.z; cs.w += v.w;
    }
    red[threadIdx.x] = cs;
    __syncthreads();
    if (rg == 0) {
        for (int g2 = 1; g2 < groups; ++g2) { const float4 o = red[g2 * tpr + c4]; cs.x += o.x; cs.y += o.y; cs.z += o.z; cs.w += o.w; }
        reinterpret_cast<float4*>(part + (long long)blockIdx.x * N)[c4] = cs;
    }
}

// ============================================================================ epilogues
// ---- plain store: C = alpha*acc (+bias) (act) (+C)   or raw split-K partials
// Eight epilogue warps: warp e drains TMEM lane quarter e % 4 (32 accumulator rows) and column half e / 4 of the
// tile, 32 columns at a time.  Each 32 x 32 block goes registers -> XOR-swizzled shared staging -> registers, so
// that global traffic is row-contiguous: one warp instruction moves 4 rows x 128 B (full lines) instead of 32
// rows x 16 B.  Bias / activation / accumulate run after the transpose, where aux and C reads coalesce too.
struct StoreParams {
    float* C; long long ldc; const float* bias; int act; float* aux_out; const float* aux_in; float alpha;
    int accumulate; float* partial;
    float* out_hi; float* out_lo;        // optional: also emit the result as a split-TF32 (hi, lo) pair [same ldc]
    int partial_accumulate;              // split-K partial slabs are running sums (added to, not overwritten)
    float* colpart;                      // optional (PRE epilogue, N % 4 == 0): column sums of the stored values per 32-row
                                         // group, [4 * m_tiles, N] -> reduce_partials gives the bias gradient
    // Segmented output (plain epilogue only, seg_cols % 32 == 0): columns [s * seg_cols, (s + 1) * seg_cols) of the result
    // are the separate [M, seg_cols] tensor C + s * seg_stride (pitch ldc) with its own bias — one launch computes
    // q, k and v from one read of the shared A operand.
    int seg_cols = 0; long long seg_stride = 0; const float* bias1 = nullptr; const float* bias2 = nullptr;
};
// PRE: the epilogue reads global memory (GELU pre-activation and / or the running sum) and prefetches those reads.
// CS: it also leaves per-32-row-group column sums of what it stores (StoreParams::colpart).
// ACT >= 0: the activation is a compile-time constant and nothing accumulates (the two FFN GEMMs with an activation in the
// epilogue get their own instantiations: the generic body carries all four activations inside its 8-fold unrolled store
// loop, and ncu showed those launches stalled on branch resolution and instruction fetch — profiles/r3i_*)
template <int BN, bool PRE = false, bool CS = false, int ACT = -1>
struct EpiStore {
    using Params = StoreParams;
    static constexpr int WARPS = 8;
    static constexpr int SMEM_BYTES = WARPS * 32 * 32 * 4;       // one 32 x 32 fp32 staging block per warp
    const Params& p; const GemmCore& g;
    float4* stg; int quarter, half, lane;
    __device__ EpiStore(const Params& p_, const GemmCore& g_, unsigned char* smem, int epi_warp, int lane_)
        : p(p_), g(g_), quarter(epi_warp & 3), half(epi_warp >> 2), lane(lane_) {
        stg = reinterpret_cast<float4*>(smem) + epi_warp * 256;
    }
    __device__ void tile(uint32_t acc, const TileCoord& tc) {
        const long long row0 = (long long)tc.m * BM + quarter * 32;
        const int n0 = tc.n * BN;
        const int sub = lane >> 3, c4 = lane & 7;                // store phase: 4 rows x 8 float4 per instruction
        const int act = ACT >= 0 ? ACT : p.act;
        const bool accum = ACT >= 0 ? false : (p.accumulate != 0);
        float* base; long long ld; bool plain = p.partial == nullptr;
        if (plain) { base = p.C; ld = p.ldc; }
        else { base = p.partial + (long long)tc.split * g.M * g.N; ld = g.N; }
        float* const base0 = base;
        bool vec_ok = ((ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
#pragma unroll 1
        for (int c0 = half * (BN / 2); c0 < (half + 1) * (BN / 2); c0 += 32) {
            int segc = 0;                                        // first column of this chunk's output segment
            const float* bias_p = p.bias;
            if (!PRE && plain && p.seg_cols) {                   // warp-uniform: a 32-column chunk never straddles segments
                const int seg = (n0 + c0) / p.seg_cols;
                segc = seg * p.seg_cols;
                base = base0 + (long long)seg * p.seg_stride;
                bias_p = seg == 0 ? p.bias : (seg == 1 ? p.bias1 : p.bias2);
                vec_ok = ((ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
            }
            // The epilogue's global reads (GELU pre-activation, running sum) are issued before the accumulator leaves
            // TMEM: inside the store loop each one would wait out its full latency behind the previous row's stores.
            const int col = n0 + c0 + c4 * 4;
            const bool full = vec_ok && (col + 3 < g.N);
            float4 aux[8], old[8];
            const bool pre_aux = PRE && plain && full && (act == 2 || act == 4), pre_old = PRE && plain && full && accum;
            if (pre_aux || pre_old) {
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const long long row = row0 + it * 4 + sub;
                    if (row < g.M) {
                        if (pre_aux) aux[it] = __ldg(reinterpret_cast<const float4*>(p.aux_in + row * ld + col));
                        if (pre_old) old[it] = *reinterpret_cast<const float4*>(base + row * ld + col);
                    }
                }
            }
            float v[32];
            tmem_ld32(acc + c0, v);
            float4 csum = make_float4(0.f, 0.f, 0.f, 0.f);
            const bool want_cs = CS && plain && p.colpart != nullptr;
            if (want_cs && row0 >= g.M && sub == 0 && full)      // a 32-row group past the last row still owns its partial row
                *reinterpret_cast<float4*>(p.colpart + ((long long)tc.m * 4 + quarter) * g.N + col) = csum;
            if (n0 + c0 >= g.N || row0 >= g.M) continue;         // warp-uniform
#pragma unroll
            for (int j = 0; j < 8; ++j)
                stg[lane * 8 + (j ^ (lane & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (plain && bias_p) {
                const float* bp = bias_p + (col - segc);
                if (col + 3 < g.N) bias4 = __ldg(reinterpret_cast<const float4*>(bp));
                else {
                    if (col < g.N) bias4.x = __ldg(bp);
                    if (col + 1 < g.N) bias4.y = __ldg(bp + 1);
                    if (col + 2 < g.N) bias4.z = __ldg(bp + 2);
                }
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int r = it * 4 + sub;
                const long long row = row0 + r;
                float4 x = stg[r * 8 + (c4 ^ (r & 7))];
                if (row >= g.M || col >= g.N) continue;
                const long long off = row * ld + (col - segc);
                if (!plain) {                                    // split-K partials: raw (or running) sums
                    if (full) {
                        if (p.partial_accumulate) { const float4 o = *reinterpret_cast<const float4*>(base + off); x.x += o.x; x.y += o.y; x.z += o.z; x.w += o.w; }
                        *reinterpret_cast<float4*>(base + off) = x;
                    } else {
                        const float xs[4] = {x.x, x.y, x.z, x.w};
                        for (int e = 0; e < 4; ++e) if (col + e < g.N) base[off + e] = xs[e] + (p.partial_accumulate ? base[off + e] : 0.f);
                    }
                    continue;
                }
                x.x = p.alpha * x.x + bias4.x; x.y = p.alpha * x.y + bias4.y; x.z = p.alpha * x.z + bias4.z; x.w = p.alpha * x.w + bias4.w;
                if (full) {
                    if (act == 1) {
                        if (p.aux_out) *reinterpret_cast<float4*>(p.aux_out + off) = x;
                        x.x = gelu_erf(x.x); x.y = gelu_erf(x.y); x.z = gelu_erf(x.z); x.w = gelu_erf(x.w);
                    } else if (PRE && act == 2) {      // activations 2 - 4 are dispatched to the PRE instantiation only
                        const float4 a = PRE ? aux[it] : *reinterpret_cast<const float4*>(p.aux_in + off);
                        x.x *= gelu_erf_grad(a.x); x.y *= gelu_erf_grad(a.y); x.z *= gelu_erf_grad(a.z); x.w *= gelu_erf_grad(a.w);
                    } else if (PRE && act == 3) {
                        float4 gr;
                        x.x = gelu_erf_both(x.x, gr.x); x.y = gelu_erf_both(x.y, gr.y); x.z = gelu_erf_both(x.z, gr.z); x.w = gelu_erf_both(x.w, gr.w);
                        if (p.aux_out) *reinterpret_cast<float4*>(p.aux_out + off) = gr;
                    } else if (PRE && act == 4) {
                        const float4 a = PRE ? aux[it] : *reinterpret_cast<const float4*>(p.aux_in + off);
                        x.x *= a.x; x.y *= a.y; x.z *= a.z; x.w *= a.w;
                    }
                    if (accum) { const float4 o = PRE ? old[it] : *reinterpret_cast<const float4*>(base + off); x.x += o.x; x.y += o.y; x.z += o.z; x.w += o.w; }
                    if (want_cs) { csum.x += x.x; csum.y += x.y; csum.z += x.z; csum.w += x.w; }
                    if (base) *reinterpret_cast<float4*>(base + off) = x;      // C == nullptr: only the split pair leaves
                    if (p.out_hi) {
                        float4 h, l;
                        split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y); split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
                        *reinterpret_cast<float4*>(p.out_hi + off) = h;
                        *reinterpret_cast<float4*>(p.out_lo + off) = l;
                    }
                } else {
                    float xs[4] = {x.x, x.y, x.z, x.w};
                    for (int e = 0; e < 4; ++e) if (col + e < g.N) {
                        float y = xs[e];
                        if (act == 1) { if (p.aux_out) p.aux_out[off + e] = y; y = gelu_erf(y); }
                        else if (PRE && act == 2) y *= gelu_erf_grad(p.aux_in[off + e]);
                        else if (PRE && act == 3) { float gr; y = gelu_erf_both(y, gr); if (p.aux_out) p.aux_out[off + e] = gr; }
                        else if (PRE && act == 4) y *= p.aux_in[off + e];
                        if (accum) y += base[off + e];
                        if (base) base[off + e] = y;
                        if (p.out_hi) { float h, l; split_tf32(y, h, l); p.out_hi[off + e] = h; p.out_lo[off + e] = l; }
                    }
                }
            }
            if (want_cs) {                                       // lanes c4, c4 + 8, c4 + 16, c4 + 24 hold the same 4 columns
#pragma unroll
                for (int o = 8; o <= 16; o <<= 1) {
                    csum.x += __shfl_xor_sync(0xffffffffu, csum.x, o); csum.y += __shfl_xor_sync(0xffffffffu, csum.y, o);
                    csum.z += __shfl_xor_sync(0xffffffffu, csum.z, o); csum.w += __shfl_xor_sync(0xffffffffu, csum.w, o);
                }
                if (sub == 0 && full) *reinterpret_cast<float4*>(p.colpart + ((long long)tc.m * 4 + quarter) * g.N + col) = csum;
            }
            __syncwarp();                                        // staging block is reused by the next chunk
        }
    }
    __device__ void finish() {}
};

// ---- online log-sum-exp over the CTA's strip of item tiles (thread = user row)
// With `logits` set the fp32 logits tile is also written out, row-contiguous through a per-warp staging block (4 rows x
// 128 B per store instruction): the backward pass reads it back instead of recomputing s E^T (ce_bwd_logits.cuh).
template <int BN>
struct EpiLse {
    struct Params {
        const long long* answers; float* part_m; float* part_s; float* tgt;
        int store_logits;                 // 1: also write the tile through `logits_map`
        CUtensorMap logits_map;           // [M, N] fp32 (pitch ld), box {32 cols, 32 rows}, 128-B swizzle
    };
    static constexpr int WARPS = 8;                      // two warps per TMEM lane quarter, each owning half of the tile's columns
    static constexpr int SMEM_BYTES = WARPS * 4096 + 1024;      // one 32 x 32 fp32 box per warp, 1024-B aligned (swizzle atom)
    const Params& p; const GemmCore& g; int rit;
    unsigned char* box; int quarter, half, lane;
    float m = -INFINITY, s = 0.f;
    long long ans = -1;
    __device__ EpiLse(const Params& p_, const GemmCore& g_, unsigned char* smem, int epi_warp, int lane_)
        : p(p_), g(g_), rit((epi_warp & 3) * 32 + lane_), quarter(epi_warp & 3), half(epi_warp >> 2), lane(lane_) {
        box = (unsigned char*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023) + epi_warp * 4096;
        const long long row = (long long)(blockIdx.x % g.m_tiles) * BM + rit;
        if (row < g.M) ans = p.answers[row];
    }
    __device__ __forceinline__ void chunk(float (&v)[32], const TileCoord& tc, int n0, int c0) {
        if (n0 + c0 >= g.N) return;
        if (p.store_logits) {
            // this warp's 32 x 32 block leaves through a TMA store: the row goes into the box as 8 swizzled 16-byte chunks
            // (conflict-free), one lane issues the bulk copy, the LSU never sees the global write and ragged edges are clipped
            if (lane == 0) tma_store_wait_read<0>();             // the previous block's copy has finished reading the box
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(box + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&p.logits_map, box, n0 + c0, tc.m * BM + quarter * 32);
                tma_store_commit();
            }
        }
        float cm = -INFINITY;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (n0 + c0 + j >= g.N) v[j] = -INFINITY;
            cm = fmaxf(cm, v[j]);
        }
        const float mn = fmaxf(m, cm);
        float acc_s = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) acc_s += fast_exp(v[j] - mn);
        s = s * fast_exp(m - mn) + acc_s;      // m == -inf on the first chunk: ex2(-inf) = 0
        m = mn;
        const long long a = ans - (n0 + c0);
        if (a >= 0 && a < 32) {
            float t = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) if (j == (int)a) t = v[j];
            const long long row = (long long)tc.m * BM + rit;
            p.tgt[row] = t;
        }
    }
    __device__ void tile(uint32_t acc, const TileCoord& tc) {
        static_assert(BN == 128, "EpiLse: two 32-column chunks per warp");
        const int n0 = tc.n * BN, ca = half * (BN / 2), cb = ca + 32;
        uint32_t ra[32], rb[32];
        tmem_ld32_issue(acc + ca, ra);                 // both chunks in flight before the single wait
        tmem_ld32_issue(acc + cb, rb);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(ra[j]);
        chunk(v, tc, n0, ca);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(rb[j]);
        chunk(v, tc, n0, cb);
    }
    __device__ void finish() {       // one (max, sum) partial per (strip, column half) and user
        if (p.store_logits && lane == 0) tma_store_wait<0>();        // every store of this warp has landed before the CTA exits
        const int m_tile = blockIdx.x % g.m_tiles, strip = blockIdx.x / g.m_tiles;
        const long long row = (long long)m_tile * BM + rit;
        if (row < g.M) { p.part_m[(long long)(strip * 2 + half) * g.M + row] = m; p.part_s[(long long)(strip * 2 + half) * g.M + row] = s; }
    }
};

// ---- per-user running top-k over the CTA's strip (thread = user row; sorted list in shared memory)
template <int BN>
struct EpiTopk {
    struct Params {
        const long long* indptr; const long long* indices; const long long* hrows; long long offset; int k;
        float* cand_val; long long* cand_idx;      // [nstrips][B][k]
        const float* seed_val;                     // optional [B][seed_ld]: seed_val[row*seed_ld + k-1] is a proven lower
        int seed_ld;                               //   bound of the row's final k-th score (a k-th best over a sample)
    };
    static constexpr int KMAX = 32;
    static constexpr int WARPS = 4;
    static constexpr int SMEM_BYTES = KMAX * 128 * 8;
    const Params& p; const GemmCore& g; int rit;
    float* lv; int* li;
    float thr = -INFINITY; int thr_i = 0x7fffffff;
    long long hlo = 0, hhi = 0, hnext = 0x7fffffffffffffffll;     // cursor into the row's history (sorted global ids)
    __device__ EpiTopk(const Params& p_, const GemmCore& g_, unsigned char* smem, int epi_warp, int lane) : p(p_), g(g_), rit(epi_warp * 32 + lane) {
        lv = reinterpret_cast<float*>(smem);
        li = reinterpret_cast<int*>(smem + KMAX * 128 * 4);
        for (int i = 0; i < p.k; ++i) { lv[i * 128 + rit] = -INFINITY; li[i * 128 + rit] = 0x7fffffff; }
        const long long row = (long long)(blockIdx.x % g.m_tiles) * BM + rit;
        if (row < g.M && p.indptr) {
            const long long hr = p.hrows ? p.hrows[row] : row;      // CSR row of this batch row (hist_rows indirection)
            hlo = p.indptr[hr]; hhi = p.indptr[hr + 1]; if (hlo < hhi) hnext = p.indices[hlo];
        }
        // seeded threshold: (thr, INT_MAX) admits every score >= thr, so the sample's own k-th item re-enters
        if (row < g.M && p.seed_val) thr = p.seed_val[row * p.seed_ld + p.k - 1];
    }
    __device__ __forceinline__ bool before(float va, int ia, float vb, int ib) const { return va > vb || (va == vb && ia < ib); }
    __device__ void insert(float v, int id) {
        // history items are scored 0.0 elsewhere: skip them here.  Candidate ids arrive in ascending order within a
        // strip, so a cursor over the row's sorted history replaces a binary search (one cached load per step).
        const long long gid = p.offset + id;
        while (hnext < gid) {
            ++hlo;
            hnext = hlo < hhi ? p.indices[hlo] : 0x7fffffffffffffffll;
        }
        if (hnext == gid) return;
        int q = p.k - 1;
        while (q > 0 && before(v, id, lv[(q - 1) * 128 + rit], li[(q - 1) * 128 + rit])) {
            lv[q * 128 + rit] = lv[(q - 1) * 128 + rit];
            li[q * 128 + rit] = li[(q - 1) * 128 + rit];
            --q;
        }
        lv[q * 128 + rit] = v; li[q * 128 + rit] = id;
        const float l20 = lv[(p.k - 1) * 128 + rit];
        if (!(l20 < thr)) { thr = l20; thr_i = li[(p.k - 1) * 128 + rit]; }      // a seeded threshold stays until the list beats it
    }
    __device__ void tile(uint32_t acc, const TileCoord& tc) {
        const int n0 = tc.n * BN;
        const bool row_ok = (long long)tc.m * BM + rit < g.M;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(acc + c0, v);
            if (!row_ok || n0 + c0 >= g.N) continue;
            // once the list is warm almost no chunk holds a candidate: one max over the chunk decides
            float cm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 32; ++j) cm = fmaxf(cm, (n0 + c0 + j < g.N) ? v[j] : -INFINITY);
            if (cm < thr) continue;                     // ties with thr still take the exact path below
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int id = n0 + c0 + j;
                if (id < g.N && before(v[j], id, thr, thr_i)) insert(v[j], id);
            }
        }
    }
    __device__ void finish() {
        const int m_tile = blockIdx.x % g.m_tiles, strip = blockIdx.x / g.m_tiles;
        const long long row = (long long)m_tile * BM + rit;
        if (row >= g.M) return;
        for (int i = 0; i < p.k; ++i) {
            const long long o = ((long long)strip * g.M + row) * p.k + i;
            const int id = li[i * 128 + rit];
            p.cand_val[o] = lv[i * 128 + rit];
            p.cand_idx[o] = id == 0x7fffffff ? -1 : p.offset + id;
        }
    }
};

// ---- CE gradient of one logits chunk: G = (softmax - onehot) * d_loss / B, emitted as a split-TF32 (hi, lo) pair
//      [B][ldg] (items contiguous).  The same arrays feed both gradient GEMMs: ds = G E reads them as a K-major A
//      operand, dE = G^T s as an MN-major A operand.  Same staged, coalesced store scheme as EpiStore.
template <int BN>
struct EpiCeGrad {
    struct Params {
        const float* lse; const long long* answers; const float* d_loss; long long n_base;
        float *g_hi, *g_lo; long long ldg;          // [B][ldg]   (cols = chunk items), ldg % 4 == 0
    };
    static constexpr int WARPS = 8;
    static constexpr int SMEM_BYTES = WARPS * 32 * 32 * 4;
    const Params& p; const GemmCore& g;
    float4* stg; int quarter, half, lane;
    __device__ EpiCeGrad(const Params& p_, const GemmCore& g_, unsigned char* smem, int epi_warp, int lane_)
        : p(p_), g(g_), quarter(epi_warp & 3), half(epi_warp >> 2), lane(lane_) {
        stg = reinterpret_cast<float4*>(smem) + epi_warp * 256;
    }
    __device__ void tile(uint32_t acc, const TileCoord& tc) {
        const long long row0 = (long long)tc.m * BM + quarter * 32;
        const int n0 = tc.n * BN;
        const int sub = lane >> 3, c4 = lane & 7;
        const float scale = __ldg(p.d_loss) / (float)g.M;
        float l[8]; long long a[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const long long row = row0 + it * 4 + sub;
            const bool ok = row < g.M;
            l[it] = ok ? __ldg(p.lse + row) : 0.f;
            a[it] = ok ? __ldg(p.answers + row) - p.n_base : -1;
        }
#pragma unroll 1
        for (int c0 = half * (BN / 2); c0 < (half + 1) * (BN / 2); c0 += 32) {
            float v[32];
            tmem_ld32(acc + c0, v);
            if (n0 + c0 >= g.N || row0 >= g.M) continue;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                stg[lane * 8 + (j ^ (lane & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            const int col = n0 + c0 + c4 * 4;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int r = it * 4 + sub;
                const long long row = row0 + r;
                const float4 x = stg[r * 8 + (c4 ^ (r & 7))];
                if (row >= g.M || col >= g.N) continue;
                float pr[4] = {expf(x.x - l[it]), expf(x.y - l[it]), expf(x.z - l[it]), expf(x.w - l[it])};
                const long long da = a[it] - col;
#pragma unroll
                for (int e = 0; e < 4; ++e) if (da == e) pr[e] -= 1.0f;
                float4 h, lo;
                split_tf32(pr[0] * scale, h.x, lo.x); split_tf32(pr[1] * scale, h.y, lo.y);
                split_tf32(pr[2] * scale, h.z, lo.z); split_tf32(pr[3] * scale, h.w, lo.w);
                const long long off = row * p.ldg + col;
                if (col + 3 < g.N) {
                    *reinterpret_cast<float4*>(p.g_hi + off) = h;
                    *reinterpret_cast<float4*>(p.g_lo + off) = lo;
                } else {
                    const float hs[4] = {h.x, h.y, h.z, h.w}, ls[4] = {lo.x, lo.y, lo.z, lo.w};
                    for (int e = 0; e < 4; ++e) if (col + e < g.N) { p.g_hi[off + e] = hs[e]; p.g_lo[off + e] = ls[e]; }
                }
            }
            __syncwarp();
        }
    }
    __device__ void finish() {}
};

}  // namespace
}  // namespace bsa
#include "ce_grad_ds.cuh"
#include "ce_bwd_logits.cuh"
namespace bsa {
namespace {

// ============================================================================ host side
struct Operand {           // a (hi, lo) pair as stored in global memory
    const float* hi; const float* lo;
    long long rows, cols, pitch;      // row-major [rows, cols]
    bool mn;                          // true: rows index K (contraction), cols index M/N
};

int make_core(GemmCore& g, int M, int N, int K, const Operand& A, const Operand& B, int BN, int sched, int splits) {
    g.M = M; g.N = N; g.K = K;
    g.m_tiles = cdiv(M, BM); g.n_tiles = cdiv(N, BN);
    g.k_blocks_total = cdiv(K, BK);
    if (splits < 1) splits = 1;
    if (splits > g.k_blocks_total) splits = g.k_blocks_total;
    g.k_blocks_per_split = cdiv(g.k_blocks_total, splits);
    g.splits = cdiv(g.k_blocks_total, g.k_blocks_per_split);      // every split owns >= 1 k-block
    g.sched = sched;
    int rc;
    if ((rc = make_tmap_2d(&g.a_hi, A.hi, A.rows, A.cols, A.pitch, A.mn ? BK : BM, A.mn))) return rc;
    if ((rc = make_tmap_2d(&g.a_lo, A.lo, A.rows, A.cols, A.pitch, A.mn ? BK : BM, A.mn))) return rc;
    if ((rc = make_tmap_2d(&g.b_hi, B.hi, B.rows, B.cols, B.pitch, B.mn ? BK : BN, B.mn))) return rc;
    if ((rc = make_tmap_2d(&g.b_lo, B.lo, B.rows, B.cols, B.pitch, B.mn ? BK : BN, B.mn))) return rc;
    return 0;
}

template <int BN, int STAGES, bool A_MN, bool B_MN, class Epi>
int launch(const GemmCore& g, const typename Epi::Params& ep, int grid, cudaStream_t st) {
    constexpr int smem = gemm_smem_bytes<BN, STAGES>(Epi::SMEM_BYTES);
    static_assert(smem <= 227 * 1024, "shared memory budget exceeded");
    static bool configured = false;
    auto* kern = umma_gemm_kernel<BN, STAGES, A_MN, B_MN, Epi>;
    if (!configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = true;
    }
    kern<<<grid, gemm_threads(Epi::WARPS), smem, st>>>(g, ep);
    BSA_LAUNCH_CHECK();
    return 0;
}

int num_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

template <int BN, int STAGES, class Epi>
int launch_major(const GemmCore& g, const typename Epi::Params& ep, bool a_mn, bool b_mn, int grid, cudaStream_t st) {
    if (!a_mn && !b_mn) return launch<BN, STAGES, false, false, Epi>(g, ep, grid, st);
    if (!a_mn && b_mn) return launch<BN, STAGES, false, true, Epi>(g, ep, grid, st);
    if (a_mn && !b_mn) return launch<BN, STAGES, true, false, Epi>(g, ep, grid, st);
    return launch<BN, STAGES, true, true, Epi>(g, ep, grid, st);
}

// seg_rows > 0: rows [s * seg_rows, (s + 1) * seg_rows) of the result are the separate tensor C / C1 / C2 (pitch ldc) — the
// three weight gradients of the q / k / v projections leave one GEMM
__global__ void splitk_reduce_kernel2(const float* __restrict__ part, float* __restrict__ C, long long MN, int N, long long ldc,
                                      int splits, float alpha, int accumulate, int seg_rows, float* __restrict__ C1,
                                      float* __restrict__ C2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= MN) return;
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s += part[(long long)k * MN + i];
    long long m = i / N;
    const long long n = i - m * N;
    float* dst = C;
    if (seg_rows) {
        const int seg = (int)(m / seg_rows);
        m -= (long long)seg * seg_rows;
        dst = seg == 0 ? C : (seg == 1 ? C1 : C2);
    }
    float v = alpha * s;
    if (accumulate) v += dst[m * ldc + n];
    dst[m * ldc + n] = v;
}

int do_split(const float* x, float* hi, float* lo, long long n, cudaStream_t st) {
    if (n == 0) return 0;
    long long blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    split_tf32_kernel<<<(int)blocks, 256, 0, st>>>(x, hi, lo, n);
    BSA_LAUNCH_CHECK();
    return 0;
}

int do_split_lo(const float* x, float* lo, long long n, cudaStream_t st) {
    if (n == 0) return 0;
    long long blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    split_lo_trunc_kernel<<<(int)blocks, 256, 0, st>>>(x, lo, n);
    BSA_LAUNCH_CHECK();
    return 0;
}
bool e_trunc() {
    static const bool on = [] { const char* e = getenv("BSA_E_TRUNC"); return !(e && atoi(e) == 0); }();
    return on;
}

inline size_t al(size_t x) { return (x + 1023) / 1024 * 1024; }

}  // namespace

// ============================================================================ generic GEMM (plain store epilogue)
// C[M,N] = alpha * A B^T (+bias)(act)(+C).  Operands are given UNSPLIT; (hi, lo) pairs are produced into `ws`.
size_t umma_gemm_workspace_bytes(long long a_elems, long long b_elems, int M, int N, int K) {
    const int splits_max = 64;
    return 2 * al((size_t)a_elems * 4) + 2 * al((size_t)b_elems * 4) + al((size_t)splits_max * M * N * 4) + 4096;
}

int umma_split(const float* x, float* hi, float* lo, long long n, cudaStream_t st) { return do_split(x, hi, lo, n, st); }

int umma_split_lo(const float* x, float* lo, long long n, cudaStream_t st) { return do_split_lo(x, lo, n, st); }

int umma_split_multi(int count, const float* const* x, float* const* hi, float* const* lo, const long long* n, cudaStream_t st) {
    BSA_REQUIRE(count >= 0 && count <= 8, "umma_split_multi: at most 8 tensors");
    if (count == 0) return 0;
    SplitMultiP t;
    long long nmax = 0;
    for (int i = 0; i < 8; ++i) {
        const int j = i < count ? i : 0;
        t.x[i] = x[j]; t.hi[i] = hi[j]; t.lo[i] = lo[j]; t.n[i] = i < count ? n[j] : 0;
        if (i < count) {
            BSA_REQUIRE(((reinterpret_cast<uintptr_t>(x[j]) | reinterpret_cast<uintptr_t>(hi[j]) | reinterpret_cast<uintptr_t>(lo[j])) & 15) == 0,
                        "umma_split_multi: tensors must be 16-byte aligned");
            if (n[j] > nmax) nmax = n[j];
        }
    }
    if (nmax == 0) return 0;
    long long bx = (nmax / 4 + 255) / 256;
    if (bx > 128) bx = 128;
    if (bx < 1) bx = 1;
    split_tf32_multi_kernel<<<dim3((unsigned)bx, (unsigned)count), 256, 0, st>>>(t);
    BSA_LAUNCH_CHECK();
    return 0;
}

bool umma_split_colsum_ok(int N, const float* x, const float* hi, const float* lo) {
    const int tpr = N >> 2;
    return (N & 3) == 0 && tpr >= 1 && tpr <= 256 && 256 % tpr == 0 &&
           ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(hi) | reinterpret_cast<uintptr_t>(lo)) & 15) == 0;
}
int umma_split_colsum_rows(long long M) { return (int)((M + SPLIT_CS_ROWS - 1) / SPLIT_CS_ROWS); }
// part: [umma_split_colsum_rows(M), N]
int umma_split_colsum(const float* x, float* hi, float* lo, long long M, int N, float* part, cudaStream_t st, long long out_pitch,
                      long long in_pitch) {
    BSA_REQUIRE(umma_split_colsum_ok(N, x, hi, lo), "umma_split_colsum: unsupported width %d or alignment", N);
    if (out_pitch == 0) out_pitch = N;
    if (in_pitch == 0) in_pitch = N;
    BSA_REQUIRE(out_pitch >= N && (out_pitch & 3) == 0 && in_pitch >= N && (in_pitch & 3) == 0, "umma_split_colsum: bad pitch");
    if (M == 0) return 0;
    split_colsum_kernel<<<umma_split_colsum_rows(M), 256, 0, st>>>(x, hi, lo, M, N, part, out_pitch >> 2, in_pitch >> 2);
    BSA_LAUNCH_CHECK();
    return 0;
}

size_t umma_gemm_partial_bytes(int M, int N) { return al((size_t)64 * M * N * 4) + 1024; }

// K-splits a plain-epilogue umma_gemm_ps launch uses when it is given split-K scratch (1 = none)
int umma_gemm_splits(int M, int N, int K) {
    const int BN = N > 128 ? 256 : 128;
    const int tiles = cdiv(M, BM) * cdiv(N, BN);
    if (tiles >= num_sms()) return 1;
    int splits = num_sms() / tiles;
    const int kb = cdiv(K, BK);
    if (splits > kb / 4) splits = kb / 4;
    if (splits > 64) splits = 64;
    return splits < 1 ? 1 : splits;
}

// Operands already split: a = (a_hi, a_lo), b = (b_hi, b_lo).  `partial` backs split-K (may be null: no split-K).
int umma_gemm_ps(int M, int N, int K, const float* a_hi, const float* a_lo, long long a_rows, long long a_cols,
                 long long lda, bool a_mn, const float* b_hi, const float* b_lo, long long b_rows, long long b_cols,
                 long long ldb, bool b_mn, float* C, long long ldc, const float* bias, int act, float* aux_out,
                 const float* aux_in, float alpha, int accumulate, float* out_hi, float* out_lo, float* partial,
                 size_t partial_bytes, cudaStream_t st, float* colpart, int seg_cols, long long seg_stride, const float* bias1,
                 const float* bias2, int seg_rows, float* C1, float* C2) {
    BSA_REQUIRE(M > 0 && N > 0 && K > 0, "umma_gemm: empty problem");
    BSA_REQUIRE(!seg_rows || (umma_gemm_splits(M, N, K) > 1 && partial && C && C1 && C2 && !bias && act == 0 && !out_hi && !colpart &&
                              M % seg_rows == 0 && M / seg_rows == 3),
                "umma_gemm: row-segmented output needs the split-K path (3 segments, plain epilogue)");
    if (seg_cols) {
        BSA_REQUIRE(C && act == 0 && !accumulate && !colpart && !out_hi && seg_cols % 32 == 0 && N % seg_cols == 0 &&
                    N / seg_cols <= 3, "umma_gemm: segmented output needs a plain epilogue and segments of a multiple of 32 columns");
        partial = nullptr;
    }
    BSA_REQUIRE(!colpart || ((N & 3) == 0 && (ldc & 3) == 0), "umma_gemm: fused column sums need N %% 4 == 0");
    if (colpart) partial = nullptr;                   // the column sums come from complete tiles
    BSA_REQUIRE(C || (out_hi && out_lo && !accumulate), "umma_gemm: C may be omitted only when the split pair is emitted");
    if (!C) partial = nullptr;                        // no split-K without an fp32 destination
    int rc;
    Operand oa{a_hi, a_lo, a_rows, a_cols, lda, a_mn}, ob{b_hi, b_lo, b_rows, b_cols, ldb, b_mn};
    const int BN = N > 128 ? 256 : 128;
    int splits = 1;
    if (!bias && act == 0 && !out_hi && partial) splits = umma_gemm_splits(M, N, K);
    GemmCore g;
    if ((rc = make_core(g, M, N, K, oa, ob, BN, SCHED_N_FASTEST, splits))) return rc;
    StoreParams ep{C, ldc, bias, act, aux_out, aux_in, alpha, accumulate, nullptr, out_hi, out_lo, 0, colpart};
    ep.seg_cols = seg_cols; ep.seg_stride = seg_stride; ep.bias1 = bias1; ep.bias2 = bias2;
    BSA_REQUIRE(!seg_rows || g.splits > 1, "umma_gemm: row-segmented output needs more than one K-split");
    if (g.splits > 1) {
        BSA_REQUIRE((size_t)g.splits * M * N * 4 <= partial_bytes, "umma_gemm: workspace too small for split-K");
        ep.partial = partial;
    }
    const long long total = (long long)g.m_tiles * g.n_tiles * g.splits;
    const int grid = (int)(total < num_sms() ? total : num_sms());
    {
        ProfScope prof(PROF_LINEAR_UMMA, st, 2.0 * M * N * (double)K);
        const bool pre = act >= 2 || accumulate || colpart;
        if (BN == 256 && !a_mn && !b_mn && act == 3 && !accumulate && !colpart && !seg_cols)          // FFN d -> 4d, GELU
            rc = launch<256, 2, false, false, EpiStore<256, true, false, 3>>(g, ep, grid, st);
        else if (BN == 256 && !a_mn && !b_mn && act == 1 && !accumulate && !colpart && !seg_cols)     // the same in inference mode
            rc = launch<256, 2, false, false, EpiStore<256, false, false, 1>>(g, ep, grid, st);
        else if (BN == 256 && !a_mn && b_mn && act == 4 && !accumulate && colpart)                   // FFN dX 4d <- d, x GELU'
            rc = launch<256, 2, false, true, EpiStore<256, true, true, 4>>(g, ep, grid, st);
        else if (BN == 256) rc = colpart ? launch_major<256, 2, EpiStore<256, true, true>>(g, ep, a_mn, b_mn, grid, st)
                            : pre   ? launch_major<256, 2, EpiStore<256, true>>(g, ep, a_mn, b_mn, grid, st)
                                    : launch_major<256, 2, EpiStore<256>>(g, ep, a_mn, b_mn, grid, st);
        else rc = colpart ? launch_major<128, 3, EpiStore<128, true, true>>(g, ep, a_mn, b_mn, grid, st)
                  : pre   ? launch_major<128, 3, EpiStore<128, true>>(g, ep, a_mn, b_mn, grid, st)
                          : launch_major<128, 3, EpiStore<128>>(g, ep, a_mn, b_mn, grid, st);
    }
    if (rc) return rc;
    if (g.splits > 1) {
        const long long MN = (long long)M * N;
        splitk_reduce_kernel2<<<cdiv(MN, 256), 256, 0, st>>>(partial, C, MN, N, ldc, g.splits, alpha, accumulate, seg_rows, C1, C2);
        BSA_LAUNCH_CHECK();
    }
    return 0;
}

int umma_gemm(int M, int N, int K, const float* A, long long a_rows, long long a_cols, long long lda, bool a_mn,
              const float* B, long long b_rows, long long b_cols, long long ldb, bool b_mn, float* C, long long ldc,
              const float* bias, int act, float* aux_out, const float* aux_in, float alpha, int accumulate, void* ws,
              size_t ws_bytes, cudaStream_t st) {
    BSA_REQUIRE(M > 0 && N > 0 && K > 0, "umma_gemm: empty problem");
    char* p = (char*)ws;
    const size_t ab = al((size_t)a_rows * lda * 4), bb = al((size_t)b_rows * ldb * 4);
    float* a_hi = (float*)p; p += ab;
    float* a_lo = (float*)p; p += ab;
    float* b_hi = (float*)p; p += bb;
    float* b_lo = (float*)p; p += bb;
    float* partial = (float*)p;
    BSA_REQUIRE((size_t)(p - (char*)ws) <= ws_bytes, "umma_gemm: workspace too small");
    int rc;
    if ((rc = do_split(A, a_hi, a_lo, a_rows * lda, st))) return rc;
    if ((rc = do_split(B, b_hi, b_lo, b_rows * ldb, st))) return rc;
    return umma_gemm_ps(M, N, K, a_hi, a_lo, a_rows, a_cols, lda, a_mn, b_hi, b_lo, b_rows, b_cols, ldb, b_mn, C, ldc, bias,
                        act, aux_out, aux_in, alpha, accumulate, nullptr, nullptr, partial, ws_bytes - (size_t)(p - (char*)ws), st);
}

// ============================================================================ cross-entropy
namespace {
constexpr int CE_BN = 128;
constexpr int CE_STAGES = 3;

struct CeWs {
    float *s_hi, *s_lo, *e_hi, *e_lo, *part_m, *part_s, *tgt;
    float *logits, *lse_pad;      // stored logits S [B, ld] (forward -> backward) and lse padded to a multiple of 128; null if not eligible
    long long ld;
    float *g_hi, *g_lo, *partial;
    int nstrips, chunk;
    size_t bytes;
};

// the logits-storing CE (forward writes S, backward reads it: ce_bwd_logits.cuh) needs the 256-column accumulator shape
bool ce_logits_eligible(int B, int d, long long N) {
    return d > 128 && d <= CBL_ND && (d & 3) == 0 && N < (1ll << 31) - 256 && B >= 1;
}

int ce_strips(int B) {
    const int m_tiles = cdiv(B, BM);
    int s = num_sms() / m_tiles;
    return s < 1 ? 1 : s;
}

int ce_bwd_chunk(int B, long long N) {
    // one wave of 128 x 256 logits tiles per chunk, and small enough that G, G^T (hi, lo) stay in L2
    const int m_tiles = cdiv(B, BM);
    long long per_wave = (long long)(num_sms() / m_tiles > 0 ? num_sms() / m_tiles : 1) * 256;
    static long long g_mib = [] { const char* e = getenv("BSA_CE_G_MIB"); long long v = e ? atoll(e) : 0; return v > 0 ? v : 640ll; }();
    long long cap = (g_mib << 20) / (8ll * (B > 0 ? B : 1));         // G (hi, lo): 8 B per element
    long long c = per_wave;
    while (c * 2 <= cap) c *= 2;
    if (c > cap && cap >= 256) c = cap / 256 * 256;
    if (c > N) c = (N + 31) / 32 * 32;
    if (c < 32) c = 32;
    return (int)c;
}

CeWs carve_ce(void* ws, int B, int d, long long N, bool bwd) {
    CeWs w;
    char* p = (char*)ws;
    auto take = [&](size_t bytes) { char* r = p; p += al(bytes); return (float*)r; };
    w.s_hi = take((size_t)B * d * 4); w.s_lo = take((size_t)B * d * 4);
    w.e_hi = take((size_t)N * d * 4); w.e_lo = take((size_t)N * d * 4);
    w.nstrips = ce_strips(B);
    w.part_m = take((size_t)2 * w.nstrips * B * 4); w.part_s = take((size_t)2 * w.nstrips * B * 4);      // (strip, column half) partials
    w.tgt = take((size_t)B * 4);
    w.logits = w.lse_pad = nullptr;
    w.ld = (N + 31) / 32 * 32;
    if (ce_logits_eligible(B, d, N)) {
        w.logits = take((size_t)B * w.ld * 4);
        w.lse_pad = take((size_t)((B + 127) / 128 * 128) * 4);
    }
    w.chunk = ce_bwd_chunk(B, N);
    if (bwd) {
        w.g_hi = take((size_t)B * w.chunk * 4); w.g_lo = take((size_t)B * w.chunk * 4);
        w.partial = take((size_t)64 * B * d * 4);
    }
    w.bytes = (size_t)(p - (char*)ws) + 1024;
    return w;
}

// Row r's log-sum-exp from the per-strip (max, sum) partials: one WARP per row, lanes over the strips (the first version
// ran one thread per row in a single block: 2 x 74 dependent strided loads per thread, 43 us warm / 81 us cold at C3).
__global__ void __launch_bounds__(256)
ce_combine_kernel(const float* __restrict__ part_m, const float* __restrict__ part_s, int nstrips, int B, float* __restrict__ lse) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= B) return;
    float m = -INFINITY;
    for (int s = lane; s < nstrips; s += 32) m = fmaxf(m, part_m[(long long)s * B + r]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float sum = 0.f;
    for (int s = lane; s < nstrips; s += 32) {
        const float ps = part_s[(long long)s * B + r];
        if (ps > 0.f) sum += ps * expf(part_m[(long long)s * B + r] - m);
    }
    sum = warp_sum(sum);                                   // xor butterfly: every lane holds the same bits
    if (lane == 0) lse[r] = m + logf(sum);
}
// loss = mean_r (lse[r] - tgt[r]) in a fixed order (one block)
__global__ void ce_loss_kernel(const float* __restrict__ lse, const float* __restrict__ tgt, int B, float* __restrict__ loss) {
    __shared__ float sh[8];
    float acc = 0.f;
    for (int r = threadIdx.x; r < B; r += blockDim.x) acc += lse[r] - tgt[r];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += sh[i];
        *loss = t / (float)B;
    }
}
}  // namespace

// 2: forward stores the logits, backward = two A-from-TMEM kernels over them (ce_bwd_logits.cuh), where d allows and the
//    backward receives the forward's workspace; 1: logits recomputed, gradient tile + ds in one kernel (ce_grad_ds.cuh);
// 0: logits recomputed, separate GEMMs
int g_ce_fused = 2;
void set_ce_fused(int mode) { g_ce_fused = mode < 0 ? 0 : (mode > 2 ? 2 : mode); }
int get_ce_fused() { return g_ce_fused; }

namespace {
// ds[u,:] -= scale * E[ans[u],:]  — the "- onehot" part of (softmax - onehot) for the user-side gradient (one block per user)
__global__ void ce_onehot_fix_ds_kernel(float* __restrict__ ds, const float* __restrict__ E, const long long* __restrict__ answers,
                                        const float* __restrict__ d_loss, int B, int d, long long N) {
    const int u = blockIdx.x;
    const long long a = answers[u];
    if (a < 0 || a >= N) return;
    const float scale = __ldg(d_loss) / (float)B;
    for (int c = threadIdx.x; c < d; c += blockDim.x) ds[(long long)u * d + c] -= scale * E[a * d + c];
}
// dE[ans[u],:] -= scale * s[u,:]  — the item-side part.  Users that share an answer are handled by the block of the FIRST
// such user, in ascending user order, so the result does not depend on scheduling (no atomics).
__global__ void ce_onehot_fix_de_kernel(float* __restrict__ dE, const float* __restrict__ s, const long long* __restrict__ answers,
                                        const float* __restrict__ d_loss, int B, int d, long long N) {
    const int u = blockIdx.x;
    const long long a = answers[u];
    if (a < 0 || a >= N) return;
    int earlier = 0;
    for (int v = threadIdx.x; v < u; v += blockDim.x) earlier |= (answers[v] == a);
    if (__syncthreads_or(earlier)) return;
    const float scale = __ldg(d_loss) / (float)B;
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
        float acc = 0.f;
        for (int v = u; v < B; ++v) if (answers[v] == a) acc += s[(long long)v * d + c];
        dE[a * d + c] -= scale * acc;
    }
}
__global__ void pad_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, int n, int n_pad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_pad) dst[i] = i < n ? src[i] : 0.f;
}
}  // namespace

bool umma_ce_supported(int B, int d, long long N) { return B >= 1 && d % 4 == 0 && d >= 4 && N >= 1; }

size_t umma_ce_workspace_bytes(int B, int d, long long N) { return carve_ce(nullptr, B, d, N, true).bytes; }

int umma_ce_fwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, float* loss,
                float* lse, void* ws, size_t ws_bytes, cudaStream_t st, float* tgt_out) {
    CeWs w = carve_ce(ws, B, d, N, false);
    BSA_REQUIRE(ws_bytes >= w.bytes, "bsa_ce_fwd: workspace too small (%zu < %zu)", ws_bytes, w.bytes);
    int rc;
    if (tgt_out) {        // one row shard of a vocabulary-parallel CE: targets outside the shard (answer < 0) contribute 0
        BSA_CHECK_CUDA(cudaMemsetAsync(w.tgt, 0, (size_t)B * 4, st));
        if (!loss) loss = (float*)((char*)ws + w.bytes - 512);
    }
    if ((rc = do_split(s, w.s_hi, w.s_lo, (long long)B * d, st))) return rc;
    if (e_trunc()) {
        w.e_hi = const_cast<float*>(E);
        if ((rc = do_split_lo(E, w.e_lo, N * d, st))) return rc;
    } else if ((rc = do_split(E, w.e_hi, w.e_lo, N * d, st))) return rc;
    Operand oa{w.s_hi, w.s_lo, B, d, d, false}, ob{w.e_hi, w.e_lo, N, d, d, false};
    GemmCore g;
    if ((rc = make_core(g, B, (int)N, d, oa, ob, CE_BN, SCHED_STRIP, 1))) return rc;
    int nstrips = w.nstrips;
    if (nstrips > g.n_tiles) nstrips = g.n_tiles;
    EpiLse<CE_BN>::Params ep;
    ep.answers = answers; ep.part_m = w.part_m; ep.part_s = w.part_s; ep.tgt = w.tgt;
    ep.store_logits = (g_ce_fused == 2 && w.logits) ? 1 : 0;
    if (ep.store_logits) { if ((rc = make_tmap_2d(&ep.logits_map, w.logits, B, N, w.ld, 32, false))) return rc; }
    else memset(&ep.logits_map, 0, sizeof(ep.logits_map));
    {
        ProfScope prof(PROF_CE_FWD_UMMA, st, 2.0 * B * (double)N * d);
        if ((rc = launch<CE_BN, CE_STAGES, false, false, EpiLse<CE_BN>>(g, ep, nstrips * g.m_tiles, st))) return rc;
    }
    ce_combine_kernel<<<cdiv(B, 8), 256, 0, st>>>(w.part_m, w.part_s, 2 * nstrips, B, lse);
    BSA_LAUNCH_CHECK();
    ce_loss_kernel<<<1, 256, 0, st>>>(lse, w.tgt, B, loss);
    BSA_LAUNCH_CHECK();
    if (tgt_out) BSA_CHECK_CUDA(cudaMemcpyAsync(tgt_out, w.tgt, (size_t)B * 4, cudaMemcpyDeviceToDevice, st));
    return 0;
}

int umma_ce_bwd(int B, int d, long long N, const float* s, const float* E, const long long* answers, const float* lse,
                const float* d_loss, float* ds, float* dE, void* ws, size_t ws_bytes, bool reuse_split, cudaStream_t st) {
    CeWs w = carve_ce(ws, B, d, N, true);
    BSA_REQUIRE(ws_bytes >= w.bytes, "bsa_ce_bwd: workspace too small (%zu < %zu)", ws_bytes, w.bytes);
    int rc;
    if (e_trunc()) w.e_hi = const_cast<float*>(E);
    if (!reuse_split) {       // else: left there by umma_ce_fwd (same carve order for the forward-only part)
        if ((rc = do_split(s, w.s_hi, w.s_lo, (long long)B * d, st))) return rc;
        if (e_trunc()) { if ((rc = do_split_lo(E, w.e_lo, N * d, st))) return rc; }
        else if ((rc = do_split(E, w.e_hi, w.e_lo, N * d, st))) return rc;
    }
    ProfScope prof(PROF_CE_BWD_UMMA, st, 4.0 * B * (double)N * d);
    if (g_ce_fused == 2 && reuse_split && w.logits) {
        // ---- logits left by the forward pass: ds and dE each by one A-from-TMEM kernel, then the onehot rows
        static bool configured = false;
        if (!configured) {
            BSA_CHECK_CUDA(cudaFuncSetAttribute(ce_bwd_logits_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CBL_SMEM));
            BSA_CHECK_CUDA(cudaFuncSetAttribute(ce_bwd_logits_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CBL_SMEM));
            configured = true;
        }
        const int b_pad = (B + 127) / 128 * 128;
        pad_copy_kernel<<<cdiv(b_pad, 256), 256, 0, st>>>(lse, w.lse_pad, B, b_pad);
        BSA_LAUNCH_CHECK();
        const int m_tiles = cdiv(B, BM), n_tiles = (int)((N + BM - 1) / BM);
        int nstrips = ce_strips(B);
        if (nstrips > 64) nstrips = 64;
        if (nstrips > n_tiles) nstrips = n_tiles;
        CeBwdLogitsParams up;        // user owner: ds partial slabs
        if ((rc = make_tmap_2d(&up.s_map, w.logits, B, N, w.ld, BM, false))) return rc;
        if ((rc = make_tmap_2d(&up.b_hi, w.e_hi, N, d, d, BK, true))) return rc;
        if ((rc = make_tmap_2d(&up.b_lo, w.e_lo, N, d, d, BK, true))) return rc;
        up.B = B; up.N = (int)N; up.d = d; up.owner_tiles = m_tiles; up.n_tiles = n_tiles; up.nstrips = nstrips;
        up.lse = w.lse_pad; up.d_loss = d_loss; up.out = w.partial;
        ce_bwd_logits_kernel<false><<<nstrips * m_tiles, CBL_THREADS, CBL_SMEM, st>>>(up);
        BSA_LAUNCH_CHECK();
        CeBwdLogitsParams ip;        // item owner: dE tiles
        if ((rc = make_tmap_2d(&ip.s_map, w.logits, B, N, w.ld, CBL_SUB, false))) return rc;
        if ((rc = make_tmap_2d(&ip.b_hi, w.s_hi, B, d, d, BK, true))) return rc;
        if ((rc = make_tmap_2d(&ip.b_lo, w.s_lo, B, d, d, BK, true))) return rc;
        ip.B = B; ip.N = (int)N; ip.d = d; ip.owner_tiles = n_tiles; ip.n_tiles = n_tiles; ip.nstrips = 1;
        ip.lse = w.lse_pad; ip.d_loss = d_loss; ip.out = dE;
        ce_bwd_logits_kernel<true><<<n_tiles < num_sms() ? n_tiles : num_sms(), CBL_THREADS, CBL_SMEM, st>>>(ip);
        BSA_LAUNCH_CHECK();
        const long long MN = (long long)B * d;
        splitk_reduce_kernel2<<<cdiv(MN, 256), 256, 0, st>>>(w.partial, ds, MN, d, d, nstrips, 1.f, 0, 0, nullptr, nullptr);
        BSA_LAUNCH_CHECK();
        ce_onehot_fix_ds_kernel<<<B, 256, 0, st>>>(ds, E, answers, d_loss, B, d, N);
        BSA_LAUNCH_CHECK();
        ce_onehot_fix_de_kernel<<<B, 256, 0, st>>>(dE, s, answers, d_loss, B, d, N);
        BSA_LAUNCH_CHECK();
        return 0;
    }
    const int BNd = d > 128 ? 256 : 128;
    // ds = G E: partial sums stay in `partial` ACROSS chunks and are reduced once at the end
    int ds_splits = num_sms() / (cdiv(B, BM) * cdiv(d, BNd));
    if (ds_splits > 64) ds_splits = 64;
    if (ds_splits < 1) ds_splits = 1;
    const bool fused = g_ce_fused >= 1 && d > 128 && d <= CGD_ND && (d & 3) == 0;     // gradient tile + ds in one kernel
    int used_splits = 0;
    bool first = true;
    if (fused) {
        static bool configured = false;
        if (!configured) {
            BSA_CHECK_CUDA(cudaFuncSetAttribute(ce_grad_ds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CGD_SMEM));
            configured = true;
        }
        used_splits = ce_strips(B);
        if (used_splits > 64) used_splits = 64;
        BSA_CHECK_CUDA(cudaMemsetAsync(w.partial, 0, (size_t)used_splits * B * d * 4, st));
    }
    for (long long n0 = 0; n0 < N; n0 += w.chunk) {
        const int cols = (int)((N - n0) < w.chunk ? (N - n0) : w.chunk);
        if (fused) {
            CeGradDsParams fp;
            if ((rc = make_tmap_2d(&fp.s_hi, w.s_hi, B, d, d, BM, false))) return rc;
            if ((rc = make_tmap_2d(&fp.s_lo, w.s_lo, B, d, d, BM, false))) return rc;
            if ((rc = make_tmap_2d(&fp.e_hi, w.e_hi + n0 * d, cols, d, d, CGD_BN, false))) return rc;
            if ((rc = make_tmap_2d(&fp.e_lo, w.e_lo + n0 * d, cols, d, d, CGD_BN, false))) return rc;
            if ((rc = make_tmap_2d(&fp.em_hi, w.e_hi + n0 * d, cols, d, d, BK, true))) return rc;
            if ((rc = make_tmap_2d(&fp.em_lo, w.e_lo + n0 * d, cols, d, d, BK, true))) return rc;
            fp.B = B; fp.cols = cols; fp.d = d;
            fp.m_tiles = cdiv(B, BM); fp.n_tiles = cdiv(cols, CGD_BN);
            fp.nstrips = used_splits < fp.n_tiles ? used_splits : fp.n_tiles;
            fp.k1_blocks = cdiv(d, BK);
            fp.lse = lse; fp.answers = answers; fp.d_loss = d_loss; fp.n_base = n0;
            fp.g_hi = w.g_hi; fp.g_lo = w.g_lo; fp.ldg = w.chunk;
            fp.partial = w.partial;
            ce_grad_ds_kernel<<<fp.nstrips * fp.m_tiles, CGD_THREADS, CGD_SMEM, st>>>(fp);
            BSA_LAUNCH_CHECK();
        } else {
        {   // logits chunk -> G (hi, lo)
            Operand oa{w.s_hi, w.s_lo, B, d, d, false}, ob{w.e_hi + n0 * d, w.e_lo + n0 * d, cols, d, d, false};
            GemmCore g;
            if ((rc = make_core(g, B, cols, d, oa, ob, 256, SCHED_M_FASTEST, 1))) return rc;
            EpiCeGrad<256>::Params ep{lse, answers, d_loss, n0, w.g_hi, w.g_lo, (long long)w.chunk};
            const long long total = (long long)g.m_tiles * g.n_tiles;
            if ((rc = launch<256, 2, false, false, EpiCeGrad<256>>(g, ep, (int)(total < num_sms() ? total : num_sms()), st))) return rc;
        }
        {   // ds partials (+)= G E_chunk : A = G [B, cols] K-major ; B operand = E_chunk [cols (K), d (N)] MN-major
            Operand oa{w.g_hi, w.g_lo, B, cols, (long long)w.chunk, false}, ob{w.e_hi + n0 * d, w.e_lo + n0 * d, cols, d, d, true};
            GemmCore g;
            if ((rc = make_core(g, B, d, cols, oa, ob, BNd, SCHED_N_FASTEST, ds_splits))) return rc;
            // a short last chunk may use fewer splits: its tiles still add into the first g.splits partial slabs
            if (g.splits > used_splits) {
                if (!first) {      // slabs never written before must start from zero
                    BSA_CHECK_CUDA(cudaMemsetAsync(w.partial + (size_t)used_splits * B * d, 0, (size_t)(g.splits - used_splits) * B * d * 4, st));
                }
                used_splits = g.splits;
            }
            const long long total = (long long)g.m_tiles * g.n_tiles * g.splits;
            const int grid = (int)(total < num_sms() ? total : num_sms());
            StoreParams ep{ds, d, nullptr, 0, nullptr, nullptr, 1.f, 0, w.partial, nullptr, nullptr, first ? 0 : 1};
            if (BNd == 256) { if ((rc = launch<256, 2, false, true, EpiStore<256>>(g, ep, grid, st))) return rc; }
            else            { if ((rc = launch<128, 3, false, true, EpiStore<128>>(g, ep, grid, st))) return rc; }
        }
        }
        {   // dE_chunk = G^T s : A = G [B (K), cols (M)] MN-major ; B operand = s [B (K), d (N)] MN-major
            Operand oa{w.g_hi, w.g_lo, B, cols, (long long)w.chunk, true}, ob{w.s_hi, w.s_lo, B, d, d, true};
            GemmCore g;
            if ((rc = make_core(g, cols, d, B, oa, ob, BNd, SCHED_N_FASTEST, 1))) return rc;
            StoreParams ep{dE + n0 * d, d, nullptr, 0, nullptr, nullptr, 1.f, 0, nullptr};
            const long long total = (long long)g.m_tiles * g.n_tiles;
            const int grid = (int)(total < num_sms() ? total : num_sms());
            if (BNd == 256) { if ((rc = launch<256, 2, true, true, EpiStore<256>>(g, ep, grid, st))) return rc; }
            else            { if ((rc = launch<128, 3, true, true, EpiStore<128>>(g, ep, grid, st))) return rc; }
        }
        first = false;
    }
    {
        const long long MN = (long long)B * d;
        splitk_reduce_kernel2<<<cdiv(MN, 256), 256, 0, st>>>(w.partial, ds, MN, d, d, used_splits, 1.f, 0, 0, nullptr, nullptr);
        BSA_LAUNCH_CHECK();
    }
    return 0;
}

// ============================================================================ scoring + top-k
namespace {
__global__ void hist_list_kernel(const long long* __restrict__ indptr, const long long* __restrict__ indices,
                                 const long long* __restrict__ hrows, long long offset, long long rows, int k, int B, float* __restrict__ val,
                                 long long* __restrict__ idx) {
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= B) return;
    int n = 0;
    const long long hr = hrows ? hrows[row] : row;
    const long long lo = indptr ? indptr[hr] : 0, hi = indptr ? indptr[hr + 1] : 0;
    for (long long q = lo; q < hi && n < k; ++q) {
        const long long gid = indices[q];
        if (gid >= offset && gid < offset + rows) { val[(long long)row * k + n] = 0.0f; idx[(long long)row * k + n] = gid; ++n; }
    }
    for (; n < k; ++n) { val[(long long)row * k + n] = -INFINITY; idx[(long long)row * k + n] = -1; }
}

constexpr long long TOPK_SEED_ROWS = 32768;      // sample scored ahead of the main pass to seed the per-user thresholds

int topk_strips(int B, int k) {
    int s = ce_strips(B);
    const int cap = 1024 / k - 1;
    if (s > cap) s = cap;
    return s < 1 ? 1 : s;
}
}  // namespace

bool umma_topk_supported(int B, int d, int k, long long rows) {
    return B >= 1 && d % 4 == 0 && d >= 4 && k >= 1 && k <= 32 && rows >= 1 && rows < (1ll << 31);
}

size_t umma_topk_workspace_bytes(int B, int d, int k, long long rows) {
    const int ns = topk_strips(B, k) + 1;
    return 2 * al((size_t)B * d * 4) + 2 * al((size_t)rows * d * 4) + al((size_t)ns * B * k * 4) + al((size_t)ns * B * k * 8) +
           al((size_t)B * k * 4) + al((size_t)B * k * 8) + 4096;
}

int umma_score_topk(int B, int d, int k, long long rows, long long offset, const float* s, const float* E,
                    const long long* hist_indptr, const long long* hist_indices, const long long* hist_rows, float* top_val,
                    long long* top_idx, void* ws, size_t ws_bytes, bool reuse_split, cudaStream_t st) {
    BSA_REQUIRE(ws_bytes >= umma_topk_workspace_bytes(B, d, k, rows), "bsa_score_topk: workspace too small");
    char* p = (char*)ws;
    auto take = [&](size_t bytes) { char* r = p; p += al(bytes); return r; };
    float* s_hi = (float*)take((size_t)B * d * 4); float* s_lo = (float*)take((size_t)B * d * 4);
    float* e_hi = (float*)take((size_t)rows * d * 4); float* e_lo = (float*)take((size_t)rows * d * 4);
    int nstrips = topk_strips(B, k);
    float* cand_val = (float*)take((size_t)(nstrips + 1) * B * k * 4);
    long long* cand_idx = (long long*)take((size_t)(nstrips + 1) * B * k * 8);
    int rc;
    if ((rc = do_split(s, s_hi, s_lo, (long long)B * d, st))) return rc;
    if (e_trunc()) {          // the shard is its own hi half: only lo = E - truncate_tf32(E) is kept (the e_hi slot stays unused)
        e_hi = const_cast<float*>(E);
        if (!reuse_split && (rc = do_split_lo(E, e_lo, rows * d, st))) return rc;
    } else if (!reuse_split && (rc = do_split(E, e_hi, e_lo, rows * d, st))) return rc;
    Operand oa{s_hi, s_lo, B, d, d, false}, ob{e_hi, e_lo, rows, d, d, false};
    GemmCore g;
    if ((rc = make_core(g, B, (int)rows, d, oa, ob, CE_BN, SCHED_STRIP, 1))) return rc;
    if (nstrips > g.n_tiles) nstrips = g.n_tiles;
    float* seed_val = (float*)take((size_t)B * k * 4);
    long long* seed_idx = (long long*)take((size_t)B * k * 8);
    // Threshold seeding.  Every strip starts its per-user list cold, and a cold list is where the epilogue's time goes
    // (a warp takes the serial insertion path whenever ANY of its 32 users has a candidate).  So the first TOPK_SEED_ROWS
    // rows are scored once ahead of time; each user's k-th best there (history excluded, as in the main pass) is a lower
    // bound of the final k-th score, and the main pass skips everything below it from its first tile on.  The sample
    // rows are scored again by the main pass, so the result does not depend on the seed.
    const float* seed = nullptr;
    if (rows >= 8 * TOPK_SEED_ROWS) {
        GemmCore gs;
        Operand obs{e_hi, e_lo, TOPK_SEED_ROWS, d, d, false};
        if ((rc = make_core(gs, B, (int)TOPK_SEED_ROWS, d, oa, obs, CE_BN, SCHED_STRIP, 1))) return rc;
        int ss = nstrips < gs.n_tiles ? nstrips : gs.n_tiles;
        EpiTopk<CE_BN>::Params sp{hist_indptr, hist_indices, hist_rows, offset, k, cand_val, cand_idx, nullptr, 0};
        ProfScope prof(PROF_TOPK_UMMA, st, 0.0);     // time counted, no work credited: the sample is scored again below
        if ((rc = launch<CE_BN, CE_STAGES, false, false, EpiTopk<CE_BN>>(gs, sp, ss * gs.m_tiles, st))) return rc;
        if ((rc = topk_merge(cand_val, cand_idx, ss, B, k, seed_val, seed_idx, st))) return rc;
        seed = seed_val;
    }
    EpiTopk<CE_BN>::Params ep{hist_indptr, hist_indices, hist_rows, offset, k, cand_val, cand_idx, seed, k};
    {
        ProfScope prof(PROF_TOPK_UMMA, st, 2.0 * B * (double)rows * d);
        if ((rc = launch<CE_BN, CE_STAGES, false, false, EpiTopk<CE_BN>>(g, ep, nstrips * g.m_tiles, st))) return rc;
    }
    // history items enter as one more candidate list with score exactly 0.0 (trainers.py:134)
    hist_list_kernel<<<cdiv(B, 128), 128, 0, st>>>(hist_indptr, hist_indices, hist_rows, offset, rows, k, B,
                                                  cand_val + (size_t)nstrips * B * k, cand_idx + (size_t)nstrips * B * k);
    BSA_LAUNCH_CHECK();
    return topk_merge(cand_val, cand_idx, nstrips + 1, B, k, top_val, top_idx, st);
}

}  // namespace bsa

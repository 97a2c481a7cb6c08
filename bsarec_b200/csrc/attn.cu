// Fused attention forward/backward for one BSARecBlock (see attn.cuh).
//
// Work split: one CTA = (sequence b, head h, 64 "owner" rows), 4 warps x 16 rows.  The other side of the L x L score
// matrix is streamed through shared memory in blocks of 64 rows (cp.async, double buffered).  All contractions run
// on the tensor cores as error-compensated TF32 (mma.sync m16n8k8, hi*hi + hi*lo + lo*hi with the split done in
// registers), so results stay within fp32 rounding of the reference's sgemm + softmax.  Row max / row sum are
// quad shuffles over the accumulator fragments (warp-shuffle softmax).
//
// Mask semantics are the reference's, literally: every (query i, key j) pair gets +0 if (ids[j] > 0 and j <= i) else
// -10000 before the softmax, so a query row with no permitted key (padding positions under
// left padding) attends to ALL L keys.  Key tiles whose every pair is masked for rows that do have a permitted key
// contribute exp(-10000 - max) == 0 exactly and are skipped.
#include "attn.cuh"
#include "umma_engine.cuh"
#include <cstdlib>

namespace bsa {
namespace {

constexpr int AT_WARPS = 4;
constexpr int AT_THREADS = AT_WARPS * 32;
constexpr int AT_ROWS = 16 * AT_WARPS;     // owner rows per CTA
constexpr int AT_BLK = 64;                 // streamed rows per shared-memory block

__device__ __forceinline__ void split3(float x, uint32_t& hi, uint32_t& lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xFFFFE000u;  // round to nearest tf32, ties away (== cvt.rna), on the int pipe
    lo = __float_as_uint(x - __uint_as_float(hi));      // exact; the tensor core reads its top 19 bits
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// c += a * b with fp32 operands a[4] (A fragment) and b0, b1 (B fragment): small terms first
__device__ __forceinline__ void mma3(float (&c)[4], const float (&a)[4], float b0, float b1) {
    uint32_t ah[4], al[4], bh0, bl0, bh1, bl1;
#pragma unroll
    for (int i = 0; i < 4; ++i) split3(a[i], ah[i], al[i]);
    split3(b0, bh0, bl0);
    split3(b1, bh1, bl1);
    mma_tf32(c, al, bh0, bh1);
    mma_tf32(c, ah, bl0, bl1);
    mma_tf32(c, ah, bh0, bh1);
}

// same with the A fragment already split
__device__ __forceinline__ void mma3s(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], float b0, float b1) {
    uint32_t bh0, bl0, bh1, bl1;
    split3(b0, bh0, bl0);
    split3(b1, bh1, bl1);
    mma_tf32(c, al, bh0, bh1);
    mma_tf32(c, ah, bl0, bl1);
    mma_tf32(c, ah, bh0, bh1);
}

// Shared tile [64 rows][DH floats], 16-byte chunks XOR-swizzled by (row & 7): both fragment access patterns below
// are bank-conflict free.
template <int DH>
__device__ __forceinline__ int swz(int row, int col) { return row * DH + (col ^ ((row & 7) << 2)); }

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// rows [r0, r0+64) of one head of src [B*L, d] -> swizzled tile; rows >= L are zero-filled
template <int DH>
__device__ __forceinline__ void load_block(float* tile, const float* __restrict__ src, long long seq_base, int r0, int L,
                                           int d, int col0) {
    constexpr int CH = DH / 4;                       // 16-byte chunks per row
    for (int i = threadIdx.x; i < AT_BLK * CH; i += AT_THREADS) {
        const int r = i / CH, c = i - r * CH;
        float* dst = tile + r * DH + ((c ^ (r & 7)) << 2);
        if (r0 + r < L) cp_async16(dst, src + (seq_base + r0 + r) * d + col0 + c * 4);
        else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// A fragments of 16 rows x DH columns straight from global memory (rows >= L read as zero)
template <int DH>
__device__ __forceinline__ void load_a_frags(float (&a)[DH / 8][4], const float* __restrict__ src, long long seq_base,
                                             int i0, int L, int d, int col0, int g, int t) {
    const bool ok0 = i0 + g < L, ok1 = i0 + g + 8 < L;
    const float* p0 = src + (seq_base + i0 + g) * d + col0 + t;
    const float* p1 = p0 + 8ll * d;
#pragma unroll
    for (int ks = 0; ks < DH / 8; ++ks) {
        a[ks][0] = ok0 ? __ldg(p0 + ks * 8) : 0.f;
        a[ks][1] = ok1 ? __ldg(p1 + ks * 8) : 0.f;
        a[ks][2] = ok0 ? __ldg(p0 + ks * 8 + 4) : 0.f;
        a[ks][3] = ok1 ? __ldg(p1 + ks * 8 + 4) : 0.f;
    }
}

// dropout multipliers of two consecutive elements idx, idx + 1 of the [B,H,L,L] probability tensor
__device__ __forceinline__ void dropout_mult2(const Dropout& dr, unsigned long long idx, float& m0, float& m1) {
    if (dr.thr == 0u) { m0 = m1 = 1.0f; return; }
    const unsigned long long w = splitmix64(dropout_key(dr) + (idx >> 2));
    const unsigned int s = (unsigned int)(idx & 3ull);
    m0 = (((unsigned int)(w >> (16 * s)) & 0xFFFFu) >= dr.thr) ? dr.scale : 0.0f;
    if (s != 3u) m1 = (((unsigned int)(w >> (16 * (s + 1))) & 0xFFFFu) >= dr.thr) ? dr.scale : 0.0f;
    else m1 = dropout_mult(dr, idx + 1);
}

// first index with ids > 0 (L if none); also fills valid[] (1 = item, 0 = padding).  All threads must call.
__device__ __forceinline__ int scan_valid(const long long* __restrict__ idb, int L, unsigned char* valid, int* s_first) {
    if (threadIdx.x == 0) *s_first = L;
    __syncthreads();
    int first = L;
    for (int j = threadIdx.x; j < L; j += AT_THREADS) {
        const bool v = idb[j] > 0;
        valid[j] = v ? 1 : 0;
        if (v && j < first) first = j;
    }
    if (first < L) atomicMin(s_first, first);
    __syncthreads();
    return *s_first;
}

// ============================================================================ forward
// smem: K tiles [2][64*DH], V tiles [2][64*DH], valid[L], first-valid index
template <int DH>
constexpr size_t attn_smem_bytes(int L) { return (size_t)4 * AT_BLK * DH * 4 + (size_t)((L + 15) / 16 * 16) + 16; }

template <int DH>
__global__ void __launch_bounds__(AT_THREADS)
attn_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                const long long* __restrict__ ids, float* __restrict__ ctx, float* __restrict__ lse, int H, int L, int d,
                float scale, Dropout dr) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) unsigned char at_smem[];
    float* Ks = reinterpret_cast<float*>(at_smem);
    float* Vs = Ks + 2 * AT_BLK * DH;
    unsigned char* valid = reinterpret_cast<unsigned char*>(Vs + 2 * AT_BLK * DH);
    int* s_first = reinterpret_cast<int*>(valid + (L + 15) / 16 * 16);

    const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
    const int r0 = blockIdx.y * AT_ROWS;
    const long long seq_base = (long long)b * L;
    const int col0 = h * DH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int first_valid = scan_valid(ids + seq_base, L, valid, s_first);

    const int i0 = r0 + warp * 16;
    const bool active = i0 < L;
    const int i1 = min(i0 + 15, L - 1);
    const bool warp_all = i0 < first_valid;            // some row has no permitted key: it sees every key
    const int ig0 = i0 + g, ig1 = i0 + g + 8;

    // CTA-uniform range of key blocks
    const int nblk = (L + AT_BLK - 1) / AT_BLK;
    const int r1 = min(r0 + AT_ROWS - 1, L - 1);
    int kb_lo = 0, kb_hi = nblk;
    if (r0 >= first_valid) { kb_lo = first_valid / AT_BLK; kb_hi = r1 / AT_BLK + 1; }

    uint32_t qh[DH / 8][4], ql[DH / 8][4];
    {
        float a[DH / 8][4];
        load_a_frags<DH>(a, q, seq_base, i0, L, d, col0, g, t);
#pragma unroll
        for (int ks = 0; ks < DH / 8; ++ks)
#pragma unroll
            for (int i = 0; i < 4; ++i) split3(a[ks][i], qh[ks][i], ql[ks][i]);
    }
    float o[DH / 8][4];
#pragma unroll
    for (int n = 0; n < DH / 8; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const unsigned long long pbase0 = ((unsigned long long)bh * L + ig0) * L, pbase1 = ((unsigned long long)bh * L + ig1) * L;

    load_block<DH>(Ks, k, seq_base, kb_lo * AT_BLK, L, d, col0);
    load_block<DH>(Vs, v, seq_base, kb_lo * AT_BLK, L, d, col0);
    cp_async_commit();
    for (int kb = kb_lo; kb < kb_hi; ++kb) {
        const int buf = (kb - kb_lo) & 1;
        if (kb + 1 < kb_hi) {
            load_block<DH>(Ks + (buf ^ 1) * AT_BLK * DH, k, seq_base, (kb + 1) * AT_BLK, L, d, col0);
            load_block<DH>(Vs + (buf ^ 1) * AT_BLK * DH, v, seq_base, (kb + 1) * AT_BLK, L, d, col0);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float* Kt = Ks + buf * AT_BLK * DH;
        const float* Vt = Vs + buf * AT_BLK * DH;
        const int jb = kb * AT_BLK;
        // which 8-key tiles of this block matter to this warp
        unsigned live = 0;
        if (active) {
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const int j0 = jb + nt * 8;
                if (j0 < L && (warp_all || (j0 <= i1 && j0 + 7 >= first_valid))) live |= 1u << nt;
            }
        }
        if (live) {
            float s[8][4];
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
                if (live >> nt & 1) {
#pragma unroll
                    for (int ks = 0; ks < DH / 8; ++ks) {
                        const float b0 = Kt[swz<DH>(nt * 8 + g, ks * 8 + t)];
                        const float b1 = Kt[swz<DH>(nt * 8 + g, ks * 8 + t + 4)];
                        mma3s(s[nt], qh[ks], ql[ks], b0, b1);
                    }
                    const int j = jb + nt * 8 + 2 * t;
                    const bool v0 = j < L && valid[j], v1 = j + 1 < L && valid[j + 1];
                    s[nt][0] = s[nt][0] * scale + ((v0 && j <= ig0) ? 0.f : -10000.f);
                    s[nt][1] = s[nt][1] * scale + ((v1 && j + 1 <= ig0) ? 0.f : -10000.f);
                    s[nt][2] = s[nt][2] * scale + ((v0 && j <= ig1) ? 0.f : -10000.f);
                    s[nt][3] = s[nt][3] * scale + ((v1 && j + 1 <= ig1) ? 0.f : -10000.f);
                    if (j >= L) { s[nt][0] = -INFINITY; s[nt][2] = -INFINITY; }
                    if (j + 1 >= L) { s[nt][1] = -INFINITY; s[nt][3] = -INFINITY; }
                } else {
                    s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = -INFINITY;
                }
            }
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
                mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);      // finite: a live tile has a finite entry per row
            const float c0 = expf(m0 - mn0), c1 = expf(m1 - mn1);        // exp(-inf) = 0 on the first block
            m0 = mn0; m1 = mn1;
            l0 *= c0; l1 *= c1;
#pragma unroll
            for (int n = 0; n < DH / 8; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                if (!(live >> nt & 1)) continue;
                float p[4];
                p[0] = expf(s[nt][0] - m0); p[1] = expf(s[nt][1] - m0);
                p[2] = expf(s[nt][2] - m1); p[3] = expf(s[nt][3] - m1);
                l0 += p[0] + p[1]; l1 += p[2] + p[3];
                if (dr.thr) {
                    const int j = jb + nt * 8 + 2 * t;
                    float d0, d1, d2, d3;
                    dropout_mult2(dr, pbase0 + j, d0, d1);
                    dropout_mult2(dr, pbase1 + j, d2, d3);
                    p[0] *= d0; p[1] *= d1; p[2] *= d2; p[3] *= d3;
                }
                // accumulator fragment -> A fragment: k-slot t <-> key 2t, k-slot t+4 <-> key 2t+1
                uint32_t ah[4], al[4];
                split3(p[0], ah[0], al[0]); split3(p[2], ah[1], al[1]); split3(p[1], ah[2], al[2]); split3(p[3], ah[3], al[3]);
#pragma unroll
                for (int n = 0; n < DH / 8; ++n) {
                    const float b0 = Vt[swz<DH>(nt * 8 + 2 * t, n * 8 + g)];
                    const float b1 = Vt[swz<DH>(nt * 8 + 2 * t + 1, n * 8 + g)];
                    mma3s(o[n], ah, al, b0, b1);
                }
            }
        }
        __syncthreads();
    }
    if (!active) return;
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = 1.0f / l0, inv1 = 1.0f / l1;
    if (ig0 < L) {
        float* dst = ctx + (seq_base + ig0) * d + col0 + 2 * t;
#pragma unroll
        for (int n = 0; n < DH / 8; ++n) *reinterpret_cast<float2*>(dst + n * 8) = make_float2(o[n][0] * inv0, o[n][1] * inv0);
        if (t == 0) *reinterpret_cast<float2*>(lse + 2 * ((long long)bh * L + ig0)) = make_float2(m0, logf(l0));
    }
    if (ig1 < L) {
        float* dst = ctx + (seq_base + ig1) * d + col0 + 2 * t;
#pragma unroll
        for (int n = 0; n < DH / 8; ++n) *reinterpret_cast<float2*>(dst + n * 8) = make_float2(o[n][2] * inv1, o[n][3] * inv1);
        if (t == 0) *reinterpret_cast<float2*>(lse + 2 * ((long long)bh * L + ig1)) = make_float2(m1, logf(l1));
    }
}

// ============================================================================ backward
// One kernel body, two roles.  KEY_OWNER == false: the CTA owns 64 query rows, streams key/value blocks and produces
// dQ (and delta = rowsum(dctx * ctx)).  KEY_OWNER == true: the CTA owns 64 key rows, streams query/dctx blocks and
// produces dK and dV.  Both recompute P = exp(S' - lse) tile by tile; nothing of size L x L touches memory and no
// atomics are used (each output row has exactly one owner -> bit-reproducible).
//   P  = softmax(S'),  Pd = P o M (dropout),  ctx = Pd V
//   dPd = dctx V^T,  dS = P o (dPd o M - delta),  dQ = scale dS K,  dK = scale dS^T Q,  dV = Pd^T dctx
template <int DH>
constexpr size_t attn_bwd_smem_bytes(int L) { return attn_smem_bytes<DH>(L) + 6 * AT_BLK * 4; }

template <int DH, bool KEY_OWNER>
__global__ void __launch_bounds__(AT_THREADS)
attn_bwd_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                const long long* __restrict__ ids, const float* __restrict__ ctx, const float* __restrict__ lse,
                const float* __restrict__ dctx, float* __restrict__ out1, float* __restrict__ out2, float* __restrict__ delta,
                int H, int L, int d, float scale, Dropout dr, float* __restrict__ pd_out = nullptr,
                float* __restrict__ ds_out = nullptr, int NK = 0) {
    dr = dropout_resolve(dr);
    extern __shared__ __align__(16) unsigned char at_smem[];
    float* Us = reinterpret_cast<float*>(at_smem);            // streamed operand of S:  K blocks (query owner) / Q blocks
    float* Ws = Us + 2 * AT_BLK * DH;                         // streamed operand of dP: V blocks (query owner) / dctx blocks
    float2* lse_s = reinterpret_cast<float2*>(Ws + 2 * AT_BLK * DH);   // [2][64] (row max, log row sum) (key owner only)
    float* del_s = reinterpret_cast<float*>(lse_s + 2 * AT_BLK);       // [2][64]
    unsigned char* valid = reinterpret_cast<unsigned char*>(del_s + 2 * AT_BLK);
    int* s_first = reinterpret_cast<int*>(valid + (L + 15) / 16 * 16);

    const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
    const int r0 = blockIdx.y * AT_ROWS;
    const long long seq_base = (long long)b * L;
    const int col0 = h * DH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int first_valid = scan_valid(ids + seq_base, L, valid, s_first);

    const int i0 = r0 + warp * 16;                            // first owner row of this warp
    const bool active = i0 < L;
    const int i1 = min(i0 + 15, L - 1);
    const int ig0 = i0 + g, ig1 = i0 + g + 8;
    const int r1 = min(r0 + AT_ROWS - 1, L - 1);
    const int nblk = (L + AT_BLK - 1) / AT_BLK;
    const float* usrc = KEY_OWNER ? q : k;
    const float* wsrc = KEY_OWNER ? dctx : v;

    // owner-side A fragments: X feeds S (q or k rows), Y feeds dP (dctx or v rows)
    float X[DH / 8][4], Y[DH / 8][4];
    load_a_frags<DH>(X, KEY_OWNER ? k : q, seq_base, i0, L, d, col0, g, t);
    load_a_frags<DH>(Y, KEY_OWNER ? v : dctx, seq_base, i0, L, d, col0, g, t);
    float2 lse0 = make_float2(0.f, 0.f), lse1 = lse0;         // (row max, log row sum) per owner row (query owner only)
    float del0 = 0.f, del1 = 0.f;
    if (!KEY_OWNER) {
        float O[DH / 8][4];
        load_a_frags<DH>(O, ctx, seq_base, i0, L, d, col0, g, t);
#pragma unroll
        for (int ks = 0; ks < DH / 8; ++ks) {
            del0 += O[ks][0] * Y[ks][0] + O[ks][2] * Y[ks][2];
            del1 += O[ks][1] * Y[ks][1] + O[ks][3] * Y[ks][3];
        }
        del0 += __shfl_xor_sync(0xffffffffu, del0, 1); del0 += __shfl_xor_sync(0xffffffffu, del0, 2);
        del1 += __shfl_xor_sync(0xffffffffu, del1, 1); del1 += __shfl_xor_sync(0xffffffffu, del1, 2);
        if (ig0 < L) { lse0 = *reinterpret_cast<const float2*>(lse + 2 * ((long long)bh * L + ig0)); if (t == 0) delta[(long long)bh * L + ig0] = del0; }
        if (ig1 < L) { lse1 = *reinterpret_cast<const float2*>(lse + 2 * ((long long)bh * L + ig1)); if (t == 0) delta[(long long)bh * L + ig1] = del1; }
    }
    float acc1[DH / 8][4], acc2[KEY_OWNER ? DH / 8 : 1][4];
#pragma unroll
    for (int n = 0; n < DH / 8; ++n) { acc1[n][0] = acc1[n][1] = acc1[n][2] = acc1[n][3] = 0.f; }
#pragma unroll
    for (int n = 0; n < (KEY_OWNER ? DH / 8 : 1); ++n) { acc2[n][0] = acc2[n][1] = acc2[n][2] = acc2[n][3] = 0.f; }

    // streamed blocks this CTA needs (CTA-uniform)
    auto needed = [&](int kb) -> bool {
        const int c0 = kb * AT_BLK, c1 = min(c0 + AT_BLK - 1, L - 1);
        if (!KEY_OWNER) return r0 < first_valid || (c0 <= r1 && c1 >= first_valid);
        return c0 < first_valid || (c1 >= r0 && r1 >= first_valid);
    };
    auto next_needed = [&](int kb) -> int {
        for (++kb; kb < nblk && !needed(kb); ++kb) {}
        return kb;
    };
    auto issue = [&](int kb, int buf) {
        load_block<DH>(Us + buf * AT_BLK * DH, usrc, seq_base, kb * AT_BLK, L, d, col0);
        load_block<DH>(Ws + buf * AT_BLK * DH, wsrc, seq_base, kb * AT_BLK, L, d, col0);
        if (KEY_OWNER && threadIdx.x < AT_BLK) {
            const int i = kb * AT_BLK + threadIdx.x;
            lse_s[buf * AT_BLK + threadIdx.x] = i < L ? *reinterpret_cast<const float2*>(lse + 2 * ((long long)bh * L + i)) : make_float2(0.f, 0.f);
            del_s[buf * AT_BLK + threadIdx.x] = i < L ? delta[(long long)bh * L + i] : 0.f;
        }
        cp_async_commit();
    };

    int kb = next_needed(-1), buf = 0;
    if (kb < nblk) issue(kb, 0);
    while (kb < nblk) {
        const int nxt = next_needed(kb);
        if (nxt < nblk) { issue(nxt, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();
        const float* Ut = Us + buf * AT_BLK * DH;
        const float* Wt = Ws + buf * AT_BLK * DH;
        const int cb = kb * AT_BLK;
        if (active) {
#pragma unroll 1
            for (int nt = 0; nt < AT_BLK / 16; ++nt) {                 // 16 streamed rows per step
                const int c0 = cb + nt * 16;
                bool live;
                if (!KEY_OWNER) live = c0 < L && (i0 < first_valid || (c0 <= i1 && c0 + 15 >= first_valid));
                else            live = c0 < L && (c0 < first_valid || (c0 + 15 >= i0 && i1 >= first_valid));
                if (!live) continue;
                float sf[2][4], df[2][4];
#pragma unroll
                for (int u = 0; u < 2; ++u) { sf[u][0] = sf[u][1] = sf[u][2] = sf[u][3] = 0.f; df[u][0] = df[u][1] = df[u][2] = df[u][3] = 0.f; }
#pragma unroll
                for (int ks = 0; ks < DH / 8; ++ks) {
                    uint32_t xh[4], xl[4], yh[4], yl[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { split3(X[ks][i], xh[i], xl[i]); split3(Y[ks][i], yh[i], yl[i]); }
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int rr = nt * 16 + u * 8 + g;
                        mma3s(sf[u], xh, xl, Ut[swz<DH>(rr, ks * 8 + t)], Ut[swz<DH>(rr, ks * 8 + t + 4)]);
                        mma3s(df[u], yh, yl, Wt[swz<DH>(rr, ks * 8 + t)], Wt[swz<DH>(rr, ks * 8 + t + 4)]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int c = c0 + u * 8 + 2 * t;                  // streamed index of fragment columns c, c + 1
                    float p[4], ds[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int own = (e & 2) ? ig1 : ig0;           // owner index of this element's row
                        const int str = c + (e & 1);                   // streamed index of this element's column
                        const int qi = KEY_OWNER ? str : own, kj = KEY_OWNER ? own : str;
                        float2 ls; float dl;
                        if (!KEY_OWNER) { ls = (e & 2) ? lse1 : lse0; dl = (e & 2) ? del1 : del0; }
                        else { ls = lse_s[buf * AT_BLK + (str - cb)]; dl = del_s[buf * AT_BLK + (str - cb)]; }
                        const bool in = str < L && own < L;
                        const bool ok = in && kj <= qi && valid[kj];
                        const float sp = sf[u][e] * scale + (ok ? 0.f : -10000.f);
                        float pr = in ? expf((sp - ls.x) - ls.y) : 0.f;     // sp - max is exact even when both sit near -10000
                        float mlt = 1.0f;
                        if (dr.thr) mlt = dropout_mult(dr, ((unsigned long long)bh * L + (in ? qi : 0)) * L + (in ? kj : 0));
                        ds[e] = pr * (df[u][e] * mlt - dl);
                        p[e] = pr * mlt;
                    }
                    if (!KEY_OWNER && pd_out) {
                        // leave Pd = P o M and dS for the tcgen05 key-owner kernel (attn_kv_umma.cu): [B*H*L, NK] fp32, zeroed by
                        // the launcher; a fragment row holds columns c, c + 1 -> one float2 per row, 32 contiguous bytes per quad
                        if (ig0 < L) {
                            *reinterpret_cast<float2*>(pd_out + ((long long)bh * L + ig0) * NK + c) = make_float2(p[0], p[1]);
                            *reinterpret_cast<float2*>(ds_out + ((long long)bh * L + ig0) * NK + c) = make_float2(ds[0], ds[1]);
                        }
                        if (ig1 < L) {
                            *reinterpret_cast<float2*>(pd_out + ((long long)bh * L + ig1) * NK + c) = make_float2(p[2], p[3]);
                            *reinterpret_cast<float2*>(ds_out + ((long long)bh * L + ig1) * NK + c) = make_float2(ds[2], ds[3]);
                        }
                    }
                    uint32_t ah[4], al[4];
                    split3(ds[0], ah[0], al[0]); split3(ds[2], ah[1], al[1]); split3(ds[1], ah[2], al[2]); split3(ds[3], ah[3], al[3]);
                    const int rr = nt * 16 + u * 8 + 2 * t;
#pragma unroll
                    for (int n = 0; n < DH / 8; ++n)
                        mma3s(acc1[n], ah, al, Ut[swz<DH>(rr, n * 8 + g)], Ut[swz<DH>(rr + 1, n * 8 + g)]);
                    if (KEY_OWNER) {
                        split3(p[0], ah[0], al[0]); split3(p[2], ah[1], al[1]); split3(p[1], ah[2], al[2]); split3(p[3], ah[3], al[3]);
#pragma unroll
                        for (int n = 0; n < DH / 8; ++n)
                            mma3s(acc2[n], ah, al, Wt[swz<DH>(rr, n * 8 + g)], Wt[swz<DH>(rr + 1, n * 8 + g)]);
                    }
                }
            }
        }
        __syncthreads();
        kb = nxt; buf ^= 1;
    }
    if (!active) return;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int row = half ? ig1 : ig0;
        if (row >= L) continue;
        float* d1 = out1 + (seq_base + row) * d + col0 + 2 * t;
#pragma unroll
        for (int n = 0; n < DH / 8; ++n)
            *reinterpret_cast<float2*>(d1 + n * 8) = make_float2(acc1[n][2 * half] * scale, acc1[n][2 * half + 1] * scale);
        if (KEY_OWNER) {
            float* d2 = out2 + (seq_base + row) * d + col0 + 2 * t;
#pragma unroll
            for (int n = 0; n < DH / 8; ++n)
                *reinterpret_cast<float2*>(d2 + n * 8) = make_float2(acc2[n][2 * half], acc2[n][2 * half + 1]);
        }
    }
}

template <int DH>
int launch_fwd(const float* q, const float* k, const float* v, const long long* ids, float* ctx, float* lse, int B, int H,
               int L, int d, Dropout dr, cudaStream_t st) {
    const size_t smem = attn_smem_bytes<DH>(L);
    static size_t configured = 0;
    if (smem > configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const dim3 grid(B * H, (L + AT_ROWS - 1) / AT_ROWS);
    ProfScope prof(PROF_ATTN_FWD, st, 4.0 * B * H * (double)L * L * DH);
    attn_fwd_kernel<DH><<<grid, AT_THREADS, smem, st>>>(q, k, v, ids, ctx, lse, H, L, d, 1.0f / sqrtf((float)DH), dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

template <int DH>
int launch_bwd(const float* q, const float* k, const float* v, const long long* ids, const float* ctx, const float* lse,
               const float* dctx, float* dq, float* dk, float* dv, float* delta, int B, int H, int L, int d, Dropout dr,
               cudaStream_t st, int ldo) {
    const size_t smem = attn_bwd_smem_bytes<DH>(L);
    BSA_REQUIRE(ldo == 0 || ldo == d || attn_bwd_strided_ok(L, d, H), "attn_bwd: strided outputs need the tcgen05 backward");
    static size_t configured = 0;
    if (smem > configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<DH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<DH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const dim3 grid(B * H, (L + AT_ROWS - 1) / AT_ROWS);
    const float scale = 1.0f / sqrtf((float)DH);
    ProfScope prof(PROF_ATTN_BWD, st, 8.0 * B * H * (double)L * L * DH);      // canonical: dP, dV, dQ, dK (no credit for the S / dP recomputation)
    if (g_attn_bwd_umma && g_attn_fwd_umma && attn_umma_supported(L, d, H)) {   // query owner on tcgen05 (dQ, delta)
        if (g_attn_bwd_umma >= 2) {      // ... which also leaves Pd and dS for the tcgen05 key owner (dK, dV): no recomputation
            const long long NK = (L + 15) / 16 * 16, bhl = (long long)B * H * L;
            float* pd = delta + (bhl + 63) / 64 * 64;
            float* ds = pd + bhl * NK;
            int rc = attn_bwd_q_umma(q, k, v, ids, ctx, lse, dctx, dq, delta, pd, ds, B, H, L, d, dr, st, ldo);
            if (rc) return rc;
            return attn_bwd_kv_umma(q, dctx, pd, ds, dk, dv, B, H, L, d, st, ldo);
        }
        const int rc = attn_bwd_q_umma(q, k, v, ids, ctx, lse, dctx, dq, delta, nullptr, nullptr, B, H, L, d, dr, st);
        if (rc) return rc;
    } else if (g_attn_bwd_umma >= 2 && L > 208) {
        // long sequences (C5): the mma.sync query owner (the whole key axis does not fit one TMEM accumulator) also leaves
        // the Pd / dS tiles, and the key owner runs on tcgen05 from them — no second recomputation of the probabilities
        const long long NK = (L + 15) / 16 * 16, bhl = (long long)B * H * L;
        float* pd = delta + (bhl + 63) / 64 * 64;
        float* ds = pd + bhl * NK;
        BSA_CHECK_CUDA(cudaMemsetAsync(pd, 0, (size_t)(2 * bhl * NK) * sizeof(float), st));      // causal / padding blocks are skipped by the writer
        attn_bwd_kernel<DH, false><<<grid, AT_THREADS, smem, st>>>(q, k, v, ids, ctx, lse, dctx, dq, nullptr, delta, H, L, d, scale, dr,
                                                                   pd, ds, (int)NK);
        BSA_LAUNCH_CHECK();
        return attn_bwd_kv_umma(q, dctx, pd, ds, dk, dv, B, H, L, d, st);
    } else {
        attn_bwd_kernel<DH, false><<<grid, AT_THREADS, smem, st>>>(q, k, v, ids, ctx, lse, dctx, dq, nullptr, delta, H, L, d, scale, dr);
        BSA_LAUNCH_CHECK();
    }
    attn_bwd_kernel<DH, true><<<grid, AT_THREADS, smem, st>>>(q, k, v, ids, ctx, lse, dctx, dk, dv, delta, H, L, d, scale, dr);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace

// 1: the forward goes to the tcgen05 kernel where its shape allows (BSA_ATTN_UMMA=0 keeps the mma.sync kernel)
int g_attn_fwd_umma = [] { const char* e = getenv("BSA_ATTN_UMMA"); return e ? atoi(e) : 1; }();
// 1: the query-owner half of the backward goes to the tcgen05 kernel too (BSA_ATTN_BWD_UMMA=0 keeps mma.sync)
// 2 (default): the key-owner half too, from the Pd / dS tiles the query owner stores (attn_kv_umma.cu); 1: key owner on mma.sync
int g_attn_bwd_umma = [] { const char* e = getenv("BSA_ATTN_BWD_UMMA"); return e ? atoi(e) : 2; }();

long long attn_bwd_scratch_floats(int B, int H, int L, int d) {
    const long long bhl = (long long)B * H * L;
    const bool fused = attn_fused_supported(L, d, H);
    if (g_attn_bwd_umma >= 2 && fused && ((g_attn_fwd_umma && attn_umma_supported(L, d, H)) || L > 208)) {
        const long long NK = (L + 15) / 16 * 16;
        return (bhl + 63) / 64 * 64 + 2 * bhl * NK + 64;
    }
    return bhl;
}

bool attn_fused_supported(int L, int d, int H) {
    if (H <= 0 || d % H != 0 || d % 4 != 0 || L < 1) return false;
    const int dh = d / H;
    return dh == 32 || dh == 64;
}

int attn_fwd(const float* q, const float* k, const float* v, const long long* ids, float* ctx, float* lse, int B, int H,
             int L, int d, Dropout dr, cudaStream_t st, float* ctx_lo) {
    BSA_REQUIRE(attn_fused_supported(L, d, H), "attn_fwd: unsupported head size %d", H ? d / H : 0);
    if (B == 0) return 0;
    if (g_attn_fwd_umma && attn_umma_supported(L, d, H)) return attn_fwd_umma(q, k, v, ids, ctx, lse, B, H, L, d, dr, st, ctx_lo);
    const int rc = d / H == 64 ? launch_fwd<64>(q, k, v, ids, ctx, lse, B, H, L, d, dr, st) : launch_fwd<32>(q, k, v, ids, ctx, lse, B, H, L, d, dr, st);
    if (rc || !ctx_lo) return rc;
    return umma_split_lo(ctx, ctx_lo, (long long)B * L * d, st);       // the mma.sync kernel does not emit it
}

int attn_bwd(const float* q, const float* k, const float* v, const long long* ids, const float* ctx, const float* lse,
             const float* dctx, float* dq, float* dk, float* dv, float* delta, int B, int H, int L, int d, Dropout dr,
             cudaStream_t st, int ldo) {
    BSA_REQUIRE(attn_fused_supported(L, d, H), "attn_bwd: unsupported head size %d", H ? d / H : 0);
    if (B == 0) return 0;
    if (d / H == 64) return launch_bwd<64>(q, k, v, ids, ctx, lse, dctx, dq, dk, dv, delta, B, H, L, d, dr, st, ldo);
    return launch_bwd<32>(q, k, v, ids, ctx, lse, dctx, dq, dk, dv, delta, B, H, L, d, dr, st, ldo);
}

bool attn_bwd_strided_ok(int L, int d, int H) {
    return g_attn_bwd_umma >= 2 && g_attn_fwd_umma && attn_fused_supported(L, d, H) && attn_umma_supported(L, d, H);
}

}  // namespace bsa

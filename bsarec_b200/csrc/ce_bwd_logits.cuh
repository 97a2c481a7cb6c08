// Cross-entropy backward from the STORED logits (bsarec.py:33-35; the forward's EpiLse epilogue leaves S = s E^T in
// HBM): no logits recomputation, no gradient tile G in HBM.  One kernel template, instantiated twice:
//
//   USER owner (ds):  a CTA owns 128 users and walks a strip of item tiles;   ds[u,:]  += sum_i P[u,i] E[i,:]
//   ITEM owner (dE):  a CTA owns 128 items and walks ALL users, tile by tile;  dE[i,:]   = sum_u P[u,i] s[u,:]
//
// with P = softmax(S) * dloss / B = exp(S - lse) * scale.  The "- onehot" part of (softmax - onehot) touches one item
// per user and is applied afterwards by two row kernels (ce_onehot_fix_*), so both contractions here are plain P-weighted
// sums.  Per 64 "stream" elements (items for the USER owner, users for the ITEM owner):
//   TMA brings the S sub-tile into shared memory; the 8 converter warps (thread = owner row = TMEM lane) read their
//   row (USER) / column (ITEM) of it, form P, split it into a TF32 (hi, lo) pair and store the pair into TMEM;
//   one elected thread issues  acc[128 x d] += P[128 x 64] * Bop[64 x d]  as tcgen05.mma with the A operand in TMEM
//   (3 MMAs per 8-wide k-step: lo*hi + hi*lo + hi*hi) and Bop = the stream rows of E (USER) / s (ITEM) as an MN-major
//   split pair streamed through a 2-stage TMA ring.
// TMEM (512 columns): two P buffers of (64 hi + 64 lo) columns — sub-tile j+1 is converted while sub-tile j is
// multiplied — and the 256-column accumulator, which stays resident for the owner tile's whole walk: it leaves once,
// with plain coalesced stores, as the strip's partial slab (USER; the slabs are summed afterwards in fixed order) or
// as the finished dE tile (ITEM) -> no atomics anywhere, bit-reproducible.
// Work executed = the canonical 4 B N d FLOP of SURVEY.md §8(d); HBM traffic = S read twice + E + dE.
#pragma once
#include "umma_gemm.cuh"

namespace bsa {
namespace umma {

struct CeBwdLogitsParams {
    CUtensorMap s_map;          // logits [B, ld] fp32, 128-B swizzle: box {32 cols, 128 rows} (USER) / {32 cols, 64 rows} (ITEM)
    CUtensorMap b_hi, b_lo;     // MN-major B operand (32-B swizzle atom), box {32 cols, 32 rows}: E [N, d] (USER) / s [B, d] (ITEM)
    int B, N, d;
    int owner_tiles;            // USER: ceil(B / 128); ITEM: ceil(N / 128)
    int n_tiles;                // ceil(N / 128)
    int nstrips;                // USER: strips of item tiles per user tile
    const float* lse;           // [B rounded up to 128], pad = 0
    const float* d_loss;
    float* out;                 // USER: partial [nstrips][B][d]; ITEM: dE [N, d]  (both overwritten)
};

constexpr int CBL_SUB = 64;                 // stream elements per P buffer
constexpr int CBL_ND = 256;                 // accumulator columns (d <= 256)
constexpr int CBL_RING = 2;
constexpr int CBL_STAGE_BYTES = 64 * 1024;  // 32 stream rows x 256 columns x (hi, lo)
constexpr int CBL_SBUF_BYTES = 32 * 1024;   // one S sub-tile
constexpr int CBL_EPI_WARPS = 8;
constexpr int CBL_THREADS = 128 + 32 * CBL_EPI_WARPS;
constexpr int CBL_SMEM = 1024 + CBL_RING * CBL_STAGE_BYTES + 2 * CBL_SBUF_BYTES + CBL_EPI_WARPS * 4096 + 256;

template <bool ITEM_OWNER>
struct CblWalk {
    int jobs, nsub, strip;
    __device__ explicit CblWalk(const CeBwdLogitsParams& p) {
        if (ITEM_OWNER) {
            strip = 0;
            jobs = (int)blockIdx.x < p.owner_tiles ? (p.owner_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
            nsub = (p.B + CBL_SUB - 1) / CBL_SUB;
        } else {
            strip = blockIdx.x / p.owner_tiles;
            jobs = 1;
            nsub = strip < p.n_tiles ? 2 * ((p.n_tiles - 1 - strip) / p.nstrips + 1) : 0;
        }
    }
    __device__ int owner0(const CeBwdLogitsParams& p, int job) const {
        return ITEM_OWNER ? ((int)blockIdx.x + job * (int)gridDim.x) * BM : ((int)blockIdx.x % p.owner_tiles) * BM;
    }
    __device__ int stream0(const CeBwdLogitsParams& p, int j) const {
        return ITEM_OWNER ? j * CBL_SUB : (strip + (j >> 1) * p.nstrips) * BM + (j & 1) * CBL_SUB;
    }
};

template <bool ITEM_OWNER>
__global__ void __launch_bounds__(CBL_THREADS, 1)
ce_bwd_logits_kernel(const __grid_constant__ CeBwdLogitsParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* ring = smem;
    unsigned char* sbuf = smem + CBL_RING * CBL_STAGE_BYTES;
    float4* stg_all = reinterpret_cast<float4*>(sbuf + 2 * CBL_SBUF_BYTES);
    uint64_t* full_bar = (uint64_t*)(sbuf + 2 * CBL_SBUF_BYTES + CBL_EPI_WARPS * 4096);
    uint64_t* empty_bar = full_bar + CBL_RING;
    uint64_t* s_full = empty_bar + CBL_RING;       // S sub-tile landed                 (TMA -> converters)
    uint64_t* s_empty = s_full + 2;                // S sub-tile read into registers    (converters -> TMA)
    uint64_t* g_ready = s_empty + 2;               // P pair stored into TMEM           (converters -> MMA)
    uint64_t* g_free = g_ready + 2;                // MMAs reading the P buffer retired (MMA -> converters)
    uint64_t* acc_full = g_free + 2;               // owner tile's accumulator complete (MMA -> converters)
    uint64_t* acc_empty = acc_full + 1;            // accumulator drained               (converters -> MMA)
    uint32_t* tmem_ptr = (uint32_t*)(acc_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const CblWalk<ITEM_OWNER> walk(p);

    if (warp == 0 && elect_one()) { prefetch_tmap(&p.s_map); prefetch_tmap(&p.b_hi); prefetch_tmap(&p.b_lo); }
    if (warp == 1 && elect_one()) {
        for (int i = 0; i < CBL_RING; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], CBL_EPI_WARPS);
            mbar_init(&g_ready[i], CBL_EPI_WARPS); mbar_init(&g_free[i], 1);
        }
        mbar_init(acc_full, 1); mbar_init(acc_empty, CBL_EPI_WARPS);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t t_acc = tmem_base + 256;
    const long long total = (long long)walk.jobs * walk.nsub;

    if (warp == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            auto load_s = [&](long long g) {
                const int job = (int)(g / walk.nsub), j = (int)(g % walk.nsub), b = (int)(g & 1);
                const int o0 = walk.owner0(p, job), s0 = walk.stream0(p, j);
                mbar_wait(&s_empty[b], (uint32_t)(((g >> 1) & 1) ^ 1));
                mbar_expect_tx(&s_full[b], CBL_SBUF_BYTES);
                unsigned char* dst = sbuf + b * CBL_SBUF_BYTES;
                if (ITEM_OWNER) {            // rows = users (stream), cols = items (owner): 4 boxes of 32 items x 64 users
#pragma unroll
                    for (int q = 0; q < 4; ++q) tma_load_2d(dst + q * 8192, &p.s_map, &s_full[b], o0 + q * 32, s0);
                } else {                     // rows = users (owner), cols = items (stream): 2 boxes of 32 items x 128 users
#pragma unroll
                    for (int q = 0; q < 2; ++q) tma_load_2d(dst + q * 16384, &p.s_map, &s_full[b], s0 + q * 32, o0);
                }
            };
            if (total > 0) load_s(0);
            if (total > 1) load_s(1);
            int cr = 0;
            for (long long g = 0; g < total; ++g) {
                const int s0 = walk.stream0(p, (int)(g % walk.nsub));
                for (int kb = 0; kb < CBL_SUB / BK; ++kb, ++cr) {       // the sub-tile's stream rows as MN-major B operand
                    const int s = cr % CBL_RING;
                    mbar_wait(&empty_bar[s], ((cr / CBL_RING) & 1) ^ 1);
                    unsigned char* st = ring + s * CBL_STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], CBL_STAGE_BYTES);
#pragma unroll
                    for (int q = 0; q < CBL_ND / 32; ++q) {
                        tma_load_2d(st + q * (BK * 128), &p.b_hi, &full_bar[s], q * 32, s0 + kb * BK);
                        tma_load_2d(st + 32768 + q * (BK * 128), &p.b_lo, &full_bar[s], q * 32, s0 + kb * BK);
                    }
                }
                if (g + 2 < total) load_s(g + 2);
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer
        if (elect_one()) {
            constexpr uint32_t IDESC = idesc_tf32(BM, CBL_ND, false, true);
            long long g = 0;
            int cr = 0;
            for (int job = 0; job < walk.jobs; ++job) {
                if (job > 0) {                               // the converters still drain the previous owner tile
                    mbar_wait(acc_empty, (uint32_t)((job - 1) & 1));
                    tc_fence_after();
                }
                uint32_t accumulate = 0;
                for (int j = 0; j < walk.nsub; ++j, ++g) {
                    const int b = (int)(g & 1);
                    mbar_wait(&g_ready[b], (uint32_t)((g >> 1) & 1));
                    tc_fence_after();
                    const uint32_t t_hi = tmem_base + (uint32_t)(b * 2 * CBL_SUB), t_lo = t_hi + CBL_SUB;
                    for (int kb = 0; kb < CBL_SUB / BK; ++kb, ++cr) {
                        const int s = cr % CBL_RING;
                        mbar_wait(&full_bar[s], (cr / CBL_RING) & 1);
                        tc_fence_after();
                        const uint32_t b_hi = smem_u32(ring + s * CBL_STAGE_BYTES);
                        const uint32_t b_lo = b_hi + 32768;
#pragma unroll
                        for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                            const uint64_t dbh = smem_desc(b_hi + ks * 1024, BK * 128, 512, LAYOUT_SW128_BASE32B);
                            const uint64_t dbl = smem_desc(b_lo + ks * 1024, BK * 128, 512, LAYOUT_SW128_BASE32B);
                            const uint32_t col = (uint32_t)(kb * BK + ks * UMMA_K);
                            mma_tf32_ts(t_acc, t_lo + col, dbh, IDESC, accumulate);
                            mma_tf32_ts(t_acc, t_hi + col, dbl, IDESC, 1u);
                            mma_tf32_ts(t_acc, t_hi + col, dbh, IDESC, 1u);
                            accumulate = 1u;
                        }
                        mma_commit(&empty_bar[s]);
                    }
                    mma_commit(&g_free[b]);
                }
                mma_commit(acc_full);
            }
        }
    } else if (warp >= 4) {
        // ===================================================================== converters / epilogue
        const int ew = warp - 4, quarter = ew & 3, half = ew >> 2;
        float4* stg = stg_all + ew * 256;
        const uint32_t lane_base = ((uint32_t)(quarter * 32) << 16);
        const int r = quarter * 32 + lane;                 // owner row inside the tile = TMEM lane
        const float scale = __ldg(p.d_loss) / (float)p.B;
        const int sub = lane >> 3, c4 = lane & 7;
        long long g = 0;
        for (int job = 0; job < walk.jobs; ++job) {
            const int o0 = walk.owner0(p, job);
            const float l_own = ITEM_OWNER ? 0.f : __ldg(p.lse + o0 + r);      // lse is padded to a multiple of 128
            for (int j = 0; j < walk.nsub; ++j, ++g) {
                const int b = (int)(g & 1);
                const int s0 = walk.stream0(p, j);
                mbar_wait(&s_full[b], (uint32_t)((g >> 1) & 1));
                const unsigned char* sb = sbuf + b * CBL_SBUF_BYTES;
                uint32_t hi[32], lo[32];
                if (ITEM_OWNER) {
                    // this thread's item is column `lane` of box `quarter`; users half*32 .. +32 are its 32 K-columns
                    const unsigned char* base = sb + quarter * 8192 + ((lane & 3) << 2);
                    const float4* lp = reinterpret_cast<const float4*>(p.lse + s0 + half * 32);
#pragma unroll
                    for (int i4 = 0; i4 < 8; ++i4) {
                        const float4 l4 = __ldg(lp + i4);
                        const float ls[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int u = half * 32 + i4 * 4 + e;
                            const float v = *reinterpret_cast<const float*>(base + u * 128 + ((((lane >> 2) ^ (u & 7))) << 4));
                            const float gv = fast_exp(v - ls[e]) * scale;
                            float h, lw;
                            split_tf32(gv, h, lw);
                            hi[i4 * 4 + e] = __float_as_uint(h); lo[i4 * 4 + e] = __float_as_uint(lw);
                        }
                    }
                } else {
                    const unsigned char* base = sb + half * 16384 + r * 128;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 x = *reinterpret_cast<const float4*>(base + ((c ^ (r & 7)) << 4));
                        const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float gv = fast_exp(xs[e] - l_own) * scale;
                            float h, lw;
                            split_tf32(gv, h, lw);
                            hi[c * 4 + e] = __float_as_uint(h); lo[c * 4 + e] = __float_as_uint(lw);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_empty[b]);               // the values are in registers: S buffer may be refilled
                mbar_wait(&g_free[b], (uint32_t)(((g >> 1) & 1) ^ 1));  // MMAs of sub-tile g-2 no longer read this P buffer
                tc_fence_after();
                const uint32_t t_hi = tmem_base + (uint32_t)(b * 2 * CBL_SUB + half * 32) + lane_base;
                tmem_st32(t_hi, hi);
                tmem_st32(t_hi + CBL_SUB, lo);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&g_ready[b]);
            }
            // ---- the owner tile's accumulator: TMEM -> staging -> global
            mbar_wait(acc_full, (uint32_t)(job & 1));
            tc_fence_after();
            const long long row0 = (long long)o0 + quarter * 32;
            const long long row_lim = ITEM_OWNER ? p.N : p.B;
            float* out = ITEM_OWNER ? p.out : p.out + (long long)walk.strip * p.B * p.d;
#pragma unroll 1
            for (int cc = 0; cc < CBL_ND / 64; ++cc) {
                const int col0 = half * (CBL_ND / 2) + cc * 32;
                float v[32];
                tmem_ld32(t_acc + lane_base + col0, v);
                if (col0 >= p.d || row0 >= row_lim) continue;
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) stg[lane * 8 + (jj ^ (lane & 7))] = make_float4(v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]);
                __syncwarp();
                const int col = col0 + c4 * 4;
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const int rr = t * 4 + sub;
                    const float4 x = stg[rr * 8 + (c4 ^ (rr & 7))];
                    if (row0 + rr < row_lim && col < p.d) {           // d % 4 == 0: whole float4 in range
                        float* dst = out + (row0 + rr) * p.d + col;
                        *reinterpret_cast<float4*>(dst) = x;
                    }
                }
                __syncwarp();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace umma
}  // namespace bsa

// Cross-entropy backward, fused stage: one kernel recomputes a chunk of logits, turns it into the gradient tile
// G = (softmax - onehot) * dloss / B, and contracts it with the item rows on the spot:
//     S  [128 users x 128 items] = s E_tile^T            (tcgen05, both operands from shared memory)
//     G  = epilogue(S)  -> written back into TMEM as a split-TF32 pair (hi over S, lo beside it)  [and to global for dE]
//     ds [128 users x d] += G E_tile                     (tcgen05 with the A operand read from TMEM)
// The ds accumulator lives in TMEM for the CTA's whole strip of item tiles, so the separate split-K `ds = G E` GEMM
// (4 output tiles, 37 splits, prologue-bound) and its reads of G disappear.  TMEM: [0,128) S/G_hi, [128,256) G_lo,
// [256,512) ds.  Same warp roles and stage ring as umma_gemm_kernel; phase 1 stages hold (s, E) K-major tiles, phase 2
// stages hold E_tile as an MN-major B operand (K = items, N = d).
#pragma once
#include "umma_gemm.cuh"

namespace bsa {
namespace umma {

struct CeGradDsParams {
    CUtensorMap s_hi, s_lo;       // [B, d], box 32 x 128
    CUtensorMap e_hi, e_lo;       // chunk rows of E [cols, d] as K-major operand, box 32 x 128
    CUtensorMap em_hi, em_lo;     // the same rows as MN-major operand (32-byte swizzle atom), box 32 x 32
    int B, cols, d;
    int m_tiles, n_tiles, nstrips, k1_blocks;
    const float* lse; const long long* answers; const float* d_loss; long long n_base;
    float *g_hi, *g_lo; long long ldg;
    float* partial;               // [nstrips][B][d] running ds sums (added to)
};

constexpr int CGD_BN = 128;                 // items per tile
constexpr int CGD_ND = 256;                 // ds accumulator columns (d <= 256)
constexpr int CGD_STAGES = 3;
constexpr int CGD_STAGE_BYTES = 64 * 1024;
constexpr int CGD_EPI_WARPS = 8;
constexpr int CGD_THREADS = 128 + 32 * CGD_EPI_WARPS;
constexpr int CGD_SMEM = 1024 + CGD_STAGES * CGD_STAGE_BYTES + CGD_EPI_WARPS * 4096 + 256;

__global__ void __launch_bounds__(CGD_THREADS, 1)
ce_grad_ds_kernel(const __grid_constant__ CeGradDsParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* stage_base = smem;
    float4* stg_all = reinterpret_cast<float4*>(smem + CGD_STAGES * CGD_STAGE_BYTES);
    uint64_t* full_bar = (uint64_t*)(smem + CGD_STAGES * CGD_STAGE_BYTES + CGD_EPI_WARPS * 4096);
    uint64_t* empty_bar = full_bar + CGD_STAGES;
    uint64_t* s_full = empty_bar + CGD_STAGES;     // S accumulator complete          (MMA -> epilogue)
    uint64_t* g_ready = s_full + 1;                // G pair written back to TMEM      (epilogue -> MMA)
    uint64_t* final_bar = g_ready + 1;             // last ds MMA retired              (MMA -> epilogue)
    uint64_t* g_consumed = final_bar + 1;          // G copied out of TMEM to global   (epilogue -> MMA: S may be overwritten)
    uint32_t* tmem_ptr = (uint32_t*)(g_consumed + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tile = blockIdx.x % p.m_tiles, strip = blockIdx.x / p.m_tiles;
    constexpr int K2_BLOCKS = CGD_BN / BK;          // 4 k-blocks of 32 items

    if (warp == 0 && elect_one()) {
        prefetch_tmap(&p.s_hi); prefetch_tmap(&p.s_lo); prefetch_tmap(&p.e_hi); prefetch_tmap(&p.e_lo);
        prefetch_tmap(&p.em_hi); prefetch_tmap(&p.em_lo);
    }
    if (warp == 1 && elect_one()) {
        for (int i = 0; i < CGD_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        mbar_init(s_full, 1); mbar_init(g_ready, CGD_EPI_WARPS); mbar_init(final_bar, 1); mbar_init(g_consumed, CGD_EPI_WARPS);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t t_s = tmem_base, t_lo = tmem_base + CGD_BN, t_ds = tmem_base + 2 * CGD_BN;

    if (warp == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            int c = 0;
            for (int n = strip; n < p.n_tiles; n += p.nstrips) {
                for (int kb = 0; kb < p.k1_blocks; ++kb, ++c) {          // phase 1: (s, E) K-major tiles
                    const int s = c % CGD_STAGES;
                    mbar_wait(&empty_bar[s], ((c / CGD_STAGES) & 1) ^ 1);
                    unsigned char* st = stage_base + s * CGD_STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], CGD_STAGE_BYTES);
                    tma_load_2d(st, &p.s_hi, &full_bar[s], kb * BK, m_tile * BM);
                    tma_load_2d(st + 16384, &p.s_lo, &full_bar[s], kb * BK, m_tile * BM);
                    tma_load_2d(st + 32768, &p.e_hi, &full_bar[s], kb * BK, n * CGD_BN);
                    tma_load_2d(st + 49152, &p.e_lo, &full_bar[s], kb * BK, n * CGD_BN);
                }
                for (int kb = 0; kb < K2_BLOCKS; ++kb, ++c) {            // phase 2: E rows as MN-major B (K = items)
                    const int s = c % CGD_STAGES;
                    mbar_wait(&empty_bar[s], ((c / CGD_STAGES) & 1) ^ 1);
                    unsigned char* st = stage_base + s * CGD_STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], CGD_STAGE_BYTES);
#pragma unroll
                    for (int q = 0; q < CGD_ND / 32; ++q) {
                        tma_load_2d(st + q * (BK * 128), &p.em_hi, &full_bar[s], q * 32, n * CGD_BN + kb * BK);
                        tma_load_2d(st + 32768 + q * (BK * 128), &p.em_lo, &full_bar[s], q * 32, n * CGD_BN + kb * BK);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer
        if (elect_one()) {
            constexpr uint32_t IDESC1 = idesc_tf32(BM, CGD_BN, false, false);
            constexpr uint32_t IDESC2 = idesc_tf32(BM, CGD_ND, false, true);
            int c = 0, it = 0;
            uint32_t ds_acc = 0;
            for (int n = strip; n < p.n_tiles; n += p.nstrips, ++it) {
                uint32_t accumulate = 0;
                if (it > 0) {                               // the epilogue still copies tile it-1's G out of TMEM
                    mbar_wait(g_consumed, (it - 1) & 1);
                    tc_fence_after();
                }
                for (int kb = 0; kb < p.k1_blocks; ++kb, ++c) {
                    const int s = c % CGD_STAGES;
                    mbar_wait(&full_bar[s], (c / CGD_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(stage_base + s * CGD_STAGE_BYTES);
                    const uint32_t a_lo = a_hi + 16384, b_hi = a_hi + 32768, b_lo = a_hi + 49152;
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                        const uint64_t dah = smem_desc(a_hi + ks * 32, 16, 1024, LAYOUT_SW128);
                        const uint64_t dal = smem_desc(a_lo + ks * 32, 16, 1024, LAYOUT_SW128);
                        const uint64_t dbh = smem_desc(b_hi + ks * 32, 16, 1024, LAYOUT_SW128);
                        const uint64_t dbl = smem_desc(b_lo + ks * 32, 16, 1024, LAYOUT_SW128);
                        mma_tf32(t_s, dal, dbh, IDESC1, accumulate);
                        mma_tf32(t_s, dah, dbl, IDESC1, 1u);
                        mma_tf32(t_s, dah, dbh, IDESC1, 1u);
                        accumulate = 1u;
                    }
                    mma_commit(&empty_bar[s]);
                }
                mma_commit(s_full);
                mbar_wait(g_ready, it & 1);                 // epilogue has replaced S by (G_hi, G_lo) in TMEM
                tc_fence_after();
                for (int kb = 0; kb < K2_BLOCKS; ++kb, ++c) {
                    const int s = c % CGD_STAGES;
                    mbar_wait(&full_bar[s], (c / CGD_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t b_hi = smem_u32(stage_base + s * CGD_STAGE_BYTES);
                    const uint32_t b_lo = b_hi + 32768;
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                        const uint64_t dbh = smem_desc(b_hi + ks * 1024, BK * 128, 512, LAYOUT_SW128_BASE32B);
                        const uint64_t dbl = smem_desc(b_lo + ks * 1024, BK * 128, 512, LAYOUT_SW128_BASE32B);
                        const uint32_t col = (uint32_t)(kb * BK + ks * UMMA_K);
                        mma_tf32_ts(t_ds, t_lo + col, dbh, IDESC2, ds_acc);
                        mma_tf32_ts(t_ds, t_s + col, dbl, IDESC2, 1u);
                        mma_tf32_ts(t_ds, t_s + col, dbh, IDESC2, 1u);
                        ds_acc = 1u;
                    }
                    mma_commit(&empty_bar[s]);
                }
            }
            mma_commit(final_bar);
        }
    } else if (warp >= 4) {
        // ===================================================================== epilogue
        const int ew = warp - 4, quarter = ew & 3, half = ew >> 2;
        float4* stg = stg_all + ew * 256;
        const uint32_t lane_base = ((uint32_t)(quarter * 32) << 16);
        const long long row = (long long)m_tile * BM + quarter * 32 + lane;       // this thread's user
        const bool row_ok = row < p.B;
        const float l = row_ok ? __ldg(p.lse + row) : 0.f;
        const long long ans = row_ok ? __ldg(p.answers + row) - p.n_base : -1;
        const float scale = __ldg(p.d_loss) / (float)p.B;
        const int sub = lane >> 3, c4 = lane & 7;
        const long long row0 = (long long)m_tile * BM + quarter * 32;
        int it = 0;
        for (int n = strip; n < p.n_tiles; n += p.nstrips, ++it) {
            mbar_wait(s_full, it & 1);
            tc_fence_after();
#pragma unroll 1
            for (int cc = 0; cc < 2; ++cc) {            // S -> (G_hi, G_lo) in place: the only work on the MMA's critical path
                const int col0 = half * 64 + cc * 32;
                const int item0 = n * CGD_BN + col0;
                float v[32];
                tmem_ld32(t_s + lane_base + col0, v);
                uint32_t hi[32], lo[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float g = 0.f;
                    if (row_ok && item0 + j < p.cols) {
                        g = expf(v[j] - l);
                        if (ans == item0 + j) g -= 1.0f;
                        g *= scale;
                    }
                    float h, lw;
                    split_tf32(g, h, lw);
                    hi[j] = __float_as_uint(h); lo[j] = __float_as_uint(lw);
                }
                tmem_st32(t_s + lane_base + col0, hi);
                tmem_st32(t_lo + lane_base + col0, lo);
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(g_ready);
            // while the ds MMAs run: copy the pair out to global memory (row-contiguous through the staging block) for
            // the dE = G^T s GEMM
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                const int col0 = half * 64 + (cc & 1) * 32;
                const int item0 = n * CGD_BN + col0;
                float* gdst = (cc >> 1) ? p.g_lo : p.g_hi;
                float v[32];
                tmem_ld32(((cc >> 1) ? t_lo : t_s) + lane_base + col0, v);
#pragma unroll
                for (int j = 0; j < 8; ++j) stg[lane * 8 + (j ^ (lane & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                __syncwarp();
                const int col = item0 + c4 * 4;
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const int r = t * 4 + sub;
                    const float4 x = stg[r * 8 + (c4 ^ (r & 7))];
                    if (row0 + r < p.B && col < p.cols) {
                        float* dst = gdst + (row0 + r) * p.ldg + col;
                        if (col + 3 < p.cols) *reinterpret_cast<float4*>(dst) = x;
                        else { const float xs[4] = {x.x, x.y, x.z, x.w}; for (int e = 0; e < 4; ++e) if (col + e < p.cols) dst[e] = xs[e]; }
                    }
                }
                __syncwarp();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(g_consumed);
        }
        // ---- the strip's ds partial: TMEM -> staging -> running sums in global memory
        mbar_wait(final_bar, 0);
        tc_fence_after();
        float* slab = p.partial + (long long)strip * p.B * p.d;
#pragma unroll 1
        for (int cc = 0; cc < CGD_ND / 64; ++cc) {
            const int col0 = half * (CGD_ND / 2) + cc * 32;
            float v[32];
            tmem_ld32(t_ds + lane_base + col0, v);
            if (col0 >= p.d || row0 >= p.B) continue;
#pragma unroll
            for (int j = 0; j < 8; ++j) stg[lane * 8 + (j ^ (lane & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            const int col = col0 + c4 * 4;
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const int r = t * 4 + sub;
                float4 x = stg[r * 8 + (c4 ^ (r & 7))];
                if (row0 + r < p.B && col < p.d) {           // d % 4 == 0: whole float4 in range
                    // fire-and-forget vector reduction: no read latency on the critical path.  Each element gets exactly
                    // one add per launch and launches are stream-ordered, so the summation order stays fixed.
                    float* dst = slab + (row0 + r) * p.d + col;
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
                }
            }
            __syncwarp();
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

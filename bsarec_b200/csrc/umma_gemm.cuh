// Warp-specialised tcgen05 GEMM with error-compensated TF32 ("3xTF32": hi*hi + hi*lo + lo*hi, fp32 accumulate
// in TMEM) and a pluggable epilogue.  fp32 parity (<=1e-4 on activations/gradients, identical top-k sets) rules
// out single-pass TF32/bf16, so every operand is supplied as a (hi, lo) pair of fp32 arrays
// with tf32-representable values and each k-step issues three tcgen05.mma instructions.
//
//   D[M, N] = sum_k A(m,k) * B(n,k)
//   A_MN == false: A global is [M rows, K cols] (K contiguous)  -> K-major smem tile  [128 rows x 128 B]
//   A_MN == true : A global is [K rows, M cols] (M contiguous)  -> MN-major smem slabs [32 K-rows x 128 B] x 4
//   B likewise with N in place of M.
//
// Roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warp 2 = TMEM allocator,
// warps 4-7 = epilogue (thread t owns accumulator row t of the 128-row tile; TMEM lane quarter = warp % 4).
// Pipelines: smem full/empty ring (TMA <-> MMA) and a double-buffered TMEM accumulator (MMA <-> epilogue).
// Persistent: each CTA walks the tile schedule; see Sched below.
#pragma once
#include "umma.cuh"
#include "common.cuh"

namespace bsa {
namespace umma {

constexpr int BM = 128;          // UMMA M (cta_group::1)
constexpr int BK = 32;           // fp32 elements per k-block = one 128-byte swizzle row
constexpr int UMMA_K = 8;        // tf32 elements per tcgen05.mma
constexpr int GEMM_THREADS = 256;          // 4 control warps + 4 epilogue warps (Epi::WARPS == 4)
constexpr int gemm_threads(int epi_warps) { return 128 + 32 * epi_warps; }

enum Sched : int {
    SCHED_N_FASTEST = 0,   // tile t -> (m = t / n_tiles, n = t % n_tiles): neighbouring CTAs share the A tile
    SCHED_M_FASTEST = 1,   // tile t -> (m = t % m_tiles, n = t / m_tiles): neighbouring CTAs share the B tile
    SCHED_STRIP = 2,       // CTA keeps m = cta % m_tiles and walks n = strip, strip + nstrips, ... (stateful epilogues)
};

struct GemmCore {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
    int M, N, K;
    int m_tiles, n_tiles;
    int k_blocks_total;      // ceil(K / BK)
    int splits;              // split-K factor (>= 1)
    int k_blocks_per_split;
    int sched;
};

template <int BN>
struct StageLayout {
    static constexpr int A_BYTES = BM * BK * 4;          // 16 KiB
    static constexpr int B_BYTES = BN * BK * 4;
    static constexpr int BYTES = 2 * A_BYTES + 2 * B_BYTES;
};

struct TileCoord { int m, n, split; };

__device__ __forceinline__ bool tile_of(const GemmCore& g, int it, TileCoord& tc) {
    if (g.sched == SCHED_STRIP) {
        const int nstrips = gridDim.x / g.m_tiles;
        tc.m = blockIdx.x % g.m_tiles;
        tc.n = blockIdx.x / g.m_tiles + it * nstrips;
        tc.split = 0;
        return tc.n < g.n_tiles;
    }
    const long long t = (long long)blockIdx.x + (long long)it * gridDim.x;
    const long long per_split = (long long)g.m_tiles * g.n_tiles;
    if (t >= per_split * g.splits) return false;
    tc.split = (int)(t / per_split);
    const int r = (int)(t - (long long)tc.split * per_split);
    if (g.sched == SCHED_N_FASTEST) { tc.m = r / g.n_tiles; tc.n = r % g.n_tiles; }
    else                            { tc.m = r % g.m_tiles; tc.n = r / g.m_tiles; }
    return true;
}

// Epi contract:
//   struct Params;  static constexpr int SMEM_BYTES;
//   static constexpr int WARPS;    // epilogue warps: 4 (thread = accumulator row) or 8 (two warps per TMEM lane
//                                  // quarter, each owning half of the tile's columns)
//   __device__ Epi(const Params&, const GemmCore&, unsigned char* smem, int epi_warp, int lane);
//   __device__ void tile(uint32_t acc_taddr, const TileCoord&);   // acc_taddr already carries this warp's lane base
//   __device__ void finish();                                      // once, after the CTA's last tile
template <int BN, int STAGES, bool A_MN, bool B_MN, class Epi>
__global__ void __launch_bounds__(gemm_threads(Epi::WARPS), 1)
umma_gemm_kernel(const __grid_constant__ GemmCore g, const __grid_constant__ typename Epi::Params ep) {
    using SL = StageLayout<BN>;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // carve: [stages][barriers][tmem ptr][epilogue scratch]
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* stage_base = smem;
    uint64_t* full_bar = (uint64_t*)(smem + STAGES * SL::BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_ptr = (uint32_t*)(tempty_bar + 2);
    unsigned char* epi_smem = (unsigned char*)(tmem_ptr + 4);

    const int warp = threadIdx.x >> 5;
    constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));

    if (warp == 0 && elect_one()) {
        prefetch_tmap(&g.a_hi); prefetch_tmap(&g.a_lo); prefetch_tmap(&g.b_hi); prefetch_tmap(&g.b_lo);
    }
    if (warp == 1 && elect_one()) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], Epi::WARPS); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            int c = 0;       // global k-block counter (ring position)
            TileCoord tc;
            for (int it = 0; tile_of(g, it, tc); ++it) {
                const int kb0 = tc.split * g.k_blocks_per_split;
                const int kb1 = min(g.k_blocks_total, kb0 + g.k_blocks_per_split);
                for (int kb = kb0; kb < kb1; ++kb, ++c) {
                    const int s = c % STAGES;
                    mbar_wait(&empty_bar[s], ((c / STAGES) & 1) ^ 1);
                    unsigned char* st = stage_base + s * SL::BYTES;
                    mbar_expect_tx(&full_bar[s], SL::BYTES);
                    const int k0 = kb * BK;
                    if (!A_MN) {
                        tma_load_2d(st, &g.a_hi, &full_bar[s], k0, tc.m * BM);
                        tma_load_2d(st + SL::A_BYTES, &g.a_lo, &full_bar[s], k0, tc.m * BM);
                    } else {
#pragma unroll
                        for (int q = 0; q < BM / 32; ++q) {
                            tma_load_2d(st + q * (BK * 128), &g.a_hi, &full_bar[s], tc.m * BM + q * 32, k0);
                            tma_load_2d(st + SL::A_BYTES + q * (BK * 128), &g.a_lo, &full_bar[s], tc.m * BM + q * 32, k0);
                        }
                    }
                    unsigned char* sb = st + 2 * SL::A_BYTES;
                    if (!B_MN) {
                        tma_load_2d(sb, &g.b_hi, &full_bar[s], k0, tc.n * BN);
                        tma_load_2d(sb + SL::B_BYTES, &g.b_lo, &full_bar[s], k0, tc.n * BN);
                    } else {
#pragma unroll
                        for (int q = 0; q < BN / 32; ++q) {
                            tma_load_2d(sb + q * (BK * 128), &g.b_hi, &full_bar[s], tc.n * BN + q * 32, k0);
                            tma_load_2d(sb + SL::B_BYTES + q * (BK * 128), &g.b_lo, &full_bar[s], tc.n * BN + q * 32, k0);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer
        if (elect_one()) {
            constexpr uint32_t IDESC = idesc_tf32(BM, BN, A_MN, B_MN);
            int c = 0;
            TileCoord tc;
            for (int it = 0; tile_of(g, it, tc); ++it) {
                const int buf = it & 1;
                mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t acc = tmem_base + (uint32_t)(buf * BN);
                const int kb0 = tc.split * g.k_blocks_per_split;
                const int kb1 = min(g.k_blocks_total, kb0 + g.k_blocks_per_split);
                uint32_t accumulate = 0;
                for (int kb = kb0; kb < kb1; ++kb, ++c) {
                    const int s = c % STAGES;
                    mbar_wait(&full_bar[s], (c / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(stage_base + s * SL::BYTES);
                    const uint32_t a_lo = a_hi + SL::A_BYTES;
                    const uint32_t b_hi = a_hi + 2 * SL::A_BYTES;
                    const uint32_t b_lo = b_hi + SL::B_BYTES;
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                        // K-major  (SWIZZLE_128B): advance 32 B inside the 128-B row; SBO = 1024 B between 8-row groups.
                        // MN-major (SWIZZLE_128B_BASE32B): one MMA eats 8 K-rows = 1024 B; LBO = stride between 32-wide
                        // MN slabs (BK*128 B), SBO = 512 B between 4-row swizzle groups.
                        const uint32_t a_off = A_MN ? ks * 1024 : ks * 32;
                        const uint32_t b_off = B_MN ? ks * 1024 : ks * 32;
                        constexpr uint32_t a_lbo = A_MN ? BK * 128 : 16, b_lbo = B_MN ? BK * 128 : 16;
                        constexpr uint32_t a_sbo = A_MN ? 512 : 1024, b_sbo = B_MN ? 512 : 1024;
                        constexpr uint32_t a_lt = A_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
                        constexpr uint32_t b_lt = B_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
                        const uint64_t dah = smem_desc(a_hi + a_off, a_lbo, a_sbo, a_lt);
                        const uint64_t dal = smem_desc(a_lo + a_off, a_lbo, a_sbo, a_lt);
                        const uint64_t dbh = smem_desc(b_hi + b_off, b_lbo, b_sbo, b_lt);
                        const uint64_t dbl = smem_desc(b_lo + b_off, b_lbo, b_sbo, b_lt);
                        mma_tf32(acc, dal, dbh, IDESC, accumulate);      // small terms first
                        mma_tf32(acc, dah, dbl, IDESC, 1u);
                        mma_tf32(acc, dah, dbh, IDESC, 1u);
                        accumulate = 1u;
                    }
                    mma_commit(&empty_bar[s]);          // frees the smem stage when these MMAs retire
                }
                if (kb1 <= kb0) {
                    // empty K range (K == 0): nothing was issued; the epilogue must still be released
                }
                mma_commit(&tfull_bar[buf]);            // accumulator complete
            }
        }
    } else if (warp >= 4) {
        // ===================================================================== epilogue
        const int quarter = warp & 3;          // TMEM lanes a warp may touch: 32 * (warp id % 4) ...
        Epi epi(ep, g, epi_smem, warp - 4, threadIdx.x & 31);
        TileCoord tc;
        for (int it = 0; tile_of(g, it, tc); ++it) {
            const int buf = it & 1;
            mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
            tc_fence_after();
            const uint32_t acc = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(quarter * 32) << 16);
            epi.tile(acc, tc);
            tc_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(&tempty_bar[buf]);
        }
        epi.finish();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

template <int BN, int STAGES>
constexpr int gemm_smem_bytes(int epi_bytes) {
    return 1024 /*align slack*/ + STAGES * StageLayout<BN>::BYTES + (2 * STAGES + 4) * 8 + 16 + epi_bytes;
}

}  // namespace umma
}  // namespace bsa

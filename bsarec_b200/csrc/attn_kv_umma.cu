// Key-owner half of the attention backward on tcgen05 (dK, dV), for the shapes of attn_umma.cu (head sizes 64 / 32, L <= 208).
//
// dV = Pd^T dO and dK = scale dS^T Q need the TRANSPOSED probability / score-gradient tiles as A operands.  The
// query-owner kernel (attn_bwd_q_umma_kernel) has both in registers — queries on the TMEM lanes — when it forms dS, so it
// leaves them in HBM as plain fp32 [B*H*L, NK] arrays (2 x 0.34 GB at C3, written once and read once: ~0.2 ms of HBM
// time against the 1.27 ms the mma.sync key-owner kernel took to RECOMPUTE them), and this kernel is the pattern of
// ce_bwd_logits.cuh's item owner:
//   one CTA = (sequence, head, 128 keys); keys sit on the TMEM lanes;
//   per 32 queries: TMA brings the [32 queries x 128 keys] tiles of Pd and dS (128-B swizzle) and the 32 rows of dO and Q
//   (MN-major B operands, 32-B swizzle atom, raw fp32); the 8 converter warps read their key's COLUMN of the two tiles,
//   split it into TF32 (hi, lo) pairs and store them into TMEM, and split the dO / Q rows in place (hi) / beside (lo);
//   one elected thread issues dV[128 x 64] += Pd^T[128 x 32] dO[32 x 64] and dK[128 x 64] += dS^T Q (A from TMEM, 3 MMAs
//   per 8-wide k-step).  Three stages of shared memory and three TMEM pair buffers: blocks j+1, j+2 are loaded / converted
//   while block j is multiplied (the per-block chain TMA -> convert -> MMA is latency-, not throughput-bound).
//   The accumulators leave once, row-contiguous.  No recomputation of exp / dropout here, no atomics.
// TMEM: buffer b at columns [128 b, 128 b + 128) = Pd hi | Pd lo | dS hi | dS lo (32 each); dV at [384, 448), dK at [448, 512).
#include "attn.cuh"
#include "umma.cuh"

namespace bsa {

using namespace umma;

namespace {

constexpr int KV_DH = 64;                       // largest head size (stage layout); instantiated for 64 and 32
constexpr int KV_QB = 32;                       // queries per block (K of one MMA batch)
constexpr int KV_STAGE = 64 * 1024;             // Pd tile 16 K | dS tile 16 K | dO hi 8 K | Q hi 8 K | dO lo 8 K | Q lo 8 K
constexpr int KV_OFF_DS = 16384, KV_OFF_DO = 32768, KV_OFF_Q = 40960, KV_OFF_LO = 16384;    // lo = hi + 16 K
constexpr int KV_TX = 32768 + 16384;            // bytes TMA delivers per stage (both tiles + the two raw row blocks)
constexpr int KV_EPI_WARPS = 8;
constexpr int KV_THREADS = 128 + 32 * KV_EPI_WARPS;
constexpr int KV_NST = 3;                       // pipeline depth (shared-memory stages == TMEM pair buffers)
constexpr int KV_SMEM = 1024 + KV_NST * KV_STAGE + 256;

struct AttnBwdKvParams {
    CUtensorMap pd, ds;           // [B*H*L, NK] fp32, box {32 keys, 32 queries}, 128-B swizzle
    CUtensorMap dout, q;          // [B*L, d] fp32, box {32 cols, 32 rows}, 32-B swizzle atom (MN-major B operands)
    float* dk; float* dv;
    int ldo;                      // row pitch of dk / dv (d, or 3d when dq | dk | dv are one packed [B*L, 3d] tensor)
    int H, L, d;
    float scale;
};

__device__ __forceinline__ void kv_split(float x, float& hi, float& lo) {
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    lo = x - hi;
}

template <int DH>
__global__ void __launch_bounds__(KV_THREADS, 1)
attn_bwd_kv_umma_kernel(const __grid_constant__ AttnBwdKvParams p) {
    constexpr int NCH = DH / 32;                // 32-column slabs of the dO / Q rows
    extern __shared__ __align__(1024) unsigned char kv_smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)kv_smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + KV_NST * KV_STAGE);   // stage landed                              (TMA -> converters)
    uint64_t* g_ready = full_bar + KV_NST;                        // pairs in TMEM, dO / Q rows split           (converters -> MMA)
    uint64_t* free_bar = g_ready + KV_NST;                        // MMAs reading stage / TMEM buffer retired   (MMA -> TMA)
    uint64_t* acc_full = free_bar + KV_NST;
    uint32_t* tmem_ptr = (uint32_t*)(acc_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.x, b = bh / p.H, h = bh - b * p.H;
    const int kt = blockIdx.y;
    const long long seq_base = (long long)b * p.L;
    const int col0 = h * DH;
    const int nblk = (p.L + KV_QB - 1) / KV_QB;

    if (warp == 0 && elect_one()) { prefetch_tmap(&p.pd); prefetch_tmap(&p.ds); prefetch_tmap(&p.dout); prefetch_tmap(&p.q); }
    if (warp == 1 && elect_one()) {
        for (int i = 0; i < KV_NST; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&g_ready[i], KV_EPI_WARPS); mbar_init(&free_bar[i], 1); }
        mbar_init(acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t t_dv = tmem_base + 128 * KV_NST, t_dk = t_dv + KV_DH;      // fixed columns for both head sizes

    if (warp == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            for (int j = 0; j < nblk; ++j) {
                const int s = j % KV_NST;
                mbar_wait(&free_bar[s], (uint32_t)(((j / KV_NST) & 1) ^ 1));
                unsigned char* st = smem + s * KV_STAGE;
                mbar_expect_tx(&full_bar[s], 32768 + 2 * NCH * 4096);
                const int grow = (int)((long long)bh * p.L + j * KV_QB);        // row of query j*32 in the [B*H*L, NK] arrays
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    tma_load_2d(st + q * 4096, &p.pd, &full_bar[s], kt * 128 + q * 32, grow);
                    tma_load_2d(st + KV_OFF_DS + q * 4096, &p.ds, &full_bar[s], kt * 128 + q * 32, grow);
                }
                const int xrow = (int)(seq_base + j * KV_QB);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    tma_load_2d(st + KV_OFF_DO + c * 4096, &p.dout, &full_bar[s], col0 + c * 32, xrow);
                    tma_load_2d(st + KV_OFF_Q + c * 4096, &p.q, &full_bar[s], col0 + c * 32, xrow);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer
        if (elect_one()) {
            constexpr uint32_t IDESC = idesc_tf32(128, DH, false, true);
            uint32_t acc = 0;
            for (int j = 0; j < nblk; ++j) {
                const int s = j % KV_NST;
                mbar_wait(&g_ready[s], (uint32_t)((j / KV_NST) & 1));
                tc_fence_after();
                const uint32_t tb = tmem_base + (uint32_t)(s * 128);
                const uint32_t st = smem_u32(smem + s * KV_STAGE);
                const uint32_t do_hi = st + KV_OFF_DO, do_lo = do_hi + KV_OFF_LO, q_hi = st + KV_OFF_Q, q_lo = q_hi + KV_OFF_LO;
#pragma unroll
                for (int ks = 0; ks < KV_QB / 8; ++ks) {
                    const uint32_t col = (uint32_t)(ks * 8);
                    const uint64_t doh = smem_desc(do_hi + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
                    const uint64_t dol = smem_desc(do_lo + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
                    const uint64_t qh = smem_desc(q_hi + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
                    const uint64_t ql = smem_desc(q_lo + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
                    mma_tf32_ts(t_dv, tb + 32 + col, doh, IDESC, acc);        // Pd lo * dO hi
                    mma_tf32_ts(t_dv, tb + col, dol, IDESC, 1u);              // Pd hi * dO lo
                    mma_tf32_ts(t_dv, tb + col, doh, IDESC, 1u);              // Pd hi * dO hi
                    mma_tf32_ts(t_dk, tb + 96 + col, qh, IDESC, acc);         // dS lo * Q hi
                    mma_tf32_ts(t_dk, tb + 64 + col, ql, IDESC, 1u);
                    mma_tf32_ts(t_dk, tb + 64 + col, qh, IDESC, 1u);
                    acc = 1u;
                }
                mma_commit(&free_bar[s]);
            }
            mma_commit(acc_full);
        }
    } else if (warp >= 4) {
        // ===================================================================== converters / epilogue
        const int ew = warp - 4, quarter = ew & 3, half = ew >> 2;        // half 0: Pd -> dV, half 1: dS -> dK
        const uint32_t lane_base = ((uint32_t)(quarter * 32) << 16);
        const int ct = (warp - 4) * 32 + lane;                           // 0..255 over the converter threads
        for (int j = 0; j < nblk; ++j) {
            const int s = j % KV_NST;
            mbar_wait(&full_bar[s], (uint32_t)((j / KV_NST) & 1));
            unsigned char* st = smem + s * KV_STAGE;
            // this thread's key is column `lane` of box `quarter`; the block's 32 queries are its 32 K-columns
            const unsigned char* base = st + half * KV_OFF_DS + quarter * 4096 + ((lane & 3) << 2);
            const int nq = p.L - j * KV_QB;                              // queries of this block inside the sequence
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int u = 0; u < 32; ++u) {
                float v = *reinterpret_cast<const float*>(base + u * 128 + ((((lane >> 2) ^ (u & 7))) << 4));
                if (u >= nq) v = 0.f;                                    // rows past L belong to the next (sequence, head)
                float hh, lw;
                kv_split(v, hh, lw);
                hi[u] = __float_as_uint(hh); lo[u] = __float_as_uint(lw);
            }
            // dO and Q rows: raw fp32 stays as hi, lo 16 KiB further (same swizzled positions); 16 KiB / 256 threads
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                unsigned char* ptr = st + KV_OFF_DO + k * 4096 + ct * 16;
                const float4 x = *reinterpret_cast<const float4*>(ptr);
                // the raw words stay as the hi half (the MMA reads their upper 19 bits); lo = x - truncate_tf32(x)
                auto lo1 = [](float t) { return t - __uint_as_float(__float_as_uint(t) & 0xFFFFE000u); };
                *reinterpret_cast<float4*>(ptr + KV_OFF_LO) = make_float4(lo1(x.x), lo1(x.y), lo1(x.z), lo1(x.w));
            }
            fence_proxy_async_smem();
            const uint32_t tb = tmem_base + (uint32_t)(s * 128 + half * 64) + lane_base;
            tmem_st32(tb, hi);
            tmem_st32(tb + 32, lo);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&g_ready[s]);
        }
        // ---- accumulators: half 0 stores dV, half 1 stores scale * dK; thread = key row, 256 contiguous bytes
        mbar_wait(acc_full, 0);
        tc_fence_after();
        const int jkey = kt * 128 + quarter * 32 + lane;
        const float mul = half ? p.scale : 1.0f;
        float* dst = (half ? p.dk : p.dv) + (seq_base + (jkey < p.L ? jkey : 0)) * p.ldo + col0;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
            float v[32];
            tmem_ld32((half ? t_dk : t_dv) + lane_base + c * 32, v);
            if (jkey < p.L) {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    *reinterpret_cast<float4*>(dst + c * 32 + e * 4) =
                        make_float4(v[4 * e] * mul, v[4 * e + 1] * mul, v[4 * e + 2] * mul, v[4 * e + 3] * mul);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

int attn_bwd_kv_umma(const float* q, const float* dctx, const float* pd, const float* ds, float* dk, float* dv, int B, int H,
                     int L, int d, cudaStream_t st, int ldo) {
    if (B == 0) return 0;
    static bool configured = false;
    if (!configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kv_umma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, KV_SMEM));
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kv_umma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, KV_SMEM));
        configured = true;
    }
    const int dh = d / H;
    AttnBwdKvParams p;
    const int NK = (L + 15) / 16 * 16;
    const long long rows = (long long)B * L, grows = (long long)B * H * L;
    int rc;
    if ((rc = make_tmap_2d(&p.pd, pd, grows, NK, NK, KV_QB, false))) return rc;
    if ((rc = make_tmap_2d(&p.ds, ds, grows, NK, NK, KV_QB, false))) return rc;
    if ((rc = make_tmap_2d(&p.dout, dctx, rows, d, d, KV_QB, true))) return rc;
    if ((rc = make_tmap_2d(&p.q, q, rows, d, d, KV_QB, true))) return rc;
    p.dk = dk; p.dv = dv; p.H = H; p.L = L; p.d = d; p.ldo = ldo ? ldo : d;
    p.scale = 1.0f / sqrtf((float)dh);
    const dim3 grid(B * H, (L + 127) / 128);
    if (dh == 64) attn_bwd_kv_umma_kernel<64><<<grid, KV_THREADS, KV_SMEM, st>>>(p);
    else          attn_bwd_kv_umma_kernel<32><<<grid, KV_THREADS, KV_SMEM, st>>>(p);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

// Causal multi-head attention forward on the 5th-generation tensor cores (tcgen05 + TMEM), for sequences whose whole
// key axis fits one accumulator: L <= 208 (C1: L = 50, C2/C3: L = 200).  Same semantics as attn.cu (the mma.sync
// kernel, which stays for longer sequences and for the backward pass): model/_modules.py:108-133 with the literal
// additive -10000 mask of _abstract_model.py:53-69, dropout on the probabilities, error-compensated TF32.
//
// One CTA = (sequence b, head h, 128 query rows).
//   TMA   : Q tile, K rows, V rows of the head arrive as raw fp32 (128-byte swizzle; V in the 32-byte-atom variant
//           that MN-major tf32 operands need); every thread then splits them IN PLACE into the (hi, lo) pair of the
//           3xTF32 scheme — elementwise, so the swizzled layout is untouched — and publishes with fence.proxy.async.
//   MMA 1 : S[128, NK] = Q K^T          (both operands from shared memory, accumulator in TMEM columns [0, NK))
//   all 8 warps (two per TMEM lane quarter, alternate 32-column chunks): two sweeps over the row in TMEM — max, then
//           exp / sum / dropout — and the probabilities go BACK into TMEM as a split pair (hi over S, lo beside it).
//   MMA 2 : O[128, dh] = P V            (A operand read from TMEM, B = V rows as MN-major operand)
//   epilogue: O / rowsum -> ctx, (rowmax, log rowsum) -> lse for the backward pass.
// TMEM columns: [0,224) S then P_hi, [224,448) P_lo, [448,512) O.
#include "attn.cuh"
#include "umma.cuh"

namespace bsa {
#ifdef BSA_ATTN_STAMPS
// Phase-stamp instrumentation (profiling builds only): one mid-grid CTA records globaltimer at each phase boundary.
__device__ unsigned long long g_attn_stamps[64];      // [q-tile][16 forward stamps, 16 backward stamps]
#define BSA_STAMP(i) do { if (threadIdx.x == 0 && blockIdx.x == gridDim.x / 2 && blockIdx.y < 2) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_attn_stamps[blockIdx.y * 32 + (i)] = t_; } } while (0)
#else
#define BSA_STAMP(i) do {} while (0)
#endif

using namespace umma;

namespace {

constexpr int AU_DH = 64;                  // largest head size (the shared-memory plan); kernels are instantiated for 64 and 32
constexpr int AU_NK_MAX = 208;             // keys per sequence (UMMA N, multiple of 16)
constexpr int AU_SCOLS = 224;              // TMEM columns reserved for S / P_hi (7 chunks of 32)
constexpr int AUF_PARTS = 4;                   // 4 warps per TMEM lane quarter (the elementwise passes are latency-bound)
constexpr int AUF_THREADS = 128 * AUF_PARTS;
constexpr int AU_Q_BYTES = 128 * AU_DH * 4;                 // 32 KiB: 2 k-blocks x [128 rows x 128 B]
constexpr int AU_K_BYTES = AU_NK_MAX * AU_DH * 4;           // 52 KiB: 2 k-blocks x [NK rows x 128 B]
constexpr int AU_V_BYTES = 7 * 2 * 4096;                    // 56 KiB: 7 k-blocks x 2 column chunks x [32 rows x 128 B]
constexpr int AU_OFF_QHI = 0;
constexpr int AU_OFF_QLO = AU_OFF_QHI + AU_Q_BYTES;
constexpr int AU_OFF_KHI = AU_OFF_QLO + AU_Q_BYTES;         // after MMA 1: V_lo (56 KiB, runs on into K_lo)
constexpr int AU_OFF_KLO = AU_OFF_KHI + AU_K_BYTES;
constexpr int AU_OFF_V = AU_OFF_KLO + AU_K_BYTES;
constexpr int AU_OFF_BAR = AU_OFF_V + AU_V_BYTES;
constexpr int AU_OFF_DELTA = AU_OFF_BAR + 128 + 256;         // backward: delta of the CTA's 128 query rows
constexpr int AU_SMEM = 1024 + AU_OFF_DELTA + 512;           // alignment slack + barriers + valid[] + delta[]
static_assert(AU_SMEM <= 232448, "attention tile does not fit shared memory");
static_assert(AU_V_BYTES <= 2 * AU_K_BYTES, "V_lo must fit the retired K region");

struct AttnUmmaParams {
    CUtensorMap q, k, v;          // [B*L, d] fp32: box 32 x 128, box 32 x NK/2, box 32 x 32 (atom32)
    const long long* ids;
    float* ctx; float* lse;
    float* ctx_lo;                // optional: ctx - truncate_tf32(ctx), the lo half of the output projection's A operand
    int H, L, d, NK;
    float scale;
    Dropout dr;
};

__device__ __forceinline__ void split_tf32_dev(float x, float& hi, float& lo) {
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    lo = x - hi;
}

// raw fp32 region -> the raw words stay as the hi half (a tf32 MMA reads only their upper 19 bits), lo = x - truncate_tf32(x)
// goes to +lo_off (same swizzled positions): one shared-memory store per element instead of two
__device__ __forceinline__ float lo_trunc_dev(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ void split_region(unsigned char* base, int bytes, int lo_off, int tid, int nthreads) {
    for (int o = tid * 16; o < bytes; o += nthreads * 16) {
        const float4 x = *reinterpret_cast<const float4*>(base + o);
        *reinterpret_cast<float4*>(base + o + lo_off) = make_float4(lo_trunc_dev(x.x), lo_trunc_dev(x.y), lo_trunc_dev(x.z), lo_trunc_dev(x.w));
    }
}

// bits [0, width) of: columns j0 + b that are <= limit  (limit - j0 >= width - 1: all ones; < 0: none)
__device__ __forceinline__ uint32_t low_bits_upto(int limit, int j0) {
    const int rel = limit - j0;
    return rel >= 31 ? 0xffffffffu : (rel < 0 ? 0u : ((2u << rel) - 1u));
}
// valid[] bytes -> one bit per key, 8 words in shared memory (one ballot per word by warp 0)
__device__ __forceinline__ void pack_valid_bits(const unsigned char* valid, uint32_t* vbits, int warp, int lane) {
    if (warp == 0) {
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            const uint32_t word = __ballot_sync(0xffffffffu, valid[w * 32 + lane] != 0);
            if (lane == 0) vbits[w] = word;
        }
    }
}

template <int DH>
__global__ void __launch_bounds__(AUF_THREADS, 1)
attn_fwd_umma_kernel(const __grid_constant__ AttnUmmaParams p) {
    constexpr int NKB = DH / 32;                 // 32-float k-blocks of the head dimension
    constexpr int VKB = NKB * 4096;              // bytes of one 32-key block of the MN-major V operand
    extern __shared__ __align__(1024) unsigned char au_smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)au_smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bar_qk = (uint64_t*)(smem + AU_OFF_BAR);       // Q and K tiles landed              (TMA -> all)
    uint64_t* bar_v = bar_qk + 1;                            // V rows landed                     (TMA -> warps 0-3)
    uint64_t* s_full = bar_qk + 2;                           // S complete in TMEM                (MMA -> all)
    uint64_t* o_full = bar_qk + 5;                           // O complete in TMEM                (MMA -> warps 4-7)
    uint32_t* tmem_ptr = (uint32_t*)(bar_qk + 6);
    unsigned char* valid = smem + AU_OFF_BAR + 128;          // [AU_SCOLS + 32]: 1 = key j is an item, 0 = padding / j >= L
    uint32_t* vbits = (uint32_t*)(smem + AU_OFF_BAR + 64);   // the same as 8 x 32 bits (the mask test is a bit test per score)
    int* s_first = (int*)(bar_qk + 7);                       // first position holding an item (L if none)
    float* xch = (float*)(smem + AU_OFF_QHI);                // [2][AUF_PARTS][128]: row max / row sum parts (Q is dead after MMA 1)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.x, b = bh / p.H, h = bh - b * p.H;
    const int qt = blockIdx.y;
    const long long seq_base = (long long)b * p.L;
    const int NK = p.NK, nkb = (NK + 31) / 32;
    const int col0 = h * DH;
    const int k_kb_bytes = NK * 128;                         // one 32-float k-block of the K operand

    BSA_STAMP(7);
    // The operand loads are issued first: the ids reads and the TMEM allocation below run under their latency.
    if (warp == 0 && elect_one()) {
        mbar_init(bar_qk, 1); mbar_init(bar_v, 1); mbar_init(s_full, 1);
        mbar_init(o_full, 1);
        fence_barrier_init();
        mbar_expect_tx(bar_qk, (uint32_t)(NKB * (16384 + k_kb_bytes)));
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
            tma_load_2d(smem + AU_OFF_QHI + kb * 16384, &p.q, bar_qk, col0 + kb * 32, (int)(seq_base + qt * 128));
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes, &p.k, bar_qk, col0 + kb * 32, (int)seq_base);
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes + (NK / 2) * 128, &p.k, bar_qk, col0 + kb * 32, (int)seq_base + NK / 2);
        }
        mbar_expect_tx(bar_v, (uint32_t)(nkb * VKB));
        for (int kb = 0; kb < nkb; ++kb) {
#pragma unroll
            for (int c = 0; c < NKB; ++c)
                tma_load_2d(smem + AU_OFF_V + kb * VKB + c * 4096, &p.v, bar_v, col0 + c * 32, (int)seq_base + kb * 32);
        }
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    if (threadIdx.x == 0) *s_first = p.L;
    for (int j = threadIdx.x; j < AU_SCOLS + 32; j += AUF_THREADS)
        valid[j] = (j < p.L && p.ids[seq_base + j] > 0) ? 1 : 0;
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < p.L && valid[threadIdx.x]) atomicMin(s_first, (int)threadIdx.x);   // L <= 208 < blockDim
    pack_valid_bits(valid, vbits, warp, threadIdx.x & 31);
    __syncthreads();
    BSA_STAMP(0);
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t t_s = tmem_base, t_lo = tmem_base + AU_SCOLS, t_o = tmem_base + 2 * AU_SCOLS;

    // ---- Q, K: raw fp32 -> (hi, lo), all 8 warps
    mbar_wait(bar_qk, 0);
    BSA_STAMP(1);
    split_region(smem + AU_OFF_QHI, NKB * 16384, AU_OFF_QLO - AU_OFF_QHI, threadIdx.x, AUF_THREADS);
    split_region(smem + AU_OFF_KHI, NKB * k_kb_bytes, AU_OFF_KLO - AU_OFF_KHI, threadIdx.x, AUF_THREADS);
    fence_proxy_async_smem();
    __syncthreads();
    BSA_STAMP(2);

    if (warp == 1 && elect_one()) {
        // ===================================================================== MMA 1: S = Q K^T
        const uint32_t idesc = idesc_tf32(128, NK, false, false);
        const uint32_t qh = smem_u32(smem + AU_OFF_QHI), ql = smem_u32(smem + AU_OFF_QLO);
        const uint32_t kh = smem_u32(smem + AU_OFF_KHI), kl = smem_u32(smem + AU_OFF_KLO);
        uint32_t acc = 0;
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint64_t dah = smem_desc(qh + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dal = smem_desc(ql + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbh = smem_desc(kh + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbl = smem_desc(kl + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                mma_tf32(t_s, dal, dbh, idesc, acc);
                mma_tf32(t_s, dah, dbl, idesc, 1u);
                mma_tf32(t_s, dah, dbh, idesc, 1u);
                acc = 1u;
            }
        }
        mma_commit(s_full);
    }
    __syncwarp();
    // ---- V: raw fp32 -> hi in place, lo into the K region once MMA 1 has retired
    mbar_wait(bar_v, 0);
    mbar_wait(s_full, 0);
    BSA_STAMP(3);
    tc_fence_after();
    split_region(smem + AU_OFF_V, nkb * VKB, AU_OFF_KHI - AU_OFF_V, threadIdx.x, AUF_THREADS);
    fence_proxy_async_smem();

    // ========================================================================= softmax: thread = (query row, column part)
    // Two warps share each TMEM lane quarter and take alternate 32-column chunks; row max and row sum are exchanged
    // through shared memory.  A chunk every pair of which is masked for rows that do have a permitted key contributes
    // exp(-10000 - max) == 0 exactly: it is not read, P = 0 is stored.
    const int quarter = warp & 3, part = warp >> 2;        // 4 warps per TMEM lane quarter: chunks c = part, part + 4, ...
    const uint32_t lane_base = ((uint32_t)(quarter * 32) << 16);
    const int i0 = qt * 128 + quarter * 32;
    const int i = i0 + lane;                                       // position in the sequence
    const bool row_ok = i < p.L;
    const int nch = (NK + 31) / 32;
    const int first_valid = *s_first;
    const bool can_skip = i0 >= first_valid;                       // every row of this warp has a permitted key
    const unsigned long long pbase = ((unsigned long long)bh * p.L + (unsigned long long)(row_ok ? i : 0)) * p.L;
    unsigned live = 0;
    for (int c = part; c < nch; c += AUF_PARTS)
        if (!(can_skip && (c * 32 > i0 + 31 || c * 32 + 31 < first_valid))) live |= 1u << c;
    float m = -INFINITY;
#pragma unroll 1
    for (int c = part; c < nch; c += AUF_PARTS) {                  // sweep 1: row max of the masked, scaled scores
        if (!(live >> c & 1)) continue;
        float v[32];
        tmem_ld32(t_s + lane_base + c * 32, v);
        const uint32_t allowed = vbits[c] & low_bits_upto(i, c * 32), in_l = low_bits_upto(p.L - 1, c * 32);
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
            const float s = v[jj] * p.scale + ((allowed >> jj & 1u) ? 0.f : -10000.f);
            if (in_l >> jj & 1u) m = fmaxf(m, s);
        }
    }
    xch[part * 128 + quarter * 32 + lane] = m;
    asm volatile("bar.sync %0, 128;" ::"r"(quarter + 1) : "memory");
#pragma unroll
    for (int q = 0; q < AUF_PARTS; ++q) m = fmaxf(m, xch[q * 128 + quarter * 32 + lane]);
    float l = 0.f;
    const bool fast_rng = (pbase & 3ull) == 0ull;
    const Dropout drr = dropout_resolve(p.dr);                     // the site key, formed once per thread
#pragma unroll 1
    for (int c = part; c < nch; c += AUF_PARTS) {                  // sweep 2: exp, sum, dropout, split, back into TMEM
        uint32_t hi[32], lo[32];
        if (live >> c & 1) {
            float v[32];
            tmem_ld32(t_s + lane_base + c * 32, v);
            const uint32_t allowed = vbits[c] & low_bits_upto(i, c * 32);
            const uint32_t in_l = row_ok ? low_bits_upto(p.L - 1, c * 32) : 0u;
#pragma unroll
            for (int g4 = 0; g4 < 8; ++g4) {
                float mult[4];
                const int j0 = c * 32 + g4 * 4;
                if (fast_rng) dropout_mult4(drr, pbase + j0, mult);
                else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) mult[e] = dropout_mult(drr, pbase + j0 + e);
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int jj = g4 * 4 + e;
                    const float s = v[jj] * p.scale + ((allowed >> jj & 1u) ? 0.f : -10000.f);
                    float pr = 0.f;
                    if (in_l >> jj & 1u) {
                        pr = fast_exp(s - m);
                        l += pr;
                        pr *= mult[e];
                    }
                    float h, lw;
                    split_tf32_dev(pr, h, lw);
                    hi[g4 * 4 + e] = __float_as_uint(h); lo[g4 * 4 + e] = __float_as_uint(lw);
                }
            }
        } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) { hi[e] = 0u; lo[e] = 0u; }
        }
        tmem_st32(t_s + lane_base + c * 32, hi);
        tmem_st32(t_lo + lane_base + c * 32, lo);
    }
    xch[AUF_PARTS * 128 + part * 128 + quarter * 32 + lane] = l;
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();                                               // P pair in TMEM, V pair in smem, row sums in xch
    BSA_STAMP(4);
    l = 0.f;
#pragma unroll
    for (int q = 0; q < AUF_PARTS; ++q) l += xch[AUF_PARTS * 128 + q * 128 + quarter * 32 + lane];

    if (warp == 1 && elect_one()) {
        // ===================================================================== MMA 2: O = P V
        tc_fence_after();
        const uint32_t idesc = idesc_tf32(128, DH, false, true);
        const uint32_t vh = smem_u32(smem + AU_OFF_V), vl = smem_u32(smem + AU_OFF_KHI);
        uint32_t acc = 0;
        for (int kstep = 0; kstep < NK / 8; ++kstep) {
            const int kb = kstep >> 2, ks = kstep & 3;
            const uint64_t dbh = smem_desc(vh + kb * VKB + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
            const uint64_t dbl = smem_desc(vl + kb * VKB + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
            const uint32_t col = (uint32_t)(kstep * 8);
            mma_tf32_ts(t_o, t_lo + col, dbh, idesc, acc);
            mma_tf32_ts(t_o, t_s + col, dbl, idesc, 1u);
            mma_tf32_ts(t_o, t_s + col, dbh, idesc, 1u);
            acc = 1u;
        }
        mma_commit(o_full);
    }
    __syncwarp();
    // ========================================================================= epilogue: parts 0 and 1 store O columns [32 part, 32 part + 32)
    mbar_wait(o_full, 0);
    BSA_STAMP(5);
    tc_fence_after();
    if (part < NKB) {
        const int half = part;
        const float inv = row_ok ? 1.0f / l : 0.f;
        float* dst = p.ctx + (seq_base + (row_ok ? i : 0)) * p.d + col0 + half * 32;
        float v[32];
        tmem_ld32(t_o + lane_base + half * 32, v);
        if (row_ok) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float4 o = make_float4(v[4 * e] * inv, v[4 * e + 1] * inv, v[4 * e + 2] * inv, v[4 * e + 3] * inv);
                *reinterpret_cast<float4*>(dst + e * 4) = o;
                if (p.ctx_lo)
                    *reinterpret_cast<float4*>(p.ctx_lo + (dst - p.ctx) + e * 4) =
                        make_float4(lo_trunc_dev(o.x), lo_trunc_dev(o.y), lo_trunc_dev(o.z), lo_trunc_dev(o.w));
            }
            if (half == 0) *reinterpret_cast<float2*>(p.lse + 2 * ((long long)bh * p.L + i)) = make_float2(m, logf(l));
        }
    }
    tc_fence_before();
    __syncthreads();
    BSA_STAMP(6);
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}


// ============================================================================ backward, query owner (dQ and delta)
// Same tile and the same TMEM plan as the forward: S = Q K^T -> columns [0,224); dPd = dO V^T -> columns [224,448); the
// saved (row max, log row sum) turn S into P, dS = P o (dPd o M - delta) replaces S / dPd as a split pair, and
// dQ = scale dS K is the third MMA (A from TMEM, B = the K rows as an MN-major operand) into columns [448,512).
// Shared memory is reused phase by phase: region 1 holds Q then dO, region 2 holds K then V (K-major), region 3 the K
// rows in the MN-major layout (its lo half moves into region 2 once the dPd MMA has retired).  dK / dV stay on the
// key-owner mma.sync kernel (they need P^T / dS^T as A operands), which reads the delta written here.
struct AttnBwdQParams {
    CUtensorMap q, k, dout, vk, km;    // q, dout: box 32 x 128; k, vk: box 32 x NK/2; km: K rows, box 32 x 32 (atom32)
    const long long* ids;
    const float* ctx; const float* dctx; const float* lse;
    float* dq; float* delta;
    float* pd_out; float* ds_out;      // optional [B*H*L, NK] fp32: Pd = P o M and dS, left for the key-owner kernel (attn_kv_umma.cu)
    int ldo;                           // row pitch of dq (d, or 3d when dq | dk | dv are one packed [B*L, 3d] tensor)
    int H, L, d, NK;
    float scale;
    Dropout dr;
};

template <int DH>
__global__ void __launch_bounds__(AUF_THREADS, 1)
attn_bwd_q_umma_kernel(const __grid_constant__ AttnBwdQParams p) {
    constexpr int NKB = DH / 32;
    constexpr int VKB = NKB * 4096;
    extern __shared__ __align__(1024) unsigned char au_smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)au_smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bar_qk = (uint64_t*)(smem + AU_OFF_BAR);       // Q and K tiles landed
    uint64_t* bar_km = bar_qk + 1;                           // K rows (MN-major layout) landed
    uint64_t* s_full = bar_qk + 2;                           // S complete in TMEM
    uint64_t* bar_dov = bar_qk + 3;                          // dO and V tiles landed
    uint64_t* dp_full = bar_qk + 4;                          // dPd complete in TMEM
    uint64_t* o_full = bar_qk + 5;                           // dQ complete in TMEM
    uint32_t* tmem_ptr = (uint32_t*)(bar_qk + 6);
    unsigned char* valid = smem + AU_OFF_BAR + 128;
    uint32_t* vbits = (uint32_t*)(smem + AU_OFF_BAR + 64);
    int* s_first = (int*)(bar_qk + 7);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.x, b = bh / p.H, h = bh - b * p.H;
    const int qt = blockIdx.y;
    const long long seq_base = (long long)b * p.L;
    const int NK = p.NK, nkb = (NK + 31) / 32;
    const int col0 = h * DH;
    const int k_kb_bytes = NK * 128;

    BSA_STAMP(31);
    if (warp == 0 && elect_one()) {
        mbar_init(bar_qk, 1); mbar_init(bar_km, 1); mbar_init(s_full, 1); mbar_init(bar_dov, 1); mbar_init(dp_full, 1);
        mbar_init(o_full, 1);
        fence_barrier_init();
        mbar_expect_tx(bar_qk, (uint32_t)(NKB * (16384 + k_kb_bytes)));
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
            tma_load_2d(smem + AU_OFF_QHI + kb * 16384, &p.q, bar_qk, col0 + kb * 32, (int)(seq_base + qt * 128));
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes, &p.k, bar_qk, col0 + kb * 32, (int)seq_base);
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes + (NK / 2) * 128, &p.k, bar_qk, col0 + kb * 32, (int)seq_base + NK / 2);
        }
        mbar_expect_tx(bar_km, (uint32_t)(nkb * VKB));
        for (int kb = 0; kb < nkb; ++kb) {
#pragma unroll
            for (int c = 0; c < NKB; ++c)
                tma_load_2d(smem + AU_OFF_V + kb * VKB + c * 4096, &p.km, bar_km, col0 + c * 32, (int)seq_base + kb * 32);
        }
        prefetch_tmap(&p.dout); prefetch_tmap(&p.vk);
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    if (threadIdx.x == 0) *s_first = p.L;
    for (int j = threadIdx.x; j < AU_SCOLS + 32; j += AUF_THREADS)
        valid[j] = (j < p.L && p.ids[seq_base + j] > 0) ? 1 : 0;
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < p.L && valid[threadIdx.x]) atomicMin(s_first, (int)threadIdx.x);
    pack_valid_bits(valid, vbits, warp, threadIdx.x & 31);
    __syncthreads();
    BSA_STAMP(16);
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t t_s = tmem_base, t_lo = tmem_base + AU_SCOLS, t_o = tmem_base + 2 * AU_SCOLS;

    mbar_wait(bar_qk, 0);
    BSA_STAMP(17);
    split_region(smem + AU_OFF_QHI, NKB * 16384, AU_OFF_QLO - AU_OFF_QHI, threadIdx.x, AUF_THREADS);
    split_region(smem + AU_OFF_KHI, NKB * k_kb_bytes, AU_OFF_KLO - AU_OFF_KHI, threadIdx.x, AUF_THREADS);
    fence_proxy_async_smem();
    __syncthreads();
    BSA_STAMP(18);

    const uint32_t idesc_s = idesc_tf32(128, NK, false, false);
    if (warp == 1 && elect_one()) {
        // ===================================================================== MMA 1: S = Q K^T
        const uint32_t qh = smem_u32(smem + AU_OFF_QHI), ql = smem_u32(smem + AU_OFF_QLO);
        const uint32_t kh = smem_u32(smem + AU_OFF_KHI), kl = smem_u32(smem + AU_OFF_KLO);
        uint32_t acc = 0;
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint64_t dah = smem_desc(qh + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dal = smem_desc(ql + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbh = smem_desc(kh + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbl = smem_desc(kl + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                mma_tf32(t_s, dal, dbh, idesc_s, acc);
                mma_tf32(t_s, dah, dbl, idesc_s, 1u);
                mma_tf32(t_s, dah, dbh, idesc_s, 1u);
                acc = 1u;
            }
        }
        mma_commit(s_full);
    }
    __syncwarp();

    // ---- per-thread row state while MMA 1 runs: thread = (query row, column half)
    const int quarter = warp & 3, part = warp >> 2;        // 4 warps per TMEM lane quarter, 16-column pieces c = part, part + 4, ...
    const uint32_t lane_base = ((uint32_t)(quarter * 32) << 16);
    const int i0 = qt * 128 + quarter * 32;
    const int i = i0 + lane;
    const bool row_ok = i < p.L;
    const int nch = NK / 16;                                       // 16-column pieces (NK is a multiple of 16)
    const int first_valid = *s_first;
    const bool can_skip = i0 >= first_valid;
    const unsigned long long pbase = ((unsigned long long)bh * p.L + (unsigned long long)(row_ok ? i : 0)) * p.L;
    unsigned live = 0;
    for (int c = part; c < nch; c += AUF_PARTS)
        if (!(can_skip && (c * 16 > i0 + 31 || c * 16 + 15 < first_valid))) live |= 1u << c;
    float2 ls = make_float2(0.f, 0.f);
    if (row_ok) ls = *reinterpret_cast<const float2*>(p.lse + 2 * ((long long)bh * p.L + i));
    {   // delta[row] = <ctx[row], dctx[row]> over the head's columns: half a warp per row (coalesced 256-byte rows),
        // 8 rows per warp; a row-per-thread read of the two tensors costs 5 - 15 us of exposed latency per CTA
        float* s_delta = reinterpret_cast<float*>(smem + AU_OFF_DELTA);
        const int hl = lane >> 4, c4 = lane & 15;
        float4 a[4], g[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int ri = qt * 128 + warp * 8 + it * 2 + hl;
            a[it] = g[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ri < p.L && c4 < DH / 4) {
                a[it] = __ldg(reinterpret_cast<const float4*>(p.ctx + (seq_base + ri) * p.d + col0) + c4);
                g[it] = __ldg(reinterpret_cast<const float4*>(p.dctx + (seq_base + ri) * p.d + col0) + c4);
            }
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            float dsum = a[it].x * g[it].x + a[it].y * g[it].y + a[it].z * g[it].z + a[it].w * g[it].w;
#pragma unroll
            for (int o = 8; o >= 1; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
            const int r = warp * 8 + it * 2 + hl, ri = qt * 128 + r;
            if (c4 == 0) {
                s_delta[r] = dsum;
                if (ri < p.L) p.delta[(long long)bh * p.L + ri] = dsum;
            }
        }
    }

    // ---- Q and K are dead once S is complete: dO and V take their place
    mbar_wait(s_full, 0);
    BSA_STAMP(19);
    tc_fence_after();
    if (warp == 0 && elect_one()) {
        fence_proxy_async_smem();
        mbar_expect_tx(bar_dov, (uint32_t)(NKB * (16384 + k_kb_bytes)));
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
            tma_load_2d(smem + AU_OFF_QHI + kb * 16384, &p.dout, bar_dov, col0 + kb * 32, (int)(seq_base + qt * 128));
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes, &p.vk, bar_dov, col0 + kb * 32, (int)seq_base);
            tma_load_2d(smem + AU_OFF_KHI + kb * k_kb_bytes + (NK / 2) * 128, &p.vk, bar_dov, col0 + kb * 32, (int)seq_base + NK / 2);
        }
    }
    __syncwarp();
    mbar_wait(bar_dov, 0);
    BSA_STAMP(20);
    split_region(smem + AU_OFF_QHI, NKB * 16384, AU_OFF_QLO - AU_OFF_QHI, threadIdx.x, AUF_THREADS);
    split_region(smem + AU_OFF_KHI, NKB * k_kb_bytes, AU_OFF_KLO - AU_OFF_KHI, threadIdx.x, AUF_THREADS);
    fence_proxy_async_smem();
    __syncthreads();
    BSA_STAMP(21);

    if (warp == 1 && elect_one()) {
        // ===================================================================== MMA 2: dPd = dO V^T
        const uint32_t gh = smem_u32(smem + AU_OFF_QHI), gl = smem_u32(smem + AU_OFF_QLO);
        const uint32_t vh = smem_u32(smem + AU_OFF_KHI), vl = smem_u32(smem + AU_OFF_KLO);
        uint32_t acc = 0;
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint64_t dah = smem_desc(gh + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dal = smem_desc(gl + kb * 16384 + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbh = smem_desc(vh + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                const uint64_t dbl = smem_desc(vl + kb * k_kb_bytes + ks * 32, 16, 1024, LAYOUT_SW128);
                mma_tf32(t_lo, dal, dbh, idesc_s, acc);
                mma_tf32(t_lo, dah, dbl, idesc_s, 1u);
                mma_tf32(t_lo, dah, dbh, idesc_s, 1u);
                acc = 1u;
            }
        }
        mma_commit(dp_full);
    }
    __syncwarp();
    mbar_wait(bar_km, 0);
    mbar_wait(dp_full, 0);
    BSA_STAMP(22);
    tc_fence_after();
    // K rows (MN-major layout): hi in place, lo into region 2 (V is dead now)
    split_region(smem + AU_OFF_V, nkb * VKB, AU_OFF_KHI - AU_OFF_V, threadIdx.x, AUF_THREADS);
    fence_proxy_async_smem();

    // ========================================================================= dS = P o (dPd o M - delta), back into TMEM
    const float delta = reinterpret_cast<const float*>(smem + AU_OFF_DELTA)[quarter * 32 + lane];
    const bool fast_rng = (pbase & 3ull) == 0ull;
    const Dropout drr = dropout_resolve(p.dr);                     // the site key, formed once per thread
    float* pd_row = (p.pd_out && row_ok) ? p.pd_out + ((long long)bh * p.L + i) * NK : nullptr;
    float* ds_row = (p.ds_out && row_ok) ? p.ds_out + ((long long)bh * p.L + i) * NK : nullptr;
#pragma unroll 1
    for (int c = part; c < nch; c += AUF_PARTS) {
        uint32_t hi[16], lo[16];
        if (live >> c & 1) {
            float vs[16], vd[16];
            tmem_ld16(t_s + lane_base + c * 16, vs);
            tmem_ld16(t_lo + lane_base + c * 16, vd);
            const uint32_t allowed = (vbits[(c * 16) >> 5] >> ((c * 16) & 31)) & low_bits_upto(i, c * 16) & 0xffffu;
            const uint32_t in_l = row_ok ? (low_bits_upto(p.L - 1, c * 16) & 0xffffu) : 0u;
#pragma unroll
            for (int g4 = 0; g4 < 4; ++g4) {
                float mult[4];
                const int j0 = c * 16 + g4 * 4;
                if (fast_rng) dropout_mult4(drr, pbase + j0, mult);
                else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) mult[e] = dropout_mult(drr, pbase + j0 + e);
                }
                float pd4[4], ds4[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int jj = g4 * 4 + e;
                    float ds = 0.f, pd = 0.f;
                    if (in_l >> jj & 1u) {
                        const float sp = vs[jj] * p.scale + ((allowed >> jj & 1u) ? 0.f : -10000.f);
                        const float pr = fast_exp((sp - ls.x) - ls.y);
                        pd = pr * mult[e];
                        ds = pr * (vd[g4 * 4 + e] * mult[e] - delta);
                    }
                    pd4[e] = pd; ds4[e] = ds;
                    float hh, lw;
                    split_tf32_dev(ds, hh, lw);
                    hi[g4 * 4 + e] = __float_as_uint(hh); lo[g4 * 4 + e] = __float_as_uint(lw);
                }
                if (pd_row) {        // 16 contiguous floats per piece: two full 32-byte sectors per row and matrix
                    *reinterpret_cast<float4*>(pd_row + j0) = make_float4(pd4[0], pd4[1], pd4[2], pd4[3]);
                    *reinterpret_cast<float4*>(ds_row + j0) = make_float4(ds4[0], ds4[1], ds4[2], ds4[3]);
                }
            }
        } else {
#pragma unroll
            for (int e = 0; e < 16; ++e) { hi[e] = 0u; lo[e] = 0u; }
            if (pd_row) {
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int g4 = 0; g4 < 4; ++g4) {
                    *reinterpret_cast<float4*>(pd_row + c * 16 + g4 * 4) = z;
                    *reinterpret_cast<float4*>(ds_row + c * 16 + g4 * 4) = z;
                }
            }
        }
        tmem_st16(t_s + lane_base + c * 16, hi);
        tmem_st16(t_lo + lane_base + c * 16, lo);
    }
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();
    BSA_STAMP(23);

    if (warp == 1 && elect_one()) {
        // ===================================================================== MMA 3: dQ = dS K
        tc_fence_after();
        const uint32_t idesc = idesc_tf32(128, DH, false, true);
        const uint32_t kh = smem_u32(smem + AU_OFF_V), kl = smem_u32(smem + AU_OFF_KHI);
        uint32_t acc = 0;
        for (int kstep = 0; kstep < NK / 8; ++kstep) {
            const int kb = kstep >> 2, ks = kstep & 3;
            const uint64_t dbh = smem_desc(kh + kb * VKB + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
            const uint64_t dbl = smem_desc(kl + kb * VKB + ks * 1024, 4096, 512, LAYOUT_SW128_BASE32B);
            const uint32_t col = (uint32_t)(kstep * 8);
            mma_tf32_ts(t_o, t_lo + col, dbh, idesc, acc);
            mma_tf32_ts(t_o, t_s + col, dbl, idesc, 1u);
            mma_tf32_ts(t_o, t_s + col, dbh, idesc, 1u);
            acc = 1u;
        }
        mma_commit(o_full);
    }
    __syncwarp();
    mbar_wait(o_full, 0);
    BSA_STAMP(24);
    tc_fence_after();
    if (part < DH / 16) {
        float v[16];
        tmem_ld16(t_o + lane_base + part * 16, v);                 // part -> dQ columns [16 part, 16 part + 16)
        if (row_ok) {
            float* dst = p.dq + (seq_base + i) * p.ldo + col0 + part * 16;
#pragma unroll
            for (int e = 0; e < 4; ++e)
                *reinterpret_cast<float4*>(dst + e * 4) =
                    make_float4(v[4 * e] * p.scale, v[4 * e + 1] * p.scale, v[4 * e + 2] * p.scale, v[4 * e + 3] * p.scale);
        }
    }
    tc_fence_before();
    __syncthreads();
    BSA_STAMP(25);
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

bool attn_umma_supported(int L, int d, int H) {
    if (H <= 0 || d % H != 0 || (d / H != 64 && d / H != 32) || d % 4 != 0) return false;
    return L >= 1 && L <= AU_NK_MAX;
}

int attn_fwd_umma(const float* q, const float* k, const float* v, const long long* ids, float* ctx, float* lse, int B, int H,
                  int L, int d, Dropout dr, cudaStream_t st, float* ctx_lo) {
    BSA_REQUIRE(attn_umma_supported(L, d, H), "attn_fwd_umma: unsupported shape (L=%d d=%d H=%d)", L, d, H);
    if (B == 0) return 0;
    static bool configured = false;
    if (!configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_umma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM));
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_umma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM));
        configured = true;
    }
    const int dh = d / H;
    AttnUmmaParams p;
    const int NK = (L + 15) / 16 * 16;
    const long long rows = (long long)B * L;
    int rc;
    if ((rc = make_tmap_2d(&p.q, q, rows, d, d, 128, false))) return rc;
    if ((rc = make_tmap_2d(&p.k, k, rows, d, d, NK / 2, false))) return rc;
    if ((rc = make_tmap_2d(&p.v, v, rows, d, d, 32, true))) return rc;
    p.ids = ids; p.ctx = ctx; p.lse = lse; p.H = H; p.L = L; p.d = d; p.NK = NK;
    p.ctx_lo = ctx_lo;
    p.scale = 1.0f / sqrtf((float)dh);
    p.dr = dr;
    const dim3 grid(B * H, (L + 127) / 128);
    ProfScope prof(PROF_ATTN_FWD, st, 4.0 * B * H * (double)L * L * dh);
    if (dh == 64) attn_fwd_umma_kernel<64><<<grid, AUF_THREADS, AU_SMEM, st>>>(p);
    else          attn_fwd_umma_kernel<32><<<grid, AUF_THREADS, AU_SMEM, st>>>(p);
    BSA_LAUNCH_CHECK();
    return 0;
}

int attn_bwd_q_umma(const float* q, const float* k, const float* v, const long long* ids, const float* ctx, const float* lse,
                    const float* dctx, float* dq, float* delta, float* pd_out, float* ds_out, int B, int H, int L, int d,
                    Dropout dr, cudaStream_t st, int ldo) {
    BSA_REQUIRE(attn_umma_supported(L, d, H), "attn_bwd_q_umma: unsupported shape (L=%d d=%d H=%d)", L, d, H);
    if (B == 0) return 0;
    static bool configured = false;
    if (!configured) {
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_q_umma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM));
        BSA_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_q_umma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM));
        configured = true;
    }
    const int dh = d / H;
    AttnBwdQParams p;
    const int NK = (L + 15) / 16 * 16;
    const long long rows = (long long)B * L;
    int rc;
    if ((rc = make_tmap_2d(&p.q, q, rows, d, d, 128, false))) return rc;
    if ((rc = make_tmap_2d(&p.k, k, rows, d, d, NK / 2, false))) return rc;
    if ((rc = make_tmap_2d(&p.dout, dctx, rows, d, d, 128, false))) return rc;
    if ((rc = make_tmap_2d(&p.vk, v, rows, d, d, NK / 2, false))) return rc;
    if ((rc = make_tmap_2d(&p.km, k, rows, d, d, 32, true))) return rc;
    p.ids = ids; p.ctx = ctx; p.dctx = dctx; p.lse = lse; p.dq = dq; p.delta = delta;
    p.pd_out = pd_out; p.ds_out = ds_out;
    p.ldo = ldo ? ldo : d;
    p.H = H; p.L = L; p.d = d; p.NK = NK;
    p.scale = 1.0f / sqrtf((float)dh);
    p.dr = dr;
    const dim3 grid(B * H, (L + 127) / 128);
    if (dh == 64) attn_bwd_q_umma_kernel<64><<<grid, AUF_THREADS, AU_SMEM, st>>>(p);
    else          attn_bwd_q_umma_kernel<32><<<grid, AUF_THREADS, AU_SMEM, st>>>(p);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace bsa

#ifdef BSA_ATTN_STAMPS
extern "C" __attribute__((visibility("default"))) int bsa_debug_attn_stamps(unsigned long long* out32) {
    return (int)cudaMemcpyFromSymbol(out32, bsa::g_attn_stamps, sizeof(unsigned long long) * 64);
}
#endif

// fp32 FFMA GEMM: 128 x BN x 16 tiles, 256 threads, (2x4) x (CGx4) register tile per thread,
// register-staged double buffering.  See gemm_simt.cuh for the operand conventions.
#include "gemm_simt.cuh"

namespace bsa {

namespace {

// BM = 128: (2x4) x (CGx4) register tile per thread.  BM = 32: 2 x (CGx4) — the small-problem variant: a [512, 256]
// output (the last-row path's [B, d] linears) is 8 CTAs at 128 x 128 but 64 at 32 x 64, and those launches are bound by
// how many SMs they reach, not by FMA:LDS ratio.
// BK = 16 with two shared-memory buffers, or BK = 64 with one (BM = 32 only): the small launches are a chain of k-tiles
// that each wait out one global-load latency (profiles/r2u launch list: 17 us for K = 256, 52 us for K = 1024 — one
// microsecond per 16-wide tile), so the deep variant keeps 4x the bytes in flight per round trip.
template <int BM, int BN, int BK, bool AK, bool BKM>
__global__ void __launch_bounds__(256) gemm_kernel(GemmP p) {
    constexpr int LDSA = BM + 4, LDSB = BN + 4;
    constexpr int NBUF = BK == 16 ? 2 : 1;
    constexpr int TM = BM / 16;            // rows per thread
    constexpr int CG = BN / 64;            // column groups of 4 per thread
    constexpr int NAL = BM * BK / 256;     // A elements loaded per thread per k-tile
    constexpr int NBL = BN * BK / 256;     // B elements loaded per thread per k-tile
    __shared__ __align__(16) float As[NBUF][BK][LDSA];
    __shared__ __align__(16) float Bs[NBUF][BK][LDSB];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    int z = blockIdx.z, split = 0;
    if (p.splitk > 1) { split = z; z = 0; }
    const int zo = z / p.inner, zi = z - zo * p.inner;
    const float* __restrict__ A = p.A + zo * p.sAo + zi * p.sAi;
    const float* __restrict__ B = p.B + zo * p.sBo + zi * p.sBi;

    int kbeg = 0, kend = p.K;
    if (p.splitk > 1) {
        int kc = ((p.K + p.splitk - 1) / p.splitk + BK - 1) / BK * BK;
        kbeg = split * kc;
        kend = min(p.K, kbeg + kc);
    }

    float acc[TM][CG * 4];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < CG * 4; ++j) acc[i][j] = 0.f;
    auto row_of = [&](int i) { return BM == 128 ? (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4)) : ty * TM + i; };

    float ra[NAL], rb[NBL];

    auto load_tiles = [&](int k0) {
#pragma unroll
        for (int i = 0; i < NAL; ++i) {
            const int e = tid + 256 * i;
            int m, k;
            if (AK) { k = e % BK; m = e / BK; }
            else    { m = e % BM; k = e / BM; }
            const int gm = m0 + m, gk = k0 + k;
            float v = 0.f;
            if (gm < p.M && gk < kend)
                v = AK ? __ldg(A + (long long)gm * p.lda + gk) : __ldg(A + (long long)gk * p.lda + gm);
            ra[i] = v;
        }
#pragma unroll
        for (int i = 0; i < NBL; ++i) {
            const int e = tid + 256 * i;
            int n, k;
            if (BKM) { k = e % BK; n = e / BK; }
            else     { n = e % BN; k = e / BN; }
            const int gn = n0 + n, gk = k0 + k;
            float v = 0.f;
            if (gn < p.N && gk < kend)
                v = BKM ? __ldg(B + (long long)gn * p.ldb + gk) : __ldg(B + (long long)gk * p.ldb + gn);
            rb[i] = v;
        }
    };
    auto store_tiles = [&](int buf) {
#pragma unroll
        for (int i = 0; i < NAL; ++i) {
            const int e = tid + 256 * i;
            int m, k;
            if (AK) { k = e % BK; m = e / BK; }
            else    { m = e % BM; k = e / BM; }
            As[buf][k][m] = ra[i];
        }
#pragma unroll
        for (int i = 0; i < NBL; ++i) {
            const int e = tid + 256 * i;
            int n, k;
            if (BKM) { k = e % BK; n = e / BK; }
            else     { n = e % BN; k = e / BN; }
            Bs[buf][k][n] = rb[i];
        }
    };

    const int ntiles = (kend > kbeg) ? (kend - kbeg + BK - 1) / BK : 0;
    if (ntiles > 0) {
        load_tiles(kbeg);
        store_tiles(0);
    }
    __syncthreads();
    for (int t = 0; t < ntiles; ++t) {
        const int buf = NBUF == 2 ? (t & 1) : 0;
        if (t + 1 < ntiles) load_tiles(kbeg + (t + 1) * BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[TM], b[CG * 4];
            if constexpr (BM == 128) {
                const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
                const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
                a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
                a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
            } else {
#pragma unroll
                for (int i = 0; i < TM; ++i) a[i] = As[buf][k][ty * TM + i];
            }
#pragma unroll
            for (int g = 0; g < CG; ++g) {
                const float4 bv = *reinterpret_cast<const float4*>(&Bs[buf][k][g * 64 + tx * 4]);
                b[g * 4 + 0] = bv.x; b[g * 4 + 1] = bv.y; b[g * 4 + 2] = bv.z; b[g * 4 + 3] = bv.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < CG * 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (NBUF == 1) __syncthreads();                  // single buffer: everyone has finished reading it
        if (t + 1 < ntiles) store_tiles(NBUF == 2 ? (buf ^ 1) : 0);
        __syncthreads();
    }

    // ------------------------------------------------------------------ epilogue
    if (p.splitk > 1) {
        float* __restrict__ part = p.partial + (long long)split * p.M * p.N;
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int gm = m0 + row_of(i);
            if (gm >= p.M) continue;
#pragma unroll
            for (int j = 0; j < CG * 4; ++j) {
                const int gn = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
                if (gn < p.N) part[(long long)gm * p.N + gn] = acc[i][j];
            }
        }
        return;
    }
    float* __restrict__ C = p.C + zo * p.sCo + zi * p.sCi;
    float* __restrict__ auxo = p.aux_out ? p.aux_out + zo * p.sCo + zi * p.sCi : nullptr;
    const float* __restrict__ auxi = p.aux_in ? p.aux_in + zo * p.sCo + zi * p.sCi : nullptr;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int gm = m0 + row_of(i);
        if (gm >= p.M) continue;
#pragma unroll
        for (int j = 0; j < CG * 4; ++j) {
            const int gn = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
            if (gn >= p.N) continue;
            const long long off = (long long)gm * p.ldc + gn;
            float v = p.alpha * acc[i][j];
            if (p.bias) v += __ldg(p.bias + gn);
            if (p.act == ACT_GELU_SAVE_PRE) { auxo[off] = v; v = gelu_erf(v); }
            else if (p.act == ACT_MUL_GELU_GRAD) v *= gelu_erf_grad(auxi[off]);
            else if (p.act == ACT_GELU_SAVE_GRAD) { float gr; v = gelu_erf_both(v, gr); if (auxo) auxo[off] = gr; }
            else if (p.act == ACT_MUL_AUX) v *= auxi[off];
            if (p.accumulate) v += C[off];
            C[off] = v;
        }
    }
}

// fixed-order sum of the split-K slabs, then the same epilogue as the unsplit kernel (bias, activation, accumulate)
__global__ void splitk_reduce_kernel(GemmP p, long long MN) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= MN) return;
    float s = 0.f;
    for (int k = 0; k < p.splitk; ++k) s += p.partial[(long long)k * MN + i];
    const long long m = i / p.N, n = i - m * p.N;
    const long long off = m * p.ldc + n;
    float v = p.alpha * s;
    if (p.bias) v += __ldg(p.bias + n);
    if (p.act == ACT_GELU_SAVE_PRE) { p.aux_out[off] = v; v = gelu_erf(v); }
    else if (p.act == ACT_MUL_GELU_GRAD) v *= gelu_erf_grad(p.aux_in[off]);
    else if (p.act == ACT_GELU_SAVE_GRAD) { float gr; v = gelu_erf_both(v, gr); if (p.aux_out) p.aux_out[off] = gr; }
    else if (p.act == ACT_MUL_AUX) v *= p.aux_in[off];
    if (p.accumulate) v += p.C[off];
    p.C[off] = v;
}

template <int BM, int BN, int BK = 16>
int dispatch(const GemmP& p, dim3 grid, cudaStream_t st) {
    if (p.a_kmajor && p.b_kmajor) gemm_kernel<BM, BN, BK, true, true><<<grid, 256, 0, st>>>(p);
    else if (p.a_kmajor && !p.b_kmajor) gemm_kernel<BM, BN, BK, true, false><<<grid, 256, 0, st>>>(p);
    else if (!p.a_kmajor && p.b_kmajor) gemm_kernel<BM, BN, BK, false, true><<<grid, 256, 0, st>>>(p);
    else gemm_kernel<BM, BN, BK, false, false><<<grid, 256, 0, st>>>(p);
    BSA_LAUNCH_CHECK();
    return 0;
}

}  // namespace

int choose_splitk(int M, int N, int K) {
    const int tiles = cdiv(M, 128) * cdiv(N, N <= 64 ? 64 : 128);
    if (tiles < 96 && K <= 4096) {
        // the [B, d]-sized problems of the last-row path run on 32 x 64 tiles (launch_gemm's `small`): one CTA per SM walks
        // its k-tiles as a chain of load -> store -> barrier -> compute phases with nothing to overlap them (17 - 50 us for
        // 0.07 - 0.27 GFLOP).  Splitting K down to 64 per CTA puts 2 - 4 independent CTAs on every SM instead.
        const int t32 = cdiv(M, 32) * cdiv(N, 64);
        int s = (148 * 2 + t32 - 1) / t32;
        const int maxs = K / 64 > 1 ? K / 64 : 1;
        if (s > maxs) s = maxs;
        if (s > 1) { const int kc = ((K + s - 1) / s + 63) / 64 * 64; s = (K + kc - 1) / kc; }      // no empty split
        return s < 1 ? 1 : s;
    }
    int s = (148 * 2 + tiles - 1) / tiles;
    const int maxs = (K + 255) / 256;      // at least 256 of K per split
    if (s > maxs) s = maxs;
    if (s < 1) s = 1;
    if (s > 512) s = 512;
    return s;
}

int launch_gemm(const GemmP& p, cudaStream_t st) {
    if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return 0;
    BSA_REQUIRE(p.splitk >= 1 && (p.splitk == 1 || (p.batch == 1 && p.partial)), "launch_gemm: invalid split-K configuration");
    BSA_REQUIRE(p.inner >= 1 && p.batch % p.inner == 0, "launch_gemm: batch %% inner != 0");
    int BN = p.N <= 64 ? 64 : 128;
    const int nz = p.splitk > 1 ? p.splitk : p.batch;
    const bool small = (long long)cdiv(p.M, 128) * cdiv(p.N, BN) * nz < 96;      // would leave a third of the SMs idle
    if (small) BN = 64;
    dim3 grid(cdiv(p.M, small ? 32 : 128), cdiv(p.N, BN), nz);
    BSA_REQUIRE(grid.y <= 65535 && grid.z <= 65535, "launch_gemm: grid too large (M=%d batch=%d)", p.M, p.batch);
    int rc;
    {
        ProfScope prof(PROF_GEMM_SIMT, st, 2.0 * p.M * p.N * (double)p.K * p.batch);
        const int kspan = p.splitk > 1 ? (p.K + p.splitk - 1) / p.splitk : p.K;
        rc = small ? (kspan >= 64 ? dispatch<32, 64, 64>(p, grid, st) : dispatch<32, 64>(p, grid, st))
                   : BN == 64 ? dispatch<128, 64>(p, grid, st) : dispatch<128, 128>(p, grid, st);
    }
    if (rc) return rc;
    if (p.splitk > 1) {
        const long long MN = (long long)p.M * p.N;
        splitk_reduce_kernel<<<cdiv(MN, 256), 256, 0, st>>>(p, MN);
        BSA_LAUNCH_CHECK();
    }
    return 0;
}

}  // namespace bsa

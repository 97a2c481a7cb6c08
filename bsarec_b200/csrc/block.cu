// bsa_block_fwd / bsa_block_bwd: one BSARecBlock (model/bsarec.py:62-104, model/_modules.py:61-140).
#include "../../include/bsarec_b200.h"
#include "common.cuh"
#include <cstdlib>
#include "gemm_simt.cuh"
#include "rowwise.cuh"
#include "umma_engine.cuh"
#include "attn.cuh"
#include "lastrow.cuh"

#include <cmath>

namespace bsa {

int linear_impl();
int attn_impl();

namespace {

struct Bump {               // bump allocator over caller-provided memory, 256-byte aligned
    char* base; size_t off = 0, cap;
    Bump(void* p, size_t bytes) : base((char*)p), cap(bytes) {}
    float* f(long long n) {
        size_t bytes = ((size_t)n * sizeof(float) + 255) / 256 * 256;
        float* r = (float*)(base + off);
        off += bytes;
        return r;
    }
    bool ok() const { return off <= cap; }
};

struct Saved {              // layout of the per-layer saved-activation buffer
    float *z_f, *mean_f, *rstd_f, *q, *k, *v, *P, *lse, *ctx, *z_a, *mean_a, *rstd_a, *y, *h_pre, *h_act, *z_o, *mean_o,
        *rstd_o;
    // tcgen05 path: split-TF32 (hi, lo) copies of the four linear inputs, written once in forward and reused by the
    // weight-gradient GEMMs in backward
    float *x_hi, *x_lo, *ctx_hi, *ctx_lo, *y_hi, *y_lo, *h_hi, *h_lo;
    // ... and of the six weight matrices, [wq | wk | wv | wo | w1 | w2] = 12 d^2 floats each: split by ONE launch at the top of
    // the forward pass and read again by the backward pass (the weights do not change in between)
    float *w_hi, *w_lo;
    bool fused;             // attention ran in the fused kernel: lse [B,H,L] is saved instead of P [B,H,L,L]
    size_t bytes;
};

Saved carve_saved(void* p, const bsa_block_desc& c) {
    const long long M = (long long)c.B * c.L, d = c.d;
    Bump b(p, ~(size_t)0);
    Saved s;
    s.z_f = b.f(M * d); s.mean_f = b.f(M); s.rstd_f = b.f(M);
    s.q = b.f(M * d); s.k = b.f(M * d); s.v = b.f(M * d);
    s.fused = attn_fused_supported(c.L, c.d, c.H) && attn_impl() == 2;
    s.P = s.fused ? nullptr : b.f((long long)c.B * c.H * c.L * c.L);
    s.lse = b.f(2ll * c.B * c.H * c.L);
    s.ctx = b.f(M * d);
    s.z_a = b.f(M * d); s.mean_a = b.f(M); s.rstd_a = b.f(M);
    s.y = b.f(M * d);
    s.h_pre = b.f(M * 4 * d); s.h_act = b.f(M * 4 * d);
    s.z_o = b.f(M * d); s.mean_o = b.f(M); s.rstd_o = b.f(M);
    s.x_hi = s.x_lo = s.ctx_hi = s.ctx_lo = s.y_hi = s.y_lo = s.h_hi = s.h_lo = s.w_hi = s.w_lo = nullptr;
    if (linear_impl() == 2) {
        s.x_lo = b.f(M * d); s.ctx_lo = b.f(M * d);       // x, ctx, y are their own hi halves: only lo = t - truncate_tf32(t) is kept
        s.y_lo = b.f(M * d); s.h_hi = b.f(M * 4 * d); s.h_lo = b.f(M * 4 * d);
        s.w_hi = b.f(12 * d * d); s.w_lo = b.f(12 * d * d);
    }
    s.bytes = b.off;
    return s;
}

// floats of split-K scratch that covers every weight-gradient GEMM of the block
long long splitk_floats(long long M, long long d) {
    long long a = (long long)choose_splitk((int)(4 * d), (int)d, (int)M) * 4 * d * d;
    long long b = (long long)choose_splitk((int)d, (int)(4 * d), (int)M) * 4 * d * d;
    long long c = (long long)choose_splitk((int)d, (int)d, (int)M) * d * d;
    long long m = a > b ? a : b;
    return (m > c ? m : c) + 1024;
}

// an activation operand: the fp32 tensor and (tcgen05 path) its split pair
struct Act { const float* x; const float* hi; const float* lo; };

struct Lin {               // where the linears run: tcgen05 3xTF32 engine (default) or the fp32 FFMA kernel
    bool umma;
    float *w_hi, *w_lo;    // [4d*d] scratch for the weight's split pair
    float* partial; size_t partial_bytes;      // split-K slabs (weight gradients)
    bool inference = false;                    // forward only: nothing is kept for a backward pass
    // lazy operand splits (the [B, d] last-row path): an Act that arrives without its pair is split here on the fly
    float *s1_hi = nullptr, *s1_lo = nullptr, *s2_hi = nullptr, *s2_lo = nullptr;
    // FFMA path, small problems: split-K slabs for the forward / input-gradient GEMMs (their epilogue runs in the reduction)
    float* sk = nullptr; long long sk_floats = 0;
};
// floats of split-K scratch that covers any [B, <= 4d] output of the last-row path: s <= 296 / tiles + 1 slabs of M x N
long long small_sk_floats(long long B, long long d) { return 296ll * 32 * 64 + 2 * B * 4 * d + 1024; }
void ffma_splitk(const Lin& ln, GemmP& p) {
    if (!ln.sk) return;
    const int s = choose_splitk(p.M, p.N, p.K);
    if (s > 1 && (long long)s * p.M * p.N <= ln.sk_floats) { p.splitk = s; p.partial = ln.sk; }
}

// the (hi, lo) pair of an activation operand: its own, or a split made now into the Lin's scratch
int act_pair(const Lin& ln, const Act& a, long long n, float* s_hi, float* s_lo, const float*& hi, const float*& lo, cudaStream_t st);

int act_pair(const Lin& ln, const Act& a, long long n, float* s_hi, float* s_lo, const float*& hi, const float*& lo, cudaStream_t st) {
    (void)ln;
    if (a.hi) { hi = a.hi; lo = a.lo; return 0; }
    BSA_REQUIRE(s_hi && s_lo, "tcgen05 linear: operand has no split pair and no scratch was provided");
    hi = s_hi; lo = s_lo;
    return umma_split(a.x, s_hi, s_lo, n, st);
}

long long lin_ws_floats(long long d) { return 2 * (4 * d * d + 256) + (long long)(umma_gemm_partial_bytes((int)(4 * d), (int)d) / 4); }

Lin make_lin(Bump& b, long long d) {
    Lin ln;
    ln.umma = linear_impl() == 2;
    ln.w_hi = b.f(4 * d * d + 256); ln.w_lo = b.f(4 * d * d + 256);
    ln.partial_bytes = umma_gemm_partial_bytes((int)(4 * d), (int)d);
    ln.partial = b.f((long long)(ln.partial_bytes / 4));
    return ln;
}

size_t workspace_bytes(const bsa_block_desc& c) {
    const long long M = (long long)c.B * c.L, d = c.d;
    const long long PL = (long long)c.B * c.H * c.L * c.L;
    Bump b(nullptr, ~(size_t)0);
    // backward is the larger user: dyb, dd, dzf, dq, dk, dv, dctx [M,d]; dh [M,4d]; Pd, dP [B,H,L,L]
    for (int i = 0; i < 7; ++i) b.f(M * d);
    b.f(M * 4 * d);
    if (attn_fused_supported(c.L, c.d, c.H) && attn_impl() == 2) b.f(attn_bwd_scratch_floats(c.B, c.H, c.L, c.d));   // delta (+ Pd, dS tiles)
    else { b.f(PL); b.f(PL); }
    // partials: LN gamma/beta, column sums, split-K, beta2
    for (int i = 0; i < 9; ++i) b.f((long long)rowwise_partial_blocks(M) * d);     // three LayerNorm sites x (gamma, beta, column) partials
    for (int i = 0; i < 3; ++i) b.f((long long)colsum_partial_rows(M) * d);        // dq / dk / dv column partials
    b.f((long long)colsum_partial_rows(M) * 4 * d);
    b.f(splitk_floats(M, d));
    b.f((long long)c.B * d);
    // tcgen05 linears: weight split, split-K slabs, split pairs of the upstream gradients ([M,d] and [M,4d])
    b.f(lin_ws_floats(d));
    b.f(M * d); b.f(M * d); b.f(M * 4 * d); b.f(M * 4 * d);
    return b.off + 8192;
}

// offsets (in units of d^2 floats) of the weight pairs inside Saved::w_hi / w_lo
enum WeightSlot : int { W_Q = 0, W_K = 1, W_V = 2, W_O = 3, W_1 = 4, W_2 = 8 };

// all six weight matrices of the block -> their split pairs in the saved buffer, one launch
int split_block_weights(const Saved& s, const bsa_block_desc& c, const bsa_block_weights& w, cudaStream_t st) {
    const long long dd = (long long)c.d * c.d;
    const float* src[6]; float* hi[6]; float* lo[6]; long long n[6];
    int k = 0;
    auto add = [&](const float* W, int slot, long long cnt) {
        if (!W) return;
        src[k] = W; hi[k] = s.w_hi + slot * dd; lo[k] = s.w_lo + slot * dd; n[k] = cnt; ++k;
    };
    if (c.kind == 0) { add(w.wq, W_Q, dd); add(w.wk, W_K, dd); add(w.wv, W_V, dd); add(w.wo, W_O, dd); }
    add(w.w1, W_1, 4 * dd); add(w.w2, W_2, 4 * dd);
    return umma_split_multi(k, src, hi, lo, n, st);
}

// x [M, K] @ W[N, K]^T + bias -> out [M, N]      (nn.Linear forward); out_hi/out_lo: also emit out's split pair
// pw_hi / pw_lo (tcgen05 path, optional): W's split pair made earlier (split_block_weights) — no split launch here
int linear_fwd(const Lin& ln, const Act& a, const float* W, const float* bias, float* out, long long M, int N, int K, int act,
               float* aux_out, float* out_hi, float* out_lo, cudaStream_t st, const float* pw_hi = nullptr, const float* pw_lo = nullptr) {
    if (ln.umma) {
        int rc = 0;
        if (!pw_hi) { rc = umma_split(W, ln.w_hi, ln.w_lo, (long long)N * K, st); pw_hi = ln.w_hi; pw_lo = ln.w_lo; }
        if (rc) return rc;
        // when the epilogue emits the result's split pair, that pair is all the tcgen05 path reads afterwards: no fp32 copy
        if (ln.inference) { aux_out = nullptr; if (act == ACT_GELU_SAVE_GRAD) act = ACT_GELU_SAVE_PRE; }    // nothing kept: plain GELU
        const float *ah_, *al_;
        if ((rc = act_pair(ln, a, M * K, ln.s1_hi, ln.s1_lo, ah_, al_, st))) return rc;
        return umma_gemm_ps((int)M, N, K, ah_, al_, M, K, K, false, pw_hi, pw_lo, N, K, K, false, out_hi ? nullptr : out, N, bias,
                            act, aux_out, nullptr, 1.f, 0, out_hi, out_lo, nullptr, 0, st);
    }
    GemmP p;
    p.A = a.x; p.B = W; p.C = out; p.M = (int)M; p.N = N; p.K = K; p.lda = K; p.ldb = K; p.ldc = N;
    p.bias = bias; p.act = act; p.aux_out = aux_out;
    ffma_splitk(ln, p);
    return launch_gemm(p, st);
}

// q, k, v = x Wq^T + bq, x Wk^T + bk, x Wv^T + bv in ONE tcgen05 launch (N = 3d, the three weights split into one [3d, d]
// pair): the shared A operand is read once and the epilogue sends each column segment to its own [M, d] tensor
// (q + s * seg_stride).  BSA_QKV_FUSE=0 restores the three launches.
bool qkv_fuse_enabled() {
    static const bool on = [] { const char* e = getenv("BSA_QKV_FUSE"); return !(e && atoi(e) == 0); }();
    return on;
}
int linear_fwd_qkv(const Lin& ln, const Act& a, const bsa_block_weights& w, float* q, long long seg_stride, long long M, int d,
                   cudaStream_t st, const float* pw_hi, const float* pw_lo) {      // pw: the [wq | wk | wv] pair, [3d, d]
    (void)ln;
    return umma_gemm_ps((int)M, 3 * d, d, a.hi, a.lo, M, d, d, false, pw_hi, pw_lo, 3 * d, d, d, false, q, d, w.bq, ACT_NONE,
                        nullptr, nullptr, 1.f, 0, nullptr, nullptr, nullptr, 0, st, nullptr, d, seg_stride, w.bk, w.bv);
}

// dx [M, K] (+)= dy [M, N] @ W[N, K]              (nn.Linear input gradient)
int linear_bwd_input(const Lin& ln, const Act& dy, const float* W, float* dx, long long M, int N, int K, int accumulate, int act,
                     const float* aux_in, float* out_hi, float* out_lo, cudaStream_t st, float* colpart = nullptr,
                     const float* pw_hi = nullptr, const float* pw_lo = nullptr) {
    if (ln.umma) {
        int rc = 0;
        if (!pw_hi) { rc = umma_split(W, ln.w_hi, ln.w_lo, (long long)N * K, st); pw_hi = ln.w_hi; pw_lo = ln.w_lo; }
        if (rc) return rc;
        // colpart (tcgen05 path only): the epilogue also leaves column sums of dx per 32-row group, [4 * ceil(M / 128), K]
        const float *gh_, *gl_;
        if ((rc = act_pair(ln, dy, M * N, ln.s1_hi, ln.s1_lo, gh_, gl_, st))) return rc;
        return umma_gemm_ps((int)M, K, N, gh_, gl_, M, N, N, false, pw_hi, pw_lo, N, K, K, true, dx, K, nullptr, act, nullptr,
                            aux_in, 1.f, accumulate, out_hi, out_lo, nullptr, 0, st, colpart);
    }
    GemmP p;
    p.A = dy.x; p.B = W; p.C = dx; p.M = (int)M; p.N = K; p.K = N; p.lda = N; p.ldb = K; p.ldc = K;
    p.b_kmajor = false; p.accumulate = accumulate; p.act = act; p.aux_in = aux_in;
    ffma_splitk(ln, p);
    return launch_gemm(p, st);
}

// dW [N, K] = dy[M, N]^T @ x [M, K] ; db [N] = colsum(dy)   (deterministic split-K + column sums)
int linear_bwd_weight(const Lin& ln, const Act& dy, const Act& x, float* dW, float* db, long long M, int N, int K, float* part_sk,
                      float* part_cs, cudaStream_t st, long long ld_dy = 0) {
    if (ln.umma) {
        const float *gh_, *gl_, *xh_, *xl_;
        int rc;
        if ((rc = act_pair(ln, dy, M * N, ln.s1_hi, ln.s1_lo, gh_, gl_, st))) return rc;
        if ((rc = act_pair(ln, x, M * K, ln.s2_hi, ln.s2_lo, xh_, xl_, st))) return rc;
        // ld_dy: dy's pair is a column window of a wider [M, ld_dy] buffer (the packed dq | dk | dv pair)
        rc = umma_gemm_ps(N, K, (int)M, gh_, gl_, M, N, ld_dy ? ld_dy : N, true, xh_, xl_, M, K, K, true, dW, K, nullptr, 0, nullptr,
                          nullptr, 1.f, 0, nullptr, nullptr, ln.partial, ln.partial_bytes, st);
        if (rc) return rc;
        return db ? colsum(dy.x, M, N, part_cs, db, 0, st) : 0;      // db == nullptr: the producer already summed the columns
    }
    GemmP p;
    p.A = dy.x; p.B = x.x; p.C = dW; p.M = N; p.N = K; p.K = (int)M; p.lda = N; p.ldb = K; p.ldc = K;
    p.a_kmajor = false; p.b_kmajor = false;
    p.splitk = choose_splitk(N, K, (int)M);
    if (p.splitk > 1) p.partial = part_sk;
    int rc = launch_gemm(p, st);
    if (rc) return rc;
    return db ? colsum(dy.x, M, N, part_cs, db, 0, st) : 0;
}

// the same with the tensor as its own hi half: only lo = x - truncate_tf32(x) is written (umma_split_lo)
int split_act_lo(const Lin& ln, const float* x, float* lo, long long n, Act& out, cudaStream_t st) {
    out.x = x; out.hi = ln.umma ? x : nullptr; out.lo = ln.umma ? lo : nullptr;
    if (!ln.umma) return 0;
    return umma_split_lo(x, lo, n, st);
}

// split an activation into (hi, lo) when the tcgen05 path will consume it
int split_act(const Lin& ln, const float* x, float* hi, float* lo, long long n, Act& out, cudaStream_t st) {
    out.x = x; out.hi = hi; out.lo = lo;
    if (!ln.umma) return 0;
    return umma_split(x, hi, lo, n, st);
}

}  // namespace

int g_attn_impl = 2;        // 1 = unfused GEMM + softmax kernels, 2 = fused attention kernel where the head size allows
int attn_impl() { return g_attn_impl == 2 && !g_attn_fwd_umma ? 3 : g_attn_impl; }
void set_attn_impl(int impl) {        // 1 unfused | 2 fused, forward on tcgen05 where the shape allows | 3 fused, mma.sync only
    g_attn_impl = impl == 1 ? 1 : 2;
    g_attn_fwd_umma = impl == 3 ? 0 : 1;
}
int g_linear_impl = 2;      // 1 = fp32 FFMA, 2 = tcgen05 3xTF32 (bsa_set_linear_impl)
int linear_impl() { return g_linear_impl; }
void set_linear_impl(int impl) { g_linear_impl = impl == 1 ? 1 : 2; }

int block_fwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
              const float* x, float* x_out, void* saved_p, void* ws, size_t ws_bytes, cudaStream_t st) {
    const int B = c.B, L = c.L, d = c.d, H = c.H, dh = d / H;
    const long long M = (long long)B * L, PL = (long long)B * H * L * L;
    BSA_REQUIRE(H > 0 && d % H == 0, "hidden size (%d) is not a multiple of the number of attention heads (%d)", d, H);
    BSA_REQUIRE(d % 4 == 0 && d <= 1024 && L <= 1024, "unsupported shape: d=%d (need d%%4==0, <=1024), L=%d (<=1024)", d, L);
    BSA_REQUIRE((long long)B * H <= 65535, "batch*heads = %lld exceeds 65535 per call; split the batch", (long long)B * H);
    if (M == 0) return 0;
    Saved s = carve_saved(saved_p, c);
    Bump b(ws, ws_bytes);
    float* dsp = b.f(M * d);
    float* tmp = b.f(M * d);
    const Dropout dr_f = make_dropout(c.seed, 1 + 4 * c.layer, c.p_hidden, c.training);
    const Dropout dr_p = make_dropout(c.seed, 2 + 4 * c.layer, c.p_attn, c.training);
    const Dropout dr_a = make_dropout(c.seed, 3 + 4 * c.layer, c.p_hidden, c.training);
    const Dropout dr_o = make_dropout(c.seed, 4 + 4 * c.layer, c.p_hidden, c.training);
    float* Pd = s.fused ? nullptr : (dr_p.thr ? b.f(PL) : s.P);
    Lin lin = make_lin(b, d);
    BSA_REQUIRE(b.ok(), "bsa_block_fwd: workspace too small (%zu < %zu)", ws_bytes, b.off);
    const int r = filter_rank(L, c.kept_bins);
    int rc;
    // filter branch
    // inference (bsa_block_desc::inference): the LayerNorm inputs / statistics and the GELU pre-activation are not written
    const bool inf = c.inference != 0;
    lin.inference = inf && lin.umma;
    const long long dd = (long long)d * d;
    if (lin.umma && (rc = split_block_weights(s, c, w, st))) return rc;
    auto wh = [&](int slot) -> const float* { return lin.umma ? s.w_hi + slot * dd : nullptr; };
    auto wl = [&](int slot) -> const float* { return lin.umma ? s.w_lo + slot * dd : nullptr; };
    Act ax, actx, ay, ah;
    if (c.kind == 1) {
        // FMLPRecBlock (fmlprec.py:84-118): y = LN(drop(circular_conv(x, K)) + x) — K [L,d] arrives as `sqrt_beta` — then the
        // same feed-forward; no attention
        if ((rc = circ_apply(x, w.sqrt_beta, false, nullptr, tmp, B, L, d, false, st))) return rc;
        if ((rc = residual_ln_fwd(tmp, x, w.f_ln_w, w.f_ln_b, s.y, inf ? nullptr : s.z_f, inf ? nullptr : s.mean_f,
                                  inf ? nullptr : s.rstd_f, nullptr, 0.f, 0.f, M, d, dr_f, st))) return rc;
    } else {
    if ((rc = filter_ln_fwd(x, basis, w.sqrt_beta, w.f_ln_w, w.f_ln_b, dsp, inf ? nullptr : s.z_f, inf ? nullptr : s.mean_f,
                            inf ? nullptr : s.rstd_f, B, L, d, r, dr_f, st, lin.umma ? s.x_lo : nullptr))) return rc;
    // attention branch: x is its own hi half, its lo half was left by the filter kernel (which reads all of x anyway)
    ax = Act{x, lin.umma ? x : nullptr, lin.umma ? s.x_lo : nullptr};
    if (lin.umma && qkv_fuse_enabled() && d % 32 == 0 && (s.k - s.q) == (s.v - s.k)) {
        if ((rc = linear_fwd_qkv(lin, ax, w, s.q, s.k - s.q, M, d, st, wh(W_Q), wl(W_Q)))) return rc;
    } else {
        if ((rc = linear_fwd(lin, ax, w.wq, w.bq, s.q, M, d, d, ACT_NONE, nullptr, nullptr, nullptr, st, wh(W_Q), wl(W_Q)))) return rc;
        if ((rc = linear_fwd(lin, ax, w.wk, w.bk, s.k, M, d, d, ACT_NONE, nullptr, nullptr, nullptr, st, wh(W_K), wl(W_K)))) return rc;
        if ((rc = linear_fwd(lin, ax, w.wv, w.bv, s.v, M, d, d, ACT_NONE, nullptr, nullptr, nullptr, st, wh(W_V), wl(W_V)))) return rc;
    }
    if (s.fused) {
        // ctx's lo half comes from a split pass afterwards: letting the attention epilogue store it (attn_fwd's ctx_lo argument)
        // costs that latency-bound kernel more (+0.054 ms at C3) than the pass it saves (0.034 ms) — profiles/r3n vs r3k
        if ((rc = attn_fwd(s.q, s.k, s.v, ids, s.ctx, s.lse, B, H, L, d, dr_p, st))) return rc;
    } else {
        {   // scores = q k^T / sqrt(dh)   -> written where P will live
            GemmP p;
            p.A = s.q; p.B = s.k; p.C = s.P; p.M = L; p.N = L; p.K = dh; p.lda = d; p.ldb = d; p.ldc = L;
            p.batch = B * H; p.inner = H;
            p.sAo = (long long)L * d; p.sAi = dh; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)H * L * L; p.sCi = (long long)L * L;
            p.alpha = 1.0f / sqrtf((float)dh);
            if ((rc = launch_gemm(p, st))) return rc;
        }
        if ((rc = softmax_mask_dropout_fwd(s.P, ids, s.P, Pd, B, H, L, dr_p, st))) return rc;
        {   // ctx = Pd v
            GemmP p;
            p.A = Pd; p.B = s.v; p.C = s.ctx; p.M = L; p.N = dh; p.K = L; p.lda = L; p.ldb = d; p.ldc = d;
            p.b_kmajor = false; p.batch = B * H; p.inner = H;
            p.sAo = (long long)H * L * L; p.sAi = (long long)L * L; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)L * d; p.sCi = dh;
            if ((rc = launch_gemm(p, st))) return rc;
        }
    }
    if ((rc = split_act_lo(lin, s.ctx, s.ctx_lo, M * d, actx, st))) return rc;
    if ((rc = linear_fwd(lin, actx, w.wo, w.bo, tmp, M, d, d, ACT_NONE, nullptr, nullptr, nullptr, st, wh(W_O), wl(W_O)))) return rc;
    // y = alpha*dsp + (1-alpha)*LN(drop(tmp) + x)
    // (tcgen05 path) y also leaves the LayerNorm kernel as the split pair dense_1 reads: no separate pass over y
    if ((rc = residual_ln_fwd(tmp, x, w.a_ln_w, w.a_ln_b, s.y, inf ? nullptr : s.z_a, inf ? nullptr : s.mean_a, inf ? nullptr : s.rstd_a,
                              dsp, c.alpha, 1.0f - c.alpha, M, d, dr_a, st, nullptr, lin.umma ? s.y_lo : nullptr))) return rc;
    ay = Act{s.y, lin.umma ? s.y : nullptr, lin.umma ? s.y_lo : nullptr};
    }   // kind == 0
    // feed-forward
    if (c.kind == 1 && (rc = split_act_lo(lin, s.y, s.y_lo, M * d, ay, st))) return rc;
    // dense_1's epilogue emits gelu(h) together with its split pair: dense_2 reads it without a separate pass
    // (h_pre holds gelu'(pre-activation): the backward's dX epilogue multiplies by it instead of evaluating erf + exp again)
    if ((rc = linear_fwd(lin, ay, w.w1, w.b1, s.h_act, M, 4 * d, d, ACT_GELU_SAVE_GRAD, s.h_pre, s.h_hi, s.h_lo, st, wh(W_1), wl(W_1)))) return rc;
    ah.x = s.h_act; ah.hi = s.h_hi; ah.lo = s.h_lo;
    if ((rc = linear_fwd(lin, ah, w.w2, w.b2, tmp, M, d, 4 * d, ACT_NONE, nullptr, nullptr, nullptr, st, wh(W_2), wl(W_2)))) return rc;
    if ((rc = residual_ln_fwd(tmp, s.y, w.o_ln_w, w.o_ln_b, x_out, inf ? nullptr : s.z_o, inf ? nullptr : s.mean_o,
                              inf ? nullptr : s.rstd_o, nullptr, 0.f, 0.f, M, d, dr_o, st))) return rc;
    return 0;
}

int block_bwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
              const float* x, const void* saved_p, const float* g_out, float* dx, const bsa_block_grads& g, void* ws,
              size_t ws_bytes, cudaStream_t st) {
    const int B = c.B, L = c.L, d = c.d, H = c.H, dh = d / H;
    const long long M = (long long)B * L, PL = (long long)B * H * L * L;
    BSA_REQUIRE(H > 0 && d % H == 0 && d % 4 == 0, "bsa_block_bwd: bad shape");
    if (M == 0) return 0;
    Saved s = carve_saved(const_cast<void*>(saved_p), c);
    Bump b(ws, ws_bytes);
    float* dyb = b.f(M * d);
    float* dd = b.f(M * d);
    float* dzf = b.f(M * d);
    float* dq = b.f(M * d);
    float* dk = b.f(M * d);
    float* dv = b.f(M * d);
    float* dctx = b.f(M * d);
    float* dh_ = b.f(M * 4 * d);
    float* Pd_ws = s.fused ? nullptr : b.f(PL);
    float* dP = s.fused ? b.f(attn_bwd_scratch_floats(B, H, L, d)) : b.f(PL);
    // Parameter-gradient partials: every site has its own slab and the reductions run as ONE launch at the end of the call
    // (ReduceBatch) instead of ~16 dependent 8-CTA launches between the GEMMs.
    const int nrb = rowwise_partial_blocks(M);
    float* lnp[3][3];                                      // [feed-forward LN, attention LN, filter LN][gamma, beta, column]
    for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k) lnp[i][k] = b.f((long long)nrb * d);
    float* part_qkv[3];
    for (int i = 0; i < 3; ++i) part_qkv[i] = b.f((long long)colsum_partial_rows(M) * d);
    float* part_cs = b.f((long long)colsum_partial_rows(M) * 4 * d);
    float* part_sk = b.f(splitk_floats(M, d));
    float* pbeta2 = b.f((long long)B * d);
    ReduceBatch rb;
    float *pgam = lnp[0][0], *pbet = lnp[0][1], *pcol = lnp[0][2];
    Lin lin = make_lin(b, d);
    float* g_hi = b.f(M * d); float* g_lo = b.f(M * d);            // split pair of the current [M,d] upstream gradient
    float* dh_hi = b.f(M * 4 * d); float* dh_lo = b.f(M * 4 * d);  // split pair of dh [M,4d]
    BSA_REQUIRE(b.ok(), "bsa_block_bwd: workspace too small (%zu < %zu)", ws_bytes, b.off);
    // x, ctx and y are their own hi halves (split_act_lo in the forward pass)
    const bool um = linear_impl() == 2;
    const Act ax{x, um ? x : nullptr, s.x_lo}, actx{s.ctx, um ? s.ctx : nullptr, s.ctx_lo}, ay{s.y, um ? s.y : nullptr, s.y_lo},
        ah{s.h_act, s.h_hi, s.h_lo};
    Act ag;
    const Dropout dr_f = make_dropout(c.seed, 1 + 4 * c.layer, c.p_hidden, c.training);
    const Dropout dr_p = make_dropout(c.seed, 2 + 4 * c.layer, c.p_attn, c.training);
    const Dropout dr_a = make_dropout(c.seed, 3 + 4 * c.layer, c.p_hidden, c.training);
    const Dropout dr_o = make_dropout(c.seed, 4 + 4 * c.layer, c.p_hidden, c.training);
    const int r = filter_rank(L, c.kept_bins);
    const float inv_sqrt_dh = 1.0f / sqrtf((float)dh);
    int rc, nblk = 0;
    // Small problems (one wave of 128-row tiles or less, e.g. C1): each weight-gradient GEMM runs on the side stream beside
    // the input-gradient GEMM that reads the same operands; the main stream joins before those operands are overwritten.
    Fork fk(st, lin.umma && M <= 128ll * 148);
    // the weight pairs the forward pass left in the saved buffer (split_block_weights)
    const long long dsq = (long long)d * d;
    auto wh = [&](int slot) -> const float* { return lin.umma ? s.w_hi + slot * dsq : nullptr; };
    auto wl = [&](int slot) -> const float* { return lin.umma ? s.w_lo + slot * dsq : nullptr; };

    // ---- feed-forward: x_out = LN(drop(tmp2) + y)
    // on the tcgen05 path the LN backward emits the gradient w.r.t. dense_2's output directly as a split-TF32 pair plus its
    // column partials (= b2's gradient): no fp32 copy, no separate split and column-sum passes over [M, d]
    const bool fuse_o = lin.umma && ln_bwd_split_ok(d, g_out, s.z_o, w.o_ln_w, dyb, nullptr, g_hi, g_lo);
    if (fuse_o) {
        if ((rc = ln_bwd_split(g_out, 1.0f, s.z_o, s.mean_o, s.rstd_o, w.o_ln_w, dyb, nullptr, pgam, pbet, &nblk, M, d, dr_o, g_hi, g_lo, pcol, st))) return rc;
        if ((rc = reduce_batch_add(rb, pcol, nblk, d, g.b2, nullptr, 1.f, 0, st))) return rc;
        ag = Act{nullptr, g_hi, g_lo};
    } else {
        if ((rc = ln_bwd(g_out, 1.0f, s.z_o, s.mean_o, s.rstd_o, w.o_ln_w, dyb, dr_o.thr ? dd : nullptr, pgam, pbet, &nblk, M, d, dr_o, st))) return rc;
    }
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.o_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.o_ln_b, nullptr, 1.f, 0, st))) return rc;
    if (!fuse_o) {
        const float* t2 = dr_o.thr ? dd : dyb;                       // grad wrt dense_2 output
        if ((rc = split_act(lin, t2, g_hi, g_lo, M * d, ag, st))) return rc;     // one split feeds dW2 and dh
    }
    if ((rc = linear_bwd_weight(lin, ag, ah, g.w2, fuse_o ? nullptr : g.b2, M, d, 4 * d, part_sk, part_cs, fuse_o ? fk.leaf() : st))) return rc;
    const Act adh{dh_, dh_hi, dh_lo};                                          // split pair emitted by the epilogue below
    if (lin.umma) {
        // dh leaves the GEMM only as its split pair (all dW1 / dX read) plus per-32-row column sums for db1: the fp32 copy
        // and the pass that summed it are gone; the [M,4d] dh_ buffer holds the [M/32, 4d] partial sums instead
        if ((rc = linear_bwd_input(lin, ag, w.w2, nullptr, M, d, 4 * d, 0, ACT_MUL_AUX, s.h_pre, dh_hi, dh_lo, st, dh_, wh(W_2), wl(W_2)))) return rc;
        if ((rc = reduce_batch_add(rb, dh_, 4 * (int)((M + 127) / 128), 4 * d, g.b1, nullptr, 1.f, 0, st))) return rc;
        if ((rc = linear_bwd_weight(lin, adh, ay, g.w1, nullptr, M, 4 * d, d, part_sk, part_cs, fk.leaf()))) return rc;
    } else {
        if ((rc = linear_bwd_input(lin, ag, w.w2, dh_, M, d, 4 * d, 0, ACT_MUL_AUX, s.h_pre, dh_hi, dh_lo, st))) return rc;
        if ((rc = linear_bwd_weight(lin, adh, ay, g.w1, g.b1, M, 4 * d, d, part_sk, part_cs, st))) return rc;
    }
    if ((rc = linear_bwd_input(lin, adh, w.w1, dyb, M, 4 * d, d, 1, ACT_NONE, nullptr, nullptr, nullptr, st, nullptr, wh(W_1), wl(W_1)))) return rc;   // dyb = dL/dy

    if (c.kind == 1) {
        // FMLPRecBlock: y = LN(drop(conv(x, K)) + x).  dx = d(LN input) + conv^T(K, d conv); dK = sum_b corr(d conv_b, x_b)
        pgam = lnp[2][0]; pbet = lnp[2][1];
        if ((rc = ln_bwd(dyb, 1.0f, s.z_f, s.mean_f, s.rstd_f, w.f_ln_w, dzf, dr_f.thr ? dd : nullptr, pgam, pbet, &nblk, M, d, dr_f, st))) return rc;
        if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.f_ln_w, nullptr, 1.f, 0, st))) return rc;
        if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.f_ln_b, nullptr, 1.f, 0, st))) return rc;
        const float* gconv = dr_f.thr ? dd : dzf;
        if ((rc = circ_apply(gconv, x, true, nullptr, dq, B, L, d, true, st))) return rc;              // per-sequence partials of dK
        if ((rc = reduce_batch_add(rb, dq, B, L * d, g.sqrt_beta, nullptr, 1.f, 0, st))) return rc;
        if ((rc = circ_apply(gconv, w.sqrt_beta, false, dzf, dx, B, L, d, true, st))) return rc;
        return reduce_batch_flush(rb, st);
    }
    // ---- attention branch: gsp = LN(drop(tmp) + x), weight (1-alpha)
    if ((rc = fk.join())) return rc;        // the next LayerNorm backward overwrites the gradient pair the side GEMMs read
    pgam = lnp[1][0]; pbet = lnp[1][1]; pcol = lnp[1][2];
    const bool fuse_a = lin.umma && ln_bwd_split_ok(d, dyb, s.z_a, w.a_ln_w, dx, nullptr, g_hi, g_lo);
    if (fuse_a) {
        if ((rc = ln_bwd_split(dyb, 1.0f - c.alpha, s.z_a, s.mean_a, s.rstd_a, w.a_ln_w, dx, nullptr, pgam, pbet, &nblk, M, d, dr_a, g_hi, g_lo, pcol, st))) return rc;
        if ((rc = reduce_batch_add(rb, pcol, nblk, d, g.bo, nullptr, 1.f, 0, st))) return rc;
        ag = Act{nullptr, g_hi, g_lo};
    } else {
        if ((rc = ln_bwd(dyb, 1.0f - c.alpha, s.z_a, s.mean_a, s.rstd_a, w.a_ln_w, dx, dr_a.thr ? dd : nullptr, pgam, pbet, &nblk, M, d, dr_a, st))) return rc;
    }
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.a_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.a_ln_b, nullptr, 1.f, 0, st))) return rc;
    if (!fuse_a) {
        const float* t1 = dr_a.thr ? dd : dx;                        // grad wrt dense output
        if ((rc = split_act(lin, t1, g_hi, g_lo, M * d, ag, st))) return rc;
    }
    if ((rc = linear_bwd_weight(lin, ag, actx, g.wo, fuse_a ? nullptr : g.bo, M, d, d, part_sk, part_cs, fuse_a ? fk.leaf() : st))) return rc;
    if ((rc = linear_bwd_input(lin, ag, w.wo, dctx, M, d, d, 0, ACT_NONE, nullptr, nullptr, nullptr, st, nullptr, wh(W_O), wl(W_O)))) return rc;
    const bool qkv_pack_ok = lin.umma && qkv_fuse_enabled() && d % 32 == 0 && umma_split_colsum_rows(M) <= colsum_partial_rows(M) &&
                             umma_split_colsum_ok(d, dq, dh_hi, dh_lo) && umma_split_colsum_ok(d, dk, dh_hi + d, dh_lo + d) &&
                             umma_split_colsum_ok(d, dv, dh_hi + 2 * d, dh_lo + 2 * d);
    bool pack32 = false;
    if (s.fused) {
        // dP (workspace, unused by this path) doubles as the [B,H,L] delta scratch
        // pack32: the tcgen05 kernels write dq | dk | dv straight into ONE packed [M, 3d] fp32 tensor (the three contiguous
        // [M, d] buffers seen with pitch 3d), which is then its own hi half for the packed dX / dW GEMMs below
        pack32 = qkv_pack_ok && attn_bwd_strided_ok(L, d, H) && dk == dq + M * d && dv == dk + M * d;
        if (pack32) { dk = dq + d; dv = dq + 2 * d; }
        if ((rc = attn_bwd(s.q, s.k, s.v, ids, s.ctx, s.lse, dctx, dq, dk, dv, dP, B, H, L, d, dr_p, st, pack32 ? 3 * d : 0))) return rc;
    } else {
        const float* Pd = s.P;
        if (dr_p.thr) {
            if ((rc = apply_dropout(s.P, Pd_ws, PL, dr_p, st))) return rc;
            Pd = Pd_ws;
        }
        {   // dv[j,:] = sum_i Pd[i,j] dctx[i,:]
            GemmP p;
            p.A = Pd; p.B = dctx; p.C = dv; p.M = L; p.N = dh; p.K = L; p.lda = L; p.ldb = d; p.ldc = d;
            p.a_kmajor = false; p.b_kmajor = false; p.batch = B * H; p.inner = H;
            p.sAo = (long long)H * L * L; p.sAi = (long long)L * L; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)L * d; p.sCi = dh;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        {   // dPd[i,j] = dctx[i,:] . v[j,:]
            GemmP p;
            p.A = dctx; p.B = s.v; p.C = dP; p.M = L; p.N = L; p.K = dh; p.lda = d; p.ldb = d; p.ldc = L;
            p.batch = B * H; p.inner = H;
            p.sAo = (long long)L * d; p.sAi = dh; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)H * L * L; p.sCi = (long long)L * L;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        if ((rc = softmax_bwd(s.P, dP, B, H, L, dr_p, st))) return rc;            // dP now holds dS
        {   // dq = dS k / sqrt(dh)
            GemmP p;
            p.A = dP; p.B = s.k; p.C = dq; p.M = L; p.N = dh; p.K = L; p.lda = L; p.ldb = d; p.ldc = d;
            p.b_kmajor = false; p.batch = B * H; p.inner = H; p.alpha = inv_sqrt_dh;
            p.sAo = (long long)H * L * L; p.sAi = (long long)L * L; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)L * d; p.sCi = dh;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        {   // dk = dS^T q / sqrt(dh)
            GemmP p;
            p.A = dP; p.B = s.q; p.C = dk; p.M = L; p.N = dh; p.K = L; p.lda = L; p.ldb = d; p.ldc = d;
            p.a_kmajor = false; p.b_kmajor = false; p.batch = B * H; p.inner = H; p.alpha = inv_sqrt_dh;
            p.sAo = (long long)H * L * L; p.sAi = (long long)L * L; p.sBo = (long long)L * d; p.sBi = dh;
            p.sCo = (long long)L * d; p.sCi = dh;
            if ((rc = launch_gemm(p, st))) return rc;
        }
    }
    {
        const float* grads3[3] = {dq, dk, dv};
        const float* ws3[3] = {w.wq, w.wk, w.wv};
        float* dws3[3] = {g.wq, g.wk, g.wv};
        float* dbs3[3] = {g.bq, g.bk, g.bv};
        // tcgen05 path: dq | dk | dv leave their split as ONE packed [M, 3d] pair (in the dead dh pair buffers), so the input
        // gradient is a single K = 3d GEMM against the [3d, d] weight pair instead of three accumulating launches that each
        // re-read and re-write dx; the three weight gradients read their column window of the packed pair.
        if ((rc = fk.join())) return rc;    // dWo is done before the pair buffers are reused below
        const bool pack = qkv_pack_ok;
        float* const g3_hi = pack32 ? dq : dh_hi;          // hi half of the packed pair: the packed fp32 tensor itself, or a copy
        const bool pack_dw = pack && ax.hi && umma_gemm_splits(3 * d, d, (int)M) > 1 &&
                             (size_t)umma_gemm_splits(3 * d, d, (int)M) * 3 * d * d * 4 <= lin.partial_bytes;
        if (pack) {
            for (int i = 0; i < 3; ++i) {
                float *ph = g3_hi + (long long)i * d, *pl = dh_lo + (long long)i * d;
                if ((rc = umma_split_colsum(grads3[i], pack32 ? nullptr : ph, pl, M, d, part_qkv[i], st, 3 * d, pack32 ? 3 * d : 0))) return rc;
                if ((rc = reduce_batch_add(rb, part_qkv[i], umma_split_colsum_rows(M), d, dbs3[i], nullptr, 1.f, 0, st))) return rc;
                if (!pack_dw && (rc = linear_bwd_weight(lin, Act{grads3[i], ph, pl}, ax, dws3[i], nullptr, M, d, d, part_sk, part_cs, st, 3 * d))) return rc;
            }
            // [dWq; dWk; dWv] = [dq | dk | dv]^T x: one split-K GEMM over the packed pair (x's pair read once instead of three
            // times), its reduction writes the three [d, d] gradient tensors
            if (pack_dw && (rc = umma_gemm_ps(3 * d, d, (int)M, g3_hi, dh_lo, M, 3 * d, 3 * d, true, ax.hi, ax.lo, M, d, d, true, g.wq, d,
                                              nullptr, 0, nullptr, nullptr, 1.f, 0, nullptr, nullptr, lin.partial, lin.partial_bytes, fk.leaf(),
                                              nullptr, 0, 0, nullptr, nullptr, d, g.wk, g.wv))) return rc;
            if ((rc = umma_gemm_ps((int)M, d, 3 * d, g3_hi, dh_lo, M, 3 * d, 3 * d, false, s.w_hi, s.w_lo, 3 * d, d, d, true, dx, d,
                                   nullptr, ACT_NONE, nullptr, nullptr, 1.f, 1, nullptr, nullptr, nullptr, 0, st))) return rc;
        } else
        for (int i = 0; i < 3; ++i) {
            if (lin.umma && umma_split_colsum_ok(d, grads3[i], g_hi, g_lo) && umma_split_colsum_rows(M) <= colsum_partial_rows(M)) {
                // the operand split of dq / dk / dv also sums their columns: no separate pass for the bias gradients
                ag = Act{grads3[i], g_hi, g_lo};
                if ((rc = umma_split_colsum(grads3[i], g_hi, g_lo, M, d, part_qkv[i], st))) return rc;
                if ((rc = reduce_batch_add(rb, part_qkv[i], umma_split_colsum_rows(M), d, dbs3[i], nullptr, 1.f, 0, st))) return rc;
                if ((rc = linear_bwd_weight(lin, ag, ax, dws3[i], nullptr, M, d, d, part_sk, part_cs, st))) return rc;
            } else {
                if ((rc = split_act(lin, grads3[i], g_hi, g_lo, M * d, ag, st))) return rc;
                if ((rc = linear_bwd_weight(lin, ag, ax, dws3[i], dbs3[i], M, d, d, part_sk, part_cs, st))) return rc;
            }
            if ((rc = linear_bwd_input(lin, ag, ws3[i], dx, M, d, d, 1, ACT_NONE, nullptr, nullptr, nullptr, st, nullptr, wh(W_Q + i), wl(W_Q + i)))) return rc;
        }
    }

    // ---- filter branch: dsp = LN(drop(u) + x), weight alpha
    pgam = lnp[2][0]; pbet = lnp[2][1];
    if ((rc = ln_bwd(dyb, c.alpha, s.z_f, s.mean_f, s.rstd_f, w.f_ln_w, dzf, nullptr, pgam, pbet, &nblk, M, d, dr_f, st))) return rc;
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.f_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.f_ln_b, nullptr, 1.f, 0, st))) return rc;
    if ((rc = filter_bwd(x, basis, w.sqrt_beta, dzf, dx, pbeta2, B, L, d, r, dr_f, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbeta2, B, d, g.sqrt_beta, w.sqrt_beta, 2.0f, 0, st))) return rc;
    if ((rc = reduce_batch_flush(rb, st))) return rc;
    return fk.join();
}

// ================================================================================================ last block, last row
// bsa_block_last_fwd / bsa_block_last_bwd: the block evaluated at position L-1 only (see lastrow.cu for the math).
namespace {

struct SavedLast {
    float *xl, *low, *z_f, *mean_f, *rstd_f, *ql, *gq, *p, *xs, *sp, *ctx, *z_a, *mean_a, *rstd_a, *y, *h_pre, *h_act,
        *z_o, *mean_o, *rstd_o;
    size_t bytes;
};

SavedLast carve_saved_last(void* ptr, const bsa_block_desc& c) {
    const long long B = c.B, d = c.d, H = c.H, L = c.L;
    Bump b(ptr, ~(size_t)0);
    SavedLast s;
    s.xl = b.f(B * d); s.low = b.f(B * d); s.z_f = b.f(B * d); s.mean_f = b.f(B); s.rstd_f = b.f(B);
    s.ql = b.f(B * d); s.gq = b.f(B * H * d); s.p = b.f(B * H * L); s.xs = b.f(B * H * d); s.sp = b.f(B * H);
    s.ctx = b.f(B * d); s.z_a = b.f(B * d); s.mean_a = b.f(B); s.rstd_a = b.f(B); s.y = b.f(B * d);
    s.h_pre = b.f(B * 4 * d); s.h_act = b.f(B * 4 * d); s.z_o = b.f(B * d); s.mean_o = b.f(B); s.rstd_o = b.f(B);
    s.bytes = b.off;
    return s;
}

// C[z] = A[z] op B[z] for z over heads: strides in floats
int head_gemm(const float* A, int lda, bool a_k, long long sA, const float* Bm, int ldb, bool b_k, long long sB, float* C,
              int ldc, long long sC, int M, int N, int K, int H, cudaStream_t st) {
    GemmP p;
    p.A = A; p.B = Bm; p.C = C; p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.a_kmajor = a_k; p.b_kmajor = b_k; p.batch = H; p.inner = H; p.sAi = sA; p.sBi = sB; p.sCi = sC;
    return launch_gemm(p, st);
}

size_t last_workspace_bytes(const bsa_block_desc& c) {
    const long long B = c.B, d = c.d, H = c.H;
    Bump b(nullptr, ~(size_t)0);
    for (int i = 0; i < 13; ++i) b.f(B * d);        // tmp, u, dsp | dyb, dd, dxl, dzf, dctx, dql, dlow, pb2, tmp, dd_a, tmp2
    b.f(B * 4 * d);                                 // dh
    b.f(B * H * d); b.f(B * H * d); b.f(B * H); b.f(B * H);       // dxs, dgq, dsp, dcq
    for (int i = 0; i < 6; ++i) b.f((long long)rowwise_partial_blocks(B) * d);      // three LayerNorm sites x (gamma, beta) partials
    b.f((long long)colsum_partial_rows(B) * 4 * d);
    b.f(splitk_floats(B, d));
    b.f(lin_ws_floats(d));                          // tcgen05 linears: weight pair, split-K slabs ...
    for (int i = 0; i < 4; ++i) b.f(B * 4 * d + 256);      // ... and two lazily split operand pairs
    b.f(small_sk_floats(B, d));                     // FFMA linears: split-K slabs
    return b.off + 8192;
}

// the [B, d] linears of the last-row path: fp32 FFMA GEMMs by default.  BSA_LAST_UMMA=1 sends them to the tcgen05 engine with
// operands split on the fly — measured (profiles/r2n_bench*.json): no gain, a [512, 256] x K = 256 problem is 4 output tiles,
// each launch pays the engine's prologue (TMEM allocation, tensor-map fetch) plus two split launches: C3 step 12.57 ->
// 12.90 ms, C1 1.55 -> 1.70 ms.  Kept as a switch; the lever for these launches is fusing them, not a bigger engine.
Lin make_last_lin(Bump& b, long long B, long long d) {
    static const bool use_umma = [] { const char* e = getenv("BSA_LAST_UMMA"); return e && atoi(e) == 1; }();
    if (linear_impl() != 2 || !use_umma) {
        Lin ln{false, nullptr, nullptr, nullptr, 0};
        ln.sk_floats = small_sk_floats(B, d);
        ln.sk = b.f(ln.sk_floats);
        return ln;
    }
    Lin ln = make_lin(b, d);
    ln.s1_hi = b.f(B * 4 * d + 256); ln.s1_lo = b.f(B * 4 * d + 256);
    ln.s2_hi = b.f(B * 4 * d + 256); ln.s2_lo = b.f(B * 4 * d + 256);
    return ln;
}

}  // namespace

bool block_last_supported(const bsa_block_desc& c) {
    return c.kind == 0 && c.H > 0 && c.d % c.H == 0 && c.d % 4 == 0 && lastrow_supported(c.L, c.d, c.H);
}

int block_last_fwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
                   const float* x, float* out_last, void* saved_p, void* ws, size_t ws_bytes, cudaStream_t st) {
    const int B = c.B, L = c.L, d = c.d, H = c.H, dh = d / H;
    BSA_REQUIRE(block_last_supported(c), "bsa_block_last_fwd: unsupported shape (L=%d d=%d H=%d)", L, d, H);
    if (B == 0) return 0;
    SavedLast s = carve_saved_last(saved_p, c);
    Bump b(ws, ws_bytes);
    float* tmp = b.f((long long)B * d);
    float* u = b.f((long long)B * d);
    float* dsp = b.f((long long)B * d);
    const Lin lin = make_last_lin(b, B, d);
    BSA_REQUIRE(b.ok(), "bsa_block_last_fwd: workspace too small (%zu < %zu)", ws_bytes, b.off);
    const Dropout dr_f = dropout_rows(make_dropout(c.seed, 1 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const Dropout dr_p = make_dropout(c.seed, 2 + 4 * c.layer, c.p_attn, c.training);
    const Dropout dr_a = dropout_rows(make_dropout(c.seed, 3 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const Dropout dr_o = dropout_rows(make_dropout(c.seed, 4 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const int r = filter_rank(L, c.kept_bins);
    int rc;
    if ((rc = gather_last_rows(x, s.xl, B, L, d, st))) return rc;
    const Act axl{s.xl, nullptr, nullptr};
    if ((rc = linear_fwd(lin, axl, w.wq, w.bq, s.ql, B, d, d, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;
    // gq[b,h,:] = Wk_h^T q_h
    if ((rc = head_gemm(s.ql, d, true, dh, w.wk, d, false, (long long)dh * d, s.gq, H * d, d, B, d, dh, H, st))) return rc;
    if ((rc = last_attn_fwd(x, ids, s.gq, s.ql, w.bk, basis, r, s.p, s.xs, s.sp, s.low, B, H, L, d, dr_p, st))) return rc;
    // ctx[b, h cols] = Wv_h xs_h + bv_h sp_h
    if ((rc = head_gemm(s.xs, H * d, true, d, w.wv, d, true, (long long)dh * d, s.ctx, d, dh, B, dh, d, H, st))) return rc;
    if ((rc = last_ctx_bias(s.ctx, w.bv, s.sp, B, d, H, st))) return rc;
    const Act actx{s.ctx, nullptr, nullptr};
    if ((rc = linear_fwd(lin, actx, w.wo, w.bo, tmp, B, d, d, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;
    // filter branch at row L-1
    if ((rc = last_filter_u(u, s.low, s.xl, w.sqrt_beta, B, d, st))) return rc;
    if ((rc = residual_ln_fwd(u, s.xl, w.f_ln_w, w.f_ln_b, dsp, s.z_f, s.mean_f, s.rstd_f, nullptr, 0.f, 0.f, B, d, dr_f, st))) return rc;
    if ((rc = residual_ln_fwd(tmp, s.xl, w.a_ln_w, w.a_ln_b, s.y, s.z_a, s.mean_a, s.rstd_a, dsp, c.alpha, 1.0f - c.alpha, B, d, dr_a, st))) return rc;
    const Act ay{s.y, nullptr, nullptr}, ah{s.h_act, nullptr, nullptr};
    if ((rc = linear_fwd(lin, ay, w.w1, w.b1, s.h_act, B, 4 * d, d, ACT_GELU_SAVE_PRE, s.h_pre, nullptr, nullptr, st))) return rc;
    if ((rc = linear_fwd(lin, ah, w.w2, w.b2, tmp, B, d, 4 * d, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;
    return residual_ln_fwd(tmp, s.y, w.o_ln_w, w.o_ln_b, out_last, s.z_o, s.mean_o, s.rstd_o, nullptr, 0.f, 0.f, B, d, dr_o, st);
}

// d_out_last [B,d] = gradient wrt the block's output at position L-1.  dx [B,L,d] (every position) and all of `g`
// are overwritten.
int block_last_bwd(const bsa_block_desc& c, const bsa_block_weights& w, const float* basis, const long long* ids,
                   const float* x, const void* saved_p, const float* d_out_last, float* dx, const bsa_block_grads& g,
                   void* ws, size_t ws_bytes, cudaStream_t st) {
    (void)ids;
    const int B = c.B, L = c.L, d = c.d, H = c.H, dh = d / H;
    BSA_REQUIRE(block_last_supported(c), "bsa_block_last_bwd: unsupported shape (L=%d d=%d H=%d)", L, d, H);
    if (B == 0) return 0;
    SavedLast s = carve_saved_last(const_cast<void*>(saved_p), c);
    Bump b(ws, ws_bytes);
    const long long Bd = (long long)B * d;
    float* dyb = b.f(Bd); float* dd = b.f(Bd); float* dxl = b.f(Bd); float* dzf = b.f(Bd); float* dctx = b.f(Bd);
    float* dql = b.f(Bd); float* dlow = b.f(Bd); float* pb2 = b.f(Bd); float* tmpbd = b.f(Bd);
    // The six bias gradients are column sums of [B, .] tensors: each source keeps its own buffer (dd / dd_a always receive
    // the pre-dropout gradient, tmpbd / tmpbd2) so that the sums join the one reduction launch at the end of the call.
    float* dd_a = b.f(Bd); float* tmpbd2 = b.f(Bd);
    float* dh_ = b.f(Bd * 4);
    float* dxs = b.f(Bd * H); float* dgq = b.f(Bd * H); float* dsp = b.f((long long)B * H); float* dcq = b.f((long long)B * H);
    const int nrb = rowwise_partial_blocks(B);
    float* lnp[3][2];                                      // per LayerNorm site: reduced together at the end (ReduceBatch)
    for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 2; ++k) lnp[i][k] = b.f((long long)nrb * d);
    float *pgam = lnp[0][0], *pbet = lnp[0][1];
    ReduceBatch rb;
    float* part_cs = b.f((long long)colsum_partial_rows(B) * 4 * d);
    float* part_sk = b.f(splitk_floats(B, d));
    const Lin lin = make_last_lin(b, B, d);
    BSA_REQUIRE(b.ok(), "bsa_block_last_bwd: workspace too small (%zu < %zu)", ws_bytes, b.off);
    const Dropout dr_f = dropout_rows(make_dropout(c.seed, 1 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const Dropout dr_p = make_dropout(c.seed, 2 + 4 * c.layer, c.p_attn, c.training);
    const Dropout dr_a = dropout_rows(make_dropout(c.seed, 3 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const Dropout dr_o = dropout_rows(make_dropout(c.seed, 4 + 4 * c.layer, c.p_hidden, c.training), L, L - 1);
    const int r = filter_rank(L, c.kept_bins);
    int rc, nblk = 0;
    const Act axl{s.xl, nullptr, nullptr}, actx{s.ctx, nullptr, nullptr}, ay{s.y, nullptr, nullptr}, ah{s.h_act, nullptr, nullptr};
    // weight-gradient GEMMs (leaves) run on the side stream; they share part_sk, in launch order (the BSA_LAST_UMMA variant
    // shares more scratch with the main chain and stays on one stream)
    Fork fk(st, !lin.umma);

    // ---- feed-forward
    if ((rc = ln_bwd(d_out_last, 1.0f, s.z_o, s.mean_o, s.rstd_o, w.o_ln_w, dyb, dd, pgam, pbet, &nblk, B, d, dr_o, st))) return rc;
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.o_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.o_ln_b, nullptr, 1.f, 0, st))) return rc;
    const Act at2{dd, nullptr, nullptr};
    if ((rc = linear_bwd_weight(lin, at2, ah, g.w2, nullptr, B, d, 4 * d, part_sk, part_cs, fk.leaf()))) return rc;
    if ((rc = reduce_batch_add(rb, dd, B, d, g.b2, nullptr, 1.f, 0, st))) return rc;
    if ((rc = linear_bwd_input(lin, at2, w.w2, dh_, B, d, 4 * d, 0, ACT_MUL_GELU_GRAD, s.h_pre, nullptr, nullptr, st))) return rc;
    const Act adh{dh_, nullptr, nullptr};
    if ((rc = linear_bwd_weight(lin, adh, ay, g.w1, nullptr, B, 4 * d, d, part_sk, part_cs, fk.leaf()))) return rc;
    if ((rc = reduce_batch_add(rb, dh_, B, 4 * d, g.b1, nullptr, 1.f, 0, st))) return rc;
    if ((rc = linear_bwd_input(lin, adh, w.w1, dyb, B, 4 * d, d, 1, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;   // dyb = dL/dy

    // ---- attention branch: gsp = LN(drop(tmp) + xl), weight (1-alpha); dxl starts as the residual gradient
    pgam = lnp[1][0]; pbet = lnp[1][1];
    if ((rc = ln_bwd(dyb, 1.0f - c.alpha, s.z_a, s.mean_a, s.rstd_a, w.a_ln_w, dxl, dd_a, pgam, pbet, &nblk, B, d, dr_a, st))) return rc;
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.a_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.a_ln_b, nullptr, 1.f, 0, st))) return rc;
    const Act at1{dd_a, nullptr, nullptr};
    if ((rc = linear_bwd_weight(lin, at1, actx, g.wo, nullptr, B, d, d, part_sk, part_cs, fk.leaf()))) return rc;
    if ((rc = reduce_batch_add(rb, dd_a, B, d, g.bo, nullptr, 1.f, 0, st))) return rc;
    if ((rc = linear_bwd_input(lin, at1, w.wo, dctx, B, d, d, 0, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;
    // ctx_h = Wv_h xs_h + bv_h sp_h
    //   dWv[h rows] = sum_b dctx_h^T xs_h   (M = dh, N = d, K = B per head)
    if ((rc = head_gemm(dctx, d, false, dh, s.xs, H * d, false, d, g.wv, d, (long long)dh * d, dh, d, B, H, fk.leaf()))) return rc;
    if ((rc = last_head_scale(tmpbd, dctx, s.sp, B, d, H, st))) return rc;
    if ((rc = reduce_batch_add(rb, tmpbd, B, d, g.bv, nullptr, 1.f, 0, st))) return rc;
    //   dxs_h = Wv_h^T dctx_h  ;  dsp_h = bv_h . dctx_h
    if ((rc = head_gemm(dctx, d, true, dh, w.wv, d, false, (long long)dh * d, dxs, H * d, d, B, d, dh, H, st))) return rc;
    if ((rc = last_dsp(dctx, w.bv, dsp, B, d, H, st))) return rc;

    // ---- filter branch: dsp = LN(drop(u) + xl), weight alpha
    pgam = lnp[2][0]; pbet = lnp[2][1];
    if ((rc = ln_bwd(dyb, c.alpha, s.z_f, s.mean_f, s.rstd_f, w.f_ln_w, dzf, nullptr, pgam, pbet, &nblk, B, d, dr_f, st))) return rc;
    if ((rc = reduce_batch_add(rb, pgam, nblk, d, g.f_ln_w, nullptr, 1.f, 0, st))) return rc;
    if ((rc = reduce_batch_add(rb, pbet, nblk, d, g.f_ln_b, nullptr, 1.f, 0, st))) return rc;
    if ((rc = last_filter_bwd(dzf, s.xl, s.low, w.sqrt_beta, dlow, dxl, pb2, B, d, dr_f, st))) return rc;
    if ((rc = reduce_batch_add(rb, pb2, B, d, g.sqrt_beta, w.sqrt_beta, 2.0f, 0, st))) return rc;

    // ---- per-sequence adjoint over all positions
    if ((rc = last_attn_bwd(x, s.p, s.gq, dxs, dsp, dlow, basis, r, dx, dgq, dcq, B, H, L, d, dr_p, st))) return rc;
    // gq_h = Wk_h^T q_h ; cq_h = q_h . bk_h
    if ((rc = head_gemm(dgq, H * d, true, d, w.wk, d, true, (long long)dh * d, dql, d, dh, B, dh, d, H, st))) return rc;
    if ((rc = last_add_head_bias(dql, dcq, w.bk, B, d, H, st))) return rc;
    if ((rc = head_gemm(s.ql, d, false, dh, dgq, H * d, false, d, g.wk, d, (long long)dh * d, dh, d, B, H, fk.leaf()))) return rc;
    if ((rc = last_head_scale(tmpbd2, s.ql, dcq, B, d, H, st))) return rc;
    if ((rc = reduce_batch_add(rb, tmpbd2, B, d, g.bk, nullptr, 1.f, 0, st))) return rc;
    // q = xl Wq^T + bq
    const Act adql{dql, nullptr, nullptr};
    if ((rc = linear_bwd_weight(lin, adql, axl, g.wq, nullptr, B, d, d, part_sk, part_cs, fk.leaf()))) return rc;
    if ((rc = reduce_batch_add(rb, dql, B, d, g.bq, nullptr, 1.f, 0, st))) return rc;
    if ((rc = linear_bwd_input(lin, adql, w.wq, dxl, B, d, d, 1, ACT_NONE, nullptr, nullptr, nullptr, st))) return rc;
    // value.bias / key.bias done; the row L-1 residual paths go into dx last
    if ((rc = last_add_rows(dx, dxl, B, L, d, st))) return rc;
    if ((rc = reduce_batch_flush(rb, st))) return rc;
    return fk.join();
}

size_t block_last_saved_bytes(const bsa_block_desc& c) { return carve_saved_last(nullptr, c).bytes + 256; }
size_t block_last_workspace_bytes(const bsa_block_desc& c) { return last_workspace_bytes(c); }

size_t block_saved_bytes(const bsa_block_desc& c) { return carve_saved(nullptr, c).bytes + 256; }
size_t block_workspace_bytes(const bsa_block_desc& c) { return workspace_bytes(c); }

}  // namespace bsa

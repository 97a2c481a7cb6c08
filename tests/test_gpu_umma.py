"""tcgen05 (3xTF32) engine through the C ABI: GEMM in all operand orientations, fused cross-entropy and
fused scoring/top-k, each against a torch fp64 reference and against the fp32 FFMA path of the same library."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _gemm(impl, A, B, M, N, K, a_mn, b_mn, bias=None, alpha=1.0, acc=None):
    from bsarec_b200 import _cabi
    lib = _cabi.load()
    dev = A.device
    Cout = acc.clone() if acc is not None else torch.empty(M, N, device=dev)
    lda, ldb = A.stride(0), B.stride(0)
    wsb = lib.bsa_gemm_workspace_bytes(M, N, K, lda, ldb, int(a_mn), int(b_mn))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    _cabi.check(lib.bsa_gemm(impl, M, N, K, A.data_ptr(), lda, int(a_mn), B.data_ptr(), ldb, int(b_mn), Cout.data_ptr(),
                             Cout.stride(0), None if bias is None else bias.data_ptr(), alpha, int(acc is not None),
                             ws.data_ptr(), wsb, st), "bsa_gemm")
    torch.cuda.synchronize()
    return Cout


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (128, 256, 64), (200, 130, 72), (517, 300, 260), (6, 24, 8), (1000, 64, 1024)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_umma_gemm_matches_fp64_and_ffma(M, N, K, a_mn, b_mn, cuda_device):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + 2 * a_mn + b_mn)
    # pitches are multiples of 4 elements (16 B, a TMA requirement); logical sizes stay ragged
    lda = ((M if a_mn else K) + 3) // 4 * 4
    ldb = ((N if b_mn else K) + 3) // 4 * 4
    A = torch.randn((K if a_mn else M), lda, generator=g)
    B = torch.randn((K if b_mn else N), ldb, generator=g)
    Al = (A[:, :M].t() if a_mn else A[:, :K]).double()
    Bl = (B[:, :N].t() if b_mn else B[:, :K]).double()
    ref = Al @ Bl.t()
    Ad, Bd = A.to(cuda_device), B.to(cuda_device)
    got = _gemm(2, Ad, Bd, M, N, K, a_mn, b_mn).cpu()
    ffma = _gemm(1, Ad, Bd, M, N, K, a_mn, b_mn).cpu()
    e_umma, e_ffma = rel_err(got, ref), rel_err(ffma, ref)
    assert e_umma < 5e-6, (e_umma, e_ffma)          # 3xTF32 stays at fp32-rounding level
    assert e_ffma < 5e-6


@pytest.mark.parametrize("M,N,K", [(300, 200, 96), (130, 202, 40), (1000, 1024, 64)])
def test_umma_gemm_bias_alpha_accumulate(cuda_device, M, N, K):
    # accumulate goes through the epilogue that prefetches its global reads (ragged rows, a pitch that rules out float4)
    g = torch.Generator().manual_seed(1)
    A, B = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g)
    bias, C0 = torch.randn(N, generator=g), torch.randn(M, N, generator=g)
    ref = 0.5 * (A.double() @ B.double().t()) + bias.double() + C0.double()
    got = _gemm(2, A.to(cuda_device), B.to(cuda_device), M, N, K, False, False, bias.to(cuda_device), 0.5, C0.to(cuda_device)).cpu()
    assert rel_err(got, ref) < 5e-6


def _ce(impl, s, E, ans, dloss=1.0, reuse=False):
    from bsarec_b200 import functional as Fn
    loss, lse, ws = Fn.ce_forward(s, E, ans, impl)
    dE = torch.empty_like(E)
    ds = Fn.ce_backward(s, E, ans, lse, torch.tensor(dloss, device=s.device), dE, impl, ws if reuse else None)
    torch.cuda.synchronize()
    return loss.cpu(), lse.cpu(), ds.cpu(), dE.cpu()


@pytest.mark.parametrize("B,d,N", [(6, 32, 211), (128, 64, 1000), (37, 64, 12102), (512, 256, 20000), (130, 128, 5003), (300, 192, 40000), (77, 256, 333)])
def test_fused_cross_entropy_fwd_bwd(B, d, N, cuda_device):
    g = torch.Generator().manual_seed(B + d + N)
    s = torch.randn(B, d, generator=g) * 1.5
    E = torch.randn(N, d, generator=g) * 0.3
    ans = torch.randint(1, N, (B,), generator=g)
    ans[0] = N - 1                                  # target in the ragged last tile
    sd, Ed = s.double().requires_grad_(True), E.double().requires_grad_(True)
    logits = sd @ Ed.t()
    ref_loss = torch.nn.functional.cross_entropy(logits, ans)
    (2.5 * ref_loss).backward()
    ref_lse = torch.logsumexp(logits.detach(), dim=1)
    for impl, reuse in ((1, False), (2, False), (2, True)):        # reuse: backward reads the forward's operand splits
        loss, lse, ds, dE = _ce(impl, s.to(cuda_device), E.to(cuda_device), ans.to(cuda_device), 2.5, reuse)
        assert abs(float(loss) - float(ref_loss)) < 2e-6 * abs(float(ref_loss)), impl
        assert rel_err(lse, ref_lse) < 4e-6, impl
        assert rel_err(ds, sd.grad) < 1e-5, impl
        assert rel_err(dE, Ed.grad) < 1e-5, impl


@pytest.mark.parametrize("B,d,N", [(512, 256, 20000), (77, 256, 333), (300, 192, 40000), (129, 256, 65), (64, 132, 12102),
                                   (1000, 256, 5000)])
def test_ce_backward_modes_agree(B, d, N, cuda_device):
    """The three tcgen05 cross-entropy backward paths — 2: logits stored by the forward, ds / dE by the A-from-TMEM
    kernels of ce_bwd_logits.cuh (+ onehot rows); 1: logits recomputed, fused gradient-tile + ds kernel; 0: separate
    GEMMs — against fp64, including duplicate answers (several users with the same target item), ragged user and item
    tiles, and bit-reproducibility of mode 2."""
    from bsarec_b200 import _cabi
    lib = _cabi.load()
    g = torch.Generator().manual_seed(B * 7 + d + N)
    s = torch.randn(B, d, generator=g) * 1.5
    E = torch.randn(N, d, generator=g) * 0.3
    ans = torch.randint(1, N, (B,), generator=g)
    ans[0] = N - 1
    ans[1:6] = ans[7]                               # shared target: the onehot fix must sum them deterministically
    sd, Ed = s.double().requires_grad_(True), E.double().requires_grad_(True)
    ref_loss = torch.nn.functional.cross_entropy(sd @ Ed.t(), ans)
    (1.7 * ref_loss).backward()
    old = lib.bsa_get_ce_mode()
    try:
        outs = {}
        for mode in (2, 1, 0, 2):
            lib.bsa_set_ce_mode(mode)
            loss, lse, ds, dE = _ce(2, s.to(cuda_device), E.to(cuda_device), ans.to(cuda_device), 1.7, True)
            assert abs(float(loss) - float(ref_loss)) < 2e-6 * abs(float(ref_loss)), mode
            assert rel_err(ds, sd.grad) < 1e-5, (mode, rel_err(ds, sd.grad))
            assert rel_err(dE, Ed.grad) < 1e-5, (mode, rel_err(dE, Ed.grad))
            if mode == 2 and 2 in outs:
                assert torch.equal(ds, outs[2][0]) and torch.equal(dE, outs[2][1])      # bit-reproducible
            outs[mode] = (ds, dE)
    finally:
        lib.bsa_set_ce_mode(old)


def _topk(impl, s, E, indptr, indices, k, offset=0):
    from bsarec_b200 import functional as Fn
    val, idx = Fn.score_topk(s, E, indptr, indices, k, offset, impl)
    torch.cuda.synchronize()
    return val.cpu().numpy(), idx.cpu().numpy()


def _ref_topk(s, E, indptr, indices, k, offset=0):
    """numpy: exact fp64 scores, history -> 0.0, order by (score desc, id asc)."""
    sc = (s.double() @ E.double().t()).numpy()
    out_v, out_i = [], []
    for b in range(sc.shape[0]):
        row = sc[b].copy()
        h = indices[indptr[b]:indptr[b + 1]] - offset
        h = h[(h >= 0) & (h < row.shape[0])]
        row[h] = 0.0
        order = np.lexsort((np.arange(row.shape[0]), -row))[:k]
        out_v.append(row[order])
        out_i.append(order + offset)
    return np.array(out_v), np.array(out_i)


@pytest.mark.parametrize("B,d,N,k", [(5, 32, 97, 20), (64, 64, 3417, 20), (130, 64, 12102, 10), (256, 128, 30000, 20),
                                     (37, 32, 300000, 20)])       # last: large enough for the seeded-threshold pre-pass
@pytest.mark.parametrize("negative", [False, True])
def test_fused_score_topk(B, d, N, k, negative, cuda_device):
    g = torch.Generator().manual_seed(B * 3 + N + k)
    s = torch.randn(B, d, generator=g)
    E = torch.randn(N, d, generator=g) * 0.2
    if negative:                       # most scores negative -> the 0.0 history entries must surface in the top-k
        E = -E.abs() * 0.1
        s = s.abs()
    counts = torch.randint(0, 60, (B,), generator=g).numpy()
    indptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    indices = np.concatenate([np.sort(np.random.default_rng(b).choice(np.arange(1, N), c, replace=False)) for b, c in enumerate(counts)] + [np.zeros(0, np.int64)]).astype(np.int64)
    rv, ri = _ref_topk(s, E, indptr, indices, k)
    ip = torch.from_numpy(indptr).to(cuda_device)
    ix = torch.from_numpy(indices if len(indices) else np.zeros(1, np.int64)).to(cuda_device)
    for impl in (1, 2):
        v, i = _topk(impl, s.to(cuda_device), E.to(cuda_device), ip, ix, k)
        assert np.allclose(v, rv, rtol=2e-5, atol=2e-6), impl
        diff = (i != ri)
        # an index may differ only where two fp64 scores are closer than fp32 resolution
        for b, q in zip(*np.nonzero(diff)):
            assert abs(rv[b, q] - v[b, q]) <= 2e-5 * max(1.0, abs(rv[b, q])), (impl, b, q)
        assert diff.mean() < 0.01, impl
        assert np.all(np.diff(v, axis=1) <= 0)          # descending


def test_sharded_topk_merge_equals_unsharded(cuda_device):
    from bsarec_b200 import functional as Fn
    g = torch.Generator().manual_seed(4)
    B, d, N, k, G = 70, 64, 10007, 20, 4
    s, E = torch.randn(B, d, generator=g).to(cuda_device), (torch.randn(N, d, generator=g) * 0.2).to(cuda_device)
    ids = torch.randint(1, N, (B, 30), generator=g)
    rows = [torch.unique(ids[b]) for b in range(B)]       # sorted, de-duplicated
    indptr = torch.tensor([0] + list(np.cumsum([len(r) for r in rows])), dtype=torch.int64).to(cuda_device)
    indices = torch.cat(rows).to(cuda_device)
    for impl in (1, 2):
        full_v, full_i = Fn.score_topk(s, E, indptr, indices, k, 0, impl)
        bounds = [N * j // G for j in range(G + 1)]
        cv, ci = [], []
        for j in range(G):
            v, i = Fn.score_topk(s, E[bounds[j]:bounds[j + 1]].contiguous(), indptr, indices, k, bounds[j], impl)
            cv.append(v)
            ci.append(i)
        mv, mi = Fn.topk_merge(torch.stack(cv), torch.stack(ci))
        assert torch.equal(mi, full_i) and torch.equal(mv, full_v), impl


@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_fused_adam_matches_torch_adam(wd, cuda_device):
    """bsa_adam_step (large tensors) and bsa_adam_step_multi (all the small ones in one launch, > 48 of them -> two
    launches) vs torch.optim.Adam (trainers.py:27-28) over several steps, ragged sizes included."""
    from bsarec_b200 import FusedAdam
    g = torch.Generator().manual_seed(3)
    shapes = [(1001, 64), (7,), (64, 64), (1, 1, 33), (4200, 257)] + [(3 + i, 5) for i in range(50)]
    assert 4200 * 257 > FusedAdam.SMALL >= 1001 * 64
    ref = [torch.randn(s, generator=g).to(cuda_device).requires_grad_(True) for s in shapes]
    got = [r.detach().clone().requires_grad_(True) for r in ref]
    o_ref = torch.optim.Adam(ref, lr=1e-3, betas=(0.9, 0.999), weight_decay=wd)
    o_got = FusedAdam(got, lr=1e-3, betas=(0.9, 0.999), weight_decay=wd)
    for step in range(5):
        for r, q in zip(ref, got):
            grad = (torch.randn(r.shape, generator=g) * (10.0 ** (step - 2))).to(cuda_device)
            r.grad, q.grad = grad.clone(), grad.clone()
        o_ref.step()
        o_got.step()
    for r, q in zip(ref, got):
        assert rel_err(q.detach().cpu(), r.detach().cpu()) < 1e-6
    assert o_got.state[got[0]]["step"] == 5


def test_score_topk_split_cache_follows_item_table_version(cuda_device):
    """The scoring workspace (split-TF32 copy of the item table) is reused across eval batches while the table's
    version counter stands still, and rebuilt after any in-place update — including FusedAdam's raw-pointer update."""
    from bsarec_b200 import functional as Fn, FusedAdam
    g = torch.Generator().manual_seed(8)
    B, d, N, k = 70, 64, 5003, 20
    s = torch.randn(B, d, generator=g).to(cuda_device)
    E = torch.nn.Parameter((torch.randn(N, d, generator=g) * 0.2).to(cuda_device))
    cache = {}
    v0, i0 = Fn.score_topk(s, E.detach(), None, None, k, 0, 2, cache=cache)
    ws = cache["ws"]
    v1, i1 = Fn.score_topk(s, E.detach(), None, None, k, 0, 2, cache=cache)          # reuse_split = 1
    assert cache["ws"] is ws and torch.equal(v0, v1) and torch.equal(i0, i1)
    E.grad = torch.randn(N, d, generator=g).to(cuda_device)
    FusedAdam([E], lr=0.05).step()                                                   # table changes, version bumps
    v2, i2 = Fn.score_topk(s, E.detach(), None, None, k, 0, 2, cache=cache)
    fresh_v, fresh_i = Fn.score_topk(s, E.detach(), None, None, k, 0, 2)
    assert cache["ws"] is not ws
    assert torch.equal(v2, fresh_v) and torch.equal(i2, fresh_i)
    assert not torch.equal(i2, i0)

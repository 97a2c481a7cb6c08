"""Bit-exact index work (north_star: "bit-exact for embedding gathers, scatter indices ..."): the row gather
``E[ids]`` (_abstract_model.py:17-18), the sorted scatter plan of the embedding backward (ATen's sort-based
``embedding_dense_backward``) and ``padding_idx=0`` (_abstract_model.py:10).  Everything here is compared with
``torch.equal`` / ``np.array_equal`` — no tolerance."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _id_streams(N, M, seed):
    """Heavy-duplicate, Zipf(1.05) and uniform id streams with padding zeros mixed in."""
    rng = np.random.RandomState(seed)
    zipf = np.minimum(rng.zipf(1.05, size=M), N - 1).astype(np.int64)
    heavy = rng.choice(np.array([1, 2, 3, N - 1, N // 2], dtype=np.int64), size=M)
    uni = rng.randint(1, N, size=M).astype(np.int64)
    out = {}
    for name, ids in (("zipf", zipf), ("heavy", heavy), ("uniform", uni)):
        ids = ids.copy()
        ids[rng.rand(M) < 0.3] = 0          # left padding
        out[name] = ids
    out["all_padding"] = np.zeros(M, dtype=np.int64)
    out["single"] = np.full(M, 7, dtype=np.int64)
    return out


@pytest.mark.parametrize("N,d", [(12102, 64), (1_000_000, 256), (257, 4)])
def test_gather_rows_bit_exact(N, d, cuda_device):
    from bsarec_b200 import functional as Fn
    g = torch.Generator().manual_seed(N + d)
    E = torch.randn(N, d, generator=g).to(cuda_device)
    for name, ids in _id_streams(N, 50_000, 3).items():
        idx = torch.from_numpy(ids).to(cuda_device)
        got = Fn.gather_rows(E, idx)
        assert torch.equal(got, E[idx]), name
    assert Fn.gather_rows(E, torch.zeros(0, dtype=torch.int64, device=cuda_device)).shape == (0, d)


def test_embed_kernel_gathers_exactly_the_indexed_row(cuda_device):
    """Through the real fused kernel (bsa_embed_ln_dropout_fwd): item row i carries the binary code of i as +-1 columns
    (plus zero columns), the position table is 0, LayerNorm is (1, 0).  LayerNorm keeps the sign of (x - mean), |mean| < 1,
    so the gathered row index is decoded EXACTLY from the output's sign pattern — for every position of every sequence."""
    from bsarec_b200 import _cabi
    from bsarec_b200._cabi import EmbedDesc, check
    lib = _cabi.load()
    N, d, B, L, bits = 100_003, 32, 64, 50, 17
    code = ((np.arange(N)[:, None] >> np.arange(bits)[None, :]) & 1).astype(np.float32) * 2.0 - 1.0
    E = np.zeros((N, d), dtype=np.float32)
    E[:, :bits] = code
    rng = np.random.RandomState(0)
    ids = rng.randint(1, N, size=(B, L)).astype(np.int64)
    ids[:, :10] = 0                          # padding rows gather row 0 (all bits clear), as nn.Embedding does
    ids[3] = 0
    ids[5, 20:] = N - 1
    Ed, idd = torch.from_numpy(E).to(cuda_device), torch.from_numpy(ids).to(cuda_device)
    pos = torch.zeros(L, d, device=cuda_device)
    w, b = torch.ones(d, device=cuda_device), torch.zeros(d, device=cuda_device)
    out = torch.empty(B, L, d, device=cuda_device)
    mean, rstd = torch.empty(B, L, device=cuda_device), torch.empty(B, L, device=cuda_device)
    ed = EmbedDesc(B=B, L=L, d=d, item_size=N, p_hidden=0.0, training=0, seed=0)
    check(lib.bsa_embed_ln_dropout_fwd(C.byref(ed), idd.data_ptr(), Ed.data_ptr(), pos.data_ptr(), w.data_ptr(), b.data_ptr(),
                                       out.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream), "embed fwd")
    o = out.cpu().numpy()
    decoded = ((o[:, :, :bits] > 0).astype(np.int64) << np.arange(bits)[None, None, :]).sum(-1)
    assert np.array_equal(decoded, ids)


@pytest.mark.parametrize("N", [12102, 1_000_000])
def test_sort_scatter_plan_equals_stable_sort(N, cuda_device):
    """(sorted_ids, sorted_rows) == torch.sort(ids, stable=True): ids ascending, rows of equal ids in their original
    order — the summation order of the atomic-free segmented sum, hence of the item-table gradient bits."""
    from bsarec_b200 import _cabi
    from bsarec_b200._cabi import check
    lib = _cabi.load()
    M = 512 * 200
    for name, ids in _id_streams(N, M, 11).items():
        idd = torch.from_numpy(ids).to(cuda_device)
        pb = lib.bsa_sort_scatter_plan_bytes(M)
        plan = torch.empty(pb, dtype=torch.uint8, device=cuda_device)
        sid = torch.empty(M, dtype=torch.int32, device=cuda_device)
        srow = torch.empty(M, dtype=torch.int32, device=cuda_device)
        check(lib.bsa_sort_scatter_plan(idd.data_ptr(), M, N, plan.data_ptr(), pb, sid.data_ptr(), srow.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream), "plan")
        ref_ids, ref_rows = torch.sort(idd, stable=True)
        assert torch.equal(sid.long(), ref_ids), name
        assert torch.equal(srow.long(), ref_rows), name


@pytest.mark.parametrize("N,d", [(12102, 64), (200_000, 256)])
def test_scatter_add_skips_padding_and_is_order_exact(N, d, cuda_device):
    """dE[id] += rows for id != 0 ONLY (padding_idx=0): row 0 keeps its bits; every other row equals the SEQUENTIAL fp32
    sum of its rows in original order, bit for bit (the stable plan fixes the order); untouched rows keep their bits;
    two runs give identical bits."""
    from bsarec_b200 import functional as Fn
    M = 40_000
    g = torch.Generator().manual_seed(d)
    for name, ids in _id_streams(N, M, 5).items():
        rows = torch.randn(M, d, generator=g)
        base = torch.randn(N, d, generator=g)
        dE = base.clone().to(cuda_device)
        Fn.scatter_add_rows(torch.from_numpy(ids).to(cuda_device), rows.to(cuda_device), dE)
        dE2 = base.clone().to(cuda_device)
        Fn.scatter_add_rows(torch.from_numpy(ids).to(cuda_device), rows.to(cuda_device), dE2)
        assert torch.equal(dE, dE2), name
        got = dE.cpu()
        assert torch.equal(got[0], base[0]), name                     # padding row untouched
        touched = np.unique(ids[ids != 0])
        mask = torch.ones(N, dtype=torch.bool)
        mask[torch.from_numpy(touched)] = False
        assert torch.equal(got[mask], base[mask]), name               # rows without an occurrence untouched
        # sequential fp32 reference in original row order for a sample of ids (incl. the most frequent one)
        vals, counts = np.unique(ids[ids != 0], return_counts=True)
        if len(vals) == 0:
            continue
        sample = set(vals[np.argsort(-counts)[:3]].tolist()) | set(vals[:: max(1, len(vals) // 20)].tolist())
        for i in sample:
            occ = np.nonzero(ids == i)[0]
            acc = base[i].clone()                                     # the kernel starts from dE[id] and adds the rows in order
            for r in occ:
                acc = acc + rows[r]
            assert torch.equal(got[i], acc), (name, i, len(occ))

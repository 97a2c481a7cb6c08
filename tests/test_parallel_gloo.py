"""world_size-2 checks of the multi-GPU host logic on CPU (gloo): data-parallel gradient averaging and the
item-sharded evaluation exchange.  The per-shard scorer / merger handed to
``sharded_full_sort_topk`` are numpy stand-ins with the C-ABI kernels' contract (score desc, id asc,
history items scored 0.0); the CUDA kernels themselves are covered by the gpu tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bsarec_b200.parallel import GradExchange, allreduce_gradients, shard_range, sharded_full_sort_topk

import oracle


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_rows_exactly():
    for n, w in [(12102, 8), (1_000_000, 8), (7, 4), (3, 4), (100, 1)]:
        edges = [shard_range(n, w, r) for r in range(w)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        for (a, b), (c, d) in zip(edges, edges[1:]):
            assert b == c and b >= a
        sizes = [b - a for a, b in edges]
        assert max(sizes) - min(sizes) <= 1


# ------------------------------------------------------------------------------------------- training
CFG = dict(item_size=97, hidden_size=16, max_seq_length=12, num_attention_heads=2, num_hidden_layers=1, c=3, alpha=0.6)


def _dp_worker(rank, world, out_dir):
    cfg = oracle.OracleConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **CFG)
    P = oracle.init_params(cfg, seed=3)
    g = torch.Generator().manual_seed(11)
    B = 4 * world
    ids = torch.randint(1, cfg.item_size, (B, cfg.max_seq_length), generator=g)
    ids[:, :5] = 0
    ids[1] = 0
    answers = torch.randint(1, cfg.item_size, (B,), generator=g)
    # every rank computes its shard's gradient with the oracle (stand-in for the CUDA path) ...
    lo, hi = rank * 4, rank * 4 + 4
    _, grads = oracle.loss_and_grads(P, ids[lo:hi], answers[lo:hi], cfg)
    params = []
    for k in P:
        p = torch.nn.Parameter(P[k].clone())
        p.grad = grads[k].clone()
        params.append(p)
    # ... and one bucketed all-reduce turns them into the global-batch gradient; tiny buckets force
    # both the single-tensor and the packed multi-tensor branch
    calls = allreduce_gradients(params, bucket_elems=600)
    assert calls > 1
    _, full = oracle.loss_and_grads(P, ids, answers, cfg)
    for p, k in zip(params, P):
        err = (p.grad - full[k]).abs().max().item()
        assert err <= 1e-6 * max(1.0, full[k].abs().max().item()) + 1e-9, (k, err)
    torch.save([p.grad for p in params], os.path.join(out_dir, f"g{rank}.pt"))


def test_data_parallel_allreduce_gives_global_batch_gradient(tmp_path):
    _spawn(_dp_worker, 2, str(tmp_path))
    g0 = torch.load(os.path.join(tmp_path, "g0.pt"))
    g1 = torch.load(os.path.join(tmp_path, "g1.pt"))
    for a, b in zip(g0, g1):        # replicas stay bit-identical after the collective
        assert torch.equal(a, b)


def _noop_worker(rank, world):
    p = torch.nn.Parameter(torch.ones(3))
    p.grad = torch.full((3,), 2.0)
    assert allreduce_gradients([p]) == 0
    assert torch.equal(p.grad, torch.full((3,), 2.0))


def test_allreduce_is_a_noop_on_one_rank():
    _spawn(_noop_worker, 1)


def _index_add_scatter(ids, rows, dE):
    """torch stand-in for bsa_scatter_add_rows: dE[ids[i]] += rows[i], id 0 (padding_idx) skipped, row order."""
    keep = ids != 0
    dE.index_add_(0, ids[keep], rows[keep])


def _exchange_worker(rank, world, out_dir):
    """GradExchange == all-reduce of the full per-rank gradients: sum_r (dense_r + scatter(ids_r, dz_r)), and the flat
    small-gradient buffer summed; replicas end bit-identical."""
    N, d, M = 53, 8, 40
    per_rank = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        ids = torch.randint(1, N, (5, 8), generator=g)
        for b in range(5):                                        # left padding of a different length per row and rank:
            ids[b, :int(torch.randint(0, 9, (1,), generator=g))] = 0      # ranks exchange different numbers of rows
        per_rank.append(dict(ids=ids, dense=torch.randn(N, d, generator=g), dz=torch.randn(M, d, generator=g),
                             small=torch.randn(77, generator=g)))
    me = {k: v.clone() for k, v in per_rank[rank].items()}
    ex = GradExchange()
    assert ex.world == world and ex.scale == 1.0 / world
    st = ex.prepare(me["ids"])
    ex.start_dense(st, me["dense"])
    ex.finish(st, me["small"], me["dz"], me["dense"], _index_add_scatter)
    want = torch.zeros(N, d, dtype=torch.float64)
    small = torch.zeros(77, dtype=torch.float64)
    for r in range(world):
        want += per_rank[r]["dense"].double()
        _index_add_scatter(per_rank[r]["ids"].reshape(-1), per_rank[r]["dz"].double(), want)
        small += per_rank[r]["small"].double()
    assert (me["dense"].double() - want).abs().max() < 1e-5
    assert (me["small"].double() - small).abs().max() < 1e-5
    assert ex.collectives == 5                                    # counts, dense, small, ids, rows
    torch.save([me["dense"], me["small"]], os.path.join(out_dir, f"x{rank}.pt"))


def test_overlapped_gradient_exchange_equals_full_allreduce(tmp_path):
    _spawn(_exchange_worker, 2, str(tmp_path))
    a = torch.load(os.path.join(tmp_path, "x0.pt"))
    b = torch.load(os.path.join(tmp_path, "x1.pt"))
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def _vocab_parallel_worker(rank, world):
    """Vocabulary-parallel CE host logic with torch stand-ins for the shard kernels: the combined loss equals the
    full-catalog cross-entropy of the GLOBAL batch; each rank ends with the complete gradient of ITS rows (dense part +
    the gather rows of every rank routed to it) and the summed ds of ITS users."""
    N, d, Bl, L = 48, 8, 5, 6
    g = torch.Generator().manual_seed(9)
    E = torch.randn(N, d, generator=g, dtype=torch.float64)
    s_glob = torch.randn(world * Bl, d, generator=g, dtype=torch.float64)
    ans_glob = torch.randint(1, N, (world * Bl,), generator=g)
    ids_glob = torch.randint(1, N, (world * Bl, L), generator=g)
    ids_glob[:, :2] = 0
    ids_glob[0] = 0                                              # a rank-0 row that sends nothing
    dz_glob = torch.randn(world * Bl * L, d, generator=g, dtype=torch.float64)
    small_glob = torch.randn(world, 13, generator=g, dtype=torch.float64)
    mine = slice(rank * Bl, (rank + 1) * Bl)
    ex = GradExchange(vocab_parallel=True)
    lo, hi = ex.shard(N)
    st = ex.prepare_routed(ids_glob[mine], N)
    s_all, ans_local = ex.gather_users(s_glob[mine].clone(), ans_glob[mine].clone(), N)
    assert torch.equal(s_all, s_glob)
    inside = (ans_glob >= lo) & (ans_glob < hi)
    assert torch.equal(ans_local, torch.where(inside, ans_glob - lo, torch.full_like(ans_glob, -1)))
    # stand-in for bsa_ce_fwd_partial on the shard
    logits = s_all @ E[lo:hi].T
    lse_p = torch.logsumexp(logits, 1)
    tgt_p = torch.where(ans_local >= 0, logits.gather(1, ans_local.clamp_min(0)[:, None]).squeeze(1), torch.zeros_like(lse_p))
    loss, lse = ex.combine_loss(lse_p, tgt_p)
    full = s_glob @ E.T
    want_loss = torch.nn.functional.cross_entropy(full, ans_glob)
    assert abs(float(loss) - float(want_loss)) < 1e-12
    assert (lse - torch.logsumexp(full, 1)).abs().max() < 1e-12
    # stand-in for bsa_ce_bwd on the shard: P = softmax rows of the shard / B_global, minus the onehot where it is local
    G = torch.exp(logits - lse[:, None]) / (world * Bl)
    rows_in = (ans_local >= 0).nonzero().flatten()
    G[rows_in, ans_local[rows_in]] -= 1.0 / (world * Bl)
    ds_all = G @ E[lo:hi]
    buf = torch.zeros(hi - lo + 1, d, dtype=torch.float64)
    buf[1:] = G.T @ s_all
    ds = ex.reduce_scatter_users(ds_all)
    sm = small_glob[rank].clone()
    ex.finish_routed(st, sm, dz_glob[rank * Bl * L:(rank + 1) * Bl * L].clone(), buf, lo, _index_add_scatter)
    # reference: autograd of the global loss + the scatter of every rank's rows
    Eg = E.clone().requires_grad_(True)
    sg = s_glob.clone().requires_grad_(True)
    torch.nn.functional.cross_entropy(sg @ Eg.T, ans_glob).backward()
    want_dE = Eg.grad.clone()
    _index_add_scatter(ids_glob.reshape(-1), dz_glob, want_dE)
    assert (buf[1:] - want_dE[lo:hi]).abs().max() < 1e-12
    assert buf[0].abs().max() == 0                               # the landing row of key 0 is never written
    assert (ds - sg.grad[mine]).abs().max() < 1e-12
    assert (sm - small_glob.sum(0)).abs().max() < 1e-12
    item = torch.nn.Parameter(torch.zeros(N, d))
    ex.item_param = item
    ex.publish_shard_grad(buf[1:], lo, hi)
    assert item._bsa_shard.lo == lo and item._bsa_shard.hi == hi and item._bsa_shard.grad.shape == (hi - lo, d)


def _fetch_worker(rank, world):
    """Lazy replica refresh: every rank's copy of the table is fresh only in its own shard; fetch_rows must bring exactly
    the rows this rank's batch reads (and row 0) from their owners, leaving them equal to the true table."""
    N, d, Bl, L = 48, 4, 3, 7
    g = torch.Generator().manual_seed(21)
    truth = torch.randn(N, d, generator=g)
    ids = torch.randint(1, N, (world, Bl, L), generator=g)
    ids[:, :, :2] = 0
    ex = GradExchange(vocab_parallel=True)
    lo, hi = ex.shard(N)
    table = torch.full((N, d), float("nan"))
    table[lo:hi] = truth[lo:hi]                                   # stale everywhere else
    st = ex.prepare_routed(ids[rank], N)
    ex.fetch_rows(st, table, None, None)
    need = torch.unique(ids[rank])
    assert torch.equal(table[need], truth[need])                  # includes row 0 (padding positions read it)
    untouched = torch.ones(N, dtype=torch.bool)
    untouched[need] = False
    untouched[lo:hi] = False
    assert torch.isnan(table[untouched]).all()                    # nothing else was written
    ex.item_param = torch.nn.Parameter(table.clone())
    ex.stale = True
    ex.sync_table()
    assert torch.equal(ex.item_param.data, truth) and not ex.stale


@pytest.mark.parametrize("world", [2, 3])
def test_vocab_parallel_row_fetch_and_table_sync(world):
    _spawn(_fetch_worker, world)


@pytest.mark.parametrize("world", [2, 3])
def test_vocab_parallel_exchange_equals_global_cross_entropy(world):
    _spawn(_vocab_parallel_worker, world)


def test_vocab_parallel_needs_divisible_table():
    _spawn(_vv_worker, 2)


def _vv_worker(rank, world):
    with pytest.raises(RuntimeError, match="divisible"):
        GradExchange(vocab_parallel=True).shard(7)


# ------------------------------------------------------------------------------------------- evaluation
def _np_score_topk(s, shard, indptr, indices, k, offset):
    """numpy stand-in for bsa_score_topk: global ids, score desc then id asc, history items score 0.0,
    fewer than k candidates -> (-inf, -1) padding."""
    sc = (s.double() @ shard.double().T).numpy()
    B, rows = sc.shape
    val = np.full((B, k), -np.inf, dtype=np.float32)
    idx = np.full((B, k), -1, dtype=np.int64)
    for b in range(B):
        row = sc[b].copy()
        if indptr is not None:
            h = indices[indptr[b]:indptr[b + 1]].numpy()
            h = h[(h >= offset) & (h < offset + rows)] - offset
            row[h] = 0.0
        order = np.lexsort((np.arange(rows), -row))[:k]
        val[b, :len(order)] = row[order]
        idx[b, :len(order)] = order + offset
    return torch.from_numpy(val), torch.from_numpy(idx)


def _np_merge(val, idx):
    G, B, k = val.shape
    v = val.permute(1, 0, 2).reshape(B, G * k).numpy()
    i = idx.permute(1, 0, 2).reshape(B, G * k).numpy()
    out_v = np.empty((B, k), np.float32)
    out_i = np.empty((B, k), np.int64)
    for b in range(B):
        key_i = np.where(i[b] < 0, np.iinfo(np.int64).max, i[b])
        order = np.lexsort((key_i, -v[b]))[:k]
        out_v[b], out_i[b] = v[b][order], i[b][order]
    return torch.from_numpy(out_v), torch.from_numpy(out_i)


def _eval_worker(rank, world, N, negative):
    g = torch.Generator().manual_seed(5)
    d, Bl, k = 8, 3, 20
    E = torch.randn(N, d, generator=g)
    if negative:
        E = -E.abs()                       # every unseen score negative: history zeros must surface (trainers.py:134)
    s_all = torch.randn(world * Bl, d, generator=g).abs()
    hist = [sorted(set(torch.randint(1, N, (6,), generator=g).tolist())) for _ in range(world * Bl)]
    indptr = torch.tensor(np.cumsum([0] + [len(h) for h in hist]), dtype=torch.int64)
    indices = torch.tensor([x for h in hist for x in h], dtype=torch.int64)
    lo, hi = shard_range(N, world, rank)
    top = sharded_full_sort_topk(s_all[rank * Bl:(rank + 1) * Bl], E[lo:hi], lo, indptr, indices, k,
                                 _np_score_topk, _np_merge)
    # unsharded reference: the oracle's eval branch on the full table
    scores = (s_all.double() @ E.double().T).numpy()
    masked = oracle.masked_scores(scores, indptr.numpy(), indices.numpy(), np.arange(world * Bl))
    ref = oracle.full_sort_topk(masked, k)[rank * Bl:(rank + 1) * Bl]
    mine = top.numpy()
    assert mine.shape == (Bl, k)
    for b in range(Bl):
        gb = rank * Bl + b
        # same score multiset position by position (ties between equal scores may order differently)
        assert np.allclose(masked[gb][mine[b]], masked[gb][ref[b]], rtol=0, atol=0), (rank, b)
        if not negative:
            assert list(mine[b]) == list(ref[b])


@pytest.mark.parametrize("N,negative", [(101, False), (64, False), (101, True)])
def test_item_sharded_eval_equals_unsharded(N, negative):
    _spawn(_eval_worker, 2, N, negative)

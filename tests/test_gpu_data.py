"""GPU-resident batch construction and device-side metrics through the C ABI,
bit-exact against the oracle restatement and the fixtures made from the reference's own RecDataset."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden
from oracle import dataset_oracle as do

pytestmark = pytest.mark.gpu


def _golden():
    z = np.load(os.path.join(GOLDEN, "dataset_small.npz"))
    d = {k: z[k] for k in z.files}
    ptr = d["seq_indptr"]
    return d, [d["seq_items"][ptr[u]:ptr[u + 1]].tolist() for u in range(len(ptr) - 1)]


@pytest.mark.parametrize("data_type", ["train", "valid", "test"])
def test_build_batch_matches_reference_fixture(data_type, cuda_device):
    from bsarec_b200 import SequenceStore
    d, user_seq = _golden()
    store = SequenceStore(user_seq, int(d["L"]), cuda_device)
    n = store.num_samples(data_type)
    assert n == len(d[f"{data_type}_answers"])
    g = torch.Generator().manual_seed(3)
    for idx in (torch.arange(n), torch.randperm(n, generator=g), torch.zeros(0, dtype=torch.int64)):
        users, ids, answers = store.build(data_type, idx.to(cuda_device))
        assert np.array_equal(users.cpu().numpy(), d[f"{data_type}_user_ids"][idx.numpy()])
        assert np.array_equal(ids.cpu().numpy(), d[f"{data_type}_input_ids"][idx.numpy()])
        assert np.array_equal(answers.cpu().numpy(), d[f"{data_type}_answers"][idx.numpy()])


def test_history_csr_matches_reference_rating_matrices(cuda_device):
    from bsarec_b200 import SequenceStore
    d, user_seq = _golden()
    store = SequenceStore(user_seq, int(d["L"]), cuda_device)
    for dt in ("valid", "test"):
        indptr, indices = store.history_csr(dt)
        assert np.array_equal(indptr.cpu().numpy(), d[f"{dt}_hist_indptr"])
        assert np.array_equal(indices.cpu().numpy(), d[f"{dt}_hist_indices"])


def test_build_batch_large_ragged_vs_oracle(cuda_device):
    """C3-like shape (L=200) with sequences shorter and much longer than L, 3 000 users; all three splits."""
    from bsarec_b200 import SequenceStore, DeviceBatchLoader
    rng = np.random.RandomState(9)
    L = 200
    lens = np.concatenate([[3, 4, L + 1, L + 2, L + 3, 5 * L], rng.randint(3, 3 * L, size=3000)])
    user_seq = [rng.randint(1, 1_000_000, size=int(n)).tolist() for n in lens]
    store = SequenceStore(user_seq, L, cuda_device)
    for dt in ("train", "valid", "test"):
        n = store.num_samples(dt)
        idx = torch.from_numpy(rng.randint(0, n, size=4096)).to(cuda_device)
        users, ids, answers = store.build(dt, idx)
        ou, oi, oa = do.build_batch(user_seq, L, dt, idx.cpu().tolist())
        assert np.array_equal(users.cpu().numpy(), ou) and np.array_equal(ids.cpu().numpy(), oi)
        assert np.array_equal(answers.cpu().numpy(), oa)
    # the loader covers every train sample exactly once per epoch, across 2 ranks, with a short last batch
    seen = []
    for r in range(2):
        ld = DeviceBatchLoader(store, "train", 1000, seed=1, rank=r, world_size=2, equal_shards=False)
        nb = 0
        for users, ids, answers, neg, same in ld:
            assert ids.shape[1] == L and neg.shape == answers.shape and same.shape == (answers.numel(), 0)
            seen.append(torch.stack([users, answers, (ids != 0).sum(1)], 1))
            nb += 1
        assert nb == len(ld)
    seen = torch.cat(seen).cpu().numpy()
    assert len(seen) == store.num_train_samples
    samples, users = do.expand_samples(user_seq, L, "train")
    ref = np.array([[u, s[-1], len(s) - 1] for s, u in zip(samples, users)], dtype=np.int64)
    assert np.array_equal(seen[np.lexsort(seen.T[::-1])], ref[np.lexsort(ref.T[::-1])])
    # data-parallel default (equal_shards): every rank yields the same number of batches of the same sizes — what
    # GradExchange's per-batch collectives need — and together they cover every sample, at most world-1 of them twice
    for world, bs in ((2, 1000), (3, 777), (8, 512)):
        shapes, rows = [], []
        for r in range(world):
            ld = DeviceBatchLoader(store, "train", bs, seed=5, rank=r, world_size=world)
            sizes = []
            for users, ids, answers, neg, same in ld:
                sizes.append(answers.numel())
                rows.append(torch.stack([users, answers, (ids != 0).sum(1)], 1))
            assert len(sizes) == len(ld)
            shapes.append(sizes)
        assert all(sh == shapes[0] for sh in shapes), shapes
        got = torch.cat(rows).cpu().numpy()
        assert 0 <= len(got) - len(ref) < world
        uniq = np.unique(got, axis=0)
        assert np.array_equal(uniq, np.unique(ref, axis=0))


def test_hit_rank_histogram_and_metrics_on_beauty(cuda_device):
    """Device histogram == oracle histogram on the reference's Beauty pred_list; metrics identical to
    BSARec_Beauty_best.log:258; accumulation over ragged batches; -1 (unused) slots never hit."""
    from bsarec_b200 import HitRankAccumulator
    _, data = load_golden("beauty_test")
    pred = torch.from_numpy(data["pred_list"].astype(np.int64)).to(cuda_device)
    ans = torch.from_numpy(data["answers"].astype(np.int64)).to(cuda_device)
    acc = HitRankAccumulator(20, cuda_device)
    for s in range(0, pred.shape[0], 1000):
        acc.update(pred[s:s + 1000], ans[s:s + 1000])
    ref = do.hit_rank_hist(data["answers"].tolist(), data["pred_list"].tolist(), 20)
    assert np.array_equal(acc.hist.cpu().numpy(), ref)
    recall, ndcg = acc.metrics([5, 10, 20])
    got = [recall[0], ndcg[0], recall[1], ndcg[1], recall[2], ndcg[2]]
    assert got == pytest.approx(list(data["metrics"]), abs=1e-12)
    assert ["%.4f" % v for v in got] == ["0.0736", "0.0523", "0.1008", "0.0611", "0.1373", "0.0703"]
    acc2 = HitRankAccumulator(20, cuda_device)
    acc2.update(torch.full((7, 20), -1, dtype=torch.int64, device=cuda_device), ans[:7])
    assert acc2.hist.cpu().tolist() == [0] * 20 + [7]


def test_trainer_eval_branch_on_device_loader(cuda_device):
    """bsarec_b200.Trainer's eval branch fed by DeviceBatchLoader + the device history CSR returns the reference's
    (scores, str) pair; checked against the oracle's eval step + metric loops on the same corpus."""
    import argparse
    import oracle
    from scipy.sparse import csr_matrix
    from helpers import make_args, oracle_cfg, perturbed_params
    from bsarec_b200 import BSARecModel, Trainer, SequenceStore, DeviceBatchLoader
    d, user_seq = _golden()
    L, N = int(d["L"]), int(d["item_size"])
    cfgkw = dict(item_size=N + 340, hidden_size=64, max_seq_length=L, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)
    cfg = oracle_cfg(cfgkw)
    P = perturbed_params(cfg, 3)
    args = make_args(cfgkw, batch_size=20)
    model = BSARecModel(args)
    model.load_state_dict({k: v.clone() for k, v in P.items()})
    store = SequenceStore(user_seq, L, cuda_device)
    indptr, indices = (d["test_hist_indptr"], d["test_hist_indices"])
    args.test_rating_matrix = csr_matrix((np.ones(len(indices)), indices, indptr), shape=(len(user_seq), N + 340))
    args.valid_rating_matrix = args.test_rating_matrix
    args.keep_pred_list = True
    tr = Trainer(model, None, None, DeviceBatchLoader(store, "test", 20), args, None)
    scores, text = tr.test(0)
    ids = torch.from_numpy(d["test_input_ids"])
    rows = np.arange(len(user_seq))
    with torch.no_grad():
        last = oracle.encode(P, ids, cfg, train=False)[-1][:, -1, :]
        masked = oracle.masked_scores(oracle.predict_full(last, P["item_embeddings.weight"]).numpy(), indptr, indices, rows)
    pred = oracle.full_sort_topk(masked, 20)
    assert np.array_equal(tr.answer_list, d["test_answers"])
    for u in rows:                      # identical lists except where the oracle's own scores tie (history zeros)
        if not np.array_equal(tr.pred_list[u], pred[u]):
            assert np.allclose(masked[u, tr.pred_list[u]], masked[u, pred[u]], rtol=2e-6, atol=1e-7), u
    answers = d["test_answers"].tolist()
    ref = []
    for k in (5, 10, 20):
        ref += [oracle.recall_at_k(answers, tr.pred_list, k), oracle.ndcg_k(answers, tr.pred_list, k)]
    assert scores == pytest.approx(ref, abs=1e-12)
    assert "HR@5" in text and "NDCG@20" in text


@pytest.mark.parametrize("fused_adam", [False, True])
def test_train_loop_on_device_batches_tracks_oracle_adam(fused_adam, cuda_device):
    """The reference's train branch (trainers.py:94-116) end to end: DeviceBatchLoader batches -> calculate_loss ->
    zero_grad -> backward -> Adam.step for 6 steps (dropout 0, sequential sampler), against the oracle's CPU loop on
    the batches the reference's own RecDataset yields.  Checks every step's loss and the final parameters."""
    import oracle
    from helpers import make_args, oracle_cfg, perturbed_params
    from bsarec_b200 import BSARecModel, Trainer, FusedAdam, SequenceStore, DeviceBatchLoader
    d, user_seq = _golden()
    L, N = int(d["L"]), int(d["item_size"])
    cfgkw = dict(item_size=N, hidden_size=64, max_seq_length=L, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)
    cfg = oracle_cfg(cfgkw)
    P = perturbed_params(cfg, 4)
    args = make_args(cfgkw, batch_size=64)
    model = BSARecModel(args)
    model.load_state_dict({k: v.clone() for k, v in P.items()})
    store = SequenceStore(user_seq, L, cuda_device)
    loader = DeviceBatchLoader(store, "train", 64, shuffle=False)
    tr = Trainer(model, loader, None, None, args, None)
    if fused_adam:
        tr.optim = FusedAdam(tr.model.parameters(), lr=args.lr, betas=(args.adam_beta1, args.adam_beta2), weight_decay=args.weight_decay)
    losses = []
    tr.model.train()
    for step, batch in enumerate(loader):
        if step == 6:
            break
        user_ids, input_ids, answers, neg, same = batch
        loss = tr.model.calculate_loss(input_ids, answers, neg, same, user_ids)
        tr.optim.zero_grad()
        loss.backward()
        tr.optim.step()
        losses.append(loss.item())
    ref = oracle.adam_train_step(P, cfg, lr=args.lr, betas=(args.adam_beta1, args.adam_beta2), weight_decay=args.weight_decay)
    ref_losses = []
    for step in range(6):
        sl = slice(step * 64, (step + 1) * 64)
        ref_losses.append(ref(torch.from_numpy(d["train_input_ids"][sl]), torch.from_numpy(d["train_answers"][sl])))
    assert losses == pytest.approx(ref_losses, rel=2e-5)
    got = {k: v.detach().cpu() for k, v in tr.model.state_dict().items()}
    for k, v in ref.P.items():
        if k.endswith("key.bias"):      # analytically zero gradient: Adam turns rounding noise into +-lr steps on both sides
            continue
        # Adam's first steps move every weight by ~lr regardless of gradient size: compare against the total update size
        assert (got[k] - v.detach()).abs().max().item() < 0.05 * 6 * args.lr, k

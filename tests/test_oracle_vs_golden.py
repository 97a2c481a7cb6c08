"""Pins the CPU oracle to the unmodified reference (vectors made by tests/golden/make_golden.py)."""
import os
import sys

import numpy as np
import pytest
import torch

import oracle
from oracle import OracleConfig
from conftest import load_golden, golden_params, golden_grads, rel_err

SMALL = ["small_a", "small_b", "small_c"]


def cfg_of(cfgkw):
    return OracleConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **cfgkw)


@pytest.mark.parametrize("name", SMALL)
@pytest.mark.parametrize("use_fft", [True, False])
def test_forward_layers_match_reference(name, use_fft):
    cfgkw, data = load_golden(name)
    cfg = cfg_of(cfgkw)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"])
    outs = oracle.encode(P, ids, cfg, train=False, use_fft=use_fft)
    assert len(outs) == cfg.num_hidden_layers + 1
    for i, o in enumerate(outs):
        assert rel_err(o, data[f"layer_out.{i}"]) < (2e-6 if use_fft else 1e-5), (name, i)


@pytest.mark.parametrize("name", SMALL)
def test_loss_and_all_grads_match_reference(name):
    cfgkw, data = load_golden(name)
    cfg = cfg_of(cfgkw)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"])
    answers = torch.from_numpy(data["answers"])
    loss, grads = oracle.loss_and_grads(P, ids, answers, cfg, train=True)
    assert abs(float(loss) - float(data["loss"])) < 1e-6 * abs(float(data["loss"]))
    ref = golden_grads(data)
    assert set(ref) == set(grads) == set(oracle.param_names(cfg.num_hidden_layers))
    for k in ref:
        if k.endswith("key.bias"):          # analytically zero: absolute bound
            assert grads[k].abs().max() < 1e-7 and ref[k].abs().max() < 1e-7
        else:
            assert rel_err(grads[k], ref[k]) < 2e-5, k


@pytest.mark.parametrize("name", SMALL)
def test_fp64_oracle_explicit_filter_agrees(name):
    """The L x L real-filter form in fp64 reproduces the fp32 reference to fp32 rounding."""
    cfgkw, data = load_golden(name)
    cfg = cfg_of(cfgkw)
    P = golden_params(data, torch.float64)
    loss, grads = oracle.loss_and_grads(P, torch.from_numpy(data["ids"]), torch.from_numpy(data["answers"]),
                                        cfg, train=True, use_fft=False)
    assert abs(float(loss) - float(data["loss"])) < 2e-6 * abs(float(data["loss"]))
    ref = golden_grads(data)
    for k in ref:
        if not k.endswith("key.bias"):
            # the fp32 reference's own rounding reaches ~1e-4 on cancellation-heavy grads (key.weight)
            assert rel_err(grads[k], ref[k]) < 2e-4, k


@pytest.mark.parametrize("name", SMALL)
def test_eval_branch_matches_reference(name):
    cfgkw, data = load_golden(name)
    cfg = cfg_of(cfgkw)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"])
    B = ids.shape[0]
    with torch.no_grad():
        s = oracle.encode(P, ids, cfg)[-1][:, -1, :]
        scores = oracle.predict_full(s, P["item_embeddings.weight"]).numpy()
    assert rel_err(scores, data["scores"]) < 2e-6
    pred = oracle.eval_step(P, ids, cfg, data["hist_indptr"], data["hist_indices"], np.arange(B), int(data["k_eval"]))
    masked = oracle.masked_scores(data["scores"], data["hist_indptr"], data["hist_indices"], np.arange(B))
    for b in range(B):
        # identical ordered lists unless two of the reference's own scores tie
        if not np.array_equal(pred[b], data["pred_list"][b]):
            assert np.allclose(np.sort(masked[b, pred[b]]), np.sort(masked[b, data["pred_list"][b]]), rtol=1e-6)


@pytest.mark.parametrize("L,c", [(8, 8), (8, 3), (7, 9), (50, 5), (200, 5), (25, 9), (64, 64), (1024, 32), (6, 1)])
def test_filter_matrix_equals_fft_operator(L, c):
    kept = c // 2 + 1
    x = torch.randn(3, L, 5, dtype=torch.float64, generator=torch.Generator().manual_seed(L * 100 + c))
    spec = torch.fft.rfft(x, dim=1, norm="ortho")
    spec[:, kept:, :] = 0
    low = torch.fft.irfft(spec, n=L, dim=1, norm="ortho")
    Fm = torch.from_numpy(oracle.filter_matrix(L, kept))
    assert torch.allclose(torch.einsum("ts,bsd->btd", Fm, x), low, atol=1e-12)
    C = oracle.filter_basis(L, kept)
    assert np.allclose(C.T @ C, np.eye(C.shape[1]), atol=1e-12)       # orthonormal
    assert np.allclose(Fm.numpy(), Fm.numpy().T)                       # symmetric -> self-adjoint backward


def test_metrics_match_reference_definitions():
    ans = [3, 9, 4]
    pred = np.array([[3, 1, 2, 5, 6], [1, 2, 3, 4, 5], [7, 4, 1, 2, 3]])
    assert oracle.recall_at_k(ans, pred, 5) == pytest.approx(2 / 3)
    assert oracle.ndcg_k(ans, pred, 5) == pytest.approx((1.0 + 1.0 / np.log2(3)) / 3)


def test_beauty_known_answer():
    """Shipped checkpoint + Beauty test split -> BSARec_Beauty_best.log:258, via the oracle."""
    _, data = load_golden("beauty_test")
    assert "'HR@5': '0.0736'" in str(data["info"]) and "'NDCG@20': '0.0703'" in str(data["info"])
    cfg = OracleConfig(item_size=int(data["item_size"]), hidden_size=64, max_seq_length=50, num_attention_heads=1,
                       num_hidden_layers=2, c=5, alpha=0.7)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"].astype(np.int64))
    U = 2048                                  # a bounded slice keeps the CPU suite fast; the GPU suite does all 22 363
    pred = oracle.eval_step(P, ids[:U], cfg, data["hist_indptr"].astype(np.int64),
                            data["hist_indices"].astype(np.int64), np.arange(U))
    ref = data["pred_list"][:U].astype(np.int64)
    same = (pred == ref).all(axis=1)
    assert same.mean() > 0.999
    for u in np.nonzero(~same)[0]:
        assert set(pred[u]) == set(ref[u]) or data["gap_rel"][u] < 1e-5
    m = oracle.full_sort_metrics(data["answers"][:U].tolist(), pred)
    mref = oracle.full_sort_metrics(data["answers"][:U].tolist(), ref)
    assert m == pytest.approx(mref, abs=1e-12)
    full = oracle.full_sort_metrics(data["answers"].tolist(), data["pred_list"].astype(np.int64))
    assert ["%.4f" % v for v in full] == ["0.0736", "0.0523", "0.1008", "0.0611", "0.1373", "0.0703"]


KNOWN_ANSWERS = {
    # fixture: (params fixture, c, alpha, line of the shipped log the reference run reproduces)
    "lastfm_test": ("lastfm_test", 3, 0.9, "BSARec_LastFM_best.log:237",
                    ["0.0523", "0.0344", "0.0807", "0.0435", "0.1174", "0.0526"]),
    "lastfm_valid": ("lastfm_test", 3, 0.9, "BSARec_LastFM_best.log:203 (best epoch 47's validation line)",
                     ["0.0706", "0.0479", "0.1037", "0.0582", "0.1615", "0.0727"]),
    # Beauty validation: the log's line :224 was written by a GPU run (0.1633 / 0.0852 at @20); the reference on this
    # container's CPU gives 0.1634 / 0.0853 — the one-unit cross-device difference of SURVEY.md App. B #12.  The
    # fixture holds what the unmodified reference computes HERE; index sets are what is compared.
    "beauty_valid": ("beauty_test", 5, 0.7, "BSARec_Beauty_best.log:224 (+1 ulp of the 4th decimal at @20 on CPU)",
                     ["0.0891", "0.0641", "0.1238", "0.0753", "0.1634", "0.0853"]),
}


@pytest.mark.parametrize("name", sorted(KNOWN_ANSWERS))
def test_second_checkpoint_and_validation_split_known_answers(name):
    """LastFM checkpoint (alpha 0.9, c 3 -> 2 kept bins: a different filter rank and blend) on its test and validation
    splits, and the Beauty validation split (history = all but the last two items, dataset.py:132), via the oracle."""
    pname, c, alpha, _, strings = KNOWN_ANSWERS[name]
    _, data = load_golden(name)
    _, pdata = load_golden(pname)
    cfg = OracleConfig(item_size=int(data["item_size"]), hidden_size=64, max_seq_length=50, num_attention_heads=1,
                       num_hidden_layers=2, c=c, alpha=alpha)
    P = golden_params(pdata)
    ids = torch.from_numpy(data["ids"].astype(np.int64))
    U = min(1090, ids.shape[0])
    pred = oracle.eval_step(P, ids[:U], cfg, data["hist_indptr"].astype(np.int64),
                            data["hist_indices"].astype(np.int64), np.arange(U))
    ref = data["pred_list"][:U].astype(np.int64)
    same = (pred == ref).all(axis=1)
    assert same.mean() > 0.995
    for u in np.nonzero(~same)[0]:
        assert set(pred[u]) == set(ref[u]) or data["gap_rel"][u] < 1e-5
    full = oracle.full_sort_metrics(data["answers"].tolist(), data["pred_list"].astype(np.int64))
    assert full == pytest.approx(list(data["metrics"]), abs=1e-12)
    assert ["%.4f" % v for v in full] == strings


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/model"), reason="live reference only in the build container")
def test_oracle_vs_live_reference_random_config():
    """Import the reference as it lies and compare on a fresh random config (not a stored vector)."""
    sys.path.insert(0, "/root/reference/src")
    import argparse
    from model import MODEL_DICT
    args = argparse.Namespace(item_size=150, hidden_size=32, max_seq_length=17, batch_size=4, hidden_dropout_prob=0.0,
                              attention_probs_dropout_prob=0.0, num_attention_heads=2, num_hidden_layers=1,
                              hidden_act="gelu", c=4, alpha=0.55, initializer_range=0.02)
    torch.manual_seed(5)
    m = MODEL_DICT["bsarec"](args=args)
    ids = torch.randint(0, 150, (4, 17))
    ids[0, :9] = 0
    ans = torch.randint(1, 150, (4,))
    loss = m.calculate_loss(ids, ans, None, None, None)
    loss.backward()
    cfg = OracleConfig(item_size=150, hidden_size=32, max_seq_length=17, num_attention_heads=2, num_hidden_layers=1,
                       c=4, alpha=0.55, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    P = {k: v.detach() for k, v in m.state_dict().items()}
    assert list(P) == oracle.param_names(1)
    l2, grads = oracle.loss_and_grads(P, ids, ans, cfg)
    assert abs(float(l2) - float(loss)) < 1e-6
    for n, p in m.named_parameters():
        if not n.endswith("key.bias"):
            assert rel_err(grads[n], p.grad) < 2e-5, n


def test_dropout_rng_statistics_and_determinism():
    keep = oracle.dropout_keep_mask(1234, 3, (64, 50, 32), 0.5)
    assert keep.dtype == bool and abs(keep.mean() - 0.5) < 5e-3
    assert np.array_equal(keep, oracle.dropout_keep_mask(1234, 3, (64, 50, 32), 0.5))
    assert (keep != oracle.dropout_keep_mask(1234, 4, (64, 50, 32), 0.5)).mean() > 0.4
    assert (keep != oracle.dropout_keep_mask(1235, 3, (64, 50, 32), 0.5)).mean() > 0.4
    k2 = oracle.dropout_keep_mask(9, 0, (100000,), 0.25)
    assert abs(k2.mean() - 0.75) < 5e-3
    # lanes of one 64-bit word are independent enough: no correlation between neighbours
    a = keep.reshape(-1)
    assert abs(np.corrcoef(a[:-1], a[1:])[0, 1]) < 0.01


# ------------------------------------------------------------------ batch construction / metrics
def _dataset_golden():
    from conftest import GOLDEN
    z = np.load(os.path.join(GOLDEN, "dataset_small.npz"))
    d = {k: z[k] for k in z.files}
    ptr = d["seq_indptr"]
    user_seq = [d["seq_items"][ptr[u]:ptr[u + 1]].tolist() for u in range(len(ptr) - 1)]
    return d, user_seq


@pytest.mark.parametrize("data_type", ["train", "valid", "test"])
def test_dataset_oracle_matches_reference_recdataset(data_type):
    """oracle.dataset_oracle vs the unmodified RecDataset's samples (fixture made by make_golden.dataset_case)."""
    from oracle import dataset_oracle as do
    d, user_seq = _dataset_golden()
    n = len(d[f"{data_type}_answers"])
    users, ids, answers = do.build_batch(user_seq, int(d["L"]), data_type, range(n))
    assert np.array_equal(users, d[f"{data_type}_user_ids"])
    assert np.array_equal(ids, d[f"{data_type}_input_ids"])
    assert np.array_equal(answers, d[f"{data_type}_answers"])
    samples, _ = do.expand_samples(user_seq, int(d["L"]), data_type)
    assert len(samples) == n


@pytest.mark.parametrize("data_type", ["valid", "test"])
def test_history_oracle_matches_reference_rating_matrix(data_type):
    from oracle import dataset_oracle as do
    d, user_seq = _dataset_golden()
    indptr, indices = do.history_csr(user_seq, data_type)
    assert np.array_equal(indptr, d[f"{data_type}_hist_indptr"])
    assert np.array_equal(indices, d[f"{data_type}_hist_indices"])


def test_hit_rank_hist_reproduces_reference_metrics_on_beauty():
    """Histogram form of HR/NDCG == the reference's per-user loops on the Beauty test split's pred_list."""
    from oracle import dataset_oracle as do
    from bsarec_b200.data import metrics_from_hist
    _, data = load_golden("beauty_test")
    hist = do.hit_rank_hist(data["answers"].tolist(), data["pred_list"].tolist(), 20)
    assert int(hist.sum()) == len(data["answers"])
    recall, ndcg = metrics_from_hist(hist, [5, 10, 20])
    got = [recall[0], ndcg[0], recall[1], ndcg[1], recall[2], ndcg[2]]
    assert got == pytest.approx(list(data["metrics"]), abs=1e-12)
    assert ["%.4f" % v for v in got] == ["0.0736", "0.0523", "0.1008", "0.0611", "0.1373", "0.0703"]


def test_sasrec_oracle_vs_reference_golden():
    """Sibling model: the oracle's SASRec restatement against the reference's SASRecModel —
    every layer output, the pairwise BCE loss (one user's positive is padding) and all gradients."""
    cfgkw, data = load_golden("sasrec_small")
    cfg = OracleConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **cfgkw)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"])
    pos, neg = torch.from_numpy(data["answers"]), torch.from_numpy(data["neg"])
    outs = oracle.encode_sasrec(P, ids, cfg)
    for i, o in enumerate(outs):
        assert rel_err(o.detach(), data[f"layer_out.{i}"]) < 1e-5, i
    loss, grads = oracle.sasrec_loss_and_grads(P, ids, pos, neg, cfg)
    assert abs(float(loss) - float(data["loss"])) < 1e-6 * abs(float(data["loss"]))
    ref = golden_grads(data)
    for k, g in grads.items():
        if k.endswith("key.bias"):
            assert g.abs().max() < 1e-6
            continue
        assert rel_err(g, ref[k]) < 1e-4, k


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/model"), reason="live reference only in the build container")
def test_sasrec_container_matches_live_reference_class():
    """Same state_dict keys / shapes and bit-identical seeded init as the reference's SASRecModel (host side only)."""
    sys.path.insert(0, "/root/reference/src")
    try:
        from model import MODEL_DICT as REF
        from helpers import make_args
        import bsarec_b200
        cfgkw = dict(item_size=123, hidden_size=32, max_seq_length=20, num_attention_heads=2, num_hidden_layers=3)
        torch.manual_seed(5)
        ref = REF["sasrec"](args=make_args(cfgkw, 0.5, 0.5))
        torch.manual_seed(5)
        ours = bsarec_b200.MODEL_DICT["sasrec"](make_args(cfgkw, 0.5, 0.5))
        a, b = ref.state_dict(), ours.state_dict()
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert a[k].shape == b[k].shape and torch.equal(a[k], b[k]), k
        assert [n for n, _ in ref.named_parameters()] == [n for n, _ in ours.named_parameters()]
    finally:
        sys.path.remove("/root/reference/src")
        for m in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
            sys.modules.pop(m, None)


def test_duorec_oracle_vs_reference_golden():
    """Sibling model: the oracle's DuoRec restatement (CE + lmd_sem * InfoNCE, ssl='us_x') against the reference's
    DuoRecModel — loss and all gradients."""
    cfgkw, data = load_golden("duorec_small")
    cfg = OracleConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **cfgkw)
    P = golden_params(data)
    ids, ans, sem = (torch.from_numpy(data[k]) for k in ("ids", "answers", "sem"))
    loss, grads = oracle.duorec_loss_and_grads(P, ids, ans, sem, cfg, float(data["tau"]), float(data["lmd_sem"]))
    assert abs(float(loss) - float(data["loss"])) < 1e-6 * abs(float(data["loss"]))
    ref = golden_grads(data)
    for k, g in grads.items():
        if k.endswith("key.bias"):
            assert g.abs().max() < 1e-6
            continue
        assert rel_err(g, ref[k]) < 1e-4, k


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/model"), reason="live reference only in the build container")
def test_duorec_container_matches_live_reference_class():
    """Same state_dict keys and bit-identical seeded init as the reference's DuoRecModel (which builds its embedding
    tables twice before init_weights: the RNG stream must be consumed the same way)."""
    sys.path.insert(0, "/root/reference/src")
    try:
        from model import MODEL_DICT as REF
        from helpers import make_args
        import bsarec_b200
        cfgkw = dict(item_size=123, hidden_size=32, max_seq_length=20, num_attention_heads=2, num_hidden_layers=2)
        extra = dict(tau=1.0, lmd=0.1, lmd_sem=0.1, ssl="us_x", sim="dot")
        a1, a2 = make_args(cfgkw, 0.5, 0.5), make_args(cfgkw, 0.5, 0.5)
        for a in (a1, a2):
            for k, v in extra.items():
                setattr(a, k, v)
        torch.manual_seed(5)
        ref = REF["duorec"](args=a1)
        torch.manual_seed(5)
        ours = bsarec_b200.MODEL_DICT["duorec"](a2)
        a, b = ref.state_dict(), ours.state_dict()
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert torch.equal(a[k], b[k]), k
    finally:
        sys.path.remove("/root/reference/src")
        for m in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
            sys.modules.pop(m, None)


@pytest.mark.parametrize("name", ["fmlprec_small", "fmlprec_odd"])
def test_fmlprec_oracle_vs_reference_golden(name):
    """Sibling model: the oracle's FMLPRec restatement against the reference's FMLPRecModel (even and odd L) — layer outputs,
    loss, all gradients; and the identity the kernels rest on: the learned filter == a circular convolution with taps
    irfft(W)."""
    cfgkw, data = load_golden(name)
    cfg = OracleConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_attention_heads=1, **cfgkw)
    P = golden_params(data)
    ids = torch.from_numpy(data["ids"])
    pos, neg = torch.from_numpy(data["answers"]), torch.from_numpy(data["neg"])
    outs = oracle.encode_fmlprec(P, ids, cfg)
    for i, o in enumerate(outs):
        assert rel_err(o.detach(), data[f"layer_out.{i}"]) < 1e-5, i
    loss, grads = oracle.fmlprec_loss_and_grads(P, ids, pos, neg, cfg)
    assert abs(float(loss) - float(data["loss"])) < 1e-6 * abs(float(data["loss"]))
    ref = golden_grads(data)
    for k, g in grads.items():
        assert rel_err(g, ref[k]) < 1e-4, k
    L = ids.shape[1]
    w = torch.view_as_complex(P["item_encoder.blocks.0.layer.complex_weight"].double().contiguous())[0]      # [F, d]
    taps = torch.fft.irfft(w, n=L, dim=0)                                                                       # [L, d]
    x = torch.randn(3, L, w.shape[1], dtype=torch.float64, generator=torch.Generator().manual_seed(1))
    want = torch.fft.irfft(torch.fft.rfft(x, dim=1, norm="ortho") * w, n=L, dim=1, norm="ortho")
    idx = (torch.arange(L)[:, None] - torch.arange(L)[None, :]) % L                                             # (t - s) mod L
    got = torch.einsum("tsc,bsc->btc", taps[idx], x)
    assert (got - want).abs().max() < 1e-12


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/model"), reason="live reference only in the build container")
def test_fmlprec_container_matches_live_reference_class():
    sys.path.insert(0, "/root/reference/src")
    try:
        from model import MODEL_DICT as REF
        from helpers import make_args
        import bsarec_b200
        cfgkw = dict(item_size=123, hidden_size=32, max_seq_length=21, num_hidden_layers=3, num_attention_heads=1)
        torch.manual_seed(5)
        ref = REF["fmlprec"](args=make_args(cfgkw, 0.5, 0.5))
        torch.manual_seed(5)
        ours = bsarec_b200.MODEL_DICT["fmlprec"](make_args(cfgkw, 0.5, 0.5))
        a, b = ref.state_dict(), ours.state_dict()
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert a[k].shape == b[k].shape and torch.equal(a[k], b[k]), k
    finally:
        sys.path.remove("/root/reference/src")
        for m in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
            sys.modules.pop(m, None)

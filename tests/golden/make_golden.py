"""Generate the golden fixtures in this directory from the UNMODIFIED reference.

Run in the build container only (needs /root/reference; it does not exist on the GPU box):

    python tests/golden/make_golden.py

Imports ``/root/reference/src/{model,trainers,dataset,metrics}`` as they lie, runs the
reference's own ``BSARecModel`` / ``Trainer.iteration`` on seeded inputs and stores inputs,
parameters and outputs as ``.npz``.  Nothing is copied from the reference's sources; only
its numerical outputs (and, for the Beauty known-answer case, the shipped checkpoint's
tensors and the test split's padded id matrix) are stored.
"""
import argparse
import logging
import os
import sys

import numpy as np
import torch

REF = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

from model import MODEL_DICT  # noqa: E402
from trainers import Trainer  # noqa: E402
import dataset as ref_dataset  # noqa: E402
from scipy.sparse import csr_matrix  # noqa: E402


def make_args(**kw):
    a = argparse.Namespace(
        item_size=100, hidden_size=64, max_seq_length=50, batch_size=256,
        hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_attention_heads=2,
        num_hidden_layers=2, hidden_act="gelu", c=3, alpha=0.9, initializer_range=0.02,
        no_cuda=True, lr=1e-3, adam_beta1=0.9, adam_beta2=0.999, weight_decay=0.0,
        model_type="BSARec", log_freq=1)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def synth_ids(g, B, L, N, all_pad_rows=(1,), short_rows=(2, 3)):
    ids = torch.zeros(B, L, dtype=torch.long)
    for b in range(B):
        if b in all_pad_rows:
            continue
        n = int(torch.randint(1, L + 1, (1,), generator=g))
        if b in short_rows:
            n = 1 + (b % 2)
        if b == 0:
            n = L
        ids[b, L - n:] = torch.randint(1, N, (n,), generator=g)
    answers = torch.randint(1, N, (B,), generator=g)
    return ids, answers


def perturb(model, g):
    """Move LayerNorm (1,0) and zero biases off their init so every gradient path is exercised."""
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith("LayerNorm.weight"):
                p.add_(0.2 * torch.randn(p.shape, generator=g))
            elif name.endswith(".bias"):
                p.add_(0.05 * torch.randn(p.shape, generator=g))
            elif name.endswith("weight") and "embeddings" not in name:
                p.mul_(4.0)       # 0.02-std Linear weights make attention nearly uniform


def small_case(name, seed, B, **cfgkw):
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1)
    args = make_args(**cfgkw)
    model = MODEL_DICT["bsarec"](args=args)
    perturb(model, g)
    ids, answers = synth_ids(g, B, args.max_seq_length, args.item_size)

    model.eval()
    with torch.no_grad():
        layers = model.predict(ids, None, all_sequence_output=True)
    model.train()                      # dropout probs are 0 -> deterministic
    loss = model.calculate_loss(ids, answers, None, None, None)
    model.zero_grad()
    loss.backward()

    # eval branch of Trainer.iteration on the same inputs: history = the user's own input ids
    rows, cols = [], []
    for b in range(B):
        for it in set(ids[b].tolist()) - {0}:
            rows.append(b)
            cols.append(it)
    hist = csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(B, args.item_size))
    args.train_matrix = hist
    logger = logging.getLogger("golden")
    trainer = Trainer(model, None, None, None, args, logger)
    captured = {}
    trainer.get_full_sort_score = lambda epoch, answers_, pred: captured.update(a=answers_, p=pred) or ([0] * 6, "")
    k_eval = 20
    batch = (torch.arange(B), ids, answers, torch.zeros(B, 0, dtype=torch.long), torch.zeros(B, 0, dtype=torch.long))
    trainer.iteration(0, [batch], train=False)
    with torch.no_grad():
        scores = trainer.predict_full(model.predict(ids, None)[:, -1, :]).numpy()

    out = {
        "cfg_keys": np.array(sorted(cfgkw.keys())),
        "cfg_vals": np.array([float(cfgkw[k]) for k in sorted(cfgkw.keys())]),
        "ids": ids.numpy(), "answers": answers.numpy(),
        "loss": loss.detach().numpy(),
        "scores": scores, "hist_indptr": hist.indptr.astype(np.int64), "hist_indices": hist.indices.astype(np.int64),
        "pred_list": captured["p"].astype(np.int64), "k_eval": np.array(k_eval),
    }
    for i, t in enumerate(layers):
        out[f"layer_out.{i}"] = t.numpy()
    for n, p in model.named_parameters():
        out["param." + n] = p.detach().numpy()
        out["grad." + n] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", float(loss), "params", sum(p.numel() for p in model.parameters()))


def sasrec_case(name, seed, B, **cfgkw):
    """Sibling model: the reference's SASRecModel (model/sasrec.py) — layer outputs, BCE loss with
    seeded negative ids (one user's positive is padding: it must drop out of both means), all gradients."""
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1)
    args = make_args(**cfgkw)
    model = MODEL_DICT["sasrec"](args=args)
    perturb(model, g)
    ids, answers = synth_ids(g, B, args.max_seq_length, args.item_size)
    neg = torch.randint(1, args.item_size, (B,), generator=g)
    answers[2] = 0
    model.eval()
    with torch.no_grad():
        layers = model.predict(ids, None, all_sequence_output=True)
    model.train()                      # dropout probs are 0 -> deterministic
    loss = model.calculate_loss(ids, answers, neg, None, None)
    model.zero_grad()
    loss.backward()
    out = {
        "cfg_keys": np.array(sorted(cfgkw.keys())),
        "cfg_vals": np.array([float(cfgkw[k]) for k in sorted(cfgkw.keys())]),
        "ids": ids.numpy(), "answers": answers.numpy(), "neg": neg.numpy(), "loss": loss.detach().numpy(),
    }
    for i, t in enumerate(layers):
        out[f"layer_out.{i}"] = t.numpy()
    for n, p in model.named_parameters():
        out["param." + n] = p.detach().numpy()
        out["grad." + n] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", float(loss), "params", sum(p.numel() for p in model.parameters()))


def duorec_case(name, seed, B, **cfgkw):
    """Sibling model: the reference's DuoRecModel with its default ssl='us_x' (CE + lmd_sem * InfoNCE between a second
    view of the sequence and the same-target sequence), dropout 0 so the run is deterministic: loss and all gradients."""
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1)
    args = make_args(tau=1.0, lmd=0.1, lmd_sem=0.1, ssl="us_x", sim="dot", batch_size=B, **cfgkw)
    model = MODEL_DICT["duorec"](args=args)
    perturb(model, g)
    ids, answers = synth_ids(g, B, args.max_seq_length, args.item_size)
    sem, _ = synth_ids(g, B, args.max_seq_length, args.item_size, all_pad_rows=(), short_rows=(4,))
    model.train()
    loss = model.calculate_loss(ids, answers, None, sem, None)
    model.zero_grad()
    loss.backward()
    out = {
        "cfg_keys": np.array(sorted(cfgkw.keys())),
        "cfg_vals": np.array([float(cfgkw[k]) for k in sorted(cfgkw.keys())]),
        "ids": ids.numpy(), "answers": answers.numpy(), "sem": sem.numpy(), "loss": loss.detach().numpy(),
        "tau": np.array(1.0), "lmd_sem": np.array(0.1),
    }
    for n, p in model.named_parameters():
        out["param." + n] = p.detach().numpy()
        out["grad." + n] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", float(loss), "params", sum(p.numel() for p in model.parameters()))


def fmlprec_case(name, seed, B, **cfgkw):
    """Sibling model: the reference's FMLPRecModel — layer outputs, pairwise log-sigmoid loss, all gradients (complex filter
    weights included).  The filter weights are scaled up so the learned filter is not a near-zero perturbation."""
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed + 1)
    args = make_args(**cfgkw)
    model = MODEL_DICT["fmlprec"](args=args)
    perturb(model, g)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.endswith("complex_weight"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.5)
    ids, answers = synth_ids(g, B, args.max_seq_length, args.item_size)
    neg = torch.randint(1, args.item_size, (B,), generator=g)
    answers[2] = 0
    model.eval()
    with torch.no_grad():
        layers = model.predict(ids, None, all_sequence_output=True)
    model.train()
    loss = model.calculate_loss(ids, answers, neg, None, None)
    model.zero_grad()
    loss.backward()
    out = {
        "cfg_keys": np.array(sorted(cfgkw.keys())),
        "cfg_vals": np.array([float(cfgkw[k]) for k in sorted(cfgkw.keys())]),
        "ids": ids.numpy(), "answers": answers.numpy(), "neg": neg.numpy(), "loss": loss.detach().numpy(),
    }
    for i, t in enumerate(layers):
        out[f"layer_out.{i}"] = t.numpy()
    for n, p in model.named_parameters():
        out["param." + n] = p.detach().numpy()
        out["grad." + n] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", float(loss), "params", sum(p.numel() for p in model.parameters()))


def known_answer_case(data_name, split, out_name, store_params=True, **cfgkw):
    """Known-answer: a shipped checkpoint on a real split through the reference's own Trainer.iteration.
    Beauty/test -> src/output/BSARec_Beauty_best.log:258; LastFM/test -> BSARec_LastFM_best.log:237; the valid split
    (history = all but the last two items, dataset.py:132) reproduces the best epoch's validation line of the log."""
    data_dir = REF + "/data/"
    args = make_args(hidden_size=64, max_seq_length=50, hidden_dropout_prob=0.5, attention_probs_dropout_prob=0.5,
                     batch_size=256, data_dir=data_dir, data_name=data_name, num_workers=0, **cfgkw)
    seq_dic, max_item, num_users = ref_dataset.get_seq_dic(args)
    args.item_size = max_item + 1
    args.num_users = num_users + 1
    ds = ref_dataset.RecDataset(args, seq_dic["user_seq"], data_type=split)
    valid_matrix, test_matrix = ref_dataset.get_rating_matrix(data_name, seq_dic, max_item)
    matrix = (valid_matrix if split == "valid" else test_matrix).tocsr()
    matrix.sort_indices()
    model = MODEL_DICT["bsarec"](args=args)
    sd = torch.load(REF + f"/output/BSARec_{data_name}_best.pt", map_location="cpu")
    model.load_state_dict(sd)
    args.train_matrix = matrix
    trainer = Trainer(model, None, None, None, args, logging.getLogger("golden"))
    captured = {}
    orig = trainer.get_full_sort_score

    def grab(epoch, answers, pred):
        captured.update(a=answers, p=pred)
        return orig(epoch, answers, pred)

    trainer.get_full_sort_score = grab
    U = len(ds)
    all_ids = torch.stack([ds[i][1] for i in range(U)])
    all_ans = torch.stack([ds[i][2] for i in range(U)])
    batches = []
    for s in range(0, U, 256):
        e = min(U, s + 256)
        batches.append((torch.arange(s, e), all_ids[s:e], all_ans[s:e],
                        torch.zeros(e - s, 0, dtype=torch.long), torch.zeros(e - s, 0, dtype=torch.long)))
    torch.set_grad_enabled(False)
    scores6, info = trainer.iteration(0, batches, train=False)
    # relative gap between the 20th and 21st masked score, to flag near-ties
    gaps = np.zeros(U, dtype=np.float32)
    for (uids, ids, _, _, _) in batches:
        sc = trainer.predict_full(model.predict(ids, None)[:, -1, :]).numpy().copy()
        sc[matrix[uids.numpy()].toarray() > 0] = 0
        top = -np.sort(-sc, axis=1)[:, :21]
        gaps[uids.numpy()] = (top[:, 19] - top[:, 20]) / np.maximum(np.abs(top[:, 19]), 1e-30)
    torch.set_grad_enabled(True)
    out = {
        "ids": all_ids.numpy().astype(np.int16), "answers": all_ans.numpy().astype(np.int32),
        "hist_indptr": matrix.indptr.astype(np.int32), "hist_indices": matrix.indices.astype(np.int16),
        "pred_list": captured["p"].astype(np.int16), "metrics": np.array(scores6, dtype=np.float64),
        "info": np.array(info), "gap_rel": gaps,
        "item_size": np.array(args.item_size),
        "cfg_keys": np.array(sorted(cfgkw.keys())),
        "cfg_vals": np.array([float(cfgkw[k]) for k in sorted(cfgkw.keys())]),
    }
    if store_params:
        for n, p in model.state_dict().items():
            out["param." + n] = p.numpy()
    np.savez_compressed(os.path.join(HERE, out_name + ".npz"), **out)
    print(out_name, info)


def beauty_case():
    known_answer_case("Beauty", "test", "beauty_test", num_attention_heads=1, c=5, alpha=0.7)


def dataset_case():
    """Batch construction: the reference's RecDataset / rating matrices on a seeded synthetic corpus -> dataset_small.npz."""
    rng = np.random.RandomState(5)
    L, N = 12, 60
    lens = [3, 4, 5, 13, 14, 15, 16, 40] + list(rng.randint(3, 30, size=40))
    user_seq = [list(map(int, rng.randint(1, N, size=n))) for n in lens]
    args = argparse.Namespace(max_seq_length=L, item_size=N, model_type="BSARec")
    out = {"L": L, "item_size": N, "seq_indptr": np.cumsum([0] + lens).astype(np.int64),
           "seq_items": np.concatenate(user_seq).astype(np.int64)}
    import random
    random.seed(0)
    for dt in ("train", "valid", "test"):
        ds = ref_dataset.RecDataset(args, user_seq, data_type=dt)
        rows = [ds[i] for i in range(len(ds))]
        out[f"{dt}_user_ids"] = np.array([int(r[0]) for r in rows], dtype=np.int64)
        out[f"{dt}_input_ids"] = torch.stack([r[1] for r in rows]).numpy()
        out[f"{dt}_answers"] = np.array([int(r[2]) for r in rows], dtype=np.int64)
    for dt, fn in (("valid", ref_dataset.generate_rating_matrix_valid), ("test", ref_dataset.generate_rating_matrix_test)):
        m = fn(user_seq, len(user_seq), N).tocsr()
        m.sort_indices()
        out[f"{dt}_hist_indptr"] = m.indptr.astype(np.int64)
        out[f"{dt}_hist_indices"] = m.indices.astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "dataset_small.npz"), **out)
    print("dataset_small.npz", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    if "--fmlprec" in sys.argv:
        fmlprec_case("fmlprec_small", 23, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_hidden_layers=2)
        fmlprec_case("fmlprec_odd", 29, B=5, item_size=97, hidden_size=32, max_seq_length=25, num_hidden_layers=1)
        sys.exit(0)
    if "--duorec" in sys.argv:
        duorec_case("duorec_small", 19, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_attention_heads=2,
                    num_hidden_layers=2)
        sys.exit(0)
    if "--sasrec" in sys.argv:
        sasrec_case("sasrec_small", 17, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_attention_heads=2,
                    num_hidden_layers=2)
        sys.exit(0)
    if "--known-answer-extra" in sys.argv:
        known_answer_case("LastFM", "test", "lastfm_test", num_attention_heads=1, c=3, alpha=0.9)
        known_answer_case("LastFM", "valid", "lastfm_valid", store_params=False, num_attention_heads=1, c=3, alpha=0.9)
        known_answer_case("Beauty", "valid", "beauty_valid", store_params=False, num_attention_heads=1, c=5, alpha=0.7)
        sys.exit(0)
    if "--dataset-only" in sys.argv:
        dataset_case()
        sys.exit(0)
    small_case("small_a", 7, B=6, item_size=211, hidden_size=32, max_seq_length=24, num_attention_heads=4,
               num_hidden_layers=2, c=7, alpha=0.7)
    small_case("small_b", 11, B=8, item_size=500, hidden_size=64, max_seq_length=50, num_attention_heads=1,
               num_hidden_layers=2, c=5, alpha=0.7)
    small_case("small_c", 13, B=5, item_size=97, hidden_size=64, max_seq_length=25, num_attention_heads=2,
               num_hidden_layers=3, c=9, alpha=0.3)
    beauty_case()
    # the second shipped checkpoint (alpha 0.9, c 3 -> 2 kept bins) and the validation split of the first
    known_answer_case("LastFM", "test", "lastfm_test", num_attention_heads=1, c=3, alpha=0.9)
    known_answer_case("LastFM", "valid", "lastfm_valid", store_params=False, num_attention_heads=1, c=3, alpha=0.9)
    known_answer_case("Beauty", "valid", "beauty_valid", store_params=False, num_attention_heads=1, c=5, alpha=0.7)
    sasrec_case("sasrec_small", 17, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_attention_heads=2,
                num_hidden_layers=2)
    duorec_case("duorec_small", 19, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_attention_heads=2,
                num_hidden_layers=2)
    fmlprec_case("fmlprec_small", 23, B=8, item_size=300, hidden_size=64, max_seq_length=50, num_hidden_layers=2)
    fmlprec_case("fmlprec_odd", 29, B=5, item_size=97, hidden_size=32, max_seq_length=25, num_hidden_layers=1)
    dataset_case()

"""CUDA-graph capture of the whole training / evaluation step (bsarec_b200/graphs.py): a replay must be the eager loop
body of trainers.py:103-107 / :126-149 — same kernels, same arguments — with the dropout seed and Adam's step number
read from device memory.  Replay r of a graph captured with seed s is compared with an EAGER step run with the seed
``replay_seed(s, r)``; trajectories must agree to fp32 rounding of the bias corrections (device pow vs host pow)."""
import numpy as np
import pytest
import torch

from helpers import make_args, synth_batch, oracle_cfg, perturbed_params, build_model, csr_from_ids

pytestmark = pytest.mark.gpu

CFG = dict(item_size=500, hidden_size=64, max_seq_length=50, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)


def _close(a, b, tol=2e-6):
    a, b = a.double(), b.double()
    return float((a - b).abs().max()) <= tol * max(float(b.abs().max()), 1e-30)


def _batches(n, B, seed0=10):
    return [synth_batch(seed0 + i, B, CFG["max_seq_length"], CFG["item_size"]) for i in range(n)]


def _eager_step(model, opt, ids, answers, seed, extra=(None, None, None)):
    model.set_dropout_seed(seed)
    loss = model.calculate_loss(ids, answers, *extra)
    opt.zero_grad()
    loss.backward()
    opt.step()
    model.set_dropout_seed(None)
    return loss.detach().clone()


def test_graphed_train_step_equals_eager_trajectory(cuda_device):
    from bsarec_b200 import FusedAdam, GraphedTrainStep
    from bsarec_b200.graphs import replay_seed
    P = perturbed_params(oracle_cfg(CFG, 0.5, 0.5), 3)
    B, base = 48, 12345
    batches = [(i.to(cuda_device), a.to(cuda_device)) for i, a in _batches(5, B)]
    short = tuple(t.to(cuda_device) for t in synth_batch(99, 17, CFG["max_seq_length"], CFG["item_size"]))
    # eager trajectory: steps 0, 1 with replay seeds 0, 1; a short batch (eager fallback, seed 777); steps 2, 3, 4
    m_e = build_model(CFG, P, cuda_device, 0.5, 0.5).train()
    o_e = FusedAdam(m_e.parameters(), lr=1e-2)
    losses_e = []
    for r in (0, 1):
        losses_e.append(_eager_step(m_e, o_e, *batches[r], replay_seed(base, r)))
    losses_e.append(_eager_step(m_e, o_e, *short, 777))
    for r in (2, 3, 4):
        losses_e.append(_eager_step(m_e, o_e, *batches[r], replay_seed(base, r)))
    # graphed
    m_g = build_model(CFG, P, cuda_device, 0.5, 0.5).train()
    o_g = FusedAdam(m_g.parameters(), lr=1e-2)
    step = GraphedTrainStep(m_g, o_g, seed=base)
    losses_g = []
    for r in (0, 1):
        losses_g.append(step(*batches[r]).clone())
    assert step.replays == 2 and step.launches_per_replay > 20
    m_g.set_dropout_seed(777)
    losses_g.append(step(*short).clone())               # other shape -> eager body, Adam's step number advances on the host
    m_g.set_dropout_seed(None)
    for r in (2, 3, 4):
        losses_g.append(step(*batches[r]).clone())
    assert step.replays == 5
    for i, (a, b) in enumerate(zip(losses_g, losses_e)):
        assert _close(a, b), (i, a.item(), b.item())
    for (n, pg), (_, pe) in zip(m_g.named_parameters(), m_e.named_parameters()):
        assert _close(pg.detach(), pe.detach(), 1e-5), n
        assert pg.grad is not None and _close(pg.grad, pe.grad, 1e-5), n
    sg, se = o_g.state_dict()["state"], o_e.state_dict()["state"]
    assert all(int(sg[k]["step"]) == int(se[k]["step"]) == 6 for k in sg)
    for k in sg:
        assert _close(sg[k]["exp_avg"], se[k]["exp_avg"], 1e-5) and _close(sg[k]["exp_avg_sq"], se[k]["exp_avg_sq"], 1e-5)


def test_graph_capture_leaves_parameters_and_state_untouched(cuda_device):
    """The warm-up step that precedes the capture is undone: after ONE call the model is exactly one step from its start."""
    from bsarec_b200 import FusedAdam, GraphedTrainStep
    P = perturbed_params(oracle_cfg(CFG, 0.5, 0.5), 4)
    ids, answers = (t.to(cuda_device) for t in synth_batch(5, 32, CFG["max_seq_length"], CFG["item_size"]))
    m_e = build_model(CFG, P, cuda_device, 0.5, 0.5).train()
    o_e = FusedAdam(m_e.parameters(), lr=1e-2)
    _eager_step(m_e, o_e, ids, answers, 4242)
    m_g = build_model(CFG, P, cuda_device, 0.5, 0.5).train()
    o_g = FusedAdam(m_g.parameters(), lr=1e-2)
    GraphedTrainStep(m_g, o_g, seed=4242)(ids, answers)
    for (n, pg), (_, pe) in zip(m_g.named_parameters(), m_e.named_parameters()):
        assert _close(pg.detach(), pe.detach(), 1e-6), n
    assert all(int(s["step"]) == 1 for s in o_g.state_dict()["state"].values())


def test_replays_draw_fresh_dropout_masks(cuda_device):
    """Same batch twice with lr = 0: the loss differs between replays (new masks) and repeats under the same replay seed."""
    from bsarec_b200 import FusedAdam, GraphedTrainStep
    P = perturbed_params(oracle_cfg(CFG, 0.5, 0.5), 5)
    ids, answers = (t.to(cuda_device) for t in synth_batch(6, 32, CFG["max_seq_length"], CFG["item_size"]))
    model = build_model(CFG, P, cuda_device, 0.5, 0.5).train()
    step = GraphedTrainStep(model, FusedAdam(model.parameters(), lr=0.0), seed=1)
    l = [step(ids, answers).item() for _ in range(3)]
    assert len({round(x, 7) for x in l}) == 3, l


@pytest.mark.parametrize("kind", ["sasrec", "duorec", "fmlprec"])
def test_graphed_sibling_models(kind, cuda_device):
    from bsarec_b200 import FusedAdam, GraphedTrainStep, MODEL_DICT
    from bsarec_b200.graphs import replay_seed
    a = make_args(CFG, 0.3, 0.3)
    a.tau, a.lmd, a.lmd_sem, a.ssl, a.sim = 1.0, 0.1, 0.1, "us_x", "dot"
    torch.manual_seed(0)
    m_e = MODEL_DICT[kind](a).to(cuda_device).train()
    m_g = MODEL_DICT[kind](a).to(cuda_device).train()
    m_g.load_state_dict(m_e.state_dict())
    B, L, N = 24, CFG["max_seq_length"], CFG["item_size"]
    data = []
    for i in range(3):
        ids, ans = synth_batch(40 + i, B, L, N)
        neg = torch.randint(1, N, (B,), generator=torch.Generator().manual_seed(i))
        sem, _ = synth_batch(70 + i, B, L, N)
        data.append(tuple(t.to(cuda_device) for t in (ids, ans, neg, sem)))
    o_e, o_g = FusedAdam(m_e.parameters(), lr=5e-3), FusedAdam(m_g.parameters(), lr=5e-3)
    step = GraphedTrainStep(m_g, o_g, seed=31)
    for r, (ids, ans, neg, sem) in enumerate(data):
        le = _eager_step(m_e, o_e, ids, ans, replay_seed(31, r), (neg, sem, None))
        lg = step(ids, ans, neg, sem, None)
        assert _close(lg, le), (kind, r, lg.item(), le.item())
    for (n, pg), (_, pe) in zip(m_g.named_parameters(), m_e.named_parameters()):
        assert _close(pg.detach(), pe.detach(), 1e-5), (kind, n)


def test_graphed_topk_equals_eager_and_follows_the_weights(cuda_device):
    from bsarec_b200 import FusedAdam, GraphedTopk, GraphedTrainStep
    P = perturbed_params(oracle_cfg(CFG), 6)
    model = build_model(CFG, P, cuda_device, 0.5, 0.5)
    B, L, N = 40, CFG["max_seq_length"], CFG["item_size"]
    users = 3 * B
    all_ids, all_ans = synth_batch(8, users, L, N)
    indptr, indices = csr_from_ids(all_ids)
    indptr, indices = torch.from_numpy(indptr).to(cuda_device), torch.from_numpy(indices).to(cuda_device)
    all_ids = all_ids.to(cuda_device)
    topk = GraphedTopk(model.eval(), k=20)
    for pass_no in range(2):
        for b in range(3):
            rows = torch.arange(b * B, (b + 1) * B, device=cuda_device)
            got = topk(all_ids[rows], indptr, indices, rows).clone()
            ref = model.full_sort_topk(all_ids[rows], None, indptr, indices, k=20, hist_rows=rows)
            assert torch.equal(got, ref), (pass_no, b)
        assert topk.launches_per_replay > 5
        # a graphed training step changes the table: the next evaluation must see the new weights (re-capture)
        model.train()
        step = GraphedTrainStep(model, FusedAdam(model.parameters(), lr=5e-2), seed=3)
        step(all_ids[:B], all_ans[:B].to(cuda_device))
        step(all_ids[B:2 * B], all_ans[B:2 * B].to(cuda_device))
        model.eval()


def test_trainer_with_cuda_graph_matches_eager_trainer_metrics(cuda_device):
    """bsarec_b200.Trainer with args.cuda_graph: the evaluation branch returns the eager trainer's scores exactly and the
    train branch runs (loss finite, weights move)."""
    import scipy.sparse as sp
    from bsarec_b200 import BSARecModel, Trainer
    B, L, N, users = 32, CFG["max_seq_length"], CFG["item_size"], 32 * 3 + 7
    ids, ans = synth_batch(21, users, L, N)
    indptr, indices = csr_from_ids(ids)
    mat = sp.csr_matrix((np.ones(len(indices)), indices, indptr), shape=(users, N))
    uid = torch.arange(users)
    neg = torch.randint(1, N, (users,))

    def loader(train):
        out = []
        for s in range(0, users, B):
            sl = slice(s, min(s + B, users))
            out.append((uid[sl], ids[sl], ans[sl], neg[sl], ids[sl]) if train else (uid[sl], ids[sl], ans[sl]))
        return out
    scores = {}
    for graph in (False, True):
        args = make_args(CFG, 0.5, 0.5, batch_size=B)
        args.cuda_graph = graph
        args.valid_rating_matrix = args.test_rating_matrix = args.train_matrix = mat
        torch.manual_seed(5)
        model = BSARecModel(args)
        tr = Trainer(model, loader(True), loader(False), loader(False), args, None)
        scores[graph] = tr.test(0)
        if graph:
            before = model.item_embeddings.weight.detach().clone()
            tr.train(0)
            assert not torch.equal(before, model.item_embeddings.weight.detach())
            s2, _ = tr.valid(0)
            assert all(np.isfinite(s2))
    assert scores[True] == scores[False]

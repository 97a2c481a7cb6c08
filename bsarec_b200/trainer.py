"""Host-side mirror of the reference's ``Trainer`` (src/trainers.py:8-158) for the BSARec hot path.

Same constructor and methods (train / valid / test / save / load / predict_full / get_full_sort_score /
iteration).  The train branch is the reference's loop unchanged in meaning (H2D copy, calculate_loss,
zero_grad, backward, Adam.step, loss.item()); the eval branch keeps scoring, history masking and top-20 on
the device through ``BSARecModel.full_sort_topk`` and returns the same ``(scores, str)`` pair.

``args.cuda_graph = True`` (additional, off by default): the optimizer is ``bsarec_b200.FusedAdam`` and both loop
bodies run as one CUDA-graph launch per full-size batch (``bsarec_b200/graphs.py``) — for the launch-bound shapes of
the real datasets.
"""
from __future__ import annotations

import numpy as np
import torch
from torch.optim import Adam


def hit_ranks(answers: np.ndarray, pred_list: np.ndarray) -> np.ndarray:
    """0-based rank of each user's ground-truth item in their list, or -1 (vectorised metrics.py:3-23)."""
    hit = pred_list == np.asarray(answers).reshape(-1, 1)
    return np.where(hit.any(axis=1), hit.argmax(axis=1), -1)


def recall_at_k(answers, pred_list, k: int) -> float:
    r = hit_ranks(answers, pred_list)
    return float(((r >= 0) & (r < k)).sum()) / float(len(pred_list))


def ndcg_k(answers, pred_list, k: int) -> float:
    r = hit_ranks(answers, pred_list)
    ok = (r >= 0) & (r < k)
    return float((1.0 / np.log2(r[ok] + 2.0)).sum()) / float(len(pred_list))


class BatchHistory:
    """The trainer's scipy CSR history matrix (``args.train_matrix``, dataset.py:126-160) kept on the device.  The
    scorer reads row ``user_ids[b]`` of it for batch row b (``hist_rows`` of ``bsa_score_topk``), so the eval loop
    builds no per-batch CSR and never synchronises with the host."""

    def __init__(self, csr, device):
        csr = csr.tocsr()
        csr.sort_indices()
        self.num_rows = csr.shape[0]
        self.indptr = torch.from_numpy(csr.indptr.astype(np.int64)).to(device)
        indices = csr.indices.astype(np.int64) if csr.nnz else np.zeros(1, dtype=np.int64)
        self.indices = torch.from_numpy(indices).to(device)


class Trainer:
    def __init__(self, model, train_dataloader, eval_dataloader, test_dataloader, args, logger):
        self.args = args
        self.logger = logger
        if not torch.cuda.is_available() or getattr(args, "no_cuda", False):
            raise RuntimeError("bsarec_b200.Trainer needs a CUDA device: the hot path has no CPU fallback")
        self.cuda_condition = True
        self.device = torch.device("cuda")
        self.model = model.cuda()
        self.train_dataloader = train_dataloader
        self.eval_dataloader = eval_dataloader
        self.test_dataloader = test_dataloader
        betas = (args.adam_beta1, args.adam_beta2)
        self.cuda_graph = bool(getattr(args, "cuda_graph", False))
        if self.cuda_graph:
            from .graphs import GraphedTrainStep, GraphedTopk
            from .optim import FusedAdam
            self.optim = FusedAdam(self.model.parameters(), lr=args.lr, betas=betas, weight_decay=args.weight_decay)
            self._train_step = GraphedTrainStep(self.model, self.optim)
            self._topk = GraphedTopk(self.model, k=20)
        else:
            self.optim = Adam(self.model.parameters(), lr=args.lr, betas=betas, weight_decay=args.weight_decay)
        self._hist_cache = {}
        if logger is not None:
            logger.info(f"Total Parameters: {sum([p.nelement() for p in self.model.parameters()])}")

    def train(self, epoch):
        self.iteration(epoch, self.train_dataloader, train=True)

    def valid(self, epoch):
        self.args.train_matrix = self.args.valid_rating_matrix
        return self.iteration(epoch, self.eval_dataloader, train=False)

    def test(self, epoch):
        self.args.train_matrix = self.args.test_rating_matrix
        return self.iteration(epoch, self.test_dataloader, train=False)

    def save(self, file_name):
        torch.save({k: v.cpu() for k, v in self.model.state_dict().items()}, file_name)

    def load(self, file_name):
        """Reference semantics (trainers.py:47-60) plus map_location, so GPU-saved files load anywhere."""
        state = self.model.state_dict()
        for key, value in torch.load(file_name, map_location="cpu").items():
            state[key] = value
        self.model.load_state_dict(state)

    def predict_full(self, seq_out):
        """Compat path of trainers.py:62-68: the dense [B, item_size] score matrix, via the fused scorer is
        not possible (it never materialises it), so this is a plain device matmul for callers that want it."""
        return torch.matmul(seq_out, self.model.item_embeddings.weight.detach().transpose(0, 1))

    def _post_fix(self, epoch, recall, ndcg):
        """The log line / return pair of trainers.py:70-83 from HR@k, NDCG@k at k = 5, 10, 15, 20."""
        post_fix = {
            "Epoch": epoch,
            "HR@5": '{:.4f}'.format(recall[0]), "NDCG@5": '{:.4f}'.format(ndcg[0]),
            "HR@10": '{:.4f}'.format(recall[1]), "NDCG@10": '{:.4f}'.format(ndcg[1]),
            "HR@20": '{:.4f}'.format(recall[3]), "NDCG@20": '{:.4f}'.format(ndcg[3]),
        }
        if self.logger is not None:
            self.logger.info(post_fix)
        return [recall[0], ndcg[0], recall[1], ndcg[1], recall[3], ndcg[3]], str(post_fix)

    def get_full_sort_score(self, epoch, answers, pred_list):
        ks = [5, 10, 15, 20]
        return self._post_fix(epoch, [recall_at_k(answers, pred_list, k) for k in ks], [ndcg_k(answers, pred_list, k) for k in ks])

    def get_full_sort_score_from_hist(self, epoch, hist):
        """Same return value as ``get_full_sort_score`` from the device-accumulated hit-rank histogram
        (``data.HitRankAccumulator``): no per-user host work, one read of 21 integers."""
        from .data import metrics_from_hist
        recall, ndcg = metrics_from_hist(hist, [5, 10, 15, 20])
        return self._post_fix(epoch, recall, ndcg)

    def _history(self):
        m = self.args.train_matrix
        if id(m) not in self._hist_cache:
            self._hist_cache[id(m)] = BatchHistory(m, self.device)
        return self._hist_cache[id(m)]

    def iteration(self, epoch, dataloader, train=True):
        if train:
            self.model.train()
            rec_loss, steps = 0.0, 0
            for batch in dataloader:
                user_ids, input_ids, answers, neg_answer, same_target = (t.to(self.device, non_blocking=True) for t in batch)
                if self.cuda_graph:
                    loss = self._train_step(input_ids, answers, neg_answer, same_target, user_ids)
                else:
                    loss = self.model.calculate_loss(input_ids, answers, neg_answer, same_target, user_ids)
                    self.optim.zero_grad()
                    loss.backward()
                    self.optim.step()
                rec_loss += loss.item()
                steps += 1
            post_fix = {"epoch": epoch, "rec_loss": '{:.4f}'.format(rec_loss / max(steps, 1))}
            if self.logger is not None and (epoch + 1) % getattr(self.args, "log_freq", 1) == 0:
                self.logger.info(str(post_fix))
            return None
        from .data import HitRankAccumulator
        self.model.eval()
        hist = self._history()
        acc = HitRankAccumulator(20, self.device)
        keep = getattr(self.args, "keep_pred_list", False)
        preds, answers_all = [], []
        with torch.no_grad():
            for batch in dataloader:
                user_ids, input_ids, answers = (t.to(self.device, non_blocking=True) for t in batch[:3])
                if self.cuda_graph and input_ids.shape[0] == getattr(self.args, "batch_size", -1):
                    pred = self._topk(input_ids, hist.indptr, hist.indices, user_ids)
                    if keep:
                        pred = pred.clone()          # the graph's output tensor is rewritten by the next replay
                else:
                    pred = self.model.full_sort_topk(input_ids, user_ids, hist.indptr, hist.indices, k=20, hist_rows=user_ids)
                acc.update(pred, answers)
                if keep:
                    preds.append(pred)
                    answers_all.append(answers)
        if keep:                                   # what trainers.py:151-156 accumulates, for callers that want the lists
            self.pred_list = torch.cat(preds).cpu().numpy()
            self.answer_list = torch.cat(answers_all).cpu().numpy()
        return self.get_full_sort_score_from_hist(epoch, acc.hist.cpu().numpy())

"""GPU-resident batch construction and metric accumulation.

``SequenceStore`` keeps the user sequences of ``<data_name>.txt`` (dataset.py:184-197) as one CSR on the device.
``DeviceBatchLoader`` is iterable like the reference's ``DataLoader`` over a ``RecDataset`` (dataset.py:9-117,
205-222) and yields the same 5-tuple ``(user_ids, input_ids, answers, neg_answers, same_target)`` — already on the
device, built by ``bsa_build_batch`` — so ``Trainer.iteration`` consumes it unchanged.  ``neg_answers`` is zeros:
the reference draws it with Python's ``random`` (dataset.py:66-67,120-124) and BSARec ignores it (bsarec.py:30).
``HitRankAccumulator`` keeps the hit-rank histogram HR@k / NDCG@k are sums over (metrics.py:3-23) on the device.

No CPU fallback: everything here needs the CUDA library.
"""
from __future__ import annotations

import math
from typing import Iterable, List, Sequence

import numpy as np
import torch

from . import _cabi
from ._cabi import check

MODES = {"train": 0, "valid": 1, "test": 2}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class SequenceStore:
    """CSR of user sequences on the device + the train-sample index (prefix sums of samples per user)."""

    def __init__(self, user_seq: Sequence[Sequence[int]], max_seq_length: int, device="cuda"):
        lens = np.fromiter((len(s) for s in user_seq), dtype=np.int64, count=len(user_seq))
        if len(lens) == 0:
            raise ValueError("SequenceStore: no sequences")
        if int(lens.min()) < 3:
            raise ValueError("SequenceStore: every user needs >= 3 interactions (train / valid / test targets, "
                             "dataset.py:20,26,28)")
        self.max_len = int(max_seq_length)
        self.num_users = int(len(lens))
        indptr = np.zeros(self.num_users + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        items = np.fromiter((i for s in user_seq for i in s), dtype=np.int64, count=int(indptr[-1]))
        per_user = np.minimum(np.maximum(lens - 2, 0), self.max_len)       # len(seq[-(L+2):-2]), dataset.py:20
        sample_start = np.zeros(self.num_users + 1, dtype=np.int64)
        np.cumsum(per_user, out=sample_start[1:])
        self.num_train_samples = int(sample_start[-1])
        self.device = torch.device(device)
        self.seq_indptr = torch.from_numpy(indptr).to(self.device)
        self.seq_items = torch.from_numpy(items).to(self.device)
        self.sample_start = torch.from_numpy(sample_start).to(self.device)
        self._lens_host = lens

    def num_samples(self, data_type: str) -> int:
        return self.num_train_samples if data_type == "train" else self.num_users

    def build(self, data_type: str, sample_idx: torch.Tensor):
        """``(user_ids, input_ids, answers)`` for the sampler indices ``sample_idx`` (int64, CUDA)."""
        if not sample_idx.is_cuda or sample_idx.dtype != torch.int64:
            raise RuntimeError("SequenceStore.build: sample_idx must be an int64 CUDA tensor (no CPU fallback)")
        lib = _cabi.load()
        sample_idx = sample_idx.contiguous()
        B, L = sample_idx.numel(), self.max_len
        input_ids = torch.empty(B, L, dtype=torch.int64, device=self.device)
        answers = torch.empty(B, dtype=torch.int64, device=self.device)
        user_ids = torch.empty(B, dtype=torch.int64, device=self.device)
        check(lib.bsa_build_batch(MODES[data_type], self.seq_indptr.data_ptr(), self.seq_items.data_ptr(),
                                  self.sample_start.data_ptr(), self.num_users, sample_idx.data_ptr(), B, L,
                                  input_ids.data_ptr(), answers.data_ptr(), user_ids.data_ptr(), _stream()),
              "bsa_build_batch")
        return user_ids, input_ids, answers

    def history_csr(self, data_type: str):
        """The rating matrix of dataset.py:126-160 (valid: all but the last 2 items, test: all but the last) as a
        device CSR with sorted distinct column ids — what ``Trainer``'s eval branch masks to 0.0 (trainers.py:131-134)."""
        drop = 2 if data_type == "valid" else 1
        lens = self.seq_indptr[1:] - self.seq_indptr[:-1]
        pos = torch.arange(self.seq_items.numel(), device=self.device)
        row = torch.repeat_interleave(torch.arange(self.num_users, device=self.device), lens)
        keep = (pos - self.seq_indptr[row]) < (lens[row] - drop)
        row, col = row[keep], self.seq_items[keep]
        stride = int(col.max().item()) + 1 if col.numel() else 1
        key = torch.unique(row * stride + col)                # sorted, duplicates collapse (csr_matrix sums them)
        row, col = key // stride, key % stride
        indptr = torch.zeros(self.num_users + 1, dtype=torch.int64, device=self.device)
        indptr[1:] = torch.cumsum(torch.bincount(row, minlength=self.num_users), 0)
        return indptr, col.contiguous()


class DeviceBatchLoader:
    """Iterable stand-in for ``DataLoader(RecDataset(...), sampler=Random|Sequential, batch_size=...)``."""

    def __init__(self, store: SequenceStore, data_type: str, batch_size: int, shuffle: bool | None = None, seed: int = 0,
                 rank: int = 0, world_size: int = 1, equal_shards: bool | None = None):
        """``equal_shards`` (default: on for train when world_size > 1): every rank yields the SAME number of batches of
        the SAME sizes — what the data-parallel gradient exchange needs (its collectives run once per batch and the
        mean of local means must be the global mean).  The sample order is wrap-padded to a multiple of ``world_size``
        (at most world_size-1 samples of the epoch are seen twice, as ``DistributedSampler`` does); the short last
        batch, if any, is equally short on every rank.  Off for valid/test: every user is scored exactly once and the
        ranks may differ by one sample."""
        if data_type not in MODES:
            raise ValueError(f"data_type must be one of {sorted(MODES)}")
        self.store, self.data_type, self.batch_size = store, data_type, int(batch_size)
        self.shuffle = (data_type == "train") if shuffle is None else bool(shuffle)     # dataset.py:207-208,212,217
        self.rank, self.world_size = rank, world_size
        self.equal_shards = (data_type == "train" and world_size > 1) if equal_shards is None else bool(equal_shards)
        self._gen = torch.Generator(device=store.device)
        self._gen.manual_seed(seed)
        self._zeros = {}

    def __len__(self):
        n = self._num_local()
        return (n + self.batch_size - 1) // self.batch_size

    def _num_local(self):
        n = self.store.num_samples(self.data_type)
        if self.equal_shards:
            return (n + self.world_size - 1) // self.world_size
        return (n - self.rank + self.world_size - 1) // self.world_size

    def __iter__(self):
        n = self.store.num_samples(self.data_type)
        dev = self.store.device
        order = torch.randperm(n, generator=self._gen, device=dev) if self.shuffle else torch.arange(n, device=dev)
        if self.equal_shards and n % self.world_size:
            order = torch.cat([order, order[:self.world_size - n % self.world_size]])      # wrap-pad: equal shard sizes
        order = order[self.rank::self.world_size]                 # one process per GPU: disjoint sample shards
        for lo in range(0, order.numel(), self.batch_size):
            idx = order[lo:lo + self.batch_size]
            user_ids, input_ids, answers = self.store.build(self.data_type, idx)
            B = idx.numel()
            neg = self._zeros.setdefault(B, torch.zeros(B, dtype=torch.int64, device=dev))
            same = self._zeros.setdefault(("s", B), torch.zeros(B, 0, dtype=torch.int64, device=dev))
            if self.data_type == "train":
                yield user_ids, input_ids, answers, neg, same
            else:
                yield user_ids, input_ids, answers, same, same


class HitRankAccumulator:
    """Device-side replacement for growing ``pred_list`` with ``np.append`` (trainers.py:151-156) and the per-user
    Python loops of ``recall_at_k`` / ``ndcg_k`` (metrics.py:3-23): a (k+1)-bin integer histogram of hit ranks."""

    def __init__(self, k: int = 20, device="cuda"):
        self.k = k
        self.hist = torch.zeros(k + 1, dtype=torch.int64, device=device)

    def update(self, pred: torch.Tensor, answers: torch.Tensor):
        if not (pred.is_cuda and answers.is_cuda):
            raise RuntimeError("HitRankAccumulator.update: CUDA tensors only (no CPU fallback)")
        pred, answers = pred.contiguous(), answers.contiguous()
        assert pred.dtype == torch.int64 and answers.dtype == torch.int64 and pred.shape == (answers.numel(), self.k)
        check(_cabi.load().bsa_hit_rank_hist(pred.data_ptr(), answers.data_ptr(), answers.numel(), self.k,
                                              self.hist.data_ptr(), _stream()), "bsa_hit_rank_hist")

    def all_reduce(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.hist)
        return self

    def metrics(self, ks: Iterable[int] = (5, 10, 15, 20)):
        """``([HR@k...], [NDCG@k...])`` in float64, one device→host read of k+1 integers."""
        return metrics_from_hist(self.hist.cpu().numpy(), ks)


def metrics_from_hist(hist: np.ndarray, ks: Iterable[int] = (5, 10, 15, 20)):
    users = int(hist.sum())
    recall: List[float] = []
    ndcg: List[float] = []
    for k in ks:
        recall.append(float(int(hist[:k].sum())) / float(users))
        ndcg.append(sum(int(hist[r]) * (1.0 / math.log(r + 2, 2)) for r in range(k)) / float(users))
    return recall, ndcg

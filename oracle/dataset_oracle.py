"""CPU restatement of the reference's batch construction.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows ``RecDataset`` (src/dataset.py:9-117) and the rating matrices (src/dataset.py:126-160) with plain Python
lists; pinned by ``tests/golden/dataset_small.npz`` (made by ``tests/golden/make_golden.py`` from the unmodified
reference classes) in ``tests/test_oracle_vs_golden.py``.
"""
from typing import List, Sequence, Tuple

import numpy as np


def expand_samples(user_seq: Sequence[Sequence[int]], max_len: int, data_type: str) -> Tuple[List[List[int]], List[int]]:
    """``(items per sample, user id per sample)`` — RecDataset.__init__, dataset.py:18-28."""
    samples, users = [], []
    if data_type == "train":
        for user, seq in enumerate(user_seq):
            window = list(seq[-(max_len + 2):-2])                       # dataset.py:20
            for i in range(len(window)):
                samples.append(window[:i + 1])                           # dataset.py:22
                users.append(user)
    elif data_type == "valid":
        for user, seq in enumerate(user_seq):
            samples.append(list(seq[:-1]))                               # dataset.py:26
            users.append(user)
    else:
        for user, seq in enumerate(user_seq):
            samples.append(list(seq))                                    # dataset.py:28
            users.append(user)
    return samples, users


def get_item(samples, users, index: int, max_len: int, data_type: str):
    """``(user_id, input_ids[max_len], answer)`` — RecDataset.__getitem__, dataset.py:61-72 and :74-113."""
    items = samples[index]
    input_ids, answer = items[:-1], items[-1]
    input_ids = ([0] * (max_len - len(input_ids)) + input_ids)[-max_len:]
    user_id = users[index] if data_type == "train" else index           # dataset.py:76 vs :106
    return user_id, input_ids, answer


def build_batch(user_seq, max_len: int, data_type: str, sample_idx: Sequence[int]):
    samples, users = expand_samples(user_seq, max_len, data_type)
    rows = [get_item(samples, users, int(i), max_len, data_type) for i in sample_idx]
    return (np.array([r[0] for r in rows], dtype=np.int64), np.array([r[1] for r in rows], dtype=np.int64).reshape(-1, max_len),
            np.array([r[2] for r in rows], dtype=np.int64))


def history_csr(user_seq, data_type: str):
    """Sorted distinct history ids per user: valid drops the last 2 items, test the last — dataset.py:126-160."""
    drop = 2 if data_type == "valid" else 1
    indptr, indices = [0], []
    for seq in user_seq:
        cols = sorted(set(seq[:-drop]))
        indices.extend(cols)
        indptr.append(len(indices))
    return np.array(indptr, dtype=np.int64), np.array(indices, dtype=np.int64)


def hit_rank_hist(answers, pred_list, k: int) -> np.ndarray:
    """hist[r] = users whose answer is at position r of their top-k list, hist[k] = no hit (metrics.py:3-23)."""
    hist = np.zeros(k + 1, dtype=np.int64)
    for a, row in zip(answers, pred_list):
        row = list(row[:k])
        hist[row.index(a) if a in row else k] += 1
    return hist

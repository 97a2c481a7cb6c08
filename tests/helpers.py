"""Shared helpers for the GPU parity tests."""
import argparse

import numpy as np
import torch

import oracle
from oracle import OracleConfig


def make_args(cfgkw, p_hidden=0.0, p_attn=0.0, batch_size=256):
    a = argparse.Namespace(hidden_dropout_prob=p_hidden, attention_probs_dropout_prob=p_attn, hidden_act="gelu",
                           initializer_range=0.02, batch_size=batch_size, no_cuda=False, lr=1e-3, adam_beta1=0.9,
                           adam_beta2=0.999, weight_decay=0.0, log_freq=1)
    for k, v in cfgkw.items():
        setattr(a, k, v)
    return a


def oracle_cfg(cfgkw, p_hidden=0.0, p_attn=0.0):
    return OracleConfig(hidden_dropout_prob=p_hidden, attention_probs_dropout_prob=p_attn, **cfgkw)


def build_model(cfgkw, params, device, p_hidden=0.0, p_attn=0.0):
    from bsarec_b200 import BSARecModel
    m = BSARecModel(make_args(cfgkw, p_hidden, p_attn))
    m.load_state_dict({k: v.clone() for k, v in params.items()})
    return m.to(device)


def synth_batch(seed, B, L, N, pad_every=8):
    """Seeded left-padded id matrix: lengths U{1..L}, one all-padding row per `pad_every`, ids U{1..N-1}."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.zeros(B, L, dtype=torch.long)
    for b in range(B):
        if pad_every and b % pad_every == 1:
            continue
        n = int(torch.randint(1, L + 1, (1,), generator=g))
        ids[b, L - n:] = torch.randint(1, N, (n,), generator=g)
    answers = torch.randint(1, N, (B,), generator=g)
    return ids, answers


def perturbed_params(cfg: OracleConfig, seed):
    """Oracle init pushed off its symmetric points (LayerNorm (1,0), zero biases, tiny Linear weights)."""
    P = oracle.init_params(cfg, seed)
    g = torch.Generator().manual_seed(seed + 17)
    for k in P:
        if k.endswith("LayerNorm.weight"):
            P[k] = P[k] + 0.2 * torch.randn(P[k].shape, generator=g)
        elif k.endswith(".bias"):
            P[k] = P[k] + 0.05 * torch.randn(P[k].shape, generator=g)
        elif k.endswith("weight") and "embeddings" not in k:
            P[k] = P[k] * 4.0
    return P


def csr_from_ids(ids: torch.Tensor):
    """History CSR = each row's own non-zero ids, sorted and de-duplicated."""
    indptr, indices = [0], []
    for b in range(ids.shape[0]):
        items = sorted(set(ids[b].tolist()) - {0})
        indices += items
        indptr.append(len(indices))
    return np.asarray(indptr, dtype=np.int64), np.asarray(indices, dtype=np.int64)

"""CPU restatement (torch, fp32 or fp64) of BSARec's training / full-ranking-eval hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Functional style over a plain
``dict`` of tensors keyed by the reference's 42 state_dict names (SURVEY.md
Appendix C); gradients come from torch autograd, as they do in the reference
(``src/trainers.py:105-106``).  Every function cites the reference lines it restates.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from .rng_ref import dropout_keep_mask

__all__ = [
    "OracleConfig", "DropoutPlan", "init_params", "filter_matrix", "filter_basis",
    "attention_mask", "layer_norm_tf", "embed", "frequency_layer", "attention_layer",
    "feed_forward", "encode", "calculate_loss", "loss_and_grads", "predict_full",
    "full_sort_topk", "masked_scores", "recall_at_k", "ndcg_k", "full_sort_metrics",
    "adam_train_step", "eval_step", "param_names", "encode_sasrec", "sasrec_loss", "sasrec_loss_and_grads",
    "info_nce", "duorec_loss", "duorec_loss_and_grads", "encode_fmlprec", "fmlprec_loss", "fmlprec_loss_and_grads",
]


@dataclass
class OracleConfig:
    """The fields of ``args`` the model reads."""
    item_size: int
    hidden_size: int = 64
    max_seq_length: int = 50
    num_attention_heads: int = 2
    num_hidden_layers: int = 2
    c: int = 3
    alpha: float = 0.9
    hidden_dropout_prob: float = 0.5
    attention_probs_dropout_prob: float = 0.5
    initializer_range: float = 0.02

    @property
    def kept_bins(self) -> int:           # src/model/bsarec.py:87
        return self.c // 2 + 1


class DropoutPlan:
    """Train-mode dropout with the kernels' counter RNG (oracle/rng_ref.py)."""

    def __init__(self, seed: int):
        self.seed = int(seed)

    def apply(self, x: torch.Tensor, site: int, p: float) -> torch.Tensor:
        if p <= 0.0:
            return x
        keep = torch.from_numpy(dropout_keep_mask(self.seed, site, tuple(x.shape), p))
        scale = torch.tensor(1.0, dtype=torch.float32) / torch.tensor(1.0 - p, dtype=torch.float32)
        return torch.where(keep, x * scale.to(x.dtype), torch.zeros((), dtype=x.dtype))


def _drop(x, site, p, train, plan):
    """nn.Dropout semantics: identity in eval; torch RNG if no plan is given."""
    if not train or p <= 0.0:
        return x
    if plan is None:
        return F.dropout(x, p=p, training=True)
    return plan.apply(x, site, p)


# ----------------------------------------------------------------------------- parameters

def param_names(num_layers: int) -> List[str]:
    """state_dict key order of the reference module tree."""
    names = ["item_embeddings.weight", "position_embeddings.weight", "LayerNorm.weight", "LayerNorm.bias"]
    for l in range(num_layers):
        p = f"item_encoder.blocks.{l}."
        names += [p + "layer.filter_layer.sqrt_beta",
                  p + "layer.filter_layer.LayerNorm.weight", p + "layer.filter_layer.LayerNorm.bias"]
        for lin in ("query", "key", "value", "dense"):
            names += [p + f"layer.attention_layer.{lin}.weight", p + f"layer.attention_layer.{lin}.bias"]
        names += [p + "layer.attention_layer.LayerNorm.weight", p + "layer.attention_layer.LayerNorm.bias"]
        names += [p + "feed_forward.dense_1.weight", p + "feed_forward.dense_1.bias",
                  p + "feed_forward.dense_2.weight", p + "feed_forward.dense_2.bias",
                  p + "feed_forward.LayerNorm.weight", p + "feed_forward.LayerNorm.bias"]
    return names


def init_params(cfg: OracleConfig, seed: int = 42, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Random parameters with the reference's *distributions* (not its RNG stream):
    N(0, initializer_range^2) for every Linear/Embedding weight including padding row 0,
    zero Linear biases, LayerNorm (1, 0), sqrt_beta ~ N(0,1) shared across layers
    (``_abstract_model.py:26-39``, ``bsarec.py:43-44,88``)."""
    g = torch.Generator().manual_seed(seed)
    d, L, N = cfg.hidden_size, cfg.max_seq_length, cfg.item_size
    std = cfg.initializer_range
    P: Dict[str, torch.Tensor] = {}

    def normal(*shape):
        return (torch.randn(*shape, generator=g, dtype=torch.float64) * std).to(dtype)

    P["item_embeddings.weight"] = normal(N, d)
    P["position_embeddings.weight"] = normal(L, d)
    P["LayerNorm.weight"] = torch.ones(d, dtype=dtype)
    P["LayerNorm.bias"] = torch.zeros(d, dtype=dtype)
    sqrt_beta = torch.randn(1, 1, d, generator=g, dtype=torch.float64).to(dtype)
    for l in range(cfg.num_hidden_layers):
        p = f"item_encoder.blocks.{l}."
        P[p + "layer.filter_layer.sqrt_beta"] = sqrt_beta.clone()
        P[p + "layer.filter_layer.LayerNorm.weight"] = torch.ones(d, dtype=dtype)
        P[p + "layer.filter_layer.LayerNorm.bias"] = torch.zeros(d, dtype=dtype)
        for lin in ("query", "key", "value", "dense"):
            P[p + f"layer.attention_layer.{lin}.weight"] = normal(d, d)
            P[p + f"layer.attention_layer.{lin}.bias"] = torch.zeros(d, dtype=dtype)
        P[p + "layer.attention_layer.LayerNorm.weight"] = torch.ones(d, dtype=dtype)
        P[p + "layer.attention_layer.LayerNorm.bias"] = torch.zeros(d, dtype=dtype)
        P[p + "feed_forward.dense_1.weight"] = normal(4 * d, d)
        P[p + "feed_forward.dense_1.bias"] = torch.zeros(4 * d, dtype=dtype)
        P[p + "feed_forward.dense_2.weight"] = normal(d, 4 * d)
        P[p + "feed_forward.dense_2.bias"] = torch.zeros(d, dtype=dtype)
        P[p + "feed_forward.LayerNorm.weight"] = torch.ones(d, dtype=dtype)
        P[p + "feed_forward.LayerNorm.bias"] = torch.zeros(d, dtype=dtype)
    return {k: P[k] for k in param_names(cfg.num_hidden_layers)}


# ----------------------------------------------------------------------------- building blocks

def filter_basis(L: int, kept_bins: int) -> np.ndarray:
    """Orthonormal real basis C [L, r] (float64) with ``C @ C.T`` equal to the operator
    ``irfft(zero_{k>=kept}(rfft(., norm='ortho')), n=L, norm='ortho')`` of
    ``src/model/bsarec.py:93-97``.  Column 0 is the DC vector; bin k contributes a
    cos and a sin column, except the Nyquist bin of an even L (cos only, sin == 0)."""
    kept = min(int(kept_bins), L // 2 + 1)
    t = np.arange(L, dtype=np.float64)
    cols = [np.full(L, 1.0 / math.sqrt(L))]
    for k in range(1, kept):
        w = 2.0 * math.pi * k * t / L
        if L % 2 == 0 and k == L // 2:
            cols.append(np.cos(w) / math.sqrt(L))
        else:
            cols.append(np.cos(w) * math.sqrt(2.0 / L))
            cols.append(np.sin(w) * math.sqrt(2.0 / L))
    return np.stack(cols, axis=1)


def filter_matrix(L: int, kept_bins: int) -> np.ndarray:
    """The exact real L x L low-pass operator (float64); symmetric, rank 2*kept-1."""
    C = filter_basis(L, kept_bins)
    return C @ C.T


def attention_mask(ids: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """[B,1,L,L] additive mask: 0 where key j is a real item and j <= i, else -10000
    (``_abstract_model.py:53-69``)."""
    B, L = ids.shape
    key_ok = (ids > 0)[:, None, None, :]
    # built on the host and moved, as the reference does (_abstract_model.py:61-63: triu on CPU, then .cuda())
    causal = torch.tril(torch.ones(L, L, dtype=torch.bool))[None, None].to(ids.device)
    allowed = (key_ok & causal).to(dtype)
    return (1.0 - allowed) * -10000.0


def layer_norm_tf(x, weight, bias, eps: float = 1e-12):
    """Custom LayerNorm, eps inside the sqrt, biased variance (``_modules.py:16-20``)."""
    u = x.mean(-1, keepdim=True)
    s = (x - u).pow(2).mean(-1, keepdim=True)
    return weight * ((x - u) / torch.sqrt(s + eps)) + bias


def embed(P, ids, cfg: OracleConfig, train=False, plan: Optional[DropoutPlan] = None):
    """``add_position_embedding`` (``_abstract_model.py:14-24``).  The item lookup has
    padding_idx=0: row 0 is read normally but receives no gradient from this gather."""
    L = ids.shape[1]
    item = F.embedding(ids, P["item_embeddings.weight"], padding_idx=0)
    pos = P["position_embeddings.weight"][:L].unsqueeze(0)
    h = layer_norm_tf(item + pos, P["LayerNorm.weight"], P["LayerNorm.bias"])
    return _drop(h, 0, cfg.hidden_dropout_prob, train, plan)


def frequency_layer(P, prefix, x, cfg: OracleConfig, layer: int, train=False, plan=None, use_fft=True):
    """``FrequencyLayer.forward`` (``bsarec.py:90-104``)."""
    L = x.shape[1]
    if use_fft:
        spec = torch.fft.rfft(x, dim=1, norm="ortho")
        keep = (torch.arange(spec.shape[1], device=x.device) < cfg.kept_bins).to(spec.real.dtype)[None, :, None]
        low = torch.fft.irfft(spec * keep, n=L, dim=1, norm="ortho")
    else:
        Fm = torch.from_numpy(filter_matrix(L, cfg.kept_bins)).to(device=x.device, dtype=x.dtype)
        low = torch.einsum("ts,bsd->btd", Fm, x)
    beta2 = P[prefix + "sqrt_beta"] ** 2
    mixed = low + beta2 * (x - low)
    mixed = _drop(mixed, 1 + 4 * layer, cfg.hidden_dropout_prob, train, plan)
    return layer_norm_tf(mixed + x, P[prefix + "LayerNorm.weight"], P[prefix + "LayerNorm.bias"])


def attention_layer(P, prefix, x, mask, cfg: OracleConfig, layer: int, train=False, plan=None):
    """``MultiHeadAttention.forward`` (``_modules.py:108-140``); its LayerNorm is
    ``nn.LayerNorm(eps=1e-12)`` (``_modules.py:97``)."""
    B, L, d = x.shape
    H = cfg.num_attention_heads
    dh = d // H

    def heads(t):
        return t.view(B, L, H, dh).permute(0, 2, 1, 3)

    q = heads(F.linear(x, P[prefix + "query.weight"], P[prefix + "query.bias"]))
    k = heads(F.linear(x, P[prefix + "key.weight"], P[prefix + "key.bias"]))
    v = heads(F.linear(x, P[prefix + "value.weight"], P[prefix + "value.bias"]))
    scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(dh)
    scores = scores + mask
    probs = torch.softmax(scores, dim=-1)
    probs = _drop(probs, 2 + 4 * layer, cfg.attention_probs_dropout_prob, train, plan)
    ctx = torch.matmul(probs, v).permute(0, 2, 1, 3).contiguous().view(B, L, d)
    out = F.linear(ctx, P[prefix + "dense.weight"], P[prefix + "dense.bias"])
    out = _drop(out, 3 + 4 * layer, cfg.hidden_dropout_prob, train, plan)
    return F.layer_norm(out + x, (d,), P[prefix + "LayerNorm.weight"], P[prefix + "LayerNorm.bias"], eps=1e-12)


def gelu_erf(x):
    """``FeedForward.gelu`` (``_modules.py:47-56``)."""
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def feed_forward(P, prefix, y, cfg: OracleConfig, layer: int, train=False, plan=None):
    """``FeedForward.forward`` (``_modules.py:61-69``)."""
    h = gelu_erf(F.linear(y, P[prefix + "dense_1.weight"], P[prefix + "dense_1.bias"]))
    h = F.linear(h, P[prefix + "dense_2.weight"], P[prefix + "dense_2.bias"])
    h = _drop(h, 4 + 4 * layer, cfg.hidden_dropout_prob, train, plan)
    return layer_norm_tf(h + y, P[prefix + "LayerNorm.weight"], P[prefix + "LayerNorm.bias"])


def encode(P, ids, cfg: OracleConfig, train=False, plan=None, use_fft=True) -> List[torch.Tensor]:
    """``BSARecModel.forward(all_sequence_output=True)`` (``bsarec.py:16-28,46-54,62-80``):
    list of P+1 tensors [B,L,d] (embedding output, then each block's output)."""
    dtype = P["item_embeddings.weight"].dtype
    mask = attention_mask(ids, dtype)
    x = embed(P, ids, cfg, train, plan)
    outs = [x]
    for l in range(cfg.num_hidden_layers):
        pre = f"item_encoder.blocks.{l}."
        dsp = frequency_layer(P, pre + "layer.filter_layer.", x, cfg, l, train, plan, use_fft)
        gsp = attention_layer(P, pre + "layer.attention_layer.", x, mask, cfg, l, train, plan)
        y = cfg.alpha * dsp + (1 - cfg.alpha) * gsp
        x = feed_forward(P, pre + "feed_forward.", y, cfg, l, train, plan)
        outs.append(x)
    return outs


def calculate_loss(P, ids, answers, cfg: OracleConfig, train=True, plan=None, use_fft=True):
    """``BSARecModel.calculate_loss`` (``bsarec.py:30-37``): mean cross-entropy of the
    last position's logits against all item_size columns (column 0 included)."""
    s = encode(P, ids, cfg, train, plan, use_fft)[-1][:, -1, :]
    logits = torch.matmul(s, P["item_embeddings.weight"].transpose(0, 1))
    return F.cross_entropy(logits, answers)


def loss_and_grads(P, ids, answers, cfg: OracleConfig, train=True, plan=None, use_fft=True):
    """loss (python float-able tensor) and ``{name: grad}`` for all parameters."""
    Q = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    loss = calculate_loss(Q, ids, answers, cfg, train, plan, use_fft)
    loss.backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in Q.items()}
    return loss.detach(), grads


# ----------------------------------------------------------------------------- sibling model: SASRec

def encode_sasrec(P, ids, cfg: OracleConfig, train=False, plan=None) -> List[torch.Tensor]:
    """``SASRecModel.forward(all_sequence_output=True)`` (``sasrec.py:27-40``, ``_modules.py:142-172``): the block is
    MultiHeadAttention -> FeedForward; parameter names are the reference's (``...blocks.<l>.layer.query.weight``)."""
    dtype = P["item_embeddings.weight"].dtype
    mask = attention_mask(ids, dtype)
    x = embed(P, ids, cfg, train, plan)
    outs = [x]
    for l in range(cfg.num_hidden_layers):
        pre = f"item_encoder.blocks.{l}."
        y = attention_layer(P, pre + "layer.", x, mask, cfg, l, train, plan)
        x = feed_forward(P, pre + "feed_forward.", y, cfg, l, train, plan)
        outs.append(x)
    return outs


def sasrec_loss(P, ids, pos, neg, cfg: OracleConfig, train=True, plan=None):
    """``SASRecModel.calculate_loss`` (``sasrec.py:42-63``): BCE-with-logits of the positive (label 1) and negative
    (label 0) item against the last position, each averaged over the users whose positive id is not padding."""
    s = encode_sasrec(P, ids, cfg, train, plan)[-1][:, -1, :]
    E = P["item_embeddings.weight"]
    pos_logits = (F.embedding(pos, E, padding_idx=0) * s).sum(-1)
    neg_logits = (F.embedding(neg, E, padding_idx=0) * s).sum(-1)
    idx = (pos != 0).nonzero().reshape(-1)
    loss = F.binary_cross_entropy_with_logits(pos_logits[idx], torch.ones_like(pos_logits[idx]))
    return loss + F.binary_cross_entropy_with_logits(neg_logits[idx], torch.zeros_like(neg_logits[idx]))


def sasrec_loss_and_grads(P, ids, pos, neg, cfg: OracleConfig, train=True, plan=None):
    Q = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    loss = sasrec_loss(Q, ids, pos, neg, cfg, train, plan)
    loss.backward()
    return loss.detach(), {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in Q.items()}


def info_nce(z_i: torch.Tensor, z_j: torch.Tensor, tau: float) -> torch.Tensor:
    """``DuoRecModel.info_nce`` + ``aug_nce_fct`` (``duorec.py:47-75``, sim='dot') on last-position vectors [B,d]: the
    positive of row r is its pair in the other view, the negatives every other row of the 2B but itself."""
    B = z_i.shape[0]
    n2 = 2 * B
    z = torch.cat((z_i, z_j), dim=0)
    sim = torch.mm(z, z.T) / tau
    pos = torch.cat((torch.diag(sim, B), torch.diag(sim, -B)), dim=0).reshape(n2, 1)
    mask = torch.ones((n2, n2), dtype=torch.bool)
    mask.fill_diagonal_(False)
    for i in range(B):
        mask[i, B + i] = False
        mask[B + i, i] = False
    neg = sim[mask].reshape(n2, -1)
    logits = torch.cat((pos, neg), dim=1)
    return F.cross_entropy(logits, torch.zeros(n2, dtype=torch.long))


def duorec_loss(P, ids, answers, sem_ids, cfg: OracleConfig, tau: float, lmd_sem: float, train=True, plans=(None, None, None)):
    """``DuoRecModel.calculate_loss`` (``duorec.py:91-139``) for ssl='us_x' (``sem_ids`` given) or the CE-only setting
    (``sem_ids`` None): CE of the sequence's last position over the catalog + lmd_sem * InfoNCE(second dropout view of the
    sequence, the same-target sequence).  ``plans``: one DropoutPlan per encoder pass."""
    s = encode_sasrec(P, ids, cfg, train, plans[0])[-1][:, -1, :]
    loss = F.cross_entropy(torch.matmul(s, P["item_embeddings.weight"].transpose(0, 1)), answers)
    if sem_ids is not None:
        aug = encode_sasrec(P, ids, cfg, train, plans[1])[-1][:, -1, :]
        sem = encode_sasrec(P, sem_ids, cfg, train, plans[2])[-1][:, -1, :]
        loss = loss + lmd_sem * info_nce(aug, sem, tau)
    return loss


def duorec_loss_and_grads(P, ids, answers, sem_ids, cfg, tau, lmd_sem, train=True, plans=(None, None, None)):
    Q = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    loss = duorec_loss(Q, ids, answers, sem_ids, cfg, tau, lmd_sem, train, plans)
    loss.backward()
    return loss.detach(), {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in Q.items()}


def encode_fmlprec(P, ids, cfg: OracleConfig, train=False, plan=None) -> List[torch.Tensor]:
    """``FMLPRecModel.forward(all_sequence_output=True)`` (``fmlprec.py:27-38,62-118``): per block the learned complex
    filter ``irfft(rfft(x, 'ortho') * W, 'ortho')`` + dropout + residual + LayerNorm, then the feed-forward.  Dropout sites
    as the kernels number them: 1 + 4 l for the filter layer, 4 + 4 l for the feed-forward."""
    x = embed(P, ids, cfg, train, plan)
    outs = [x]
    L = ids.shape[1]
    for l in range(cfg.num_hidden_layers):
        pre = f"item_encoder.blocks.{l}."
        w = torch.view_as_complex(P[pre + "layer.complex_weight"].contiguous())
        f = torch.fft.irfft(torch.fft.rfft(x, dim=1, norm="ortho") * w, n=L, dim=1, norm="ortho")
        h = _drop(f, 1 + 4 * l, cfg.hidden_dropout_prob, train, plan)
        y = layer_norm_tf(h + x, P[pre + "layer.LayerNorm.weight"], P[pre + "layer.LayerNorm.bias"])
        x = feed_forward(P, pre + "feed_forward.", y, cfg, l, train, plan)
        outs.append(x)
    return outs


def fmlprec_loss(P, ids, pos, neg, cfg: OracleConfig, train=True, plan=None):
    """``FMLPRecModel.calculate_loss`` (``fmlprec.py:40-60``)."""
    s = encode_fmlprec(P, ids, cfg, train, plan)[-1][:, -1, :]
    E = P["item_embeddings.weight"]
    pos_logits = (F.embedding(pos, E, padding_idx=0) * s).sum(-1)
    neg_logits = (F.embedding(neg, E, padding_idx=0) * s).sum(-1)
    return torch.mean(-torch.log(torch.sigmoid(pos_logits) + 1e-24) - torch.log(1 - torch.sigmoid(neg_logits) + 1e-24))


def fmlprec_loss_and_grads(P, ids, pos, neg, cfg: OracleConfig, train=True, plan=None):
    Q = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    loss = fmlprec_loss(Q, ids, pos, neg, cfg, train, plan)
    loss.backward()
    return loss.detach(), {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in Q.items()}


# ----------------------------------------------------------------------------- evaluation

def predict_full(seq_out: torch.Tensor, item_emb: torch.Tensor) -> torch.Tensor:
    """``Trainer.predict_full`` (``trainers.py:62-68``)."""
    return torch.matmul(seq_out, item_emb.transpose(0, 1))


def masked_scores(scores: np.ndarray, hist_indptr, hist_indices, rows) -> np.ndarray:
    """History items get score exactly 0.0, not -inf (``trainers.py:134``)."""
    out = np.array(scores, copy=True)
    for r, u in enumerate(rows):
        out[r, hist_indices[hist_indptr[u]:hist_indptr[u + 1]]] = 0.0
    return out


def full_sort_topk(scores: np.ndarray, k: int = 20) -> np.ndarray:
    """Top-k ids, descending score, via argpartition + argsort as ``trainers.py:142-149``."""
    n = len(scores)
    ind = np.argpartition(scores, -k)[:, -k:]
    sub = scores[np.arange(n)[:, None], ind]
    order = np.argsort(sub)[np.arange(n), ::-1]
    return ind[np.arange(n)[:, None], order]


def recall_at_k(actual: Sequence[int], predicted, topk: int) -> float:
    """``metrics.py:3-13`` with one ground-truth item per user."""
    hits = sum(1.0 for a, p in zip(actual, predicted) if a in set(p[:topk]))
    return hits / len(predicted)


def ndcg_k(actual: Sequence[int], predicted, topk: int) -> float:
    """``metrics.py:15-23``; idcg(1) == 1."""
    res = 0.0
    for a, p in zip(actual, predicted):
        res += sum(1.0 / math.log(j + 2, 2) for j in range(topk) if p[j] == a)
    return res / float(len(actual))


def full_sort_metrics(answers, pred_list) -> List[float]:
    """[HR@5, NDCG@5, HR@10, NDCG@10, HR@20, NDCG@20] (``trainers.py:70-83``)."""
    out = []
    for k in (5, 10, 20):
        out += [recall_at_k(answers, pred_list, k), ndcg_k(answers, pred_list, k)]
    return out


def eval_step(P, ids, cfg: OracleConfig, hist_indptr, hist_indices, rows, k: int = 20) -> np.ndarray:
    """One eval batch of ``Trainer.iteration(train=False)`` (``trainers.py:123-149``)."""
    with torch.no_grad():
        s = encode(P, ids, cfg, train=False)[-1][:, -1, :]
        scores = predict_full(s, P["item_embeddings.weight"]).cpu().numpy().copy()      # trainers.py:129: D2H of [B,N]
    return full_sort_topk(masked_scores(scores, hist_indptr, hist_indices, rows), k)


# ----------------------------------------------------------------------------- CPU baseline step

class adam_train_step:
    """The timed CPU baseline: ``calculate_loss -> zero_grad -> backward -> Adam.step
    -> loss.item()`` (``trainers.py:103-108``), torch dropout, all host threads."""

    def __init__(self, P, cfg: OracleConfig, lr=1e-3, betas=(0.9, 0.999), weight_decay=0.0):
        self.cfg = cfg
        self.P = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
        self.optim = torch.optim.Adam(list(self.P.values()), lr=lr, betas=betas, weight_decay=weight_decay)

    def __call__(self, ids, answers) -> float:
        loss = calculate_loss(self.P, ids, answers, self.cfg, train=True, plan=None)
        self.optim.zero_grad()
        loss.backward()
        self.optim.step()
        return loss.item()

"""Drop-in ``BSARecModel``: the reference's class contract over the sm_100a kernels.

Same constructor (``Cls(args=args)``), same ``calculate_loss`` / ``predict`` / ``forward`` signatures, same
``nn.Module`` parameter tree — hence the same 42 ``state_dict`` keys and shapes, the same ``parameters()``
order and, under the same ``torch.manual_seed``, bit-identical initial weights — as
``src/model/bsarec.py:7-37`` + ``src/model/_abstract_model.py:6-75``.
The modules below are parameter containers only: no torch op of theirs runs on the hot path.
"""
from __future__ import annotations

import copy
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from . import functional as Fn


class TFLayerNorm(nn.Module):
    """Container for the reference's custom LayerNorm (weight, bias; eps=1e-12 inside the sqrt)."""

    def __init__(self, hidden_size: int, eps: float = 1e-12):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.bias = nn.Parameter(torch.zeros(hidden_size))
        self.variance_epsilon = eps


class _FilterLayer(nn.Module):          # bsarec.py:82-88
    def __init__(self, args):
        super().__init__()
        self.out_dropout = nn.Dropout(args.hidden_dropout_prob)
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)
        self.c = args.c // 2 + 1
        self.sqrt_beta = nn.Parameter(torch.randn(1, 1, args.hidden_size))


class _AttentionLayer(nn.Module):       # _modules.py:76-98
    def __init__(self, args):
        super().__init__()
        if args.hidden_size % args.num_attention_heads != 0:
            raise ValueError("The hidden size (%d) is not a multiple of the number of attention heads (%d)"
                             % (args.hidden_size, args.num_attention_heads))
        self.num_attention_heads = args.num_attention_heads
        self.attention_head_size = args.hidden_size // args.num_attention_heads
        self.query = nn.Linear(args.hidden_size, args.hidden_size)
        self.key = nn.Linear(args.hidden_size, args.hidden_size)
        self.value = nn.Linear(args.hidden_size, args.hidden_size)
        self.attn_dropout = nn.Dropout(args.attention_probs_dropout_prob)
        self.dense = nn.Linear(args.hidden_size, args.hidden_size)
        self.LayerNorm = nn.LayerNorm(args.hidden_size, eps=1e-12)
        self.out_dropout = nn.Dropout(args.hidden_dropout_prob)


class _FeedForward(nn.Module):          # _modules.py:22-35
    def __init__(self, args):
        super().__init__()
        if args.hidden_act != "gelu":
            raise ValueError(f"bsarec_b200 implements hidden_act='gelu' only (got {args.hidden_act!r})")
        self.dense_1 = nn.Linear(args.hidden_size, 4 * args.hidden_size)
        self.dense_2 = nn.Linear(4 * args.hidden_size, args.hidden_size)
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(args.hidden_dropout_prob)


class _Layer(nn.Module):                # bsarec.py:67-73
    def __init__(self, args):
        super().__init__()
        self.filter_layer = _FilterLayer(args)
        self.attention_layer = _AttentionLayer(args)
        self.alpha = args.alpha


class _Block(nn.Module):                # bsarec.py:56-60
    def __init__(self, args):
        super().__init__()
        self.layer = _Layer(args)
        self.feed_forward = _FeedForward(args)


class _Encoder(nn.Module):              # bsarec.py:39-44: one block built, deep-copied P times
    def __init__(self, args):
        super().__init__()
        block = _Block(args)
        self.blocks = nn.ModuleList([copy.deepcopy(block) for _ in range(args.num_hidden_layers)])


class _TransformerBlock(nn.Module):     # _modules.py:142-151 (SASRec: the attention module IS `layer`)
    def __init__(self, args):
        super().__init__()
        self.layer = _AttentionLayer(args)
        self.feed_forward = _FeedForward(args)


class _TransformerEncoder(nn.Module):   # _modules.py:153-158
    def __init__(self, args):
        super().__init__()
        block = _TransformerBlock(args)
        self.blocks = nn.ModuleList([copy.deepcopy(block) for _ in range(args.num_hidden_layers)])


class BSARecModel(nn.Module):
    def __init__(self, args):
        super().__init__()
        self.args = args
        self.item_embeddings = nn.Embedding(args.item_size, args.hidden_size, padding_idx=0)
        self.position_embeddings = nn.Embedding(args.max_seq_length, args.hidden_size)
        self.batch_size = args.batch_size
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(args.hidden_dropout_prob)
        self.item_encoder = _Encoder(args)
        self.apply(self.init_weights)
        self._init_runtime()

    # ------------------------------------------------------------------ reference-compatible surface
    def init_weights(self, module):
        """N(0, initializer_range^2) for Linear/Embedding weights (padding row included), zero Linear
        biases, custom LayerNorm -> (1, 0); sqrt_beta and nn.LayerNorm untouched (_abstract_model.py:26-39)."""
        if isinstance(module, (nn.Linear, nn.Embedding)):
            module.weight.data.normal_(mean=0.0, std=self.args.initializer_range)
        elif isinstance(module, TFLayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)
        if isinstance(module, nn.Linear) and module.bias is not None:
            module.bias.data.zero_()

    def forward(self, input_ids, user_ids=None, all_sequence_output=False):
        """[B,L] ids -> last block output [B,L,d] (or the list of P+1 layer outputs).  Inference path:
        the result is detached; training goes through ``calculate_loss``."""
        xs, _ = Fn.encoder_forward(self._hyper(), self._tensors(), input_ids, self.training, self._next_seed(),
                                   keep_for_backward=False)
        return xs if all_sequence_output else xs[-1]

    def _init_runtime(self):
        self.ce_impl = 0          # 0 auto | 1 fp32 FFMA | 2 tcgen05 3xTF32   (bsa_ce_desc.impl)
        self.topk_impl = 0
        self._fixed_seed = None
        self._exchange = None     # parallel.GradExchange when data-parallel training is on
        self._topk_cache = {}     # scoring workspace kept while the item table is unchanged (functional.score_topk)

    def predict(self, input_ids, user_ids, all_sequence_output=False):
        return self.forward(input_ids, user_ids, all_sequence_output)

    def calculate_loss(self, input_ids, answers, neg_answers, same_target, user_ids):
        """Mean full-catalog cross-entropy at the last position (bsarec.py:30-37); neg_answers, same_target
        and user_ids are accepted and ignored, as in the reference."""
        names = [n for n, _ in self.named_parameters()]
        tensors = [p for _, p in self.named_parameters()]
        if self._exchange is not None and self._exchange.vocab_parallel and \
                getattr(self.item_embeddings.weight, "_bsa_shard", None) is not None and torch.is_grad_enabled():
            raise RuntimeError("bsarec_b200: the previous step's row-shard gradient of the item table was never applied — "
                               "vocabulary-parallel training needs bsarec_b200.FusedAdam (a stock optimizer sees no .grad "
                               "on the table and would silently leave it untrained)")
        return Fn.FusedLoss.apply(self._hyper(), names, input_ids, answers, self.training, self._next_seed(),
                                  self.ce_impl, self._exchange, *tensors)

    # ------------------------------------------------------------------ data-parallel training (additional API)
    def enable_data_parallel(self, group=None, vocab_parallel: bool = False):
        """One process per GPU, full replica, equal local batches: after this call ``loss.backward()`` leaves the
        GLOBAL-batch-mean gradient in every ``p.grad`` (NCCL exchange overlapped with the backward pass), so the
        reference's loop — ``calculate_loss -> zero_grad -> backward -> step`` (trainers.py:103-107) — runs unchanged.

        ``vocab_parallel``: the cross-entropy is sharded over the item table's rows instead (every rank scores all users
        of the global batch against its rows; ``loss`` is then the GLOBAL-batch mean on every rank).  The item table gets
        no ``.grad``: its row-shard gradient is handed to ``bsarec_b200.FusedAdam`` (required), which updates the shard
        and all-gathers the rows.  Needs ``item_size % world_size == 0`` and the tcgen05 CE path."""
        from .parallel import GradExchange
        self._exchange = GradExchange(group, vocab_parallel=vocab_parallel, item_param=self.item_embeddings.weight)
        if vocab_parallel:
            self._exchange.shard(self.args.item_size)       # raises early when the table does not divide
        return self

    def disable_data_parallel(self):
        self._exchange = None

    # ------------------------------------------------------------------ fused evaluation (additional API)
    @torch.no_grad()
    def last_hidden(self, input_ids) -> torch.Tensor:
        """[B,d] = predict(input_ids)[:, -1, :] without building the full output list on the host side."""
        xs, _ = Fn.encoder_forward(self._hyper(), self._tensors(), input_ids, False, 0, keep_for_backward=False,
                                   last_only=True)
        return xs[-1]

    @torch.no_grad()
    def full_sort_topk(self, input_ids, user_ids=None, hist_indptr=None, hist_indices=None, k: int = 20,
                       hist_rows=None):
        """Top-k item ids per user, descending score, history items scored 0.0 — the eval branch of
        ``Trainer.iteration`` (trainers.py:126-149) kept on the device.  ``hist_indptr`` / ``hist_indices`` are a CSR
        (int64, CUDA) over the batch rows, or — with ``hist_rows`` int64 [B] — over all users, row ``hist_rows[b]``
        belonging to batch row b (the trainer's rating matrix and ``user_ids``).  Returns int64 [B,k]."""
        s = self.last_hidden(input_ids)
        _, idx = Fn.score_topk(s, self.item_embeddings.weight.detach(), hist_indptr, hist_indices, k, 0, self.topk_impl,
                               cache=self._topk_cache, hist_rows=hist_rows)
        return idx

    def invalidate_eval_cache(self):
        """Forget the split copy of the item table kept for evaluation sweeps.  Called by ``train()`` and
        ``load_state_dict``; call it yourself after writing the table through ``.data`` or a raw pointer."""
        getattr(self, "_topk_cache", {}).clear()

    def train(self, mode: bool = True):
        if mode:
            self.invalidate_eval_cache()
        else:
            self.sync_item_table()           # evaluation reads the whole table: bring the replica up to date first
        return super().train(mode)

    def sync_item_table(self):
        """Vocabulary-parallel training leaves each rank's replica of the item table fresh only in its own row shard and
        the rows its batches read; this all-gathers the shards (a collective: every rank must call it).  ``eval()`` and
        ``state_dict()`` do."""
        ex = getattr(self, "_exchange", None)
        if ex is not None and ex.vocab_parallel and ex.stale:
            ex.sync_table()
            torch.autograd.graph.increment_version(self.item_embeddings.weight)
            self.invalidate_eval_cache()

    def state_dict(self, *a, **kw):
        self.sync_item_table()
        return super().state_dict(*a, **kw)

    def load_state_dict(self, *a, **kw):
        self.invalidate_eval_cache()
        ex = getattr(self, "_exchange", None)
        if ex is not None:
            ex.stale = False
        return super().load_state_dict(*a, **kw)

    # ------------------------------------------------------------------ helpers
    def set_dropout_seed(self, seed: Optional[int]):
        """Pin the per-call dropout seed (tests); ``None`` draws one per call from torch's CPU generator."""
        self._fixed_seed = seed

    def _next_seed(self) -> int:
        if self._fixed_seed is not None:
            return int(self._fixed_seed)
        if not self.training:
            return 0
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        if self._exchange is not None:          # identically seeded ranks must not share dropout masks across the global batch
            seed = (seed ^ ((self._exchange.rank + 1) * 0x9E3779B97F4A7C15)) & (2 ** 62 - 1)
        return seed

    def _hyper(self) -> Fn.Hyper:
        a = self.args
        return Fn.Hyper(item_size=a.item_size, hidden_size=a.hidden_size, max_seq_length=a.max_seq_length,
                        num_attention_heads=a.num_attention_heads, num_hidden_layers=a.num_hidden_layers,
                        kept_bins=a.c // 2 + 1, alpha=float(a.alpha), p_hidden=float(a.hidden_dropout_prob),
                        p_attn=float(a.attention_probs_dropout_prob))

    def _tensors(self) -> Dict[str, torch.Tensor]:
        return {n: p.detach() for n, p in self.named_parameters()}


class SASRecModel(BSARecModel):
    """Sibling model on the same kernels: ``SASRecModel`` of ``src/model/sasrec.py:18-63``.

    Same class contract and the reference's own ``state_dict`` keys (``item_encoder.blocks.<l>.layer.{query,key,value,
    dense,LayerNorm}.*`` — the attention module IS ``layer``, ``_modules.py:142-146`` — and ``feed_forward.*``), same
    seeded init.  The encoder is the BSARec block with the filter branch blended out: ``alpha = 0`` makes
    ``alpha * dsp + (1 - alpha) * gsp`` the attention output bit for bit (``0 * finite = 0``, ``1 * x = x``), the filter
    parameters the kernels still take are constant buffers that never reach the ``state_dict``.  The loss is the pairwise
    BCE at the last position (``sasrec.py:42-63``) through ``bsa_bce_pair_fwd/bwd``."""

    def __init__(self, args):
        nn.Module.__init__(self)
        self.args = args
        self.item_embeddings = nn.Embedding(args.item_size, args.hidden_size, padding_idx=0)
        self.position_embeddings = nn.Embedding(args.max_seq_length, args.hidden_size)
        self.batch_size = args.batch_size
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(args.hidden_dropout_prob)
        self.item_encoder = _TransformerEncoder(args)
        self.apply(self.init_weights)
        d = args.hidden_size
        self.register_buffer("_f_sqrt_beta", torch.zeros(1, 1, d), persistent=False)
        self.register_buffer("_f_ln_w", torch.ones(d), persistent=False)
        self.register_buffer("_f_ln_b", torch.zeros(d), persistent=False)
        self._init_runtime()

    def _hyper(self) -> Fn.Hyper:
        a = self.args
        return Fn.Hyper(item_size=a.item_size, hidden_size=a.hidden_size, max_seq_length=a.max_seq_length,
                        num_attention_heads=a.num_attention_heads, num_hidden_layers=a.num_hidden_layers,
                        kept_bins=1, alpha=0.0, p_hidden=float(a.hidden_dropout_prob),
                        p_attn=float(a.attention_probs_dropout_prob))

    def _kernel_named(self, detach: bool):
        """[(name the kernels' host code expects (BSARec layout), tensor)] — parameters first-class, filter dummies last."""
        out = []
        for n, p in self.named_parameters():
            if ".layer." in n:
                n = n.replace(".layer.", ".layer.attention_layer.")
            out.append((n, p.detach() if detach else p))
        for l in range(self.args.num_hidden_layers):
            pre = f"item_encoder.blocks.{l}.layer.filter_layer."
            out += [(pre + "sqrt_beta", self._f_sqrt_beta), (pre + "LayerNorm.weight", self._f_ln_w),
                    (pre + "LayerNorm.bias", self._f_ln_b)]
        return out

    def _tensors(self) -> Dict[str, torch.Tensor]:
        return dict(self._kernel_named(detach=True))

    def calculate_loss(self, input_ids, answers, neg_answers, same_target, user_ids):
        """BCE of the positive / negative item against the last position (sasrec.py:42-63); ``same_target`` and
        ``user_ids`` are accepted and ignored, as in the reference."""
        if self._exchange is not None:
            raise RuntimeError("bsarec_b200.SASRecModel: the overlapped data-parallel exchange is BSARec-only; use "
                               "parallel.allreduce_gradients after backward")
        named = self._kernel_named(detach=False)
        return Fn.FusedPairLoss.apply(self._hyper(), [n for n, _ in named], input_ids, answers, neg_answers, self.training,
                                      self._next_seed(), *[t for _, t in named])


class DuoRecModel(SASRecModel):
    """Sibling model on the same kernels: ``DuoRecModel`` of ``src/model/duorec.py:16-139``.

    SASRec's encoder (the reference's ``TransformerEncoder``, same ``state_dict`` keys), full-catalog cross-entropy at the
    last position through the fused CE kernels (``duorec.py:95-98``) and, for ``ssl = 'us_x'`` (the reference's default,
    ``utils.py:109``), the in-batch InfoNCE between a second dropout view of the sequence and the same-target sequence
    (``duorec.py:124-137``, ``sim = 'dot'``) through ``bsa_info_nce_fwd/bwd``, weight ``lmd_sem``.  The other ``ssl``
    modes of the reference stack a ``[B, d]`` and a ``[B, L, d]`` tensor (``duorec.py:54-55`` after ``:93``) and cannot run
    there either; any value outside {'us', 'un', 'su', 'us_x'} trains on the cross-entropy alone, as in the reference."""

    def __init__(self, args):
        # the reference builds both embedding tables twice (SequentialRecModel.__init__, then duorec.py:20-21): the seeded
        # RNG stream is consumed accordingly before init_weights overwrites everything
        nn.Embedding(args.item_size, args.hidden_size, padding_idx=0)
        nn.Embedding(args.max_seq_length, args.hidden_size)
        super().__init__(args)
        self.tau, self.ssl, self.sim = float(args.tau), str(args.ssl), str(args.sim)
        self.lmd_sem, self.lmd = float(args.lmd_sem), float(args.lmd)
        if self.ssl in ("us", "un", "su"):
            raise ValueError(f"ssl={self.ssl!r}: this mode of the reference concatenates [B,d] with [B,L,d] "
                             "(duorec.py:54-55) and cannot run there; use 'us_x' or a CE-only setting")
        if self.ssl == "us_x" and self.sim != "dot":
            raise ValueError("bsarec_b200.DuoRecModel implements sim='dot' (the reference's default) only")

    def calculate_loss(self, input_ids, answers, neg_answers, same_target, user_ids):
        if self._exchange is not None:
            raise RuntimeError("bsarec_b200.DuoRecModel: use parallel.allreduce_gradients after backward for data parallel")
        named = self._kernel_named(detach=False)
        base = self._next_seed()
        seeds = (base, base + 1, base + 2) if self._fixed_seed is not None else (base, self._next_seed(), self._next_seed())
        sem = same_target if self.ssl == "us_x" else None
        return Fn.FusedDuoLoss.apply(self._hyper(), [n for n, _ in named], input_ids, answers, sem, self.training, seeds,
                                     self.ce_impl, self.tau, self.lmd_sem, *[t for _, t in named])


class _FmlpLayer(nn.Module):            # fmlprec.py:95-100
    def __init__(self, args):
        super().__init__()
        self.complex_weight = nn.Parameter(torch.randn(1, args.max_seq_length // 2 + 1, args.hidden_size, 2,
                                                       dtype=torch.float32) * 0.02)
        self.out_dropout = nn.Dropout(args.hidden_dropout_prob)
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)


class _FmlpBlock(nn.Module):            # fmlprec.py:84-88
    def __init__(self, args):
        super().__init__()
        self.layer = _FmlpLayer(args)
        self.feed_forward = _FeedForward(args)


class _FmlpEncoder(nn.Module):          # fmlprec.py:62-68
    def __init__(self, args):
        super().__init__()
        block = _FmlpBlock(args)
        self.blocks = nn.ModuleList([copy.deepcopy(block) for _ in range(args.num_hidden_layers)])


class FMLPRecModel(BSARecModel):
    """Sibling model on the same kernels: ``FMLPRecModel`` of ``src/model/fmlprec.py:17-118``.

    The reference's ``state_dict`` keys (``item_encoder.blocks.<l>.layer.{complex_weight, LayerNorm.*}``,
    ``feed_forward.*``) and seeded init.  The filter layer ``irfft(rfft(x) * W)`` with its learned complex weight is run as
    what it is exactly — a per-channel circular convolution along the sequence with taps ``irfft(W)`` (``csrc/fmlp.cu``) —
    inside the block kernels' ``kind = 1`` mode (filter + feed-forward, no attention); the loss is the pairwise
    log-sigmoid of ``fmlprec.py:40-60`` (``bsa_bce_pair_*`` variant 1).  Sequences must have ``max_seq_length`` positions,
    as in the reference (the weight has ``max_seq_length // 2 + 1`` bins)."""

    def __init__(self, args):
        nn.Module.__init__(self)
        self.args = args
        self.item_embeddings = nn.Embedding(args.item_size, args.hidden_size, padding_idx=0)
        self.position_embeddings = nn.Embedding(args.max_seq_length, args.hidden_size)
        self.batch_size = args.batch_size
        self.LayerNorm = TFLayerNorm(args.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(args.hidden_dropout_prob)
        self.item_encoder = _FmlpEncoder(args)
        self.apply(self.init_weights)
        self.register_buffer("_unused", torch.zeros(4), persistent=False)      # stands in for the attention members
        self._init_runtime()

    def _hyper(self) -> Fn.Hyper:
        a = self.args
        return Fn.Hyper(item_size=a.item_size, hidden_size=a.hidden_size, max_seq_length=a.max_seq_length,
                        num_attention_heads=1, num_hidden_layers=a.num_hidden_layers, kept_bins=1, alpha=1.0,
                        p_hidden=float(a.hidden_dropout_prob), p_attn=0.0, kind=1)

    def _kernel_named(self, detach: bool):
        out = []
        for n, p in self.named_parameters():
            if ".layer." in n:
                n = n.replace(".layer.", ".layer.filter_layer.")
            out.append((n, p.detach() if detach else p))
        for l in range(self.args.num_hidden_layers):
            pre = f"item_encoder.blocks.{l}.layer.attention_layer."
            for lin in ("query", "key", "value", "dense"):
                out += [(pre + lin + ".weight", self._unused), (pre + lin + ".bias", self._unused)]
            out += [(pre + "LayerNorm.weight", self._unused), (pre + "LayerNorm.bias", self._unused)]
        return out

    def _tensors(self) -> Dict[str, torch.Tensor]:
        params = dict(self._kernel_named(detach=True))
        return Fn.fmlp_kernel_params(self._hyper(), params, self.args.max_seq_length)

    def _check_length(self, input_ids):
        if input_ids.shape[1] != self.args.max_seq_length:
            raise RuntimeError(f"FMLPRec: sequences must have max_seq_length = {self.args.max_seq_length} positions "
                               f"(got {input_ids.shape[1]}): the filter weight has max_seq_length // 2 + 1 bins (fmlprec.py:98,104)")

    def forward(self, input_ids, user_ids=None, all_sequence_output=False):
        self._check_length(input_ids)
        return super().forward(input_ids, user_ids, all_sequence_output)

    def last_hidden(self, input_ids):
        self._check_length(input_ids)
        return super().last_hidden(input_ids)

    def calculate_loss(self, input_ids, answers, neg_answers, same_target, user_ids):
        if self._exchange is not None:
            raise RuntimeError("bsarec_b200.FMLPRecModel: use parallel.allreduce_gradients after backward for data parallel")
        named = self._kernel_named(detach=False)
        return Fn.FusedFmlpLoss.apply(self._hyper(), [n for n, _ in named], input_ids, answers, neg_answers, self.training,
                                      self._next_seed(), *[t for _, t in named])

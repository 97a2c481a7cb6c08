"""Multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on the box, gloo in CPU tests).

Training is data-parallel: every rank holds a full replica, the global batch is split evenly, and one
flat fp32 all-reduce averages the gradients.  Evaluation shards the item table by rows:
user vectors are all-gathered, every rank scores all users against its shard and keeps a local top-k,
the candidates are all-gathered and merged.
"""
from __future__ import annotations

from typing import Callable, Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range [lo, hi) of rank's shard; the first n_rows % world shards get one extra row."""
    base, extra = divmod(n_rows, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_gradients(params: Iterable[torch.nn.Parameter], group=None, bucket_elems: int = 1 << 28) -> int:
    """Average ``p.grad`` over the group in flat buckets (sum / world).  Returns the number of collectives.
    Ranks must hold equally sized local batches so that the mean of local means is the global mean."""
    world = dist.get_world_size(group)
    grads = [p.grad for p in params if p.grad is not None]
    if world == 1 or not grads:
        return 0
    calls, i = 0, 0
    while i < len(grads):
        j, n = i, 0
        while j < len(grads) and (n == 0 or n + grads[j].numel() <= bucket_elems):
            n += grads[j].numel()
            j += 1
        chunk = grads[i:j]
        if len(chunk) == 1:
            flat = chunk[0].view(-1)
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            flat.div_(world)
        else:
            flat = torch.cat([g.reshape(-1) for g in chunk])
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            flat.div_(world)
            off = 0
            for g in chunk:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
        calls += 1
        i = j
    return calls


class GradExchange:
    """Data-parallel gradient averaging woven into ``functional.FusedLoss.backward`` (one instance per model).

    sum_r (dE_ce[r] + scatter(ids[r], dz[r]))  ==  all_reduce(dE_ce) + scatter(all_gather(ids), all_gather(dz)):
      * ``start_dense``: the dense cross-entropy part of the item-table gradient ([N,d], 1.02 GB at C3) is produced first
        in the backward pass; its all-reduce is issued at once on the communicator's stream and runs under the whole
        encoder backward;
      * ``finish``: the 41 small gradients are one flat buffer -> one all-reduce; the sparse gather part is exchanged as
        rows — only the positions that hold an item (padding rows carry no gradient), ids + dz rows padded to the largest
        per-rank count — and scatter-added on every rank in the same (rank, row) order, so all replicas hold
        bit-identical gradients.
    ``scale`` = 1/world is folded into d_loss (no averaging pass over the gradients).  Ranks must hold equally sized local
    batches (the mean of local means is then the global mean of ``CrossEntropyLoss``, bsarec.py:35)."""

    def __init__(self, group=None, vocab_parallel: bool = False, item_param: Optional[torch.nn.Parameter] = None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.scale = 1.0 / self.world
        self.collectives = 0
        # vocabulary-parallel cross-entropy + sharded item-table optimizer step (methods below)
        self.vocab_parallel = bool(vocab_parallel)
        self.item_param = item_param
        # vocabulary-parallel training keeps each rank's replica of the item table fresh only where it is read: the rank's own
        # row shard (updated by its FusedAdam step) and the rows of the current batch, fetched from their owners at the start
        # of the forward pass (fetch_rows).  `stale` says the rest of the replica lags behind: sync_table() all-gathers the
        # shards (BSARecModel.eval() / state_dict() call it).
        self.fetch = True
        self.stale = False

    def _issue(self, fn, *a):
        self.collectives += 1
        return fn(*a, group=self.group, async_op=True)

    def prepare(self, ids: torch.Tensor):
        """Call at the start of the forward pass with this rank's ids int64 [B,L].  Padding positions (id 0) carry no
        item-table gradient (``padding_idx=0``, _abstract_model.py:10), so only the rows of real items are exchanged.
        The number of such rows per rank (max over ranks = rows on the wire) travels to the host asynchronously here, a
        whole forward pass before ``finish`` needs it: no stream synchronisation on the backward's critical path.
        Returns the per-call state (kept on the autograd ctx, so several losses may be in flight before a backward)."""
        st = _ExchangeState()
        flat = ids.reshape(-1)
        pad = flat == 0
        # positions of the real items first, in their original order (stable), without a host round trip
        st.order = torch.argsort(pad.to(torch.int8), stable=True)
        st.flat_ids = flat
        counts = torch.empty(self.world, dtype=torch.int64, device=flat.device)
        mine = (flat.numel() - pad.sum()).reshape(1)
        dist.all_gather_into_tensor(counts, mine, group=self.group)
        self.collectives += 1
        if flat.is_cuda:
            st.cap_host = torch.empty(1, dtype=torch.int64, pin_memory=True)
            st.cap_host.copy_(counts.max().reshape(1), non_blocking=True)
            st.cap_event = torch.cuda.Event()
            st.cap_event.record()
        else:
            st.cap_host, st.cap_event = counts.max().reshape(1), None
        return st

    def start_dense(self, st: "_ExchangeState", g_item: torch.Tensor):
        st.w_dense = self._issue(dist.all_reduce, g_item.view(-1), dist.ReduceOp.SUM)

    def finish(self, st: "_ExchangeState", small_flat: torch.Tensor, dz: torch.Tensor, g_item: torch.Tensor, scatter_fn: Callable,
               gather_fn: Optional[Callable] = None):
        w_small = self._issue(dist.all_reduce, small_flat, dist.ReduceOp.SUM) if small_flat.numel() else None
        if st.cap_event is not None:
            st.cap_event.synchronize()                # recorded during the forward pass: long since complete
        cap = int(st.cap_host[0])                     # rows per rank on the wire (the largest per-rank count)
        if cap > 0:
            idx = st.order[:cap]                      # this rank's real rows, then (if it has fewer) padding rows: id 0, skipped
            ids_c = st.flat_ids[idx]
            dz_c = gather_fn(dz, idx) if gather_fn is not None else torch.index_select(dz, 0, idx)   # (gloo CPU tests: torch)
            ids_all = torch.empty(self.world * cap, dtype=torch.int64, device=dz.device)
            dz_all = torch.empty(self.world * cap, dz.shape[1], dtype=dz.dtype, device=dz.device)
            w_ids = self._issue(dist.all_gather_into_tensor, ids_all, ids_c)
            w_rows = self._issue(dist.all_gather_into_tensor, dz_all, dz_c)
            w_ids.wait()
            w_rows.wait()
        st.w_dense.wait()
        if cap > 0:
            scatter_fn(ids_all, dz_all, g_item)       # (rank, row) order, the same on every rank -> identical bits
        if w_small is not None:
            w_small.wait()


    # ------------------------------------------------------------------ vocabulary-parallel cross-entropy
    # Every rank scores ALL users of the global batch against ITS contiguous row shard of the item table:
    #   forward : all-gather the [B,d] user vectors and the answers; per-shard (log-sum-exp, target logit) partials are
    #             all-gathered ([world, 2, B_global] floats) and combined -> the GLOBAL-batch mean loss, identical on every rank;
    #   backward: each rank gets the COMPLETE gradient of its rows (dense CE part from its own kernel, gather part routed to
    #             it row by row with one all-to-all) and a partial ds for all users, which a reduce-scatter sums and hands
    #             back to the users' owners; the 41 small gradients are one all-reduce (sum: the loss is already the global
    #             mean).  No [N,d] all-reduce, no replicated Adam pass over the table: FusedAdam updates the shard and the
    #             updated rows are all-gathered (ShardedItemGrad).  Needs item_size % world == 0.
    def shard(self, n_rows: int) -> Tuple[int, int]:
        if n_rows % self.world:
            raise RuntimeError(f"vocabulary-parallel CE needs item_size ({n_rows}) divisible by the world size ({self.world})")
        per = n_rows // self.world
        return self.rank * per, (self.rank + 1) * per

    def gather_users(self, s: torch.Tensor, answers: torch.Tensor, n_rows: int):
        """-> (s_all [world*B, d], answers_local [world*B]: answer - shard offset, -1 when it lies in another shard)."""
        B = s.shape[0]
        s_all = torch.empty(self.world * B, s.shape[1], dtype=s.dtype, device=s.device)
        ans_all = torch.empty(self.world * B, dtype=answers.dtype, device=answers.device)
        dist.all_gather_into_tensor(s_all, s.contiguous(), group=self.group)
        dist.all_gather_into_tensor(ans_all, answers.contiguous(), group=self.group)
        self.collectives += 2
        lo, hi = self.shard(n_rows)
        inside = (ans_all >= lo) & (ans_all < hi)
        return s_all, torch.where(inside, ans_all - lo, torch.full_like(ans_all, -1))

    def combine_loss(self, lse_part: torch.Tensor, tgt_part: torch.Tensor):
        """Per-shard partials [B_global] -> (global-batch mean loss, global lse [B_global])."""
        mine = torch.stack([lse_part, tgt_part])                          # [2, Bg]
        parts = torch.empty(self.world, 2, lse_part.numel(), dtype=mine.dtype, device=mine.device)
        dist.all_gather_into_tensor(parts.view(self.world * 2, -1), mine, group=self.group)
        self.collectives += 1
        lse = torch.logsumexp(parts[:, 0], dim=0)
        tgt = parts[:, 1].sum(dim=0)
        return (lse - tgt).mean(), lse.contiguous()

    def reduce_scatter_users(self, ds_all: torch.Tensor) -> torch.Tensor:
        """[world*B, d] partial (this shard's contribution for every user) -> [B, d] summed over shards for MY users."""
        B = ds_all.shape[0] // self.world
        out = torch.empty(B, ds_all.shape[1], dtype=ds_all.dtype, device=ds_all.device)
        self.collectives += 1
        if ds_all.is_cuda:
            dist.reduce_scatter_tensor(out, ds_all.contiguous(), op=dist.ReduceOp.SUM, group=self.group)
        else:                                                             # gloo (CPU tests) has no reduce-scatter
            tmp = ds_all.clone()
            dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=self.group)
            out.copy_(tmp[self.rank * B:(self.rank + 1) * B])
        return out

    def prepare_routed(self, ids: torch.Tensor, n_rows: int):
        """Forward-time half of the routed row exchange: rows of real items sorted by owner shard (stable), and the
        [world, world] table of row counts on its way to the host (read in the backward pass: no sync there)."""
        st = _ExchangeState()
        flat = ids.reshape(-1)
        per = n_rows // self.world
        dest = torch.where(flat > 0, flat // per, torch.full_like(flat, self.world))      # padding sorts last, is never sent
        st.order = torch.argsort(dest, stable=True)
        st.flat_ids = flat
        mine = torch.bincount(dest, minlength=self.world + 1)[:self.world].to(torch.int64)
        counts = torch.empty(self.world, self.world, dtype=torch.int64, device=flat.device)
        dist.all_gather_into_tensor(counts.view(-1), mine, group=self.group)
        self.collectives += 1
        if flat.is_cuda:
            st.cap_host = torch.empty(self.world, self.world, dtype=torch.int64, pin_memory=True)
            st.cap_host.copy_(counts, non_blocking=True)
            st.cap_event = torch.cuda.Event()
            st.cap_event.record()
        else:
            st.cap_host, st.cap_event = counts, None
        return st

    def fetch_rows(self, st: "_ExchangeState", table: torch.Tensor, gather_fn: Optional[Callable], scatter_fn: Optional[Callable]):
        """Bring the rows of this batch's items from the ranks that own them into the local replica `table` [N, d] — the
        first thing a vocabulary-parallel forward pass does when the replica is stale.  Requests are the batch's non-zero
        ids sorted by owner (the order prepare_routed made): one all-to-all carries the ids out, the owners gather the rows
        (bsa_gather_rows), one all-to-all carries them back, bsa_scatter_rows writes them into the replica.  Row 0 (the
        padding row, read by every padded position) is broadcast from its owner.  ~L*B*(8 + 4d) bytes per rank instead of
        the whole table after every optimizer step."""
        if st.cap_event is not None:
            st.cap_event.synchronize()
        counts = st.cap_host.tolist()
        send = counts[self.rank]                                   # my requests per owner
        recv = [counts[r][self.rank] for r in range(self.world)]   # requests I serve
        n_send, n_recv = sum(send), sum(recv)
        d = table.shape[1]
        ids_req = st.flat_ids[st.order[:n_send]].contiguous()
        ids_srv = torch.empty(n_recv, dtype=torch.int64, device=table.device)
        dist.all_to_all_single(ids_srv, ids_req, recv, send, group=self.group)
        rows_srv = gather_fn(table, ids_srv) if gather_fn is not None else torch.index_select(table, 0, ids_srv)
        rows_req = torch.empty(n_send, d, dtype=table.dtype, device=table.device)
        dist.all_to_all_single(rows_req.view(-1), rows_srv.view(-1), [c * d for c in send], [c * d for c in recv], group=self.group)
        if scatter_fn is not None:
            scatter_fn(table, ids_req, rows_req)
        else:
            table.index_copy_(0, ids_req, rows_req)
        dist.broadcast(table[0], src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        self.collectives += 3

    def sync_table(self):
        """All-gather the row shards into every replica (after it: stale == False)."""
        if not (self.vocab_parallel and self.stale) or self.item_param is None:
            return
        E = self.item_param.data
        lo, hi = self.shard(E.shape[0])
        dist.all_gather_into_tensor(E.view(-1), E[lo:hi].reshape(-1).clone(), group=self.group)
        self.collectives += 1
        self.stale = False

    def finish_routed(self, st: "_ExchangeState", small_flat: torch.Tensor, dz: torch.Tensor, shard_buf: torch.Tensor, lo: int,
                      scatter_fn: Callable, gather_fn: Optional[Callable] = None):
        """shard_buf [rows + 1, d]: row 0 is a landing row (the scatter skips key 0), row 1 + r is shard row r — already
        holding the dense CE gradient.  Adds the gather-path rows of EVERY rank that fall into this shard."""
        w_small = self._issue(dist.all_reduce, small_flat, dist.ReduceOp.SUM) if small_flat.numel() else None
        if st.cap_event is not None:
            st.cap_event.synchronize()
        counts = st.cap_host.tolist()
        send = counts[self.rank]
        recv = [counts[r][self.rank] for r in range(self.world)]
        n_send, n_recv = sum(send), sum(recv)
        idx = st.order[:n_send]
        ids_s = st.flat_ids[idx]
        rows_s = gather_fn(dz, idx) if gather_fn is not None else torch.index_select(dz, 0, idx)
        ids_r = torch.empty(n_recv, dtype=torch.int64, device=dz.device)
        rows_r = torch.empty(n_recv, dz.shape[1], dtype=dz.dtype, device=dz.device)
        d = dz.shape[1]
        w1 = dist.all_to_all_single(ids_r, ids_s, recv, send, group=self.group, async_op=True)
        w2 = dist.all_to_all_single(rows_r.view(-1), rows_s.view(-1), [c * d for c in recv], [c * d for c in send],
                                    group=self.group, async_op=True)
        self.collectives += 2
        w1.wait()
        w2.wait()
        if n_recv:
            scatter_fn(ids_r - (lo - 1), rows_r, shard_buf)       # (source rank, row) order: fixed -> reproducible bits
        if w_small is not None:
            w_small.wait()

    def publish_shard_grad(self, g_shard: torch.Tensor, lo: int, hi: int):
        if self.item_param is None:
            raise RuntimeError("vocabulary-parallel exchange: item_param not set (use BSARecModel.enable_data_parallel)")
        self.item_param._bsa_shard = ShardedItemGrad(g_shard, lo, hi, self.group, self)


class ShardedItemGrad:
    """Attached to the item-table parameter as ``_bsa_shard`` by a vocabulary-parallel backward pass: the COMPLETE
    gradient of rows [lo, hi) (global-batch mean).  ``FusedAdam.step`` updates those rows and all-gathers them."""
    __slots__ = ("grad", "lo", "hi", "group", "exchange")

    def __init__(self, grad, lo, hi, group, exchange=None):
        self.grad, self.lo, self.hi, self.group, self.exchange = grad, lo, hi, group, exchange


class _ExchangeState:
    """What one forward pass leaves for its backward pass's exchange."""
    __slots__ = ("order", "flat_ids", "cap_host", "cap_event", "w_dense")


def sharded_full_sort_topk(s_local: torch.Tensor, item_shard: torch.Tensor, shard_offset: int,
                           hist_indptr_all: Optional[torch.Tensor], hist_indices_all: Optional[torch.Tensor], k: int,
                           score_fn: Callable, merge_fn: Callable, group=None, return_values: bool = False):
    """Item-sharded evaluation step.

    s_local [B_local, d]: this rank's users' last-position vectors (equal B_local on every rank).
    hist_*_all: CSR history of ALL world*B_local users of the global batch, rank-major order.
    score_fn(s, shard, indptr, indices, k, offset) -> (val [B,k], idx [B,k]); merge_fn(val [G,B,k], idx) -> (val, idx).
    Exchange: one all-gather of the user vectors, then ONE all-to-all per candidate array that routes each user's k
    candidates from every shard to the rank that owns the user (k*(4+8) bytes per user per shard on the wire — an
    all-gather would deliver every rank world times more than it merges).
    Returns int64 [B_local, k] for this rank's users (and the merged scores with ``return_values``)."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    Bl = s_local.shape[0]
    if world == 1:
        val, idx = score_fn(s_local, item_shard, hist_indptr_all, hist_indices_all, k, shard_offset)
        return (idx, val) if return_values else idx
    s_all = torch.empty(world * Bl, s_local.shape[1], dtype=s_local.dtype, device=s_local.device)
    dist.all_gather_into_tensor(s_all, s_local.contiguous(), group=group)
    val, idx = score_fn(s_all, item_shard, hist_indptr_all, hist_indices_all, k, shard_offset)
    # val / idx are [world * Bl, k] in rank-major user order: block r goes to rank r; what arrives is [world, Bl, k],
    # shard-major — this rank's users against every shard
    val_in = torch.empty(world, Bl, k, dtype=val.dtype, device=val.device)
    idx_in = torch.empty(world, Bl, k, dtype=idx.dtype, device=idx.device)
    dist.all_to_all_single(val_in.view(-1), val.contiguous().view(-1), group=group)
    dist.all_to_all_single(idx_in.view(-1), idx.contiguous().view(-1), group=group)
    top_val, top = merge_fn(val_in, idx_in)
    return (top, top_val) if return_values else top

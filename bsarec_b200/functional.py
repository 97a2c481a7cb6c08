"""Host-side orchestration of the C-ABI kernels: encoder forward/backward, fused loss, fused top-k.

PyTorch is used for device memory, streams and autograd plumbing only; every FLOP of the hot path runs in
libbsarec_sm100a.so.  Tensors handed to the library are fp32/int64, contiguous and on the same CUDA device.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _cabi
from ._cabi import BlockDesc, BlockPtrs, CeDesc, EmbedDesc, TopkDesc, check

_basis_cache: Dict[Tuple[int, int, int], torch.Tensor] = {}


@dataclass
class Hyper:
    """What the kernels need from ``args`` (SURVEY.md §8b ctor-input row)."""
    item_size: int
    hidden_size: int
    max_seq_length: int
    num_attention_heads: int
    num_hidden_layers: int
    kept_bins: int
    alpha: float
    p_hidden: float
    p_attn: float
    kind: int = 0          # bsa_block_desc.kind: 0 BSARecBlock, 1 FMLPRecBlock (sibling model)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _require_cuda(t: torch.Tensor, name: str, dtype):
    if not t.is_cuda:
        raise RuntimeError(f"bsarec_b200: `{name}` must be a CUDA tensor — this build has no CPU fallback "
                           f"(got device {t.device}). Move the model and batch to the GPU.")
    if t.dtype != dtype:
        raise RuntimeError(f"bsarec_b200: `{name}` must be {dtype}, got {t.dtype}")
    return t.contiguous()


def filter_basis(L: int, kept_bins: int, device) -> torch.Tensor:
    """C [L, r] with C C^T == the reference's rfft/low-pass/irfft operator (model/bsarec.py:93-97)."""
    lib = _cabi.load()
    key = (L, kept_bins, torch.device(device).index or 0)
    if key not in _basis_cache:
        r = lib.bsa_filter_rank(L, kept_bins)
        basis = torch.empty(L, r, dtype=torch.float32, device=device)
        check(lib.bsa_build_filter_basis(_ptr(basis), L, kept_bins, _stream()), "bsa_build_filter_basis")
        _basis_cache[key] = basis
    return _basis_cache[key]


def _block_desc(h: Hyper, B: int, L: int, layer: int, training: bool, seed: int, inference: bool = False) -> BlockDesc:
    return BlockDesc(B=B, L=L, d=h.hidden_size, H=h.num_attention_heads, kept_bins=h.kept_bins, alpha=h.alpha,
                     p_hidden=h.p_hidden, p_attn=h.p_attn, training=int(training), seed=seed, layer=layer,
                     inference=int(inference), kind=h.kind)


def _block_ptrs(tensors: Dict[str, torch.Tensor], layer: int) -> BlockPtrs:
    pre = f"item_encoder.blocks.{layer}."
    return BlockPtrs(**{f: tensors[pre + key].data_ptr() for f, key in _cabi.BLOCK_PARAM_KEYS.items()})


class EncoderState:
    """Everything one forward pass leaves behind for its backward pass."""
    __slots__ = ("ids", "B", "L", "seed", "training", "mean", "rstd", "xs", "saved", "workspace", "last_row")


LAST_ROW_PATH = True      # module switch (tests): run the last block at position L-1 only when the caller needs no more


def encoder_forward(h: Hyper, params: Dict[str, torch.Tensor], ids: torch.Tensor, training: bool, seed: int,
                    keep_for_backward: bool, last_only: bool = False) -> Tuple[List[torch.Tensor], Optional[EncoderState]]:
    """ids int64 [B,L] -> [h0, x1, ..., xP] (each [B,L,d]); mirrors BSARecModel.forward (bsarec.py:16-28).

    ``last_only``: the caller reads nothing but position L-1 of the last block (calculate_loss, bsarec.py:32; the eval
    branch, trainers.py:127).  The last list entry is then ``[B,d]`` = ``xP[:, -1, :]``, computed by
    ``bsa_block_last_fwd`` without evaluating the other L-1 rows of that block."""
    lib = _cabi.load()
    ids = _require_cuda(ids, "input_ids", torch.int64)
    B, L = ids.shape
    d, dev = h.hidden_size, ids.device
    if L > h.max_seq_length:
        raise RuntimeError(f"sequence length {L} exceeds max_seq_length {h.max_seq_length} (position table rows)")
    E = params["item_embeddings.weight"]
    with torch.cuda.device(dev):
        st = _stream()
        basis = filter_basis(L, h.kept_bins, dev)
        h0 = torch.empty(B, L, d, dtype=torch.float32, device=dev)
        mean = torch.empty(B, L, dtype=torch.float32, device=dev)
        rstd = torch.empty(B, L, dtype=torch.float32, device=dev)
        ed = EmbedDesc(B=B, L=L, d=d, item_size=E.shape[0], p_hidden=h.p_hidden, training=int(training), seed=seed)
        check(lib.bsa_embed_ln_dropout_fwd(C.byref(ed), _ptr(ids), _ptr(E), _ptr(params["position_embeddings.weight"]),
                                           _ptr(params["LayerNorm.weight"]), _ptr(params["LayerNorm.bias"]), _ptr(h0),
                                           _ptr(mean), _ptr(rstd), st), "bsa_embed_ln_dropout_fwd")
        xs = [h0]
        saved_list = []
        bd0 = _block_desc(h, B, L, 0, training, seed)
        saved_bytes = lib.bsa_block_saved_bytes(C.byref(bd0))
        ws_bytes = lib.bsa_block_workspace_bytes(C.byref(bd0))
        workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        shared_saved = None
        last_row = False
        for l in range(h.num_hidden_layers):
            bd = _block_desc(h, B, L, l, training, seed, inference=not keep_for_backward)
            if last_only and LAST_ROW_PATH and l == h.num_hidden_layers - 1 and lib.bsa_block_last_supported(C.byref(bd)):
                last_row = True
                lsaved = torch.empty(lib.bsa_block_last_saved_bytes(C.byref(bd)), dtype=torch.uint8, device=dev)
                lws_bytes = lib.bsa_block_last_workspace_bytes(C.byref(bd))
                if lws_bytes > workspace.numel():
                    workspace = torch.empty(lws_bytes, dtype=torch.uint8, device=dev)
                w = _block_ptrs(params, l)
                out_last = torch.empty(B, d, dtype=torch.float32, device=dev)
                check(lib.bsa_block_last_fwd(C.byref(bd), C.byref(w), _ptr(basis), _ptr(ids), _ptr(xs[-1]), _ptr(out_last),
                                             _ptr(lsaved), _ptr(workspace), workspace.numel(), st), "bsa_block_last_fwd")
                xs.append(out_last)
                saved_list.append(lsaved)
                break
            if keep_for_backward or shared_saved is None:
                saved = torch.empty(saved_bytes, dtype=torch.uint8, device=dev)
                shared_saved = saved
            else:
                saved = shared_saved
            w = _block_ptrs(params, l)
            x_out = torch.empty(B, L, d, dtype=torch.float32, device=dev)
            check(lib.bsa_block_fwd(C.byref(bd), C.byref(w), _ptr(basis), _ptr(ids), _ptr(xs[-1]), _ptr(x_out), _ptr(saved),
                                    _ptr(workspace), ws_bytes, st), "bsa_block_fwd")
            xs.append(x_out)
            saved_list.append(saved)
    if last_only and not last_row:          # full last block was run: hand back its last position
        xs[-1] = gather_last(xs[-1])
    if not keep_for_backward:
        return xs, None
    state = EncoderState()
    state.last_row = last_row
    state.ids, state.B, state.L, state.seed, state.training = ids, B, L, seed, training
    state.mean, state.rstd, state.xs, state.saved, state.workspace = mean, rstd, xs, saved_list, workspace
    return xs, state


def encoder_backward(h: Hyper, params: Dict[str, torch.Tensor], state: EncoderState, d_out: torch.Tensor,
                     grads: Dict[str, torch.Tensor], rows_only: bool = False) -> Optional[torch.Tensor]:
    """d_out: gradient wrt the last block's output — [B,d] (position L-1) if the forward ran with ``last_only``,
    else [B,L,d].  Writes every encoder/embedding gradient into
    ``grads`` (item table: accumulated on top of what is already there).

    ``rows_only`` (data-parallel training): the item-table scatter is left to the caller; returns ``dz`` [B*L,d], the
    gradient of the gathered rows ``E[ids]``, and does not touch ``grads["item_embeddings.weight"]``."""
    lib = _cabi.load()
    B, L, d, dev = state.B, state.L, h.hidden_size, d_out.device
    with torch.cuda.device(dev):
        st = _stream()
        basis = filter_basis(L, h.kept_bins, dev)
        ws_bytes = state.workspace.numel()
        g = d_out
        for l in reversed(range(h.num_hidden_layers)):
            bd = _block_desc(h, B, L, l, state.training, state.seed)
            w = _block_ptrs(params, l)
            gp = _block_ptrs(grads, l)
            dx = torch.empty(B, L, d, dtype=torch.float32, device=dev)
            if l == h.num_hidden_layers - 1 and state.last_row:       # g is [B,d]: gradient at position L-1
                check(lib.bsa_block_last_bwd(C.byref(bd), C.byref(w), _ptr(basis), _ptr(state.ids), _ptr(state.xs[l]),
                                             _ptr(state.saved[l]), _ptr(g), _ptr(dx), C.byref(gp), _ptr(state.workspace),
                                             ws_bytes, st), "bsa_block_last_bwd")
                g = dx
                state.saved[l] = None
                continue
            check(lib.bsa_block_bwd(C.byref(bd), C.byref(w), _ptr(basis), _ptr(state.ids), _ptr(state.xs[l]),
                                    _ptr(state.saved[l]), _ptr(g), _ptr(dx), C.byref(gp), _ptr(state.workspace), ws_bytes, st),
                  "bsa_block_bwd")
            g = dx
            state.saved[l] = None            # release the layer's activations as soon as they are consumed
        E = params["item_embeddings.weight"]
        ed = EmbedDesc(B=B, L=L, d=d, item_size=E.shape[0], p_hidden=h.p_hidden, training=int(state.training),
                       seed=state.seed)
        eb = lib.bsa_embed_bwd_workspace_bytes(C.byref(ed))
        ws = state.workspace if ws_bytes >= eb else torch.empty(eb, dtype=torch.uint8, device=dev)
        common = (_ptr(state.ids), _ptr(E), _ptr(params["position_embeddings.weight"]), _ptr(params["LayerNorm.weight"]),
                  _ptr(state.mean), _ptr(state.rstd), _ptr(g))
        tail = (_ptr(grads["position_embeddings.weight"]), _ptr(grads["LayerNorm.weight"]), _ptr(grads["LayerNorm.bias"]),
                _ptr(ws), ws.numel(), st)
        if rows_only:
            dz = torch.empty(B * L, d, dtype=torch.float32, device=dev)
            check(lib.bsa_embed_ln_dropout_bwd_rows(C.byref(ed), *common, _ptr(dz), *tail), "bsa_embed_ln_dropout_bwd_rows")
            return dz
        check(lib.bsa_embed_ln_dropout_bwd(C.byref(ed), *common, _ptr(grads["item_embeddings.weight"]), *tail),
              "bsa_embed_ln_dropout_bwd")
    return None


def scatter_add_rows(ids: torch.Tensor, rows: torch.Tensor, dE: torch.Tensor) -> None:
    """dE[ids[i]] += rows[i] for ids[i] != 0 (sorted, atomic-free; ties summed in row order).  ids int64 [M] (any shape
    with M elements), rows fp32 [M,d]."""
    lib = _cabi.load()
    M, d = rows.shape
    assert ids.numel() == M and dE.shape[1] == d
    with torch.cuda.device(rows.device):
        wsb = lib.bsa_scatter_add_workspace_bytes(M)
        ws = torch.empty(wsb, dtype=torch.uint8, device=rows.device)
        check(lib.bsa_scatter_add_rows(_ptr(ids.contiguous()), M, dE.shape[0], _ptr(rows.contiguous()), d, _ptr(dE), _ptr(ws), wsb,
                                       _stream()), "bsa_scatter_add_rows")


def gather_rows(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """src [R,d] fp32, idx int64 [M] -> [M,d] = src[idx] (bit-exact row copies; the gather of _abstract_model.py:17-18)."""
    lib = _cabi.load()
    src = _require_cuda(src, "src", torch.float32)
    idx = _require_cuda(idx, "idx", torch.int64)
    out = torch.empty(idx.numel(), src.shape[1], dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        check(lib.bsa_gather_rows(_ptr(idx), idx.numel(), src.shape[0], _ptr(src), src.shape[1], _ptr(out), _stream()),
              "bsa_gather_rows")
    return out


def scatter_rows(dst: torch.Tensor, idx: torch.Tensor, src: torch.Tensor) -> None:
    """dst[idx[i], :] = src[i, :] (duplicates in idx must carry identical rows)."""
    lib = _cabi.load()
    idx = _require_cuda(idx, "idx", torch.int64)
    src = _require_cuda(src, "src", torch.float32)
    with torch.cuda.device(dst.device):
        check(lib.bsa_scatter_rows(_ptr(idx), idx.numel(), dst.shape[0], _ptr(src), dst.shape[1], _ptr(dst), _stream()),
              "bsa_scatter_rows")


def gather_last(x: torch.Tensor) -> torch.Tensor:
    lib = _cabi.load()
    B, L, d = x.shape
    out = torch.empty(B, d, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.bsa_gather_last(_ptr(x), _ptr(out), B, L, d, _stream()), "bsa_gather_last")
    return out


def ce_forward(s: torch.Tensor, E: torch.Tensor, answers: torch.Tensor, impl: int = 0):
    """-> (loss 0-dim, lse [B]); logits never materialised (bsarec.py:32-35)."""
    lib = _cabi.load()
    B, d = s.shape
    answers = _require_cuda(answers, "answers", torch.int64)
    cd = CeDesc(B=B, d=d, item_size=E.shape[0], impl=impl, reuse_split=0)
    with torch.cuda.device(s.device):
        wsb = lib.bsa_ce_workspace_bytes(C.byref(cd))
        ws = torch.empty(wsb, dtype=torch.uint8, device=s.device)
        loss = torch.empty((), dtype=torch.float32, device=s.device)
        lse = torch.empty(B, dtype=torch.float32, device=s.device)
        check(lib.bsa_ce_fwd(C.byref(cd), _ptr(s), _ptr(E), _ptr(answers), _ptr(loss), _ptr(lse), _ptr(ws), wsb, _stream()),
              "bsa_ce_fwd")
    return loss, lse, ws


def ce_forward_partial(s_all: torch.Tensor, E_shard: torch.Tensor, answers_local: torch.Tensor):
    """One row shard of a vocabulary-parallel CE (``bsa_ce_fwd_partial``): -> (lse_part [B], tgt_part [B], workspace).
    ``answers_local`` = answer - shard offset, negative when the target lies in another shard."""
    lib = _cabi.load()
    B, d = s_all.shape
    answers_local = _require_cuda(answers_local, "answers_local", torch.int64)
    cd = CeDesc(B=B, d=d, item_size=E_shard.shape[0], impl=2, reuse_split=0)
    with torch.cuda.device(s_all.device):
        wsb = lib.bsa_ce_workspace_bytes(C.byref(cd))
        ws = torch.empty(wsb, dtype=torch.uint8, device=s_all.device)
        lse = torch.empty(B, dtype=torch.float32, device=s_all.device)
        tgt = torch.empty(B, dtype=torch.float32, device=s_all.device)
        check(lib.bsa_ce_fwd_partial(C.byref(cd), _ptr(s_all), _ptr(E_shard), _ptr(answers_local), _ptr(lse), _ptr(tgt), _ptr(ws),
                                     wsb, _stream()), "bsa_ce_fwd_partial")
    return lse, tgt, ws


def ce_backward(s, E, answers, lse, d_loss, dE: torch.Tensor, impl: int = 0, fwd_ws: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Overwrites dE [N,d] with the dense gradient; returns ds [B,d].  ``fwd_ws``: the workspace ``ce_forward``
    returned for the same (s, E) — its split copies of both operands are reused instead of being rebuilt."""
    lib = _cabi.load()
    B, d = s.shape
    cd = CeDesc(B=B, d=d, item_size=E.shape[0], impl=impl, reuse_split=int(fwd_ws is not None))
    with torch.cuda.device(s.device):
        wsb = lib.bsa_ce_workspace_bytes(C.byref(cd))
        ws = fwd_ws if fwd_ws is not None else torch.empty(wsb, dtype=torch.uint8, device=s.device)
        ds = torch.empty(B, d, dtype=torch.float32, device=s.device)
        d_loss = d_loss.to(dtype=torch.float32).contiguous()
        check(lib.bsa_ce_bwd(C.byref(cd), _ptr(s), _ptr(E), _ptr(answers), _ptr(lse), _ptr(d_loss), _ptr(ds), _ptr(dE),
                             _ptr(ws), wsb, _stream()), "bsa_ce_bwd")
    return ds


def scatter_last(ds: torch.Tensor, L: int) -> torch.Tensor:
    lib = _cabi.load()
    B, d = ds.shape
    out = torch.empty(B, L, d, dtype=torch.float32, device=ds.device)
    with torch.cuda.device(ds.device):
        check(lib.bsa_scatter_last(_ptr(ds), _ptr(out), B, L, d, _stream()), "bsa_scatter_last")
    return out


def score_topk(s: torch.Tensor, item_shard: torch.Tensor, hist_indptr: Optional[torch.Tensor],
               hist_indices: Optional[torch.Tensor], k: int = 20, shard_offset: int = 0, impl: int = 0,
               cache: Optional[dict] = None, hist_rows: Optional[torch.Tensor] = None):
    """-> (top_val fp32 [B,k], top_idx int64 [B,k]); replaces trainers.py:62-68,130-149 for one item shard.

    ``hist_rows`` int64 [B] (optional): batch row b masks CSR row ``hist_rows[b]`` — the trainer passes its whole
    rating matrix once and the batch's ``user_ids`` here (trainers.py:131-133), no per-batch CSR is built.

    ``cache``: a dict owned by the caller (the model).  It holds the split-TF32 copy of the shard (2 N d floats inside
    the workspace) for as long as the shard is the same storage at the same version counter; ``reuse_split`` is then
    set, so an evaluation sweep splits the table once, not once per batch.  The key does not depend on the batch size
    (the workspace is kept at the largest B seen), and ``cache.clear()`` — ``BSARecModel.invalidate_eval_cache()`` —
    must follow any write to the table that bypasses the version counter (``weight.data.copy_``, raw pointers)."""
    lib = _cabi.load()
    B, d = s.shape
    td = TopkDesc(B=B, d=d, k=k, shard_rows=item_shard.shape[0], shard_offset=shard_offset, impl=impl, reuse_split=0)
    with torch.cuda.device(s.device):
        wsb = lib.bsa_score_topk_workspace_bytes(C.byref(td))
        key = (item_shard.data_ptr(), item_shard._version, tuple(item_shard.shape), k, impl, s.device.index)
        # the split pair sits at a B-dependent offset inside the workspace: reusable only for the B it was made with
        if cache is not None and cache.get("key") == key and cache.get("B") == B:
            ws = cache["ws"]
            td.reuse_split = 1
        elif cache is not None and cache.get("key") == key and B in cache.get("alt", {}):
            ws = cache["alt"][B]              # e.g. the short last batch of a sweep: its own workspace, split kept too
            td.reuse_split = 1
        else:
            ws = torch.empty(wsb, dtype=torch.uint8, device=s.device)
            if cache is not None:
                if cache.get("key") == key:   # same table, new batch size: keep the main entry
                    alt = cache.setdefault("alt", {})
                    if len(alt) >= 2:
                        alt.clear()
                    alt[B] = ws
                else:
                    cache.clear()             # drop the old 2 N d buffers before keeping a new one
                    cache["key"], cache["B"], cache["ws"] = key, B, ws
        val = torch.empty(B, k, dtype=torch.float32, device=s.device)
        idx = torch.empty(B, k, dtype=torch.int64, device=s.device)
        if hist_indptr is not None:
            hist_indptr = _require_cuda(hist_indptr, "hist_indptr", torch.int64)
            hist_indices = _require_cuda(hist_indices, "hist_indices", torch.int64)
        if hist_rows is not None:
            hist_rows = _require_cuda(hist_rows, "hist_rows", torch.int64)
            if hist_rows.numel() != B:
                raise RuntimeError(f"bsarec_b200: hist_rows has {hist_rows.numel()} entries for a batch of {B}")
        check(lib.bsa_score_topk(C.byref(td), _ptr(s.contiguous()), _ptr(item_shard), _ptr(hist_indptr), _ptr(hist_indices),
                                 _ptr(hist_rows), _ptr(val), _ptr(idx), _ptr(ws), wsb, _stream()), "bsa_score_topk")
    return val, idx


def topk_merge(cand_val: torch.Tensor, cand_idx: torch.Tensor):
    """[G,B,k] candidates -> merged [B,k] (score desc, id asc)."""
    lib = _cabi.load()
    G, B, k = cand_val.shape
    val = torch.empty(B, k, dtype=torch.float32, device=cand_val.device)
    idx = torch.empty(B, k, dtype=torch.int64, device=cand_val.device)
    with torch.cuda.device(cand_val.device):
        check(lib.bsa_topk_merge(_ptr(cand_val.contiguous()), _ptr(cand_idx.contiguous()), G, B, k, _ptr(val), _ptr(idx),
                                 _stream()), "bsa_topk_merge")
    return val, idx


def alloc_grads(params: Dict[str, torch.Tensor], big: str = "item_embeddings.weight", with_big: bool = True):
    """Gradient buffers: the item table's on its own, every other tensor a view into ONE flat fp32 buffer (so a
    data-parallel run all-reduces the 41 small gradients with a single collective and no packing copies).
    -> ({name: tensor}, flat)"""
    small = [(n, t) for n, t in params.items() if n != big]
    flat = torch.empty(sum(t.numel() for _, t in small), dtype=torch.float32, device=params[big].device)
    grads, off = {big: torch.empty_like(params[big]) if with_big else None}, 0
    for n, t in small:
        grads[n] = flat[off:off + t.numel()].view_as(t)
        off += t.numel()
    return grads, flat


class FusedLoss(torch.autograd.Function):
    """calculate_loss as one autograd node: encoder fwd -> last position -> fused CE; backward in reverse.

    ``exchange`` (``parallel.GradExchange`` or None): data-parallel gradient averaging woven into the backward pass —
    the dense cross-entropy part of the item-table gradient is all-reduced WHILE the encoder backward runs, the 41 small
    gradients go in one flat all-reduce, and the sparse gather part travels as (ids, rows) and is scatter-added on
    every rank."""

    @staticmethod
    def forward(ctx, hyper: Hyper, names: Sequence[str], ids, answers, training: bool, seed: int, ce_impl: int,
                exchange, *tensors):
        params = dict(zip(names, tensors))
        if exchange is not None and exchange.vocab_parallel:
            E = params["item_embeddings.weight"]
            ctx.ex_state = exchange.prepare_routed(_require_cuda(ids, "input_ids", torch.int64), E.shape[0])
            if exchange.stale:                               # rows of this batch's items come from the ranks that own them
                exchange.fetch_rows(ctx.ex_state, E, gather_rows, scatter_rows)
        xs, state = encoder_forward(hyper, params, ids, training, seed, keep_for_backward=True, last_only=True)
        s = xs[-1]
        ctx.hyper, ctx.names, ctx.state, ctx.ce_impl, ctx.exchange = hyper, names, state, ce_impl, exchange
        if exchange is not None and exchange.vocab_parallel:
            # every rank scores ALL users of the global batch against ITS rows of the item table
            E = params["item_embeddings.weight"]
            s_all, ans_local = exchange.gather_users(s, _require_cuda(answers, "answers", torch.int64), E.shape[0])
            lo, hi = exchange.shard(E.shape[0])
            lse_p, tgt_p, ce_ws = ce_forward_partial(s_all, E[lo:hi], ans_local)
            loss, lse = exchange.combine_loss(lse_p, tgt_p)
            ctx.answers, ctx.s, ctx.lse, ctx.ce_ws = ans_local, s_all, lse, ce_ws
            ctx.save_for_backward(*tensors)
            return loss
        loss, lse, ce_ws = ce_forward(s, params["item_embeddings.weight"], answers, ce_impl)
        if exchange is not None:
            ctx.ex_state = exchange.prepare(ids)             # row counts for the sparse exchange reach the host during the forward
        ctx.answers, ctx.s, ctx.lse, ctx.ce_ws = answers, s, lse, ce_ws
        ctx.save_for_backward(*tensors)
        return loss

    @staticmethod
    def backward(ctx, d_loss):
        tensors = ctx.saved_tensors
        names, hyper, state, ex = ctx.names, ctx.hyper, ctx.state, ctx.exchange
        params = dict(zip(names, tensors))
        grads, flat = alloc_grads(params, with_big=not (ex is not None and ex.vocab_parallel))
        if state.L < params["position_embeddings.weight"].shape[0]:
            grads["position_embeddings.weight"].zero_()      # rows >= L are never touched by the kernels
        E = params["item_embeddings.weight"]
        if ex is not None and ex.vocab_parallel:
            # the item table's gradient lives only as this rank's row shard (complete for its rows); FusedAdam updates the
            # shard and all-gathers the rows (parallel.ShardedItemGrad), autograd sees no gradient for the table
            lo, hi = ex.shard(E.shape[0])
            grads["item_embeddings.weight"] = None
            buf = torch.empty(hi - lo + 1, E.shape[1], dtype=torch.float32, device=E.device)    # row 0: landing row of the
            gE_shard = buf[1:]                                                                   # scatter's skipped key 0
            ds_all = ce_backward(ctx.s, E[lo:hi], ctx.answers, ctx.lse, d_loss, gE_shard, 2, ctx.ce_ws)
            ctx.ce_ws = None
            ds = ex.reduce_scatter_users(ds_all)
            d_out = ds if state.last_row else scatter_last(ds, state.L)
            dz = encoder_backward(hyper, params, state, d_out, grads, rows_only=True)
            ex.finish_routed(ctx.ex_state, flat, dz, buf, lo, scatter_add_rows, gather_rows)
            ex.publish_shard_grad(gE_shard, lo, hi)
            ctx.state = None
            return (None,) * 8 + tuple(grads[n] for n in names)
        gE = grads["item_embeddings.weight"]
        if ex is not None:
            d_loss = d_loss * ex.scale                       # sum over ranks == mean over the global batch
        ds = ce_backward(ctx.s, E, ctx.answers, ctx.lse, d_loss, gE, ctx.ce_impl, ctx.ce_ws)
        ctx.ce_ws = None
        if ex is not None:
            ex.start_dense(ctx.ex_state, gE)                               # overlaps everything below
        d_out = ds if state.last_row else scatter_last(ds, state.L)
        dz = encoder_backward(hyper, params, state, d_out, grads, rows_only=ex is not None)
        if ex is not None:
            ex.finish(ctx.ex_state, flat, dz, gE, scatter_add_rows, gather_rows)
        ctx.state = None
        return (None,) * 8 + tuple(grads[n] for n in names)


def bce_pair_forward(s: torch.Tensor, E: torch.Tensor, pos: torch.Tensor, neg: torch.Tensor, variant: int = 0):
    """-> (loss 0-dim, workspace): pairwise BCE at the last position (sasrec.py:46-61) by ``bsa_bce_pair_fwd``."""
    lib = _cabi.load()
    B, d = s.shape
    pos = _require_cuda(pos, "answers", torch.int64)
    neg = _require_cuda(neg, "neg_answers", torch.int64)
    with torch.cuda.device(s.device):
        wsb = lib.bsa_bce_pair_workspace_bytes(B, d)
        ws = torch.empty(wsb, dtype=torch.uint8, device=s.device)
        loss = torch.empty((), dtype=torch.float32, device=s.device)
        check(lib.bsa_bce_pair_fwd(B, d, E.shape[0], _ptr(s), _ptr(E), _ptr(pos), _ptr(neg), _ptr(loss), _ptr(ws), wsb,
                                   int(variant), _stream()), "bsa_bce_pair_fwd")
    return loss, ws


def bce_pair_backward(s, E, pos, neg, d_loss, dE: torch.Tensor, ws: torch.Tensor, variant: int = 0) -> torch.Tensor:
    """Accumulates the gathered-row gradients into dE [N,d]; returns ds [B,d]."""
    lib = _cabi.load()
    B, d = s.shape
    with torch.cuda.device(s.device):
        ds = torch.empty(B, d, dtype=torch.float32, device=s.device)
        d_loss = d_loss.to(dtype=torch.float32).contiguous()
        check(lib.bsa_bce_pair_bwd(B, d, E.shape[0], _ptr(s), _ptr(E), _ptr(pos.contiguous()), _ptr(neg.contiguous()),
                                   _ptr(d_loss), _ptr(ds), _ptr(dE), _ptr(ws), ws.numel(), int(variant), _stream()), "bsa_bce_pair_bwd")
    return ds


class FusedPairLoss(torch.autograd.Function):
    """SASRec's calculate_loss (sasrec.py:42-63) as one autograd node: encoder forward (filter branch blended out) ->
    last position -> pairwise BCE; backward in reverse.  ``names`` are the kernel-side (BSARec-layout) names of
    ``tensors``; tensors that do not require grad (the constant filter stand-ins) get their gradient dropped."""

    @staticmethod
    def forward(ctx, hyper: Hyper, names: Sequence[str], ids, pos, neg, training: bool, seed: int, *tensors):
        params = dict(zip(names, tensors))
        xs, state = encoder_forward(hyper, params, ids, training, seed, keep_for_backward=True, last_only=True)
        s = xs[-1]
        loss, ws = bce_pair_forward(s, params["item_embeddings.weight"], pos, neg, variant=hyper.kind)
        ctx.hyper, ctx.names, ctx.state, ctx.s, ctx.pos, ctx.neg, ctx.ws = hyper, names, state, s, pos, neg, ws
        ctx.save_for_backward(*tensors)
        return loss

    @staticmethod
    def backward(ctx, d_loss):
        tensors = ctx.saved_tensors
        names, hyper, state = ctx.names, ctx.hyper, ctx.state
        params = dict(zip(names, tensors))
        grads, _ = alloc_grads(params)
        if state.L < params["position_embeddings.weight"].shape[0]:
            grads["position_embeddings.weight"].zero_()
        E, gE = params["item_embeddings.weight"], grads["item_embeddings.weight"]
        gE.zero_()                                       # no dense cross-entropy part here: rows are accumulated into zeros
        ds = bce_pair_backward(ctx.s, E, ctx.pos, ctx.neg, d_loss, gE, ctx.ws, variant=hyper.kind)
        d_out = ds if state.last_row else scatter_last(ds, state.L)
        encoder_backward(hyper, params, state, d_out, grads)
        ctx.state = ctx.ws = None
        return (None,) * 7 + tuple(grads[n] if t.requires_grad else None for n, t in zip(names, tensors))


def info_nce_forward(z: torch.Tensor, tau: float):
    """z [2B, d] (two views stacked) -> (loss 0-dim, workspace): DuoRec's in-batch InfoNCE (duorec.py:47-75, sim='dot')."""
    lib = _cabi.load()
    n2, d = z.shape
    B = n2 // 2
    with torch.cuda.device(z.device):
        wsb = lib.bsa_info_nce_workspace_bytes(B)
        ws = torch.empty(wsb, dtype=torch.uint8, device=z.device)
        loss = torch.empty((), dtype=torch.float32, device=z.device)
        check(lib.bsa_info_nce_fwd(B, d, float(tau), _ptr(z), _ptr(loss), _ptr(ws), wsb, _stream()), "bsa_info_nce_fwd")
    return loss, ws


def info_nce_backward(z: torch.Tensor, tau: float, d_loss: torch.Tensor, scale: float, ws: torch.Tensor) -> torch.Tensor:
    lib = _cabi.load()
    n2, d = z.shape
    with torch.cuda.device(z.device):
        dz = torch.empty_like(z)
        d_loss = d_loss.to(dtype=torch.float32).contiguous()
        check(lib.bsa_info_nce_bwd(n2 // 2, d, float(tau), _ptr(z), _ptr(d_loss), float(scale), _ptr(dz), _ptr(ws), ws.numel(),
                                   _stream()), "bsa_info_nce_bwd")
    return dz


class FusedDuoLoss(torch.autograd.Function):
    """DuoRec's calculate_loss (duorec.py:91-139, ssl = 'us_x' or CE only) as one autograd node: the sequence, a second
    dropout view of it and the same-target sequence each go through the encoder (last position only); full-catalog CE on
    the first (the fused CE kernels), in-batch InfoNCE between the other two (bsa_info_nce_*), weight lmd_sem."""

    @staticmethod
    def forward(ctx, hyper: Hyper, names: Sequence[str], ids, answers, sem_ids, training: bool, seeds, ce_impl: int,
                tau: float, lmd_sem: float, *tensors):
        params = dict(zip(names, tensors))
        E = params["item_embeddings.weight"]
        xs, st0 = encoder_forward(hyper, params, ids, training, seeds[0], keep_for_backward=True, last_only=True)
        s = xs[-1]
        loss, lse, ce_ws = ce_forward(s, E, answers, ce_impl)
        states = [st0]
        ctx.z = None
        if sem_ids is not None:
            xa, st1 = encoder_forward(hyper, params, ids, training, seeds[1], keep_for_backward=True, last_only=True)
            xm, st2 = encoder_forward(hyper, params, sem_ids, training, seeds[2], keep_for_backward=True, last_only=True)
            z = torch.cat([xa[-1], xm[-1]], dim=0).contiguous()
            nce, nce_ws = info_nce_forward(z, tau)
            loss = loss + lmd_sem * nce
            states += [st1, st2]
            ctx.z, ctx.nce_ws = z, nce_ws
        ctx.hyper, ctx.names, ctx.states, ctx.ce_impl, ctx.tau, ctx.lmd_sem = hyper, names, states, ce_impl, tau, lmd_sem
        ctx.answers, ctx.s, ctx.lse, ctx.ce_ws = answers, s, lse, ce_ws
        ctx.save_for_backward(*tensors)
        return loss

    @staticmethod
    def backward(ctx, d_loss):
        tensors = ctx.saved_tensors
        names, hyper = ctx.names, ctx.hyper
        params = dict(zip(names, tensors))
        E = params["item_embeddings.weight"]
        grads, flat = alloc_grads(params)
        gE = grads["item_embeddings.weight"]
        ds = ce_backward(ctx.s, E, ctx.answers, ctx.lse, d_loss, gE, ctx.ce_impl, ctx.ce_ws)
        outs = [ds]
        if ctx.z is not None:
            dz = info_nce_backward(ctx.z, ctx.tau, d_loss, ctx.lmd_sem, ctx.nce_ws)
            B = dz.shape[0] // 2
            outs += [dz[:B].contiguous(), dz[B:].contiguous()]
        total = None
        for state, d_out in zip(ctx.states, outs):
            g_k, flat_k = alloc_grads(params, with_big=False)
            g_k["item_embeddings.weight"] = gE                   # the gather part accumulates on top of the dense CE part
            if state.L < params["position_embeddings.weight"].shape[0]:
                g_k["position_embeddings.weight"].zero_()
            d_in = d_out if state.last_row else scatter_last(d_out, state.L)
            encoder_backward(hyper, params, state, d_in, g_k)
            total = flat_k if total is None else total.add_(flat_k)
        flat.copy_(total)
        ctx.states = None
        return (None,) * 10 + tuple(grads[n] if t.requires_grad else None for n, t in zip(names, tensors))


_fmlp_tables: Dict[Tuple[int, int], Tuple[torch.Tensor, torch.Tensor]] = {}


def fmlp_tables(L: int, device):
    """(Tc, Ts) fp32 [L, L//2+1] with irfft(W, n=L) == Tc W_re + Ts W_im (csrc/fmlp.cu), cached per (L, device)."""
    lib = _cabi.load()
    key = (L, torch.device(device).index or 0)
    if key not in _fmlp_tables:
        F_ = L // 2 + 1
        Tc = torch.empty(L, F_, dtype=torch.float32, device=device)
        Ts = torch.empty(L, F_, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            check(lib.bsa_fmlp_tables(_ptr(Tc), _ptr(Ts), L, _stream()), "bsa_fmlp_tables")
        _fmlp_tables[key] = (Tc, Ts)
    return _fmlp_tables[key]


def fmlp_kernel(complex_weight: torch.Tensor, L: int) -> torch.Tensor:
    """complex_weight [1, L//2+1, d, 2] (fmlprec.py:98) -> taps K [L, d] of the equivalent circular convolution."""
    lib = _cabi.load()
    cw = _require_cuda(complex_weight, "complex_weight", torch.float32)
    if cw.shape[1] != L // 2 + 1:
        raise RuntimeError(f"FMLPRec: sequence length {L} does not match the filter's {cw.shape[1]} frequency bins "
                           "(the reference multiplies rfft(x) by the weight elementwise: L must equal max_seq_length)")
    d = cw.shape[2]
    Tc, Ts = fmlp_tables(L, cw.device)
    K = torch.empty(L, d, dtype=torch.float32, device=cw.device)
    with torch.cuda.device(cw.device):
        check(lib.bsa_fmlp_kernel_fwd(_ptr(cw), _ptr(Tc), _ptr(Ts), _ptr(K), L, d, _stream()), "bsa_fmlp_kernel_fwd")
    return K


def fmlp_kernel_grad(dK: torch.Tensor, L: int) -> torch.Tensor:
    """dK [L, d] -> gradient of complex_weight [1, L//2+1, d, 2]."""
    lib = _cabi.load()
    d = dK.shape[1]
    Tc, Ts = fmlp_tables(L, dK.device)
    dW = torch.empty(1, L // 2 + 1, d, 2, dtype=torch.float32, device=dK.device)
    with torch.cuda.device(dK.device):
        check(lib.bsa_fmlp_kernel_bwd(_ptr(dK.contiguous()), _ptr(Tc), _ptr(Ts), _ptr(dW), L, d, _stream()), "bsa_fmlp_kernel_bwd")
    return dW


FMLP_CW = "layer.filter_layer.complex_weight"      # kernel-side name of FMLPRecLayer.complex_weight under blocks.<l>.
FMLP_K = "layer.filter_layer.sqrt_beta"            # where the block kernels expect the taps (bsa_block_desc.kind == 1)


def fmlp_kernel_params(hyper: Hyper, params: Dict[str, torch.Tensor], L: int) -> Dict[str, torch.Tensor]:
    """params with every layer's complex weight replaced by its convolution taps under the name the kernels read."""
    out = {k: v for k, v in params.items() if not k.endswith(FMLP_CW)}
    for l in range(hyper.num_hidden_layers):
        pre = f"item_encoder.blocks.{l}."
        out[pre + FMLP_K] = fmlp_kernel(params[pre + FMLP_CW], L)
    return out


class FusedFmlpLoss(torch.autograd.Function):
    """FMLPRec's calculate_loss (fmlprec.py:40-60) as one autograd node: per layer the complex filter weight becomes the
    taps of a circular convolution (bsa_fmlp_kernel_fwd), the encoder runs the kind-1 block (filter + feed-forward, no
    attention), the loss is the pairwise log-sigmoid of bsa_bce_pair_* (variant 1); backward maps dK back to d weight."""

    @staticmethod
    def forward(ctx, hyper: Hyper, names: Sequence[str], ids, pos, neg, training: bool, seed: int, *tensors):
        params = dict(zip(names, tensors))
        ids = _require_cuda(ids, "input_ids", torch.int64)
        kp = fmlp_kernel_params(hyper, params, ids.shape[1])
        xs, state = encoder_forward(hyper, kp, ids, training, seed, keep_for_backward=True, last_only=True)
        s = xs[-1]
        loss, ws = bce_pair_forward(s, kp["item_embeddings.weight"], pos, neg, variant=1)
        ctx.hyper, ctx.names, ctx.state, ctx.s, ctx.pos, ctx.neg, ctx.ws, ctx.kp = hyper, names, state, s, pos, neg, ws, kp
        ctx.save_for_backward(*tensors)
        return loss

    @staticmethod
    def backward(ctx, d_loss):
        tensors = ctx.saved_tensors
        names, hyper, state, kp = ctx.names, ctx.hyper, ctx.state, ctx.kp
        grads, _ = alloc_grads(kp)
        if state.L < kp["position_embeddings.weight"].shape[0]:
            grads["position_embeddings.weight"].zero_()
        E, gE = kp["item_embeddings.weight"], grads["item_embeddings.weight"]
        gE.zero_()
        ds = bce_pair_backward(ctx.s, E, ctx.pos, ctx.neg, d_loss, gE, ctx.ws, variant=1)
        d_out = ds if state.last_row else scatter_last(ds, state.L)
        encoder_backward(hyper, kp, state, d_out, grads)
        out = []
        for n, t in zip(names, tensors):
            if not t.requires_grad:
                out.append(None)
            elif n.endswith(FMLP_CW):
                out.append(fmlp_kernel_grad(grads[n[:-len(FMLP_CW)] + FMLP_K], state.L))
            else:
                out.append(grads[n])
        ctx.state = ctx.ws = ctx.kp = None
        return (None,) * 7 + tuple(out)

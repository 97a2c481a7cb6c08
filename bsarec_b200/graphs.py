"""Whole-step CUDA graphs for the shapes where the step is launch-bound (the real datasets: Beauty / LastFM are
L = 50, d = 64 — ~170 launches of a few microseconds each per training step).

``GraphedTrainStep`` captures the reference's iteration body (``trainers.py:103-107``)::

    loss = model.calculate_loss(input_ids, answers, neg_answer, same_target, user_ids)
    optim.zero_grad(); loss.backward(); optim.step()

into ONE graph launch per step; ``GraphedTopk`` does the same for the eval branch (``trainers.py:126-149`` ->
``BSARecModel.full_sort_topk``).  Launch arguments are frozen at capture, so the two per-step host values of the path
live in device memory instead (``bsa_set_replay_state``, include/bsarec_b200.h): word 0 is added to every dropout seed
(replay r draws exactly the masks an eager step with seed ``base + r * SEED_STRIDE`` draws — that is what the parity
test checks), word 1 to Adam's step number.  The last node of the graph advances both.

Not a tracing compiler: the graph is the same hand-written kernels in the same order, recorded once.
Single process only (the data-parallel exchanges stay eager); ``bsarec_b200.FusedAdam`` only.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import _cabi
from .optim import FusedAdam

SEED_STRIDE = 0xD1B54A32D192ED03          # odd; not the site constant of make_dropout, so (replay, site) pairs never collide
_MASK64 = (1 << 64) - 1


def _signed64(x: int) -> int:
    x &= _MASK64
    return x - (1 << 64) if x >= (1 << 63) else x


def replay_seed(base: int, replay: int) -> int:
    """The dropout seed replay number ``replay`` (0-based) of a graph captured with seed ``base`` runs with."""
    return (int(base) + replay * SEED_STRIDE) & _MASK64


class _ReplayState:
    """The two uint64 words of ``bsa_set_replay_state`` and the tensor that advances them."""

    def __init__(self, device):
        self.words = torch.zeros(2, dtype=torch.int64, device=device)
        self.inc = torch.tensor([_signed64(SEED_STRIDE), 1], dtype=torch.int64, device=device)
        self.one_step = torch.tensor([0, 1], dtype=torch.int64, device=device)

    class _Bound:
        def __init__(self, ptr):
            self.ptr = ptr

        def __enter__(self):
            lib = _cabi.load()
            self.prev = lib.bsa_get_replay_state()
            lib.bsa_set_replay_state(self.ptr)

        def __exit__(self, *exc):
            _cabi.load().bsa_set_replay_state(self.prev)

    def bound(self):
        return self._Bound(self.words.data_ptr())


def _clone_args(args: Sequence[Optional[torch.Tensor]]):
    return [a.detach().clone() if isinstance(a, torch.Tensor) else a for a in args]


class GraphedTrainStep:
    """``step = GraphedTrainStep(model, optimizer); loss = step(input_ids, answers, neg_answer, same_target, user_ids)``.

    The first call with a given batch shape runs one eager warm-up step on a scratch copy of the parameters and the
    optimizer state (undone afterwards), captures the graph, and replays it; later calls copy the batch into the
    graph's input tensors and replay.  A batch of another shape (an epoch's short last batch) runs the eager loop body.
    After a call every ``p.grad`` holds that step's gradient and ``optimizer.state[p]["step"]`` has advanced, exactly
    as after the eager body.  The returned loss is the graph's own 0-dim output tensor (valid until the next call).

    Re-captures when the optimizer's hyper-parameters change; call ``reset()`` after replacing parameter tensors."""

    def __init__(self, model, optimizer, seed: Optional[int] = None):
        if not isinstance(optimizer, FusedAdam):
            raise TypeError("GraphedTrainStep needs bsarec_b200.FusedAdam (its step number can live on the device)")
        if getattr(model, "_exchange", None) is not None:
            raise RuntimeError("GraphedTrainStep: data-parallel exchanges are not captured; use the eager step")
        self.model, self.optimizer = model, optimizer
        self.base_seed = seed
        self.reset()

    def reset(self):
        self._graph = None
        self._key = None
        self.replays = 0

    # ------------------------------------------------------------------ helpers
    def _params(self):
        return [p for g in self.optimizer.param_groups for p in g["params"]]

    def _hyper_key(self):
        return tuple((float(g["lr"]), tuple(float(b) for b in g["betas"]), float(g["eps"]), float(g["weight_decay"]))
                     for g in self.optimizer.param_groups)

    def _opt_step(self) -> int:
        for p in self._params():
            st = self.optimizer.state.get(p)
            if st:
                return int(st["step"])
        return 0

    def _eager(self, args):
        loss = self.model.calculate_loss(*args)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    # ------------------------------------------------------------------ capture
    def _capture(self, args):
        model, opt = self.model, self.optimizer
        if not model.training:
            raise RuntimeError("GraphedTrainStep: call model.train() first")
        dev = args[0].device
        params = self._params()
        self._static = _clone_args(args)
        self._state = _ReplayState(dev)
        if self.base_seed is None:
            self.base_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        prev_seed = model._fixed_seed
        # --- warm-up: one real step (lazy initialisation everywhere: optimizer state, cached operators, kernel
        # attributes), on a side stream as torch's capture recipe asks, then put parameters and state back
        with torch.no_grad():
            snap_p = [p.detach().clone() for p in params]
            snap_s = {}
            for p in params:
                st = opt.state.get(p)
                if st:
                    snap_s[p] = (int(st["step"]), st["exp_avg"].clone(), st["exp_avg_sq"].clone())
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        try:
            model.set_dropout_seed(self.base_seed)
            with torch.cuda.stream(side):
                self._eager(self._static)
            torch.cuda.current_stream(dev).wait_stream(side)
            with torch.no_grad():
                for p, s in zip(params, snap_p):
                    p.copy_(s)
                for p in params:
                    st = opt.state.get(p)
                    if not st:
                        continue
                    if p in snap_s:
                        st["step"] = snap_s[p][0]
                        st["exp_avg"].copy_(snap_s[p][1])
                        st["exp_avg_sq"].copy_(snap_s[p][2])
                    else:
                        st["step"] = 0
                        st["exp_avg"].zero_()
                        st["exp_avg_sq"].zero_()
            del snap_p, snap_s
            # --- capture
            opt.zero_grad(set_to_none=True)           # gradients are (re)allocated from the graph's pool
            steps_before = self._opt_step()
            graph = torch.cuda.CUDAGraph()
            l0 = _cabi.launch_count()
            with self._state.bound():
                with torch.cuda.graph(graph):
                    loss = model.calculate_loss(*self._static)
                    loss.backward()
                    opt.step()
                    self._state.words.add_(self._state.inc)
            self.launches_per_replay = _cabi.launch_count() - l0      # this library's kernels inside one replay
            # the captured opt.step() advanced the host-side counters although nothing has run yet
            for p in params:
                st = opt.state.get(p)
                if st:
                    st["step"] = int(st["step"]) - 1
        finally:
            model.set_dropout_seed(prev_seed)
        self._graph, self._loss = graph, loss.detach()
        self._grads = [(p, p.grad) for p in params]
        self._captured_step = steps_before + 1        # the step number baked into the Adam launches
        self._dev_step_offset = 0                     # value of word 1 on the device
        self.replays = 0

    # ------------------------------------------------------------------ call
    @staticmethod
    def _on_device(args):
        return tuple(a.cuda(non_blocking=True) if isinstance(a, torch.Tensor) and not a.is_cuda else a for a in args)

    def __call__(self, input_ids, answers, neg_answers=None, same_target=None, user_ids=None):
        args = (input_ids, answers, neg_answers, same_target, user_ids)
        if not (input_ids.is_cuda or input_ids.is_pinned()):      # pinned host batches are copied straight into the graph's inputs
            raise RuntimeError("bsarec_b200: input_ids must be a CUDA (or pinned host) tensor (no CPU fallback)")
        shapes = tuple(tuple(a.shape) if isinstance(a, torch.Tensor) else None for a in args)
        hyper = self._hyper_key()
        if self._graph is None or hyper != self._key[1]:
            self._capture(self._on_device(args))
            self._key = (shapes, hyper)
        elif shapes != self._key[0]:                  # another batch shape: the eager loop body
            return self._eager(self._on_device(args))
        with torch.no_grad():
            for dst, src in zip(self._static, args):
                if isinstance(dst, torch.Tensor):
                    dst.copy_(src, non_blocking=True)
            want = self._opt_step() + 1 - self._captured_step       # eager steps may have run in between
            if want != self._dev_step_offset:
                self._state.words.add_(self._state.one_step, alpha=want - self._dev_step_offset)
                self._dev_step_offset = want
        self._graph.replay()
        self._dev_step_offset += 1
        self.replays += 1
        opt = self.optimizer
        for p, g in self._grads:
            st = opt.state.get(p)
            if st:
                st["step"] = int(st["step"]) + 1
            if p.grad is not g:
                p.grad = g
        torch.autograd.graph.increment_version(self.model.item_embeddings.weight)      # eval's split cache keys on it
        return self._loss


class GraphedTopk:
    """``topk = GraphedTopk(model); pred = topk(input_ids, hist_indptr, hist_indices, hist_rows)`` — the eval branch
    (``BSARecModel.full_sort_topk`` with the whole-matrix history CSR and the batch's ``user_ids``) as one graph launch.
    Re-captures when the batch shape, the CSR tensors or the item table (its version counter) change; the returned
    int64 [B,k] tensor is the graph's own output (valid until the next call)."""

    def __init__(self, model, k: int = 20):
        self.model, self.k = model, k
        self._graph, self._key = None, None

    def _make_key(self, input_ids, indptr, indices):
        w = self.model.item_embeddings.weight
        return (tuple(input_ids.shape), input_ids.device.index, indptr.data_ptr(), indices.data_ptr(), w.data_ptr(), w._version)

    @torch.no_grad()
    def __call__(self, input_ids, hist_indptr, hist_indices, hist_rows):
        model = self.model
        if model.training:
            raise RuntimeError("GraphedTopk: call model.eval() first")
        key = self._make_key(input_ids, hist_indptr, hist_indices)
        if key != self._key:
            dev = input_ids.device
            self._ids, self._rows = input_ids.clone(), hist_rows.clone()
            self._indptr, self._indices = hist_indptr, hist_indices          # kept alive: the graph reads them in place
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):                                      # fills the split cache of the item table
                model.full_sort_topk(self._ids, None, hist_indptr, hist_indices, k=self.k, hist_rows=self._rows)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            l0 = _cabi.launch_count()
            with torch.cuda.graph(graph):
                out = model.full_sort_topk(self._ids, None, hist_indptr, hist_indices, k=self.k, hist_rows=self._rows)
            self.launches_per_replay = _cabi.launch_count() - l0
            self._keep = dict(getattr(model, "_topk_cache", {}))               # the split item table the graph reads
            self._graph, self._out, self._key = graph, out, key
        self._ids.copy_(input_ids, non_blocking=True)
        self._rows.copy_(hist_rows, non_blocking=True)
        self._graph.replay()
        return self._out

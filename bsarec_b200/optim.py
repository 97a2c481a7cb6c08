"""``FusedAdam``: drop-in for the ``torch.optim.Adam`` the reference's Trainer builds (``trainers.py:27-28``) with the
update done by one sm_100a kernel pass per large tensor (``bsa_adam_step``) and one launch for all the small ones
(``bsa_adam_step_multi``) instead of ~13 foreach passes.

Same constructor arguments, same update rule (no amsgrad / maximize), ``state_dict`` entries named as torch's
(``step``, ``exp_avg``, ``exp_avg_sq``).  CUDA fp32 parameters only — there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi
from ._cabi import check


class FusedAdam(torch.optim.Optimizer):
    SMALL = 1 << 18      # tensors up to this many elements are updated together, one launch per 48 of them (the grouped
                         # launch gives every tensor at most 64 blocks: Beauty's 775 k-element item table took 18 us there)

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or weight_decay < 0.0:
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _cabi.load()
        for group in self.param_groups:
            b1, b2 = group["betas"]
            small = {}                      # (device index, step) -> tensors updated together by one bsa_adam_step_multi
            for p in group["params"]:
                shard = getattr(p, "_bsa_shard", None)
                if shard is not None:            # vocabulary-parallel training: this rank owns rows [lo, hi) of the table
                    self._step_shard(lib, group, p, shard)
                    p._bsa_shard = None
                    continue
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise RuntimeError("FusedAdam: parameters and gradients must be CUDA fp32 (no CPU fallback)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] = int(st["step"]) + 1
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                if not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous")
                if p.numel() <= self.SMALL:
                    small.setdefault((p.device.index, st["step"]), []).append((p, g, st))
                    continue
                with torch.cuda.device(p.device):
                    check(lib.bsa_adam_step(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                                            p.numel(), float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                                            float(group["weight_decay"]), st["step"],
                                            torch.cuda.current_stream().cuda_stream), "bsa_adam_step")
                torch.autograd.graph.increment_version(p)     # written through a raw pointer: tell autograd / caches
            for (dev, step), items in small.items():
                n = len(items)
                arr = lambda f: (C.c_void_p * n)(*[f(it) for it in items])
                sizes = (C.c_int64 * n)(*[it[0].numel() for it in items])
                with torch.cuda.device(dev):
                    check(lib.bsa_adam_step_multi(n, arr(lambda it: it[0].data_ptr()), arr(lambda it: it[1].data_ptr()),
                                                  arr(lambda it: it[2]["exp_avg"].data_ptr()),
                                                  arr(lambda it: it[2]["exp_avg_sq"].data_ptr()), sizes, float(group["lr"]),
                                                  float(b1), float(b2), float(group["eps"]), float(group["weight_decay"]), step,
                                                  torch.cuda.current_stream().cuda_stream), "bsa_adam_step_multi")
                for it in items:
                    torch.autograd.graph.increment_version(it[0])
        return loss

    def _step_shard(self, lib, group, p, shard):
        """Adam on rows [lo, hi) of ``p`` with the complete row-shard gradient ``shard.grad`` (state kept for the shard
        only: 1/world of the memory and of the 28 B/element pass).  The other replicas are brought up to date lazily — each
        rank fetches the rows its next batch reads, `model.eval()` / `state_dict()` all-gather the shards — or, with
        `exchange.fetch = False`, by an all-gather of the updated rows right here."""
        import torch.distributed as dist
        if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
            raise RuntimeError("FusedAdam: sharded parameters must be contiguous CUDA fp32")
        b1, b2 = group["betas"]
        st = self.state[p]
        rows = p.data[shard.lo:shard.hi]
        if not st or st.get("shard") != (shard.lo, shard.hi):
            st.clear()
            st["step"] = 0
            st["shard"] = (shard.lo, shard.hi)
            st["exp_avg"] = torch.zeros_like(rows)
            st["exp_avg_sq"] = torch.zeros_like(rows)
        st["step"] = int(st["step"]) + 1
        g = shard.grad if shard.grad.is_contiguous() else shard.grad.contiguous()
        with torch.cuda.device(p.device):
            check(lib.bsa_adam_step(rows.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                                    rows.numel(), float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                                    float(group["weight_decay"]), st["step"], torch.cuda.current_stream().cuda_stream),
                  "bsa_adam_step")
        if shard.exchange is not None and shard.exchange.fetch:
            shard.exchange.stale = True          # the other ranks fetch the rows they read (GradExchange.fetch_rows / sync_table)
        else:
            dist.all_gather_into_tensor(p.data.view(-1), rows.reshape(-1).clone(), group=shard.group)
        torch.autograd.graph.increment_version(p)

#!/usr/bin/env python
"""CE forward + backward at the C3 shape (B=512, d=256, N=1M) a few times: the short command the ce kernels are profiled
under.  Prints forward / backward device times for ce mode 2 (logits stored by the forward, read by the backward) and
mode 1 (logits recomputed)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bsarec_b200 import functional as Fn, _cabi
lib = _cabi.load()
dev = torch.device("cuda:0")
B, d, N = int(os.environ.get("B", 512)), 256, int(os.environ.get("N", 1_000_000))
g = torch.Generator(device=dev).manual_seed(0)
s = torch.randn(B, d, device=dev, generator=g)
E = torch.randn(N, d, device=dev, generator=g) * 0.05
ans = torch.randint(1, N, (B,), device=dev, generator=g)
dE = torch.empty_like(E)
one = torch.ones((), device=dev)
modes = [int(m) for m in os.environ.get("MODES", "2,1").split(",")]
for mode in modes:
    lib.bsa_set_ce_mode(mode)
    for it in range(int(os.environ.get("ITERS", 2))):
        loss, lse, ws = Fn.ce_forward(s, E, ans, 2)
        Fn.ce_backward(s, E, ans, lse, one, dE, 2, ws)
    torch.cuda.synchronize()
    if os.environ.get("TIME", "1") == "1":
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        for it in range(5):
            loss, lse, ws = Fn.ce_forward(s, E, ans, 2)
        e[1].record()
        for it in range(5):
            Fn.ce_backward(s, E, ans, lse, one, dE, 2, ws)
        e[2].record(); torch.cuda.synchronize()
        print(f"ce mode {mode}: fwd {e[0].elapsed_time(e[1]) / 5:.3f} ms, bwd {e[1].elapsed_time(e[2]) / 5:.3f} ms", flush=True)

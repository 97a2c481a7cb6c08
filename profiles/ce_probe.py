#!/usr/bin/env python
"""CE forward + backward at the C3 shape (B=512, d=256, N=1M) a few times: the short command ce kernels are profiled under."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bsarec_b200 import functional as Fn
dev = torch.device("cuda:0")
B, d, N = int(os.environ.get("B", 512)), 256, int(os.environ.get("N", 1_000_000))
g = torch.Generator(device=dev).manual_seed(0)
s = torch.randn(B, d, device=dev, generator=g)
E = torch.randn(N, d, device=dev, generator=g) * 0.05
ans = torch.randint(1, N, (B,), device=dev, generator=g)
dE = torch.empty_like(E)
one = torch.ones((), device=dev)
for it in range(int(os.environ.get("ITERS", 3))):
    loss, lse, ws = Fn.ce_forward(s, E, ans, 2)
    Fn.ce_backward(s, E, ans, lse, one, dE, 2, ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for it in range(5):
    Fn.ce_backward(s, E, ans, lse, one, dE, 2, ws)
e1.record(); torch.cuda.synchronize()
print("ce_bwd ms", e0.elapsed_time(e1) / 5)

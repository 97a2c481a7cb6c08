#!/usr/bin/env python
"""Two train steps of the C3 encoder shape (L=200, d=256, 4 heads, B=512) over a small catalog: the short command the
attention kernels are profiled under (ncu -k regex:attn_)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
from bsarec_b200 import BSARecModel
from helpers import make_args

cfg = dict(bench.WORKLOADS[os.environ.get("WL", "c3")][1]); cfg["item_size"] = 20000
B = int(os.environ.get("B", 512))
dev = torch.device("cuda:0")
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, 0.5, 0.5, batch_size=B)).to(dev).train()
b = {k: v.to(dev) for k, v in bench.synth_batches(cfg, B, 1234, 1)[0].items()}
for _ in range(2):
    model.zero_grad()
    model.calculate_loss(b["ids"], b["answers"], None, None, None).backward()
torch.cuda.synchronize()
print("ok")

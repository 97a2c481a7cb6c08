#!/usr/bin/env python
"""N warm C3 train steps and nothing else (deterministic launch order for `ncu --launch-skip/--launch-count`).

    ncu --set full --clock-control none --import-source on -k regex:umma_gemm_kernel --launch-skip 52 --launch-count 26 \
        -o gpurun_out/gemms python profiles/ncu_train_steps.py 3
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import bench
from bsarec_b200 import BSARecModel, FusedAdam
from helpers import make_args

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
desc, cfg, B, _ = bench.WORKLOADS["c3"]
dev = torch.device("cuda:0")
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, bench.P_DROP, bench.P_DROP, batch_size=B)).to(dev)
optim = FusedAdam(model.parameters(), lr=1e-3)
batches = [{k: v.to(dev) for k, v in b.items()} for b in bench.synth_batches(cfg, B, 1234, 2)]
model.train(True)
for i in range(steps):
    b = batches[i % 2]
    loss = model.calculate_loss(b["ids"], b["answers"], None, None, None)
    optim.zero_grad()
    loss.backward()
    optim.step()
torch.cuda.synchronize()
print("ok", float(loss))
